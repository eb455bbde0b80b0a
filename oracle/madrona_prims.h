/*
 * madrona_prims.h — the Madrona L0 primitives the step path calls, restated.
 *
 * TEST INFRASTRUCTURE (part of oracle/).  UNVERIFIED against Madrona source:
 * the submodule shacklettbp/madrona @ 86457089d677d94e2826613c8ba62d180b50ab41
 * (reference .gitmodules:1-3, .SUBMODULES.json) is absent from /root/reference.
 * Formulas follow SURVEY.md App. C; the only independent pin is
 * scripts/viewer.py:59-65, which mirrors Quat::rotateVec, and the published
 * Random123 known-answer vectors for Threefry2x32-20.
 * Each primitive sits behind one function so it can be re-pinned later.
 */
#ifndef ORC_MADRONA_PRIMS_H
#define ORC_MADRONA_PRIMS_H

#include <stdint.h>
#include <math.h>
#include "../include/bb_math.h"

namespace orc {

/* ---- libm: deterministic spec by default, glibc with -DBB_USE_LIBM -------- */
#ifdef BB_USE_LIBM
static inline float m_sinf(float x) { return ::sinf(x); }
static inline float m_cosf(float x) { return ::cosf(x); }
static inline float m_atan2f(float y, float x) { return ::atan2f(y, x); }
static inline float m_acosf(float x) { return ::acosf(x); }
static inline float m_atanf(float x) { return ::atanf(x); }
static inline float m_erff(float x) { return ::erff(x); }
static inline float m_expf(float x) { return ::expf(x); }
#else
static inline float m_sinf(float x) { return bb_sinf(x); }
static inline float m_cosf(float x) { return bb_cosf(x); }
static inline float m_atan2f(float y, float x) { return bb_atan2f(y, x); }
static inline float m_acosf(float x) { return bb_acosf(x); }
static inline float m_atanf(float x) { return bb_atanf(x); }
static inline float m_erff(float x) { return bb_erff(x); }
static inline float m_expf(float x) { return bb_expf(x); }
#endif
static inline float m_sqrtf(float x) { return __builtin_sqrtf(x); }

constexpr float pi = 3.14159265358979323846f; /* madrona::math::pi */

/* ---- math::Vector3 -------------------------------------------------------- */
struct Vector3 {
    float x, y, z;

    float dot(const Vector3 &o) const { return x * o.x + y * o.y + z * o.z; }
    float length2() const { return x * x + y * y + z * z; }
    float length() const { return m_sqrtf(length2()); }
    float invLength() const { return 1.f / m_sqrtf(length2()); }
    Vector3 normalize() const
    {
        float inv = invLength();
        return Vector3{x * inv, y * inv, z * inv};
    }
    float &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    Vector3 &operator+=(const Vector3 &o) { x += o.x; y += o.y; z += o.z; return *this; }
    Vector3 &operator-=(const Vector3 &o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
    Vector3 &operator*=(float s) { x *= s; y *= s; z *= s; return *this; }
    static Vector3 zero() { return Vector3{0.f, 0.f, 0.f}; }
};
static inline Vector3 operator+(Vector3 a, Vector3 b) { return Vector3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline Vector3 operator-(Vector3 a, Vector3 b) { return Vector3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline Vector3 operator-(Vector3 a) { return Vector3{-a.x, -a.y, -a.z}; }
static inline Vector3 operator*(Vector3 a, float s) { return Vector3{a.x * s, a.y * s, a.z * s}; }
static inline Vector3 operator*(float s, Vector3 a) { return Vector3{a.x * s, a.y * s, a.z * s}; }
static inline float dot(Vector3 a, Vector3 b) { return a.dot(b); }
static inline Vector3 cross(Vector3 a, Vector3 b)
{
    return Vector3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

/* ---- math::Quat {w,x,y,z} -------------------------------------------------- */
struct Quat {
    float w, x, y, z;

    static Quat id() { return Quat{1.f, 0.f, 0.f, 0.f}; }
    static Quat angleAxis(float angle, Vector3 normal)
    {
        float coshalf = m_cosf(angle / 2.f);
        float sinhalf = m_sinf(angle / 2.f);
        return Quat{coshalf, normal.x * sinhalf, normal.y * sinhalf, normal.z * sinhalf};
    }
    /* v + 2 (w (p x v) + p x (p x v)); mirrored by scripts/viewer.py:59-65 */
    Vector3 rotateVec(Vector3 v) const
    {
        Vector3 pure{x, y, z};
        float scalar = w;
        Vector3 pure_x_v = cross(pure, v);
        Vector3 pure_x_pure_x_v = cross(pure, pure_x_v);
        return v + 2.f * ((pure_x_v * scalar) + pure_x_pure_x_v);
    }
};
/* Hamilton product */
static inline Quat operator*(Quat a, Quat b)
{
    return Quat{
        (a.w * b.w - a.x * b.x) - (a.y * b.y + a.z * b.z),
        (a.w * b.x + a.x * b.w) + (a.y * b.z - a.z * b.y),
        (a.w * b.y + a.y * b.w) + (a.z * b.x - a.x * b.z),
        (a.w * b.z + a.z * b.w) + (a.x * b.y - a.y * b.x),
    };
}

static inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

/* ---- rand:: (Threefry2x32-20, JAX-style split) ----------------------------- */
struct RandKey { uint32_t a, b; };

static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

static inline RandKey threefry2x32(RandKey key, uint32_t c0, uint32_t c1)
{
    const int R[8] = {13, 15, 26, 6, 17, 29, 16, 24};
    uint32_t ks[3] = {key.a, key.b, 0x1BD11BDAu ^ key.a ^ key.b};
    uint32_t x0 = c0 + ks[0], x1 = c1 + ks[1];
    for (int round = 0; round < 20; round++) {
        x0 += x1;
        x1 = rotl32(x1, R[round % 8]);
        x1 ^= x0;
        if ((round & 3) == 3) {
            int s = round / 4 + 1;
            x0 += ks[s % 3];
            x1 += ks[(s + 1) % 3] + (uint32_t)s;
        }
    }
    return RandKey{x0, x1};
}

static inline RandKey initKey(uint32_t seed, uint32_t seed_upper = 0) { return RandKey{seed, seed_upper}; }
static inline RandKey split_i(RandKey src, uint32_t idx, uint32_t idx_upper = 0) { return threefry2x32(src, idx, idx_upper); }
static inline uint32_t bits32(RandKey k) { return k.a ^ k.b; }
static inline float bitsToFloat01(uint32_t bits) { return (float)(bits >> 8) * 0x1p-24f; }

struct RNG {
    RandKey key;
    uint32_t count;
    float sampleUniform() { return bitsToFloat01(bits32(split_i(key, count++, 0))); }
};

} // namespace orc
#endif
