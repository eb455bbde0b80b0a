"""Derive the polynomial used by bb_erff (include/bb_math.h).

erfc(x) = t * exp(-x*x + P(s)),  t = 2/(2+x),  s = (2t - (1+T0)) / (1-T0)  in [-1,1]
for x in [0.5, X_MAX]  (t in [T0, 0.8]).  P is a Chebyshev interpolant converted to
monomial form in s.  Ground truth: mpmath erfc at 60 digits.
Run:  python tools/gen_erfc_coeffs.py
"""
import mpmath as mp
mp.mp.dps = 60

X_MIN, X_MAX = mp.mpf("0.5"), mp.mpf("6.5")
T_HI = 2 / (2 + X_MIN)
T_LO = 2 / (2 + X_MAX)

def f_of_t(t):
    x = 2 / t - 2
    return mp.log(mp.erfc(x) / t) + x * x

def cheb_fit(deg):
    n = deg + 1
    nodes = [mp.cos(mp.pi * (k + mp.mpf(1) / 2) / n) for k in range(n)]
    ts = [(T_HI - T_LO) / 2 * s + (T_HI + T_LO) / 2 for s in nodes]
    fs = [f_of_t(t) for t in ts]
    c = []
    for j in range(n):
        acc = mp.mpf(0)
        for k in range(n):
            acc += fs[k] * mp.cos(mp.pi * j * (k + mp.mpf(1) / 2) / n)
        c.append(acc * 2 / n)
    c[0] /= 2
    return c

def cheb_to_mono(c):
    # T0=1, T1=s, T_{n+1}=2 s T_n - T_{n-1}
    n = len(c)
    T = [[mp.mpf(0)] * n for _ in range(n)]
    T[0][0] = mp.mpf(1)
    if n > 1:
        T[1][1] = mp.mpf(1)
    for k in range(2, n):
        for i in range(n):
            T[k][i] = (2 * T[k - 1][i - 1] if i > 0 else 0) - T[k - 2][i]
    m = [mp.mpf(0)] * n
    for k in range(n):
        for i in range(n):
            m[i] += c[k] * T[k][i]
    return m

def max_err(m, samples=4000):
    worst = mp.mpf(0)
    for i in range(samples + 1):
        x = X_MIN + (X_MAX - X_MIN) * i / samples
        t = 2 / (2 + x)
        s = (2 * t - (T_HI + T_LO)) / (T_HI - T_LO)
        # evaluate in double-like rounding to be realistic
        acc = 0.0
        sd = float(s)
        for coef in reversed(m):
            acc = acc * sd + float(coef)
        approx = mp.mpf(float(t)) * mp.exp(-x * x + mp.mpf(acc))
        erf_true = 1 - mp.erfc(x)
        err = abs((1 - approx) - erf_true) / erf_true
        worst = max(worst, err)
    return worst

if __name__ == "__main__":
    for deg in (10, 12, 14, 16, 18):
        m = cheb_to_mono(cheb_fit(deg))
        print("deg", deg, "max rel err of erf:", mp.nstr(max_err(m, 800), 5))
    deg = 12
    m = cheb_to_mono(cheb_fit(deg))
    print("T_LO = %s  T_HI = %s" % (mp.nstr(T_LO, 20), mp.nstr(T_HI, 20)))
    for i, coef in enumerate(m):
        print("  P%-2d = %s," % (i, float(coef).hex()), "//", repr(float(coef)))
