/*
 * bb_fastcall.c -- a CPython callable bound to one C-ABI entry point of libbball.so and its five arguments.
 *
 * Why: a served step (bb_serve_step, include/bb_engine.h) takes ~16 us end to end at 8192 worlds; calling it through ctypes
 * costs another ~1.5 us per call (argument conversion, errno / GIL bookkeeping).  `bind5(fn, last_error, a0..a4)` returns an
 * object whose call does nothing but `fn(a0, a1, a2, a3, a4)` -- the same C-ABI call with the same pointers the ctypes path
 * makes (manager.py bind_serve_step / bind_step_host look the addresses up once) -- and raises RuntimeError with
 * bb_last_error()'s text when it returns non-zero.  Binding glue only: no compute, nothing of the step path lives here, and
 * manager.py falls back to the ctypes closure when this module is not built.
 *
 * The call keeps the GIL (it returns within microseconds; the server's own liveness checks bound it).
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stddef.h>
#include <stdint.h>

typedef int (*bb_fn5)(void *, void *, void *, void *, void *);
typedef const char *(*bb_err_fn)(void);

typedef struct {
    PyObject_HEAD
    bb_fn5 fn;
    bb_err_fn last_error;
    void *a[5];
    vectorcallfunc vectorcall;
} Bound5;

static PyObject *bound5_vectorcall(PyObject *self_, PyObject *const *args, size_t nargsf, PyObject *kwnames)
{
    Bound5 *self = (Bound5 *)self_;
    (void)args;
    if (PyVectorcall_NARGS(nargsf) != 0 || (kwnames && PyTuple_GET_SIZE(kwnames) != 0)) {
        PyErr_SetString(PyExc_TypeError, "bound call takes no arguments");
        return NULL;
    }
    const int rc = self->fn(self->a[0], self->a[1], self->a[2], self->a[3], self->a[4]);
    if (rc != 0) {
        const char *msg = self->last_error ? self->last_error() : NULL;
        PyErr_Format(PyExc_RuntimeError, "C-ABI call failed (rc %d): %s", rc, (msg && msg[0]) ? msg : "no error text");
        return NULL;
    }
    Py_RETURN_NONE;
}

static PyTypeObject Bound5Type = {
    PyVarObject_HEAD_INIT(NULL, 0)
    .tp_name = "madrona_basketball_b200._fastcall.Bound5",
    .tp_basicsize = sizeof(Bound5),
    .tp_flags = Py_TPFLAGS_DEFAULT | Py_TPFLAGS_HAVE_VECTORCALL,
    .tp_call = PyVectorcall_Call,
    .tp_vectorcall_offset = offsetof(Bound5, vectorcall),
    .tp_doc = "fn(a0, a1, a2, a3, a4) of a C-ABI entry point, bound once",
};

/* bind5(fn_addr, last_error_addr, a0, a1, a2, a3, a4): all unsigned integers (addresses; 0 = NULL) */
static PyObject *bind5(PyObject *mod, PyObject *args)
{
    unsigned long long fn, err, a0, a1, a2, a3, a4;
    (void)mod;
    if (!PyArg_ParseTuple(args, "KKKKKKK", &fn, &err, &a0, &a1, &a2, &a3, &a4)) return NULL;
    if (!fn) {
        PyErr_SetString(PyExc_ValueError, "bind5: NULL function address");
        return NULL;
    }
    Bound5 *b = PyObject_New(Bound5, &Bound5Type);
    if (!b) return NULL;
    b->fn = (bb_fn5)(uintptr_t)fn;
    b->last_error = (bb_err_fn)(uintptr_t)err;
    b->a[0] = (void *)(uintptr_t)a0; b->a[1] = (void *)(uintptr_t)a1; b->a[2] = (void *)(uintptr_t)a2;
    b->a[3] = (void *)(uintptr_t)a3; b->a[4] = (void *)(uintptr_t)a4;
    b->vectorcall = bound5_vectorcall;
    return (PyObject *)b;
}

static PyMethodDef methods[] = {
    {"bind5", bind5, METH_VARARGS, "bind5(fn_addr, last_error_addr, a0, a1, a2, a3, a4) -> zero-argument callable"},
    {NULL, NULL, 0, NULL},
};

static struct PyModuleDef moduledef = {PyModuleDef_HEAD_INIT, "_fastcall", "bound C-ABI calls (binding glue)", -1, methods, NULL, NULL, NULL, NULL};

PyMODINIT_FUNC PyInit__fastcall(void)
{
    if (PyType_Ready(&Bound5Type) < 0) return NULL;
    return PyModule_Create(&moduledef);
}
