/* _wfptcall: the one call a sampler makes millions of times -- wiener_like(x, v, sv, a, z, sz, t, st, err, ...) on a NumPy
 * array (/root/reference/src/wfpt.pyx:54-76, called from hddm/likelihoods.py:117-121) -- bound with the CPython C API
 * instead of ctypes: argument conversion through ctypes costs 4-9 microseconds of a 19-microsecond call.  The arithmetic
 * is libwfpt_b200's (wfpt_like_cached_host / wfpt_like_batch_host, include/wfpt_b200.h); nothing is computed here.
 * Returns None for anything but a C-contiguous 1-D float64 array with plain numeric arguments: the Python wrapper then
 * takes its ctypes path, which produces the reference's errors. */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <string.h>

#include "../../include/wfpt_b200.h"

static PyObject* call_wiener_like(PyObject* self, PyObject* args)
{
    PyObject* xo;
    double v, sv, a, z, sz, t, st, err, simps_err, p_outlier, w_outlier;
    int n_st, n_sz, use_adaptive, variant, cached;
    (void)self;
    if (!PyArg_ParseTuple(args, "Oddddddddiiidddii", &xo, &v, &sv, &a, &z, &sz, &t, &st, &err, &n_st, &n_sz, &use_adaptive,
                          &simps_err, &p_outlier, &w_outlier, &variant, &cached)) {
        PyErr_Clear();
        Py_RETURN_NONE;
    }
    Py_buffer view;
    if (PyObject_GetBuffer(xo, &view, PyBUF_FORMAT | PyBUF_C_CONTIGUOUS) != 0) {
        PyErr_Clear();
        Py_RETURN_NONE;
    }
    if (view.ndim != 1 || view.itemsize != 8 || !view.format || strcmp(view.format, "d") != 0) {
        PyBuffer_Release(&view);
        Py_RETURN_NONE;
    }
    const double prm[8] = {v, sv, a, z, sz, t, st, p_outlier};
    double res = 0.0;
    const int64_t n = (int64_t)view.shape[0];
    int rc;
    if (cached)
        rc = wfpt_like_cached_host((const double*)view.buf, n, prm, 1, err, n_st, n_sz, use_adaptive ? 1 : 0, simps_err, w_outlier,
                                   variant, &res);
    else
        rc = wfpt_like_batch_host((const double*)view.buf, n, prm, 1, err, n_st, n_sz, use_adaptive ? 1 : 0, simps_err, w_outlier,
                                  variant, &res);
    PyBuffer_Release(&view);
    if (rc != 0) {
        PyErr_SetString(PyExc_RuntimeError, wfpt_last_error());
        return NULL;
    }
    return PyFloat_FromDouble(res);
}

static PyMethodDef methods[] = {
    {"wiener_like", call_wiener_like, METH_VARARGS,
     "wiener_like(x, v, sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive, simps_err, p_outlier, w_outlier, variant, cached)"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef module = {PyModuleDef_HEAD_INIT, "_wfptcall", "fast binding of the wiener_like call", -1, methods,
                                    NULL, NULL, NULL, NULL};

PyMODINIT_FUNC PyInit__wfptcall(void) { return PyModule_Create(&module); }
