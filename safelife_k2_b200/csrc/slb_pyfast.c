/* slb_pyfast.c — a CPython binding of the two per-step calls of the host actor loop
 * (slb_stepper_step / slb_stepper_wait, include/safelife_b200.h).  The library itself stays a plain
 * C ABI; this module only removes the ctypes argument marshalling (about 0.5 us per call) from the
 * loop that makes two such calls per 6 us step.  Everything else keeps going through ctypes.
 *
 *   step(stepper_address, actions_address, reward_address, done_address) -> int status
 *   wait(stepper_address) -> int status
 *   dev_step(dev_stepper_address, actions, reward, done, obs_or_None, stream) -> int status (slb_dev_stepper_step,
 *   the device-buffer step behind SafeLifeEnv.step)
 *
 * Addresses are Python ints (what ctypes' c_void_p.value / tensor.data_ptr() give).  The GIL is kept:
 * both calls return within microseconds. */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

#include "../../include/safelife_b200.h"

static PyObject *py_step(PyObject *self, PyObject *const *args, Py_ssize_t nargs) {
    (void)self;
    if (nargs != 4) {
        PyErr_SetString(PyExc_TypeError, "step(stepper, actions, reward, done) takes 4 addresses");
        return NULL;
    }
    void *p[4];
    for (int k = 0; k < 4; k++) {
        if (args[k] == Py_None) {
            p[k] = NULL;
        } else {
            p[k] = PyLong_AsVoidPtr(args[k]);
            if (p[k] == NULL && PyErr_Occurred()) return NULL;
        }
    }
    int rc = slb_stepper_step((slb_stepper *)p[0], (const uint8_t *)p[1], (int32_t *)p[2], (uint8_t *)p[3]);
    return PyLong_FromLong(rc);
}

static PyObject *py_wait(PyObject *self, PyObject *arg) {
    (void)self;
    void *p = PyLong_AsVoidPtr(arg);
    if (p == NULL && PyErr_Occurred()) return NULL;
    return PyLong_FromLong(slb_stepper_wait((slb_stepper *)p));
}

/* dev_step(dev_stepper_address, actions, reward, done, obs_or_None, stream) -> int status  (slb_dev_stepper_step) */
static PyObject *py_dev_step(PyObject *self, PyObject *const *args, Py_ssize_t nargs) {
    (void)self;
    if (nargs != 6) {
        PyErr_SetString(PyExc_TypeError, "dev_step(stepper, actions, reward, done, obs, stream) takes 6 addresses");
        return NULL;
    }
    void *p[6];
    for (int k = 0; k < 6; k++) {
        if (args[k] == Py_None) {
            p[k] = NULL;
        } else {
            p[k] = PyLong_AsVoidPtr(args[k]);
            if (p[k] == NULL && PyErr_Occurred()) return NULL;
        }
    }
    int rc = slb_dev_stepper_step((slb_dev_stepper *)p[0], (const uint8_t *)p[1], (int32_t *)p[2], (uint8_t *)p[3],
                                  (uint16_t *)p[4], p[5]);
    return PyLong_FromLong(rc);
}

static PyMethodDef methods[] = {
    {"dev_step", (PyCFunction)(void (*)(void))py_dev_step, METH_FASTCALL, "slb_dev_stepper_step on raw addresses"},
    {"step", (PyCFunction)(void (*)(void))py_step, METH_FASTCALL, "slb_stepper_step on raw addresses"},
    {"wait", py_wait, METH_O, "slb_stepper_wait on a raw address"},
    {NULL, NULL, 0, NULL},
};

static struct PyModuleDef module = {PyModuleDef_HEAD_INIT, "_slbfast", "fast binding of the host-step calls", -1,
                                    methods, NULL, NULL, NULL, NULL};

PyMODINIT_FUNC PyInit__slbfast(void) { return PyModule_Create(&module); }
