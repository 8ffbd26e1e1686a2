/* b200ps._fastfeed — CPython entry points for the two calls on the per-step end-to-end path.
 *
 * HostFeeder.submit / Ticket.result are issued once per training step from the user's Python loop
 * (the `sess.run(..., feed_dict=...)` of example.py:168-170).  Through ctypes each costs 2-3 us of argument
 * marshalling — as much as the native work behind it; these wrappers take plain integers (handle, pinned
 * host pointers, stream) and return a Python int / float directly.  The module is optional: without it
 * b200ps.runtime.feeder falls back to ctypes. */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

extern int b200ps_feeder_submit(void* feeder, const void* src0, const void* src1, void* stream, int64_t* ticket);
extern int b200ps_feeder_wait(void* feeder, int64_t ticket, void* out);
extern const char* b200ps_last_error(void);

static PyObject* ff_error(const char* what, int rc) {
  PyErr_Format(PyExc_RuntimeError, "%s failed (%d): %s", what, rc, b200ps_last_error());
  return NULL;
}

/* submit(handle, src0_ptr, src1_ptr, stream_ptr) -> ticket */
static PyObject* ff_submit(PyObject* self, PyObject* const* args, Py_ssize_t nargs) {
  if (nargs != 4) { PyErr_SetString(PyExc_TypeError, "submit(handle, src0, src1, stream)"); return NULL; }
  void* h = PyLong_AsVoidPtr(args[0]);
  void* p0 = PyLong_AsVoidPtr(args[1]);
  void* p1 = PyLong_AsVoidPtr(args[2]);
  void* st = PyLong_AsVoidPtr(args[3]);
  if (PyErr_Occurred()) return NULL;
  int64_t t = 0;
  const int rc = b200ps_feeder_submit(h, p0, p1, st, &t);
  if (rc) return ff_error("feeder_submit", rc);
  return PyLong_FromLongLong((long long)t);
}

/* wait_first(handle, ticket) -> float: blocks on the unit, returns the first fp32 word of its result block */
static PyObject* ff_wait_first(PyObject* self, PyObject* const* args, Py_ssize_t nargs) {
  if (nargs != 2) { PyErr_SetString(PyExc_TypeError, "wait_first(handle, ticket)"); return NULL; }
  void* h = PyLong_AsVoidPtr(args[0]);
  const long long t = PyLong_AsLongLong(args[1]);
  if (PyErr_Occurred()) return NULL;
  float out[64];                       /* result blocks of up to 256 bytes; larger ones use the ctypes path */
  int rc;
  Py_BEGIN_ALLOW_THREADS
  rc = b200ps_feeder_wait(h, (int64_t)t, out);
  Py_END_ALLOW_THREADS
  if (rc) return ff_error("feeder_wait", rc);
  return PyFloat_FromDouble((double)out[0]);
}

static PyMethodDef methods[] = {
    {"submit", (PyCFunction)(void (*)(void))ff_submit, METH_FASTCALL, "queue one feeder unit; returns its ticket"},
    {"wait_first", (PyCFunction)(void (*)(void))ff_wait_first, METH_FASTCALL, "block on a ticket; first result word"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef moduledef = {PyModuleDef_HEAD_INIT, "_fastfeed", "fast paths of b200ps.runtime.feeder", -1, methods};

PyMODINIT_FUNC PyInit__fastfeed(void) { return PyModule_Create(&moduledef); }
