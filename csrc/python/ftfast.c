/* _ftfast: the per-call path of the Python API as a compiled CPython extension (METH_FASTCALL, no ctypes).
 *
 * Communicator.allreduce used to cross into the library through ctypes: every call built c_void_p objects, converted
 * seven arguments one by one and went through libffi with the GIL held - a few microseconds per call, 341 times per
 * GPT-2 step in the DDP hook. Here one vectorcall parses plain integers and calls ft_allreduce_ex directly; the GIL
 * is released around the call (a first call that grows the staging window runs a collective bootstrap exchange,
 * which calls back into Python on the torch path - the callback re-acquires the GIL itself).
 * The reference's API is one C call as well (MPI_Allreduce, allreduce_over_mpi/mpi_mod.hpp:1167-1221). */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

#include "flextree.h"

static int as_u64(PyObject* o, unsigned long long* out) {
  *out = PyLong_AsUnsignedLongLongMask(o);
  return !(*out == (unsigned long long)-1 && PyErr_Occurred());
}

/* allreduce_ex(comm, send_ptr, recv_ptr, count, dtype, wire_dtype, op, scale, stream) -> result code */
static PyObject* ftfast_allreduce_ex(PyObject* self, PyObject* const* args, Py_ssize_t nargs) {
  (void)self;
  if (nargs != 9) {
    PyErr_SetString(PyExc_TypeError, "allreduce_ex(comm, send, recv, count, dtype, wire_dtype, op, scale, stream)");
    return NULL;
  }
  unsigned long long comm, send, recv, count, stream;
  if (!as_u64(args[0], &comm) || !as_u64(args[1], &send) || !as_u64(args[2], &recv) || !as_u64(args[3], &count) ||
      !as_u64(args[8], &stream))
    return NULL;
  const long dt = PyLong_AsLong(args[4]), wt = PyLong_AsLong(args[5]), op = PyLong_AsLong(args[6]);
  const double scale = PyFloat_AsDouble(args[7]);
  if (PyErr_Occurred()) return NULL;
  ft_result_t rc;
  Py_BEGIN_ALLOW_THREADS
  rc = ft_allreduce_ex((const void*)(uintptr_t)send, (void*)(uintptr_t)recv, (size_t)count, (ft_dtype_t)dt, (ft_dtype_t)wt,
                       (ft_op_t)op, scale, (ft_comm_t*)(uintptr_t)comm, (void*)(uintptr_t)stream);
  Py_END_ALLOW_THREADS
  return PyLong_FromLong((long)rc);
}

static PyObject* ftfast_error_peek(PyObject* self, PyObject* const* args, Py_ssize_t nargs) {
  (void)self;
  unsigned long long comm;
  if (nargs != 1 || !as_u64(args[0], &comm)) return NULL;
  return PyLong_FromUnsignedLong(ft_comm_error_peek((const ft_comm_t*)(uintptr_t)comm));
}

static PyMethodDef methods[] = {
    {"allreduce_ex", (PyCFunction)(void (*)(void))ftfast_allreduce_ex, METH_FASTCALL, "ft_allreduce_ex on raw pointers; returns the result code"},
    {"error_peek", (PyCFunction)(void (*)(void))ftfast_error_peek, METH_FASTCALL, "ft_comm_error_peek"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef moduledef = {PyModuleDef_HEAD_INIT, "_ftfast", "flextree per-call fast path", -1, methods, NULL, NULL, NULL, NULL};

PyMODINIT_FUNC PyInit__ftfast(void) { return PyModule_Create(&moduledef); }
