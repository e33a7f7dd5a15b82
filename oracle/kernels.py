"""ctypes bindings of the CPU oracle kernels (TEST INFRASTRUCTURE).

``liboracle.so``  -- plain-C restatement, oracle/restated/kernels.c
``_ref/libspmv_ref_kernels.so`` -- the reference's csr_kernels.cpp and
openmp/csr_kernels.openmp.cpp compiled unmodified (oracle/Makefile).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def _load(path):
    if not os.path.exists(path):
        raise FileNotFoundError(
            f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`"
        )
    return C.CDLL(path)


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        _lib = _load(os.path.join(_HERE, "_build", "liboracle.so"))
        L = _lib
        L.oracle_csr_spmv_f64.argtypes = [C.c_int32, _i32p, _i32p, _f64p, C.c_double, _f64p, C.c_double, _f64p]
        L.oracle_csr_spmv_f32.argtypes = [C.c_int32, _i32p, _i32p, _f32p, C.c_float, _f32p, C.c_float, _f32p]
        L.oracle_csr_symv_f64.argtypes = [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, _f64p,
                                          C.c_double, _f64p, C.c_double, _f64p]
        L.oracle_csr_symv_f32.argtypes = [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, _f32p,
                                          C.c_float, _f32p, C.c_float, _f32p]
        L.oracle_gather_f64.argtypes = [C.c_int32, _i32p, _f64p, _f64p]
        L.oracle_scatter_add_f64.argtypes = [C.c_int32, _i32p, _f64p, _f64p]
        L.oracle_ddot.argtypes = [C.c_int32, _f64p, _f64p]
        L.oracle_ddot.restype = C.c_double
        L.oracle_daxpy.argtypes = [C.c_int32, C.c_double, _f64p, _f64p]
        L.oracle_dscal.argtypes = [C.c_int32, C.c_double, _f64p]
    return _lib


def have_ref():
    return os.path.exists(os.path.join(_HERE, "_ref", "libspmv_ref_kernels.so"))


def ref():
    global _ref
    if _ref is None:
        _ref = _load(os.path.join(_HERE, "_ref", "libspmv_ref_kernels.so"))
        R = _ref
        R.ref_csr_run_f64.argtypes = [C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_int, C.c_double, _f64p, C.c_double, _f64p]
        R.ref_omp_create.argtypes = [C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        R.ref_omp_create.restype = C.c_void_p
        R.ref_omp_run.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_double, _f64p, C.c_double, _f64p]
        R.ref_omp_destroy.argtypes = [C.c_void_p]
        R.ref_omp_set_threads.argtypes = [C.c_int]
    return _ref


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


def spmv(rowptr, colind, values, x, alpha=1.0, beta=0.0, y=None):
    """General CSR SpMV, csr_kernels.cpp:41-51.  Returns a new y."""
    dt = np.asarray(values).dtype if np.asarray(values).dtype == np.float32 else np.float64
    rowptr, colind = _c(rowptr, np.int32), _c(colind, np.int32)
    values, x = _c(values, dt), _c(x, dt)
    n = len(rowptr) - 1
    out = np.zeros(n, dt) if y is None else np.array(y, dtype=dt, copy=True)
    f = lib().oracle_csr_spmv_f32 if dt == np.float32 else lib().oracle_csr_spmv_f64
    f(n, rowptr, colind, values, alpha, x, beta, out)
    return out


def symv(rowptr, colind, values, diagonal, x, alpha=1.0, beta=0.0, y=None):
    """Symmetric CSR SpMV (strictly lower + diagonal), csr_kernels.cpp:26-40."""
    dt = np.asarray(diagonal).dtype if np.asarray(diagonal).dtype == np.float32 else np.float64
    diagonal, x = _c(diagonal, dt), _c(x, dt)
    n = len(diagonal)
    nnz = 0 if values is None else len(values)
    rowptr, colind, values = _c(rowptr, np.int32), _c(colind, np.int32), _c(values, dt)
    out = np.zeros(n, dt) if y is None else np.array(y, dtype=dt, copy=True)
    f = lib().oracle_csr_symv_f32 if dt == np.float32 else lib().oracle_csr_symv_f64
    f(n, nnz, _ptr(rowptr) if nnz else None, _ptr(colind) if nnz else None, _ptr(values) if nnz else None,
      diagonal, alpha, x, beta, out)
    return out


def gather(indices, vec):
    indices, vec = _c(indices, np.int32), _c(vec, np.float64)
    out = np.empty(len(indices), np.float64)
    lib().oracle_gather_f64(len(indices), indices, vec, out)
    return out


def ddot(x, y):
    x, y = _c(x, np.float64), _c(y, np.float64)
    return float(lib().oracle_ddot(len(x), x, y))


def daxpy(a, x, y):
    """y += a*x in place (y must be a contiguous float64 array)."""
    lib().oracle_daxpy(len(y), a, _c(x, np.float64), y)


def dscal(a, x):
    lib().oracle_dscal(len(x), a, x)


def laplace3d(nx, ny, nzg, z0, z1, stencil, symmetric=False):
    """Rows of planes [z0, z1) of the 7-/27-point Laplacian (definition: spmv_b200/synth.py::laplace3d_rows),
    generated by the plain-C oracle_laplace3d: -> (rowptr int32, global colind int32, values[, diagonal]).
    symmetric: strictly lower triangle + diagonal (Matrix.cpp:337-349 at one rank)."""
    L = lib()
    L.oracle_laplace3d.restype = C.c_longlong
    L.oracle_laplace3d.argtypes = [C.c_int, C.c_int, C.c_longlong, C.c_longlong, C.c_longlong, C.c_int, C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    rows = nx * ny * (z1 - z0)
    rp = np.zeros(rows + 1, np.int64)
    n = L.oracle_laplace3d(nx, ny, nzg, z0, z1, stencil, int(symmetric), _ptr(rp), None, None, None)
    assert n < 2 ** 31
    ci, v = np.empty(n, np.int32), np.empty(n, np.float64)
    d = np.zeros(rows, np.float64) if symmetric else None
    L.oracle_laplace3d(nx, ny, nzg, z0, z1, stencil, int(symmetric), _ptr(rp), _ptr(ci), _ptr(v), _ptr(d))
    rp = rp.astype(np.int32)
    return (rp, ci, v, d) if symmetric else (rp, ci, v)


def ref_run(rows, cols, rowptr, colind, values, diagonal, symmetric, x, alpha=1.0, beta=0.0, y=None):
    """The reference's own sequential CSRSpMV<double>::run (oracle/_ref)."""
    nnz = 0 if values is None else len(values)
    rowptr, colind, values = _c(rowptr, np.int32), _c(colind, np.int32), _c(values, np.float64)
    diagonal = _c(diagonal, np.float64)
    x = np.array(x, dtype=np.float64, copy=True)
    out = np.zeros(rows, np.float64) if y is None else np.array(y, dtype=np.float64, copy=True)
    ref().ref_csr_run_f64(rows, cols, nnz, _ptr(rowptr), _ptr(colind), _ptr(values), _ptr(diagonal),
                          int(symmetric), alpha, x, beta, out)
    return out


class RefOmp:
    """The reference's OpenMP CSRSpMV<double> (openmp/csr_kernels.openmp.cpp).

    The reference reads its thread count from the OMP_NUM_THREADS environment
    variable, not from the runtime (openmp/omp_executor.cpp:22-34); the constructor
    sets both to ``threads`` (default: OMP_NUM_THREADS, else all cores).
    """

    def __init__(self, rows, cols, rowptr, colind, values, diagonal=None, symmetric=False, threads=None):
        if threads is None:
            threads = int(os.environ.get("OMP_NUM_THREADS", "0")) or (os.cpu_count() or 1)
        self.threads = int(threads)
        os.environ["OMP_NUM_THREADS"] = str(self.threads)
        ref().ref_omp_set_threads(self.threads)
        self.rows, self.cols = rows, cols
        self.rowptr, self.colind = _c(rowptr, np.int32), _c(colind, np.int32)
        self.values, self.diagonal = _c(values, np.float64), _c(diagonal, np.float64)
        self.nnz = 0 if values is None else len(self.values)
        self.h = ref().ref_omp_create(rows, cols, self.nnz, _ptr(self.rowptr), _ptr(self.colind),
                                      _ptr(self.values), int(symmetric))

    def run(self, x, y, alpha=1.0, beta=0.0):
        ref().ref_omp_run(self.h, self.rows, self.cols, self.nnz, _ptr(self.rowptr), _ptr(self.colind),
                          _ptr(self.values), _ptr(self.diagonal), alpha, x, beta, y)

    def close(self):
        if self.h:
            ref().ref_omp_destroy(self.h)
            self.h = None
