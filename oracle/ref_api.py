"""ctypes bindings of oracle/_ref/libspmv_ref.so (TEST INFRASTRUCTURE).

That library is the reference's own host code -- L2GMap.cpp, Matrix.cpp, csr_matrix.cpp,
reference_executor.cpp, cg.cpp, read_petsc.cpp, demos/CreateA.cpp -- compiled unmodified
(oracle/Makefile) with ranks as threads.  Use ``run_ranks(n, fn)`` and call the
``RefMatrix`` builders inside ``fn``.
"""
import ctypes as C
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "libspmv_ref.so")
_p, _i32, _i64, _f64, _int = C.c_void_p, C.c_int32, C.c_int64, C.c_double, C.c_int
_pp = C.POINTER(C.c_void_p)
_RANK_FN = C.CFUNCTYPE(None, C.c_void_p)
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{_PATH} missing (built only where /root/reference exists)")
        L = C.CDLL(_PATH)
        sig = {
            "ref_run_ranks": [_int, _RANK_FN, _p],
            "ref_matrix_create": [_pp, _p, _p, _p, _i64, _i64, _p, _i64, _p, _i64, _int, _int],
            "ref_matrix_read_petsc": [_pp, C.c_char_p, _int, _int],
            "ref_matrix_create_A": [_pp, _int],
            "ref_matrix_destroy": [_p],
            "ref_matrix_info": [_p, C.POINTER(_i64)],
            "ref_matrix_block_dims": [_p, _int, C.POINTER(_i64)],
            "ref_matrix_block_copy": [_p, _int, _p, _p, _p, _p],
            "ref_l2g_get": [_p, _int, _int, _p, _i64, C.POINTER(_i64)],
            "ref_l2g_sizes": [_p, _int, C.POINTER(_i64)],
            "ref_l2g_global_to_local": [_p, _i64, C.POINTER(_i32)],
            "ref_update": [_p, _p],
            "ref_mult": [_p, _p, _p],
            "ref_cg": [_p, _p, _p, _int, _f64, C.POINTER(_int)],
            "ref_reverse_update": [_p, _p],
            "ref_demo_spmv_loop": [_p, _p, _p, _int, C.POINTER(_f64), C.POINTER(_f64)],
        }
        for n, a in sig.items():
            f = getattr(L, n)
            f.argtypes, f.restype = a, C.c_int
        L.ref_last_error.restype = C.c_char_p
        L.ref_world_rank.restype = C.c_int
        L.ref_world_size.restype = C.c_int
        L.ref_allreduce_sum.argtypes = [_f64]
        L.ref_allreduce_sum.restype = _f64
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise RuntimeError("reference: " + lib().ref_last_error().decode(errors="replace"))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def world_rank():
    return lib().ref_world_rank()


def world_size():
    return lib().ref_world_size()


def allreduce_sum(v):
    """MPI_Allreduce(SUM) of one double over the thread-ranks, folded in rank order."""
    return float(lib().ref_allreduce_sum(float(v)))


def run_ranks(n, fn):
    """fn(rank) on n thread-ranks of the reference library."""
    errors, lock = [], threading.Lock()

    def body(_):
        try:
            fn(world_rank())
        except BaseException as e:  # noqa: BLE001
            with lock:
                errors.append(e)

    rc = lib().ref_run_ranks(n, _RANK_FN(body), None)
    if errors:
        raise errors[0]
    if rc:
        raise RuntimeError("a reference thread-rank failed")


_GET = {"ranges": (0, np.int64), "ghosts": (1, np.int64), "neighbours": (2, np.int32), "send_count": (3, np.int32),
        "recv_count": (4, np.int32), "send_offset": (5, np.int32), "recv_offset": (6, np.int32),
        "indexbuf": (7, np.int32)}


class RefMatrix:
    """A spmv::Matrix<double> of the reference, on its ReferenceExecutor."""

    def __init__(self, h):
        self.h = h

    @staticmethod
    def create_matrix(rowptr, colind, values, nrows_local, ncols_local, row_ghosts, col_ghosts, symmetric=False,
                      cm=2):
        rp = np.ascontiguousarray(rowptr, dtype=np.int32)
        ci = np.ascontiguousarray(colind, dtype=np.int32)
        v = np.ascontiguousarray(values, dtype=np.float64)
        rg = np.ascontiguousarray(row_ghosts, dtype=np.int64)
        cg = np.ascontiguousarray(col_ghosts, dtype=np.int64)
        h = C.c_void_p()
        _ck(lib().ref_matrix_create(C.byref(h), _ptr(rp), _ptr(ci), _ptr(v), int(nrows_local), int(ncols_local),
                                    _ptr(rg), len(rg), _ptr(cg), len(cg), int(bool(symmetric)), int(cm)))
        return RefMatrix(h)

    @staticmethod
    def read_petsc(filename, symmetric=False, cm=2):
        h = C.c_void_p()
        _ck(lib().ref_matrix_read_petsc(C.byref(h), os.fsencode(filename), int(bool(symmetric)), int(cm)))
        return RefMatrix(h)

    @staticmethod
    def create_A(N):
        h = C.c_void_p()
        _ck(lib().ref_matrix_create_A(C.byref(h), int(N)))
        return RefMatrix(h)

    def info(self):
        out = (_i64 * 8)()
        _ck(lib().ref_matrix_info(self.h, out))
        k = ["rows", "cols", "non_zeros", "symmetric", "format_size", "has_remote", "local_nnz", "remote_nnz"]
        return dict(zip(k, list(out)))

    def block(self, which):
        dims = (_i64 * 4)()
        _ck(lib().ref_matrix_block_dims(self.h, which, dims))
        rows, cols, nnz, has_diag = list(dims)
        if nnz < 0:
            return None
        rp, ci, v = np.zeros(rows + 1, np.int32), np.zeros(nnz, np.int32), np.zeros(nnz, np.float64)
        d = np.zeros(rows, np.float64) if has_diag else None
        _ck(lib().ref_matrix_block_copy(self.h, which, _ptr(rp), _ptr(ci), _ptr(v), _ptr(d)))
        return {"rows": rows, "cols": cols, "rowptr": rp, "colind": ci, "values": v, "diagonal": d}

    def plan(self, which_map=0):
        d = {}
        for name, (what, dt) in _GET.items():
            n = _i64()
            _ck(lib().ref_l2g_get(self.h, which_map, what, None, 0, C.byref(n)))
            out = np.zeros(n.value, dtype=dt)
            _ck(lib().ref_l2g_get(self.h, which_map, what, _ptr(out), n.value, C.byref(n)))
            d[name] = out
        s = (_i64 * 8)()
        _ck(lib().ref_l2g_sizes(self.h, which_map, s))
        d.update(local_size=s[0], num_ghosts=s[1], global_size=s[2], global_offset=s[3])
        return d

    def global_to_local(self, g):
        out = _i32()
        _ck(lib().ref_l2g_global_to_local(self.h, int(g), C.byref(out)))
        return out.value

    def update(self, x):
        _ck(lib().ref_update(self.h, _ptr(x)))

    def mult(self, x, y):
        _ck(lib().ref_mult(self.h, _ptr(x), _ptr(y)))

    def reverse_update(self, x):
        """L2GMap::reverse_update on the column map (use cm = collective_blocking)."""
        _ck(lib().ref_reverse_update(self.h, _ptr(x)))

    def demo_spmv_loop(self, x, y, n_apply=100):
        """demos/spmv.cpp:81-108 -> ({"update": s, "mult": s, "copy": s}, final ||x||_2); x, y are updated."""
        t = (_f64 * 3)()
        nrm = _f64()
        _ck(lib().ref_demo_spmv_loop(self.h, _ptr(x), _ptr(y), int(n_apply), t, C.byref(nrm)))
        return {"update": t[0], "mult": t[1], "copy": t[2]}, nrm.value

    def cg(self, b, kmax, rtol):
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros(len(b))
        it = C.c_int()
        _ck(lib().ref_cg(self.h, _ptr(b), _ptr(x), int(kmax), float(rtol), C.byref(it)))
        return it.value, x

    def close(self):
        if self.h:
            lib().ref_matrix_destroy(self.h)
            self.h = None
