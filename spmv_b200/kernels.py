"""Kernel-level Python harness over the C ABI, on torch tensors.

PyTorch is plumbing here (device memory, the current stream, events); every
compute call goes through ``libb200spmv.so``.  Used by tests/ and bench.py.
"""
import ctypes as C

import torch

from ._lib import CG_STATE_DOUBLES, DOT_SCRATCH_DOUBLES, check, lib


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class CsrPlan:
    """CSRSpMV<T>::init/finalize (csr_kernels.h:71-78): the per-block analysis."""

    def __init__(self, rows, cols, rowptr, symmetric=False, nnz=None):
        self.rows, self.cols, self.symmetric = int(rows), int(cols), bool(symmetric)
        self.rowptr = rowptr
        self.nnz = int(nnz) if nnz is not None else (int(rowptr[-1].item()) if rowptr is not None else 0)
        h = C.c_void_p()
        check(lib().b200_csr_plan_create(C.byref(h), self.rows, self.cols, self.nnz, _ptr(rowptr),
                                         int(self.symmetric), _stream()))
        self.h = h

    def info(self):
        nt, mx = C.c_int32(), C.c_int32()
        hist = (C.c_int64 * 32)()
        check(lib().b200_csr_plan_info(self.h, C.byref(nt), C.byref(mx), hist))
        return {"num_tiles": nt.value, "max_row_len": mx.value, "hist": list(hist)}

    def tune(self, variant=0, tile_bytes=49152, stages=2, consumer_warps=8, ctas_per_sm=2):
        check(lib().b200_csr_plan_tune(self.h, _ptr(self.rowptr), variant, tile_bytes, stages,
                                       consumer_warps, ctas_per_sm, _stream()))

    def close(self):
        if self.h:
            check(lib().b200_csr_plan_destroy(self.h))
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def spmv(plan, rowptr, colind, values, x, y, alpha=1.0, beta=0.0):
    f = lib().b200_csr_spmv_f32 if values.dtype == torch.float32 else lib().b200_csr_spmv_f64
    check(f(plan.h, _ptr(rowptr), _ptr(colind), _ptr(values), alpha, _ptr(x), beta, _ptr(y), _stream()))


def spmv_dot(plan, rowptr, colind, values, x, y, dot):
    check(lib().b200_csr_spmv_dot_f64(plan.h, _ptr(rowptr), _ptr(colind), _ptr(values), _ptr(x), _ptr(y),
                                      _ptr(dot), _stream()))


def symv(plan, rowptr, colind, values, diag, x, y, alpha=1.0, beta=0.0):
    f = lib().b200_csr_symv_f32 if diag.dtype == torch.float32 else lib().b200_csr_symv_f64
    check(f(plan.h, _ptr(rowptr), _ptr(colind), _ptr(values), _ptr(diag), alpha, _ptr(x), beta, _ptr(y),
            _stream()))


def symv_dot(plan, rowptr, colind, values, diag, x, y, dot):
    """y = A x (lower + diagonal storage) and dot[0] = x'Ax in the same pass (b200_csr_symv_dot_f64)."""
    check(lib().b200_csr_symv_dot_f64(plan.h, _ptr(rowptr), _ptr(colind), _ptr(values), _ptr(diag), _ptr(x), _ptr(y),
                                      _ptr(dot), _stream()))


def plan_rowlist_dot_ok(plan):
    ok = C.c_int32(0)
    check(lib().b200_csr_plan_rowlist_dot_ok(plan.h, C.byref(ok)))
    return bool(ok.value)


def spmv_rowlist_dot(remote_plan, local_plan, rowptr, colind, values, x, y, dot):
    """Remote block (beta = 1) completing the x.(A x) epilogue the local block's spmv_dot / symv_dot started."""
    check(lib().b200_csr_spmv_rowlist_dot_f64(remote_plan.h, local_plan.h, _ptr(rowptr), _ptr(colind), _ptr(values),
                                              _ptr(x), _ptr(y), _ptr(dot), _stream()))


def gather(idx, vec, out):
    f = lib().b200_gather_f32 if vec.dtype == torch.float32 else lib().b200_gather_f64
    check(f(idx.numel(), _ptr(idx), _ptr(vec), _ptr(out), _stream()))


def scatter_add(idx, src, inout):
    f = lib().b200_scatter_add_f32 if src.dtype == torch.float32 else lib().b200_scatter_add_f64
    check(f(idx.numel(), _ptr(idx), _ptr(src), _ptr(inout), _stream()))


def dot(a, b, out, scratch, done_flag=None):
    check(lib().b200_cg_dot_f64(a.numel(), _ptr(a), _ptr(b), _ptr(out), _ptr(scratch), _ptr(done_flag),
                                _stream()))


def dot_static(a, b, out):
    """b200_cg_dot_static_f64: the same dot product with the library's static device scratch (allocates nothing)."""
    check(lib().b200_cg_dot_static_f64(a.numel(), _ptr(a), _ptr(b), _ptr(out), _stream()))


def new_dot_scratch(device):
    return torch.zeros(DOT_SCRATCH_DOUBLES, dtype=torch.float64, device=device)


def new_cg_state(device):
    return torch.zeros(CG_STATE_DOUBLES, dtype=torch.float64, device=device)


def cg_init_state(state, rr, rtol, kmax):
    check(lib().b200_cg_init_state(_ptr(state), _ptr(rr), rtol, kmax, _stream()))


def cg_update_xr(n, state, p, Ap, x, r):
    check(lib().b200_cg_update_xr_f64(n, _ptr(state), _ptr(p), _ptr(Ap), _ptr(x), _ptr(r), _stream()))


def cg_update_p(n, state, r, p, hist=None):
    check(lib().b200_cg_update_p_f64(n, _ptr(state), _ptr(r), _ptr(p), _ptr(hist),
                                     0 if hist is None else hist.numel(), _stream()))


def laplace3d(nx, ny, nzg, z0, z1, stencil, mode, device="cuda"):
    """Device-generated Laplacian blocks (b200_gen_laplace3d_*).

    Returns a dict with local = (rowptr, colind, values), remote = (...) or None,
    diag or None, rows, ncols_local, nghosts.
    """
    rows = nx * ny * (z1 - z0)
    rp_l = torch.empty(rows + 1, dtype=torch.int32, device=device)
    rp_r = torch.empty(rows + 1, dtype=torch.int32, device=device) if mode != 0 else None
    nl, nr = C.c_int64(), C.c_int64()
    check(lib().b200_gen_laplace3d_rowptr(nx, ny, nzg, z0, z1, stencil, mode, _ptr(rp_l), _ptr(rp_r),
                                          C.byref(nl), C.byref(nr), _stream()))
    ci_l = torch.empty(nl.value, dtype=torch.int32, device=device)
    v_l = torch.empty(nl.value, dtype=torch.float64, device=device)
    ci_r = torch.empty(nr.value, dtype=torch.int32, device=device) if mode != 0 else None
    v_r = torch.empty(nr.value, dtype=torch.float64, device=device) if mode != 0 else None
    diag = torch.empty(rows, dtype=torch.float64, device=device) if mode == 1 else None
    check(lib().b200_gen_laplace3d_fill(nx, ny, nzg, z0, z1, stencil, mode, _ptr(rp_l), _ptr(ci_l), _ptr(v_l),
                                        _ptr(diag), _ptr(rp_r), _ptr(ci_r), _ptr(v_r), _stream()))
    nghosts = (nx * ny if z0 > 0 else 0) + (nx * ny if z1 < nzg else 0)
    return {
        "local": (rp_l, ci_l, v_l),
        "remote": (rp_r, ci_r, v_r) if mode != 0 else None,
        "diag": diag,
        "rows": rows,
        "ncols_local": rows,
        "nghosts": nghosts,
    }
