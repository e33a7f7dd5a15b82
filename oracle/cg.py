"""Restatement of the reference's Conjugate Gradient with a residual history (TEST INFRASTRUCTURE).

Follows spmv/cg.cpp:21-98 statement by statement -- same scalar expressions (alpha from the
*rooted* norms squared again, cg.cpp:66; beta likewise, :78), same vector updates through the
restated BLAS-1 of oracle/restated/kernels.c (ddot in twice the working precision, Dot2) -- but keeps ||r_k|| of every
iteration, which the reference computes and throws away (cg.cpp:76-77).  `ranks` > 1 restates the
MPI version: rows are cut with owner_ranges (read_petsc.cpp:20-36), every dot product is summed
per rank and the (rounded) partial sums are folded in rank order, which is what the oracle
build's MPI shim does for MPI_Allreduce.

Pinned by tests/test_oracle.py::test_restated_cg_equals_reference against the reference's own
cg.cpp (oracle/_ref, tier 2): identical iteration counts and bit-identical solutions for a ladder
of kmax values, i.e. the same trajectory.  What stays UNPINNED is the reference's choice of BLAS
(un-vendored, CMakeLists.txt:71): the accuracy and summation order of cblas_ddot are ours
(oracle/shims/cblas.h).
"""
import numpy as np

from . import kernels as OK


def _owner_ranges(size, n):
    q, r = divmod(int(n), int(size))
    return [rank * (q + 1) if rank < r else rank * q + r for rank in range(size + 1)]


def _dot(x, y, cuts):
    s = 0.0
    for a, b in zip(cuts[:-1], cuts[1:]):
        s += OK.ddot(x[a:b], y[a:b])  # rank-local partial, then the fold in rank order
    return s


def cg(mult, b, kmax, rtol, ranks=1):
    """mult(p) -> A p (the caller's restated SpMV, e.g. oracle.kernels.spmv on the global CSR).
    Returns (iterations, x, history) with history[k-1] = ||r_k||_2 for k = 1..iterations."""
    b = np.ascontiguousarray(b, dtype=np.float64)
    n = len(b)
    cuts = _owner_ranges(ranks, n)
    r = b.copy()                                   # cg.cpp:44
    p = b.copy()                                   # cg.cpp:45
    x = np.zeros(n)                                # x0 = 0 (cg.cpp:41 leaves it to the allocator)
    rnorm0 = np.sqrt(_dot(r, r, cuts))             # cg.cpp:47-50
    rnorm_old = rnorm0
    hist = []
    k = 0
    while k < kmax:                                # cg.cpp:55
        k += 1
        Ap = mult(p)                               # cg.cpp:59-60
        pdotAp = _dot(p, Ap, cuts)                 # cg.cpp:63-65
        alpha = (rnorm_old * rnorm_old) / pdotAp   # cg.cpp:66
        OK.daxpy(alpha, p, x)                      # cg.cpp:69
        OK.daxpy(-alpha, Ap, r)                    # cg.cpp:70
        rnorm_new = np.sqrt(_dot(r, r, cuts))      # cg.cpp:73-76
        beta = (rnorm_new * rnorm_new) / (rnorm_old * rnorm_old)  # cg.cpp:77
        rnorm_old = rnorm_new
        hist.append(float(rnorm_new))
        if rnorm_new / rnorm0 < rtol:              # cg.cpp:80-81
            break
        OK.dscal(beta, p)                          # cg.cpp:84
        OK.daxpy(1.0, r, p)                        # cg.cpp:85
    return k, x, np.array(hist)
