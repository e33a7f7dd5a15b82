"""CPU: pin the oracle (plain-C restatement) against the reference's golden values
and, where oracle/_ref was built, against the reference's own compiled kernels."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import kernels as K
from tests.fixtures import (FIX_COLIND, FIX_NORM, FIX_ROWPTR, FIX_SYM_COLIND, FIX_SYM_DIAG, FIX_SYM_ROWPTR,
                            FIX_SYM_VALUES, FIX_VALUES, FIX_X, FIX_Y, essentially_equal, rel_l2, split_symmetric)

needs_ref = pytest.mark.skipif(not K.have_ref(), reason="oracle/_ref not built (no /root/reference here)")


def test_fixture_general_golden():
    y = K.spmv(FIX_ROWPTR, FIX_COLIND, FIX_VALUES, FIX_X)
    assert np.array_equal(y, FIX_Y)  # bit-exact vs Appendix D
    assert essentially_equal(np.linalg.norm(y), FIX_NORM)  # the reference's own pass criterion


def test_fixture_symmetric_golden():
    y = K.symv(FIX_SYM_ROWPTR, FIX_SYM_COLIND, FIX_SYM_VALUES, FIX_SYM_DIAG, FIX_X)
    assert np.array_equal(y, FIX_Y)
    assert essentially_equal(np.linalg.norm(y), FIX_NORM)


def test_split_symmetric_matches_fixture():
    rp, ci, v, d = split_symmetric(FIX_ROWPTR, FIX_COLIND, FIX_VALUES)
    assert np.array_equal(rp, FIX_SYM_ROWPTR) and np.array_equal(ci, FIX_SYM_COLIND)
    assert np.array_equal(v, FIX_SYM_VALUES) and np.array_equal(d, FIX_SYM_DIAG)


def test_diag_only_symmetric():
    # rank 1 of the 2-rank symmetric split: nnz == 0, rowptr never allocated (csr_matrix.cpp:35)
    x = np.array([0.5, -2.0])
    y = K.symv(None, None, None, np.array([8.0, 8.0]), x, alpha=1.0, beta=0.0)
    assert np.array_equal(y, np.array([4.0, -16.0]))


def _random_csr(n, m, density, seed):
    A = sp.random(n, m, density=density, format="csr", random_state=seed, dtype=np.float64)
    A.sort_indices()
    return A


@needs_ref
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.5, 1.0), (-2.0, 0.25)])
def test_restated_general_equals_reference(alpha, beta):
    rng = np.random.default_rng(7)
    A = _random_csr(1500, 1700, 0.01, 3)
    x, y0 = rng.standard_normal(1700), rng.standard_normal(1500)
    mine = K.spmv(A.indptr, A.indices, A.data, x, alpha, beta, y0)
    ref = K.ref_run(1500, 1700, A.indptr, A.indices, A.data, None, False, x, alpha, beta, y0)
    assert np.array_equal(mine, ref)


@needs_ref
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.5, 1.0)])
def test_restated_symmetric_equals_reference(alpha, beta):
    rng = np.random.default_rng(8)
    A = _random_csr(1200, 1200, 0.01, 5)
    S = (A + A.T + sp.eye(1200) * 3).tocsr()
    S.sort_indices()
    rp, ci, v, d = split_symmetric(S.indptr, S.indices, S.data)
    x, y0 = rng.standard_normal(1200), rng.standard_normal(1200)
    mine = K.symv(rp, ci, v, d, x, alpha, beta, y0)
    ref = K.ref_run(1200, 1200, rp, ci, v, d, True, x, alpha, beta, y0)
    assert np.array_equal(mine, ref)
    if beta == 0.0:
        full = K.spmv(S.indptr, S.indices, S.data, x, alpha, 0.0)
        assert rel_l2(mine, full) < 1e-13


@needs_ref
def test_reference_openmp_kernel_agrees():
    rng = np.random.default_rng(9)
    A = _random_csr(3000, 3000, 0.004, 11)
    x = rng.standard_normal(3000)
    want = K.spmv(A.indptr, A.indices, A.data, x)
    op = K.RefOmp(3000, 3000, A.indptr, A.indices, A.data, threads=4)
    y = np.zeros(3000)
    op.run(x.copy(), y)
    op.close()
    assert np.array_equal(y, want)  # same per-row order, rows split across threads


def test_scipy_cross_check():
    # scipy's csr matvec accumulates per row left to right like csr_kernels.cpp:42-49
    rng = np.random.default_rng(1)
    A = _random_csr(800, 900, 0.02, 2)
    x = rng.standard_normal(900)
    assert rel_l2(K.spmv(A.indptr, A.indices, A.data, x), A @ x) < 1e-15


def test_blas1_restatement():
    rng = np.random.default_rng(2)
    x, y = rng.standard_normal(1000), rng.standard_normal(1000)
    s = 0.0
    for a, b in zip(x, y):
        s += a * b
    assert K.ddot(x, y) == s
    y2 = y.copy()
    K.daxpy(0.3, x, y2)
    assert np.array_equal(y2, y + 0.3 * x)
    K.dscal(-1.5, y2)
    assert np.array_equal(y2, (y + 0.3 * x) * -1.5)
