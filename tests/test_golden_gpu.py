"""GPU parity against the committed golden vectors (tests/golden/, produced by the reference's own
compiled code -- see make_golden.py), through the reference-shaped API (spmv_b200.api).

Tolerances are BASELINE.json's: y within 1e-12 relative L2; CG iteration counts within +-1 and the
residual history within 1e-10 relative, iteration by iteration."""
import functools
import os

import numpy as np
import pytest
import torch

from spmv_b200 import api
from tests.fixtures import (FIX_COLIND, FIX_ROWPTR, FIX_VALUES, rel_l2)

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CG_GOLDEN = ["lap27_6x5x8", "lap7_12x12x12", "powerlaw_3000"]
TOL_Y = 1e-12
TOL_HIST = 1e-10


@pytest.fixture(scope="module")
def ex():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    e = api.B200Executor.create(0)
    e.use_torch_stream()
    yield e
    e.close()


def _mult(ex, A, x_host):
    cm = A.col_map()
    n = cm.local_size() + cm.num_ghosts()
    x = torch.zeros(n, dtype=torch.float64, device="cuda")
    x[: len(x_host)] = torch.from_numpy(x_host).cuda()
    y = torch.full((A.rows(),), float("nan"), dtype=torch.float64, device="cuda")
    cm.update(x)
    A.mult(x, y)
    ex.synchronize()
    return y.cpu().numpy()


@pytest.mark.parametrize("symmetric", [False, True])
def test_fixture_through_matrix_api(ex, symmetric):
    """tests/test_spmv.cpp:56-160 at one rank: create_matrix on the 5x5 fixture, update, mult."""
    g = np.load(os.path.join(GOLDEN, "fixture_5x5.npz"))
    A = api.Matrix.create_matrix(ex, FIX_ROWPTR, FIX_COLIND, FIX_VALUES, 5, 5, [], [], symmetric, api.P2P_BLOCKING)
    y = _mult(ex, A, g["x"])
    want = g["y_symmetric"] if symmetric else g["y_general"]
    assert rel_l2(y, want) < TOL_Y
    if not symmetric:
        assert np.array_equal(y, want)  # one thread per row in the reference's order
    # the reference's own pass criterion (tests/test_spmv.cpp:20-23,159-160)
    assert abs(np.linalg.norm(y) - float(g["norm"])) <= min(np.linalg.norm(y), float(g["norm"])) * 2.3e-16 * 4
    A.close()


@pytest.mark.parametrize("symmetric", [False, True])
@pytest.mark.parametrize("name", CG_GOLDEN)
def test_cg_history_against_golden(ex, name, symmetric):
    g = np.load(os.path.join(GOLDEN, f"cg_{name}.npz"))
    n = len(g["b"])
    A = api.Matrix.create_matrix(ex, g["rowptr"], g["colind"], g["values"], n, n, [], [], symmetric, api.P2P_BLOCKING)
    b = torch.from_numpy(g["b"]).cuda()
    x = torch.zeros(n, dtype=torch.float64, device="cuda")
    it, hist = api.cg(ex, A, b, x, int(g["kmax"]), float(g["rtol"]), history=True)
    want_it, want_hist = int(g["iterations"]), g["history"]
    assert abs(it - want_it) <= 1, (it, want_it)
    m = min(it, want_it)
    dev = np.abs(hist[:m] - want_hist[:m]) / want_hist[:m]
    max_row = int(np.diff(g["rowptr"]).max())
    print(f"cg {name} symmetric={symmetric}: iterations {it} (golden {want_it}), relative history deviation "
          f"max {dev.max():.3e}, first 20 iterations {dev[:20].max():.3e}")
    if not symmetric and max_row <= 32:
        # SpMV is bit-identical to the reference kernel (one thread per row, its order), the
        # vector updates are un-contracted and the dot products are order-independent (Dot2 on
        # both sides): the whole trajectory is the reference's, bit for bit
        assert it == want_it and np.array_equal(hist, want_hist)
        assert np.array_equal(x.cpu().numpy(), g["x"])
    else:
        # symmetric storage (atomic adds land in a different order) or rows walked by a warp perturb
        # y in the last bit, and CG roughly doubles a perturbation every iteration: the 1e-10 bound
        # holds over the first ~20 iterations and through the whole solve of the two stencil cases
        # (34 iterations at most); over the 57-iteration irregular solve only the iteration count and
        # the solution are reproducible
        assert dev[:20].max() < TOL_HIST, dev[:20].max()
        if name != "powerlaw_3000":
            assert dev.max() < TOL_HIST, dev.max()
    assert rel_l2(x.cpu().numpy(), g["x"]) < 1e-9
    A.close()


# ---------------------------------------------------------------------------------------------
# Matrix::mult_host: host vectors in, copies pipelined against tile ranges.  Same kernels and the
# same per-row order as mult, so general storage must be bit-identical to update + mult; the
# symmetric kernel's adds land in another order across chunks, so the BASELINE tolerance applies.
# ---------------------------------------------------------------------------------------------
@functools.lru_cache(maxsize=None)
def _host_case(kind):
    from tests import cases
    if kind == "random":  # columns all over the place: every chunk needs all of x, y rows free early
        import scipy.sparse as sp
        n = 60000
        rng = np.random.default_rng(5)
        i, j = rng.integers(0, n, 12 * n), rng.integers(0, n, 12 * n)
        A = sp.coo_matrix((rng.standard_normal(12 * n), (i, j)), shape=(n, n)).tocsr()
        A = (A + A.T + sp.eye(n) * 40.0).tocsr()
        A.sort_indices()
        return n, A.indptr.astype(np.int32), A.indices.astype(np.int32), A.data
    N, rows = cases.make_problem(kind, (48, 48, 64) if kind == "lap7" else (40, 40, 48))
    rp, col, val = rows(0, N)
    return N, rp.astype(np.int32), col.astype(np.int32), val


@pytest.mark.parametrize("chunks", [1, 3, 8, 64])
@pytest.mark.parametrize("symmetric", [False, True])
@pytest.mark.parametrize("kind", ["lap7", "lap27", "random"])
def test_mult_host_equals_mult(ex, kind, symmetric, chunks, monkeypatch):
    from oracle import kernels as K
    monkeypatch.setenv("B200_HOST_CHUNKS", str(chunks))
    n, rp, ci, v = _host_case(kind)
    A = api.Matrix.create_matrix(ex, rp, ci, v, n, n, [], [], symmetric, api.P2P_BLOCKING)
    rng = np.random.default_rng(3)
    xh = torch.from_numpy(rng.standard_normal(n)).pin_memory()
    yh = torch.full((n,), float("nan"), dtype=torch.float64).pin_memory()
    want_dev = _mult(ex, A, xh.numpy())
    for _ in range(3):  # repeated calls reuse the device vectors and the chunk table
        yh.fill_(float("nan"))
        A.mult_host(xh.numpy(), yh.numpy())
        got = yh.numpy()
        if symmetric:
            assert rel_l2(got, want_dev) < TOL_Y
        else:
            assert np.array_equal(got, want_dev)
    assert rel_l2(got, K.spmv(rp, ci, v, xh.numpy())) < TOL_Y  # and the oracle
    A.close()


def test_mult_host_pageable_and_tiny(ex):
    """Plain numpy (pageable) vectors and the 5x5 fixture (one tile, fewer tiles than chunks)."""
    g = np.load(os.path.join(GOLDEN, "fixture_5x5.npz"))
    for symmetric in (False, True):
        A = api.Matrix.create_matrix(ex, FIX_ROWPTR, FIX_COLIND, FIX_VALUES, 5, 5, [], [], symmetric, api.P2P_BLOCKING)
        y = np.full(5, np.nan)
        A.mult_host(np.ascontiguousarray(g["x"]), y)
        assert rel_l2(y, g["y_general"]) < TOL_Y
        A.close()


# ---------------------------------------------------------------------------------------------
# CG trajectory at a size where rounding has room to show: 27-point 64^3 (262 144 rows), 120
# iterations with rtol out of reach, against the reference's own cg.cpp run here through oracle/_ref
# (tier 2, one thread-rank) and against the restated CG's residual history.  General storage with
# stencil rows: iteration count, x and every ||r_k|| bit-identical.
# ---------------------------------------------------------------------------------------------
def test_cg_bit_identical_to_reference_64cubed(ex):
    from oracle import cg as OCG
    from oracle import kernels as OK
    from oracle import ref_api as R
    from spmv_b200 import synth
    n1, kmax = 64, 120
    N = n1 ** 3
    rp, col, val = synth.laplace3d_rows(n1, n1, n1, 0, n1, 27)
    rp32, c32 = rp.astype(np.int32), col.astype(np.int32)
    mult = lambda p: OK.spmv(rp32, c32, val, p)  # noqa: E731
    b_h = mult(np.ones(N))
    it_o, x_o, hist_o = OCG.cg(mult, b_h, kmax, 1e-30)
    assert it_o == kmax
    if R.available():  # the reference's own loop (cg.cpp:21-98) on its own Matrix
        res = {}

        def body(rank):
            M = R.RefMatrix.create_matrix(rp32, c32, val, N, N, [], [], False, 2)
            res["cg"] = M.cg(b_h, kmax, 1e-30)
            M.close()

        R.run_ranks(1, body)
        assert res["cg"][0] == it_o and np.array_equal(res["cg"][1], x_o)  # restatement == reference
    A = api.Matrix.create_matrix(ex, rp32, c32, val, N, N, [], [], False, api.P2P_BLOCKING)
    b = torch.from_numpy(b_h).cuda()
    x = torch.zeros(N, dtype=torch.float64, device="cuda")
    it, hist = api.cg(ex, A, b, x, kmax, 1e-30, history=True)
    assert it == it_o
    assert np.array_equal(hist, hist_o)
    assert np.array_equal(x.cpu().numpy(), x_o)
    # and the solve to the BASELINE tolerance stops at the same iteration
    it8, hist8 = api.cg(ex, A, b, x, 5000, 1e-8, history=True)
    it8_o, x8_o, hist8_o = OCG.cg(mult, b_h, 5000, 1e-8)
    assert it8 == it8_o and np.array_equal(hist8, hist8_o) and np.array_equal(x.cpu().numpy(), x8_o)
    A.close()


# ---------------------------------------------------------------------------------------------
# The CG driver's execution modes -- separate vector kernels / one fused cooperative step, stream /
# captured CUDA graph of 8 iterations, p.Ap in the SpMV epilogue or not -- are the same algorithm: same
# iteration count, same residual history and x, bit for bit (general storage, stencil rows), whatever
# kmax is relative to the graph length.  B200_CG_GRAPH_STRICT turns a capture failure into an error
# instead of a silent fall-back to the stream path.
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("graph,fused_vec,fused_dot", [(0, 0, 1), (0, 1, 1), (1, 0, 1), (1, 1, 1), (1, 1, 0), (1, 0, 0)])
def test_cg_modes_agree(ex, graph, fused_vec, fused_dot, monkeypatch):
    from oracle import cg as OCG
    from oracle import kernels as OK
    from spmv_b200 import synth
    n1 = 20
    N = n1 ** 3
    rp, col, val = synth.laplace3d_rows(n1, n1, n1, 0, n1, 27)
    rp32, c32 = rp.astype(np.int32), col.astype(np.int32)
    mult = lambda p: OK.spmv(rp32, c32, val, p)  # noqa: E731
    b_h = mult(np.ones(N))
    monkeypatch.setenv("B200_CG_GRAPH", str(graph))
    monkeypatch.setenv("B200_CG_GRAPH_STRICT", "1")
    monkeypatch.setenv("B200_CG_FUSED_VEC", str(fused_vec))
    monkeypatch.setenv("B200_CG_FUSED_DOT", str(fused_dot))
    A = api.Matrix.create_matrix(ex, rp32, c32, val, N, N, [], [], False, api.P2P_BLOCKING)
    b = torch.from_numpy(b_h).cuda()
    x = torch.zeros(N, dtype=torch.float64, device="cuda")
    for kmax, rtol in ((1, 1e-30), (7, 1e-30), (8, 1e-30), (21, 1e-30), (500, 1e-10), (3, 1e-30)):
        it_o, x_o, hist_o = OCG.cg(mult, b_h, kmax, rtol)
        x.fill_(float("nan"))
        it, hist = api.cg(ex, A, b, x, kmax, rtol, history=True)
        assert it == it_o, (kmax, it, it_o)
        assert np.array_equal(hist, hist_o), kmax
        assert np.array_equal(x.cpu().numpy(), x_o), kmax
    it, ph = api.cg_phases(ex, A, b, x, 12, 1e-30)  # profile mode: same count, sane timings
    assert it == 12 and ph["local"] > 0 and ph["vector"] > 0
    A.close()
    # a second operator at the same address range must not replay the first one's graph
    A2 = api.Matrix.create_matrix(ex, rp32, c32, 2.0 * val, N, N, [], [], False, api.P2P_BLOCKING)
    it2, hist2 = api.cg(ex, A2, b, x, 500, 1e-10, history=True)
    it2_o, x2_o, hist2_o = OCG.cg(lambda p: OK.spmv(rp32, c32, 2.0 * val, p), b_h, 500, 1e-10)
    assert it2 == it2_o and np.array_equal(hist2, hist2_o) and np.array_equal(x.cpu().numpy(), x2_o)
    A2.close()


def test_cg_symmetric_modes(ex, monkeypatch):
    """Symmetric storage (two kernels per SpMV, scatter adds) through the graph + fused step: same iteration
    count as the stream path and the oracle (+-1), history within 1e-10 over the first 20 iterations."""
    from oracle import cg as OCG
    from oracle import kernels as OK
    from spmv_b200 import synth
    n1 = 16
    N = n1 ** 3
    rp, col, val = synth.laplace3d_rows(n1, n1, n1, 0, n1, 7)
    rp32, c32 = rp.astype(np.int32), col.astype(np.int32)
    mult = lambda p: OK.spmv(rp32, c32, val, p)  # noqa: E731
    b_h = mult(np.ones(N))
    it_o, x_o, hist_o = OCG.cg(mult, b_h, 500, 1e-10)
    A = api.Matrix.create_matrix(ex, rp32, c32, val, N, N, [], [], True, api.P2P_BLOCKING)
    b = torch.from_numpy(b_h).cuda()
    x = torch.zeros(N, dtype=torch.float64, device="cuda")
    monkeypatch.setenv("B200_CG_GRAPH_STRICT", "1")
    # fd: x'Ax in the symmetric SpMV's epilogue (b200_csr_symv_dot_f64) against the separate dot kernel
    for graph, fv, fd in ((0, 0, 0), (1, 1, 0), (1, 0, 1), (0, 1, 1), (1, 1, 1), (0, 0, 1)):
        monkeypatch.setenv("B200_CG_GRAPH", str(graph))
        monkeypatch.setenv("B200_CG_FUSED_VEC", str(fv))
        monkeypatch.setenv("B200_CG_FUSED_DOT", str(fd))
        it, hist = api.cg(ex, A, b, x, 500, 1e-10, history=True)
        assert abs(it - it_o) <= 1
        m = min(20, it, it_o)
        assert np.max(np.abs(hist[:m] - hist_o[:m]) / hist_o[:m]) < TOL_HIST
        assert rel_l2(x.cpu().numpy(), x_o) < 1e-9
    A.close()
