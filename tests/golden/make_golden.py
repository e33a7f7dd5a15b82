#!/usr/bin/env python
"""make_golden.py -- regenerate the committed golden vectors (run where /root/reference exists).

Every array below is produced by the REFERENCE's own code compiled unmodified (oracle/_ref, see
oracle/Makefile), except the CG residual histories, which the reference computes but does not
keep (cg.cpp:76-77): those come from the restatement oracle/cg.py, after checking here that the
restatement and the reference's cg.cpp return the same iteration count and bit-identical x.

    python tests/golden/make_golden.py        # writes tests/golden/*.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import cg as OCG  # noqa: E402
from oracle import kernels as OK  # noqa: E402
from oracle import ref_api as R  # noqa: E402
from tests import cases  # noqa: E402
from tests.fixtures import (FIX_COLIND, FIX_ROWPTR, FIX_SYM_COLIND, FIX_SYM_DIAG, FIX_SYM_ROWPTR,  # noqa: E402
                            FIX_SYM_VALUES, FIX_VALUES, FIX_X)

PLAN_KEYS = cases.PLAN_KEYS
CG_CASES = {"lap27_6x5x8": ("lap27", (6, 5, 8)), "lap7_12x12x12": ("lap7", (12, 12, 12)),
            "powerlaw_3000": ("powerlaw_spd", 3000)}


def spd_powerlaw(n):
    """Irregular SPD operator: off-diagonal part of A + A^T of the power-law matrix, diagonal
    1 + sum |off-diagonal| of the row (strictly diagonally dominant, so CG converges steadily and
    reduction-order noise is not amplified)."""
    import scipy.sparse as sp
    from spmv_b200 import synth
    rp, col, val = synth.powerlaw_rows(n, 0, n, band=max(8, n // 16))
    A = sp.csr_matrix((val, col, rp), shape=(n, n))
    A.setdiag(0.0)
    A.eliminate_zeros()
    off = (A + A.T).tocsr()
    B = (off + sp.diags(1.0 + np.asarray(abs(off).sum(axis=1)).ravel())).tocsr()
    B.sort_indices()
    return B.indptr.astype(np.int64), B.indices.astype(np.int64), B.data.copy()


def problem(kind, n):
    if kind == "powerlaw_spd":
        rp, col, val = spd_powerlaw(n)

        def rows(r0, r1):
            return rp[r0:r1 + 1] - rp[r0], col[rp[r0]:rp[r1]], val[rp[r0]:rp[r1]]
        return n, rows
    return cases.make_problem(kind, n)


def main():
    assert R.available() and OK.have_ref(), "oracle/_ref is not built: run __graft_entry__.build() with /root/reference"
    # 1. the reference test's 5x5 fixture (tests/test_spmv.cpp:56-70), both kernels of csr_kernels.cpp
    y_gen = OK.ref_run(5, 5, FIX_ROWPTR, FIX_COLIND, FIX_VALUES, None, False, FIX_X)
    y_sym = OK.ref_run(5, 5, FIX_SYM_ROWPTR, FIX_SYM_COLIND, FIX_SYM_VALUES, FIX_SYM_DIAG, True, FIX_X)
    np.savez(os.path.join(HERE, "fixture_5x5.npz"), x=FIX_X, y_general=y_gen, y_symmetric=y_sym,
             norm=np.linalg.norm(y_gen))

    # 2. halo plans + block splits of the reference at 2 and 3 ranks (L2GMap.cpp:346-479, Matrix.cpp:164-480)
    out = {}
    for kind, n in (("fixture", None), ("create_a", 1001), ("lap7", (4, 3, 6))):
        N, rows = cases.make_problem(kind, n)
        for size in (2, 3):
            for sym, cm in ((False, 2), (True, 2), (False, 1)):
                res = {}

                def body(rank):
                    inp = cases.local_inputs(rows, N, rank, size)
                    A = R.RefMatrix.create_matrix(inp["rowptr"], inp["colind"], inp["values"], inp["nrows_local"],
                                                  inp["ncols_local"], [], inp["col_ghosts"], sym, cm)
                    res[rank] = (A.plan(0), A.block(0), A.block(1), A.info())
                    A.close()

                R.run_ranks(size, body)
                for rank in range(size):
                    plan, loc, rem, info = res[rank]
                    tag = f"{cases.problem_name(kind, n)}/s{size}/sym{int(sym)}/cm{cm}/r{rank}"
                    for k in PLAN_KEYS:
                        out[f"{tag}/plan/{k}"] = np.asarray(plan[k])
                    for name, blk in (("local", loc), ("remote", rem)):
                        if blk is None:
                            continue
                        for k in ("rowptr", "colind", "values"):
                            out[f"{tag}/{name}/{k}"] = blk[k]
                        out[f"{tag}/{name}/dims"] = np.array([blk["rows"], blk["cols"]])
                        if blk["diagonal"] is not None:
                            out[f"{tag}/{name}/diagonal"] = blk["diagonal"]
                    out[f"{tag}/non_zeros"] = np.array(info["non_zeros"])
    np.savez_compressed(os.path.join(HERE, "plans_and_blocks.npz"), **out)

    # 3. CG: iteration count and x from the reference's cg.cpp at 1 rank, history from the restatement
    for name, (kind, n) in CG_CASES.items():
        N, rows = problem(kind, n)
        rp, col, val = rows(0, N)
        rp32, c32 = rp.astype(np.int32), col.astype(np.int32)
        mult = lambda p: OK.spmv(rp32, c32, val, p)  # noqa: E731
        b = mult(np.ones(N))
        res = {}

        def body(rank):
            inp = cases.local_inputs(rows, N, rank, 1)
            A = R.RefMatrix.create_matrix(inp["rowptr"], inp["colind"], inp["values"], inp["nrows_local"],
                                          inp["ncols_local"], [], inp["col_ghosts"], False, 2)
            res["full"] = A.cg(b, 500, 1e-10)
            res["k7"] = A.cg(b, 7, 1e-30)
            A.close()

        R.run_ranks(1, body)
        it, x, hist = OCG.cg(mult, b, 500, 1e-10)
        it7, x7, _ = OCG.cg(mult, b, 7, 1e-30)
        assert it == res["full"][0] and np.array_equal(x, res["full"][1]), name
        assert it7 == res["k7"][0] and np.array_equal(x7, res["k7"][1]), name
        np.savez_compressed(os.path.join(HERE, f"cg_{name}.npz"), rowptr=rp32, colind=c32, values=val, b=b,
                            iterations=np.array(it), x=x, history=hist, kmax=np.array(500), rtol=np.array(1e-10))
        print(f"cg_{name}: N={N} iterations={it} final rel residual {hist[-1] / np.linalg.norm(b):.3e}")
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
