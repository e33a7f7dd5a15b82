"""Problem builders shared by the parity tests: a global operator is described by a
function rows(r0, r1) -> (rowptr, global colind, values); `local_inputs` turns one
rank's slice into the arguments of Matrix::create_matrix the way the reference's own
test does (tests/test_spmv.cpp:84-129)."""
import numpy as np

from spmv_b200 import synth
from tests.fixtures import FIX_COLIND, FIX_ROWPTR, FIX_VALUES


def fixture_rows(r0, r1):
    rp = (FIX_ROWPTR[r0:r1 + 1] - FIX_ROWPTR[r0]).astype(np.int64)
    sl = slice(FIX_ROWPTR[r0], FIX_ROWPTR[r1])
    return rp, FIX_COLIND[sl].astype(np.int64), FIX_VALUES[sl].copy()


def problem_name(kind, n=None):
    """Key of a problem in tests/golden/plans_and_blocks.npz."""
    if n is None:
        return kind
    return kind + "_" + ("x".join(str(v) for v in n) if isinstance(n, tuple) else str(n))


def make_problem(kind, n=None):
    """-> (N_global, rows_fn)"""
    if kind == "fixture":
        return 5, fixture_rows
    if kind in ("lap7", "lap27"):
        nx, ny, nz = n
        st = 7 if kind == "lap7" else 27
        plane = nx * ny

        def rows(r0, r1):
            # whole planes when the split allows it, else generate a cover and cut
            z0, z1 = r0 // plane, -(-r1 // plane)
            rp, col, val = synth.laplace3d_rows(nx, ny, nz, z0, max(z1, z0), st)
            a, b = r0 - z0 * plane, r1 - z0 * plane
            return (rp[a:b + 1] - rp[a]), col[rp[a]:rp[b]], val[rp[a]:rp[b]]

        return nx * ny * nz, rows
    if kind == "powerlaw":
        return n, lambda r0, r1: synth.powerlaw_rows(n, r0, r1, band=max(8, n // 16))
    if kind == "create_a":
        return n, lambda r0, r1: synth.create_a_rows(n, r0, r1)
    raise ValueError(kind)


def local_inputs(rows_fn, N, rank, size):
    rng = synth.owner_ranges(size, N)
    r0, r1 = int(rng[rank]), int(rng[rank + 1])
    rp, gcol, val = rows_fn(r0, r1)
    lcol, ghosts = synth.localize_columns(np.asarray(gcol, dtype=np.int64), r0, r1)
    return {"rowptr": np.asarray(rp, dtype=np.int32), "colind": lcol, "values": np.asarray(val, dtype=np.float64),
            "nrows_local": r1 - r0, "ncols_local": r1 - r0, "col_ghosts": ghosts, "r0": r0, "r1": r1}


def same_block(a, b):
    if a is None or b is None:
        return a is None and b is None
    ok = a["rows"] == b["rows"] and a["cols"] == b["cols"]
    for k in ("rowptr", "colind", "values"):
        ok = ok and np.array_equal(a[k], b[k])
    if a["diagonal"] is None or b["diagonal"] is None:
        return ok and a["diagonal"] is None and b["diagonal"] is None
    return ok and np.array_equal(a["diagonal"], b["diagonal"])


PLAN_KEYS = ("ranges", "ghosts", "neighbours", "send_count", "recv_count", "send_offset", "recv_offset", "indexbuf",
             "local_size", "num_ghosts", "global_size", "global_offset")


def same_plan(a, b):
    return all(np.array_equal(a[k], b[k]) for k in PLAN_KEYS)


def random_uneven_inputs(seed, size=None):
    """Seeded random sparsity on an UNEVEN row partition: per-rank create_matrix inputs (the keys of
    local_inputs) for `size` ranks (2 - 6, drawn from the seed when None).  One rank owns a single row; with
    seed % 3 == 0 the last rank ghosts nothing.  Rows are sorted and de-duplicated, values standard normal."""
    from spmv_b200 import synth
    rng = np.random.default_rng(1000 + seed)
    drawn = int(rng.integers(2, 7))
    size = drawn if size is None else size
    sizes = rng.integers(1, 300, size=size)
    sizes[rng.integers(0, size)] = 1
    ranges = np.concatenate([[0], np.cumsum(sizes)])
    n = int(ranges[-1])
    inps = []
    for r in range(size):
        r0, r1 = int(ranges[r]), int(ranges[r + 1])
        m = r1 - r0
        if seed % 3 == 0 and r == size - 1:
            lens = rng.integers(1, 4, size=m)
            gcol = np.concatenate([rng.integers(r0, r1, size=k) for k in lens])
        else:
            lens = rng.integers(0, 9, size=m)
            tot = int(lens.sum())
            near = rng.integers(max(0, r0 - 40), min(n, r1 + 40), size=tot)
            far = rng.integers(0, n, size=tot)
            gcol = np.where(rng.random(tot) < 0.8, near, far)
        rows_of = np.repeat(np.arange(m), lens)
        order = np.lexsort((gcol, rows_of))
        rows_of, gcol = rows_of[order], gcol[order]
        keep = np.ones(len(gcol), bool)
        keep[1:] = (rows_of[1:] != rows_of[:-1]) | (gcol[1:] != gcol[:-1])
        rows_of, gcol = rows_of[keep], gcol[keep]
        rowptr = np.concatenate([[0], np.cumsum(np.bincount(rows_of, minlength=m))]).astype(np.int32)
        lcol, ghosts = synth.localize_columns(gcol.astype(np.int64), r0, r1)
        inps.append({"rowptr": rowptr, "colind": lcol, "values": rng.standard_normal(len(lcol)), "nrows_local": m,
                     "ncols_local": m, "col_ghosts": ghosts, "r0": r0, "r1": r1})
    return inps
