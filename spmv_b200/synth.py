"""Deterministic synthetic inputs for the configs BASELINE.json names (host side).

The reference ships one synthetic operator, the 1-D ``create_A``
(demos/CreateA.cpp:31-103); the 3-D Laplacians and the power-law matrix of
configs 2-5 are this project's generators.  All functions
return host CSR for a slice of rows with GLOBAL column indices, sorted within
each row, ready for ``Matrix::create_matrix`` or the PETSc writer.
"""
import numpy as np


def owner_ranges(size, n):
    """read_petsc.cpp:20-36 / demos/CreateA.cpp:13-29: first n % size ranks get one more."""
    q, r = divmod(int(n), int(size))
    return np.array([rank * (q + 1) if rank < r else rank * q + r for rank in range(size + 1)], dtype=np.int64)


def bump_vector(n_global, r0, r1):
    """x_i = exp(-10 (5 (i/N - 0.5))^2), demos/spmv.cpp:63-67, tests/test_spmv.cpp:66-70."""
    i = np.arange(r0, r1, dtype=np.float64)
    z = i / float(n_global)
    return np.exp(-10.0 * np.power(5.0 * (z - 0.5), 2.0))


def create_a_rows(n, r0, r1, gamma=0.1):
    """Rows [r0, r1) of the create_A operator (demos/CreateA.cpp:52-67)."""
    rows = np.arange(r0, r1, dtype=np.int64)
    cnt = np.where((rows == 0) | (rows == n - 1), 2, 3)
    if n == 1:
        cnt[:] = 2  # the reference would index column 1 of a 1x1 matrix; not a case we use
    rowptr = np.zeros(len(rows) + 1, dtype=np.int64)
    np.cumsum(cnt, out=rowptr[1:])
    col = np.empty(rowptr[-1], dtype=np.int64)
    val = np.empty(rowptr[-1], dtype=np.float64)
    first = rows == 0
    last = rows == n - 1
    mid = ~(first | last)
    s = rowptr[:-1]
    col[s[first]], val[s[first]] = 0, 1.0 - gamma
    col[s[first] + 1], val[s[first] + 1] = 1, gamma
    col[s[last]], val[s[last]] = n - 2, gamma
    col[s[last] + 1], val[s[last] + 1] = n - 1, 1.0 - gamma
    m = rows[mid]
    col[s[mid]], val[s[mid]] = m - 1, gamma
    col[s[mid] + 1], val[s[mid] + 1] = m, 1.0 - 2.0 * gamma
    col[s[mid] + 2], val[s[mid] + 2] = m + 1, gamma
    return rowptr, col, val


def laplace3d_rows(nx, ny, nzg, z0, z1, stencil):
    """Rows of planes [z0, z1) of the 7- or 27-point Laplacian on nx*ny*nzg.

    Index i = x + nx*(y + ny*z); diagonal 6 (7-pt) or 26 (27-pt), neighbours -1,
    Dirichlet truncation at the faces.
    """
    assert stencil in (7, 27)
    z, y, x = np.meshgrid(np.arange(z0, z1), np.arange(ny), np.arange(nx), indexing="ij")
    x, y, z = x.ravel(), y.ravel(), z.ravel()
    nrows = x.size
    cols, vals, valid = [], [], []
    for dz in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                off = (dx != 0) + (dy != 0) + (dz != 0)
                if stencil == 7 and off > 1:
                    continue
                xx, yy, zz = x + dx, y + dy, z + dz
                ok = (xx >= 0) & (xx < nx) & (yy >= 0) & (yy < ny) & (zz >= 0) & (zz < nzg)
                cols.append(xx + nx * (yy + ny * zz.astype(np.int64)))
                vals.append(np.full(nrows, (6.0 if stencil == 7 else 26.0) if off == 0 else -1.0))
                valid.append(ok)
    cols, vals, valid = np.stack(cols, 1), np.stack(vals, 1), np.stack(valid, 1)
    cnt = valid.sum(1)
    rowptr = np.zeros(nrows + 1, dtype=np.int64)
    np.cumsum(cnt, out=rowptr[1:])
    return rowptr, cols[valid].astype(np.int64), vals[valid]


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def _u01(key):
    return (_splitmix64(key) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def powerlaw_rows(n, r0, r1, seed=20261104, alpha=2.0, xmin=4.0, max_len=4096, band=32768, p_global=0.1):
    """Rows [r0, r1) of the irregular power-law matrix of config 5.

    Counter-based (any rank can generate any row): row length = clamp(floor(
    Pareto(alpha, xmin)), 1, max_len) off-diagonal draws; a fraction 1-p_global of
    them uniform in a band +-`band` around the diagonal, the rest uniform over all
    columns; de-duplicated, sorted, diagonal always present; off-diagonal values
    uniform in [-1, 1), diagonal = 1 + sum |off-diagonal| (diagonally dominant).
    """
    with np.errstate(over="ignore"):
        rows = np.arange(r0, r1, dtype=np.uint64)
        base = rows * np.uint64(0x100000000) + np.uint64(seed) * np.uint64(0x9E3779B1)
        u = _u01(base)
        length = np.floor(xmin / np.power(1.0 - u, 1.0 / alpha)).astype(np.int64)
        length = np.clip(length, 1, min(max_len, max(n - 1, 1)))
        tot = int(length.sum())
        owner = np.repeat(np.arange(r1 - r0, dtype=np.int64), length)
        k = np.arange(tot, dtype=np.int64) - np.repeat(np.cumsum(length) - length, length)
        key = base[owner] + (k.astype(np.uint64) + np.uint64(1)) * np.uint64(3)
        sel, pos, valu = _u01(key), _u01(key + np.uint64(1)), _u01(key + np.uint64(2))
    grow = owner + r0
    lo = np.maximum(grow - band, 0)
    hi = np.minimum(grow + band + 1, n)
    c_band = lo + np.floor(pos * (hi - lo)).astype(np.int64)
    c_glob = np.floor(pos * n).astype(np.int64)
    col = np.where(sel < p_global, c_glob, c_band)
    col = np.clip(col, 0, n - 1)
    val = 2.0 * valu - 1.0
    keep = col != grow
    owner, col, val = owner[keep], col[keep], val[keep]
    order = np.lexsort((col, owner))
    owner, col, val = owner[order], col[order], val[order]
    first = np.ones(len(col), dtype=bool)
    first[1:] = (owner[1:] != owner[:-1]) | (col[1:] != col[:-1])
    owner, col, val = owner[first], col[first], val[first]
    # diagonal
    absum = np.bincount(owner, weights=np.abs(val), minlength=r1 - r0)
    d_owner = np.arange(r1 - r0, dtype=np.int64)
    owner = np.concatenate([owner, d_owner])
    col = np.concatenate([col, d_owner + r0])
    val = np.concatenate([val, 1.0 + absum])
    order = np.lexsort((col, owner))
    owner, col, val = owner[order], col[order], val[order]
    rowptr = np.zeros(r1 - r0 + 1, dtype=np.int64)
    np.cumsum(np.bincount(owner, minlength=r1 - r0), out=rowptr[1:])
    return rowptr, col, val


def localize_columns(col_global, c0, c1):
    """Local column ids as Matrix.cpp:294-318 / tests/test_spmv.cpp:98-119 number them:
    owned g - c0; ghosts ncols_local + rank among this rank's sorted ghosts.
    Returns (local int32 colind, sorted int64 ghosts)."""
    owned = (col_global >= c0) & (col_global < c1)
    ghosts = np.unique(col_global[~owned])
    local = np.empty(len(col_global), dtype=np.int32)
    local[owned] = (col_global[owned] - c0).astype(np.int32)
    local[~owned] = ((c1 - c0) + np.searchsorted(ghosts, col_global[~owned])).astype(np.int32)
    return local, ghosts.astype(np.int64)
