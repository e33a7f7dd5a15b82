"""Shared inputs for the tests: the reference's 5x5 fixture and helpers."""
import numpy as np

# tests/test_spmv.cpp:56-63
FIX_N = 5
FIX_ROWPTR = np.array([0, 3, 6, 9, 13, 15], dtype=np.int32)
FIX_COLIND = np.array([0, 1, 3, 0, 1, 3, 2, 3, 4, 0, 1, 2, 3, 2, 4], dtype=np.int32)
FIX_VALUES = np.array([1, -2, -3, -2, 5, 4, 6, 4, -4, -3, 4, 4, 8, -4, 8], dtype=np.float64)
# tests/test_spmv.cpp:66-70
FIX_X = np.array([np.exp(-10 * np.power(5 * (i / 5.0 - 0.5), 2.0)) for i in range(5)])
# produced by the reference's own compiled csr_kernels.cpp
FIX_Y = np.array([-0.24625499621007618, 0.32833999534154445, 0.82084998556222943,
                  0.98501998416354564, -0.32833999314207712])
FIX_NORM = 1.3857542692681533
# 1-rank symmetric split of the fixture (strictly lower + diagonal), Appendix D
FIX_SYM_ROWPTR = np.array([0, 0, 1, 1, 4, 5], dtype=np.int32)
FIX_SYM_COLIND = np.array([0, 0, 1, 2, 2], dtype=np.int32)
FIX_SYM_VALUES = np.array([-2, -3, 4, 4, -4], dtype=np.float64)
FIX_SYM_DIAG = np.array([1, 5, 6, 8, 8], dtype=np.float64)


def essentially_equal(a, b, eps=np.finfo(np.float64).eps):
    """tests/test_spmv.cpp:20-23 (Knuth)."""
    return abs(a - b) <= min(abs(a), abs(b)) * eps


def rel_l2(a, b):
    nb = np.linalg.norm(b)
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / (nb if nb > 0 else 1.0)


def split_symmetric(rowptr, colind, values):
    """Strictly-lower + diagonal split of a square CSR block (Matrix.cpp:337-349)."""
    n = len(rowptr) - 1
    rows = np.repeat(np.arange(n), np.diff(rowptr))
    lower = colind < rows
    diag = np.zeros(n)
    np.add.at(diag, rows[colind == rows], values[colind == rows])
    cnt = np.bincount(rows[lower], minlength=n)
    rp = np.zeros(n + 1, dtype=np.int32)
    np.cumsum(cnt, out=rp[1:])
    return rp, colind[lower].astype(np.int32), values[lower].astype(np.float64), diag
