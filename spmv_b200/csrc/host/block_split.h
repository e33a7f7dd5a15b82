// block_split.h -- internal: the entry-classification rule and row canonicalisation
// shared by Matrix::create_matrix and read_petsc_binary_matrix.
#pragma once

#include "Matrix.h"

#include <algorithm>
#include <cstdint>
#include <numeric>
#include <vector>

namespace spmv
{
namespace detail
{

// destination block of an entry
enum Dest : int { LOCAL = 0, REMOTE = 1, DIAG = 2, DROP = 3 };

struct Splitter {
  bool symmetric;
  bool overlap;
  std::int64_t ncols_local, row0, col0;
  // Matrix.cpp:337-358 (and the same rules for received rows, :387-406)
  Dest operator()(std::int64_t row_local, std::int32_t col_local) const
  {
    if (symmetric) {
      if (col_local < ncols_local) {
        const std::int64_t grow = row_local + row0, gcol = col_local + col0;
        return grow > gcol ? LOCAL : (grow == gcol ? DIAG : DROP);
      }
      return REMOTE;
    }
    if (overlap)
      return col_local < ncols_local ? LOCAL : REMOTE;
    return LOCAL;
  }
};

// Sort one row by column (stable) and sum duplicates in input order; returns the new
// length.  setFromTriplets semantics.
template <typename T>
inline std::int32_t canonicalise_row(std::int32_t* col, T* val, std::int32_t len, std::vector<std::int32_t>& perm,
                              std::vector<std::int32_t>& ctmp, std::vector<T>& vtmp)
{
  bool sorted = true;
  for (std::int32_t k = 1; k < len; ++k)
    if (col[k] <= col[k - 1]) {
      sorted = false;
      break;
    }
  if (sorted)
    return len;
  perm.resize(len);
  std::iota(perm.begin(), perm.end(), 0);
  std::stable_sort(perm.begin(), perm.end(), [&](std::int32_t a, std::int32_t b) { return col[a] < col[b]; });
  ctmp.assign(col, col + len);
  vtmp.assign(val, val + len);
  std::int32_t out = 0;
  for (std::int32_t k = 0; k < len; ++k) {
    const std::int32_t c = ctmp[perm[k]];
    const T v = vtmp[perm[k]];
    if (out > 0 && col[out - 1] == c)
      val[out - 1] += v;
    else {
      col[out] = c;
      val[out] = v;
      ++out;
    }
  }
  return out;
}

template <typename T>
inline void canonicalise(HostCSR<T>& m, int nthreads = 1)
{
  const std::int32_t n = m.rows;
  std::vector<std::int32_t> newlen(n);
  bool shrunk = false;
#pragma omp parallel num_threads(nthreads)
  {
    std::vector<std::int32_t> perm, ctmp;
    std::vector<T> vtmp;
#pragma omp for schedule(static) reduction(|| : shrunk)
    for (std::int32_t r = 0; r < n; ++r) {
      const std::int32_t s = m.rowptr[r], len = m.rowptr[r + 1] - s;
      newlen[r] = canonicalise_row(m.colind.data() + s, m.values.data() + s, len, perm, ctmp, vtmp);
      shrunk = shrunk || newlen[r] != len;
    }
  }
  if (!shrunk)
    return;
  std::int32_t w = 0;
  for (std::int32_t r = 0; r < n; ++r) {
    const std::int32_t s = m.rowptr[r];
    m.rowptr[r] = w;
    for (std::int32_t k = 0; k < newlen[r]; ++k) {
      m.colind[w] = m.colind[s + k];
      m.values[w] = m.values[s + k];
      ++w;
    }
  }
  m.rowptr[n] = w;
  m.colind.resize(w);
  m.values.resize(w);
}

} // namespace detail
} // namespace spmv
