// spmv_csr.cu -- fp64/fp32 CSR SpMV for sm_100a, general and symmetric storage.
//
// Replaces, behind CSRSpMV<T>::{init,run,finalize} (csr_kernels.h:71-78):
//   general   : cusparseSpMV              (cuda/csr_kernels.cuda.cu:106-151)
//   symmetric : csr_sym_gmem_atomics<T>   (cuda/csr_kernels.cuda.cu:22-47)
// Arithmetic contract: csr_kernels.cpp:20-52.
//
// Design (HBM-bound streaming kernel, no tensor cores):
//   * the plan cuts the rows into tiles of ~tile_bytes of matrix stream
//     (values 8 B + colind 4 B per entry, rowptr 4 B per row);
//   * a persistent CTA per SM slot owns tiles b, b+grid, ...; one producer lane
//     streams each tile's values / colind / rowptr slices into a shared-memory
//     ring with 1-D bulk-async copies (cp.async.bulk -> UBLKCP, completion on an
//     mbarrier), `stages` deep, so HBM requests stay in flight while the consumer
//     warps walk the previous tile;
//   * consumers walk one thread per row out of shared memory, in the reference's
//     j-ascending order with un-contracted mul/add (bit-identical to the CPU
//     kernel when the row is short); lanes of a warp hold consecutive rows, so
//     for banded / stencil matrices the x gathers coalesce and hit L1;
//   * rows longer than 32 entries are walked by a whole warp afterwards; a tile
//     that does not fit a stage (one huge row) is read directly from global
//     memory one warp per row;
//   * short-row matrices (mean row length below ~12, and every symmetric block)
//     use the two-phase "stream" walk instead: phase A is entry-parallel (each
//     thread multiplies entries tid, tid+T, ... by their gathered x and leaves
//     the products in shared memory, so every gather of a tile is independent
//     and in flight together, whatever the row lengths); phase B is one thread
//     per row summing its products in j order (same rounding sequence as the
//     row walk).  The row-length histogram taken by the plan picks the walk;
//   * symmetric storage: the transpose contribution goes out as fire-and-forget
//     RED.ADD.F64 (coalesced across the warp because lanes hold consecutive
//     rows); y is pre-scaled by beta (or zeroed) because every write is an add.
//     The tile's slices of the diagonal and of x ride in the same bulk copies,
//     and gathers that fall inside the tile's own rows are served from that
//     shared-memory copy of x (the diagonal block's near band).
#include "common.cuh"

#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>

#include <algorithm>
#include <cstdlib>
#include <vector>
#include <new>
#include <type_traits>

namespace b200
{

// Walks (template parameter WALK): one thread per row fetching 16 / 8 / 4 entries
// at a time, or the two-phase stream walk.
constexpr int WALK_ROW16 = 0, WALK_STREAM = 1, WALK_ROW8 = 2, WALK_ROW4 = 3;
// ...S: no row is longer than one chunk, so the chunk loop is gone
constexpr int WALK_ROW8S = 4, WALK_ROW4S = 5;
// segmented-scan walk for irregular rows (general storage): entry-parallel, warp-private, no CTA barrier
constexpr int WALK_SEG = 6;
// the same with L1::no_allocate x gathers
constexpr int WALK_SEGNA = 7;
__host__ __device__ constexpr bool walk_is_seg(int w) { return w == WALK_SEG || w == WALK_SEGNA; }
// split stream walk (general storage): half of the consumer warps make the products of tile i+1
// while the other half sums the rows of tile i; hand-over through mbarriers only
constexpr int WALK_SPLIT = 8;
// WALK_SEG: ints per tile in Plan::d_warp_rows (cw + 1 row offsets, padded to a 16-byte multiple for the bulk copy)
__host__ __device__ constexpr int seg_stride(int cw) { return (cw + 1 + 3) & ~3; }
constexpr int SHORT_ROW = 32;  // <= this many entries: one thread walks the row
constexpr int LONG_CAP = 256;  // long rows per tile (stage_bytes <= 96 KiB)
constexpr int MAX_STAGE_BYTES = 96 * 1024;

struct StageMeta {
  int r0;
  int nrows;   // rows of the tile
  int lo;
  int hi;      // end of the STAGED entries (rowptr[r0 + nstaged])
  int nstaged; // leading rows whose stream sits in the stage; the rest are read from global memory
  int nlong;
  int pad[2];
  int long_rows[LONG_CAP];
};

template <typename T>
struct SpmvArgs {
  const int32_t* rowptr;
  const int32_t* colind;
  const T* val;
  const T* diag;
  const T* x;
  T* y;
  T alpha;
  T beta;
  const int32_t* tile_meta;
  const int32_t* warp_rows; // WALK_SEG: seg_stride(CW) ints per tile, CW+1 row offsets: the consumer warps' segments
  int32_t flags;            // bit 0: WALK_SEG matrix stream with the evict_first L2 policy
  int32_t tile0;     // first tile of this launch (tile sub-ranges: pipelined host-vector SpMV)
  int32_t num_tiles; // end of the tile range
  int32_t stage_bytes;
  int32_t stages;
  double* partials; // dot epilogue: 2 * partials_len doubles (hi parts, then lo parts)
  int32_t partials_len;
  unsigned int* counter;
  double* dot_out;
};

__host__ __device__ __forceinline__ int round16(int b) { return (b + 15) & ~15; }

// Byte layout of one staged tile: [values | colind | rowptr], then for LAY >= 1
// the tile's slices of [diag | x] (symmetric storage) and for LAY == 2 a region
// for the stream walk's products; each loaded slice starts at the 16-byte aligned
// global element at or below the tile's first element.
constexpr int LAY_PLAIN = 0, LAY_DX = 1, LAY_DXP = 2;

template <typename T, int LAY>
struct TileLayout {
  int lo_v, lo_c, lo_r, lo_d;
  int bytes_v, bytes_c, bytes_r, bytes_d, bytes_p;
  __host__ __device__ TileLayout(int r0, int nrows, int lo, int hi)
  {
    constexpr int VA = 16 / (int)sizeof(T);
    lo_v = lo & ~(VA - 1);
    lo_c = lo & ~3;
    lo_r = r0 & ~3;
    lo_d = r0 & ~(VA - 1);
    bytes_v = round16((hi - lo_v) * (int)sizeof(T));
    bytes_c = round16((hi - lo_c) * 4);
    bytes_r = round16((r0 + nrows + 1 - lo_r) * 4);
    bytes_d = LAY >= LAY_DX ? round16((r0 + nrows - lo_d) * (int)sizeof(T)) : 0;
    bytes_p = LAY == LAY_DXP ? round16((hi - lo) * (int)sizeof(T)) : 0;
  }
  __host__ __device__ int off_c() const { return bytes_v; }
  __host__ __device__ int off_r() const { return bytes_v + bytes_c; }
  __host__ __device__ int off_d() const { return bytes_v + bytes_c + bytes_r; }
  __host__ __device__ int off_x() const { return off_d() + bytes_d; }
  __host__ __device__ int off_p() const { return off_x() + bytes_d; }
  __host__ __device__ int loaded() const { return off_p(); }
  __host__ __device__ int total() const { return off_p() + bytes_p; }
};

// 64-bit pre-check first: a tile holding one enormous row must not overflow the
// int byte counts of TileLayout.
template <typename T, int LAY>
__host__ __device__ __forceinline__ bool tile_fits(int r0, int nrows, int lo, int hi, int sb)
{
  const long long per_entry = (long long)sizeof(T) + 4 + (LAY == LAY_DXP ? (long long)sizeof(T) : 0);
  const long long per_row = 4 + (LAY >= LAY_DX ? 2ll * (long long)sizeof(T) : 0);
  if ((long long)(hi - lo) * per_entry + per_row * nrows > (long long)sb)
    return false;
  return TileLayout<T, LAY>(r0, nrows, lo, hi).total() <= sb;
}

// ---------------------------------------------------------------------------
// Row walks
// ---------------------------------------------------------------------------

// One warp, one row, entries [rs, re) addressed through `col_at` / `val_at`.
template <typename T, bool SYM, typename ColAt, typename ValAt>
__device__ __forceinline__ T warp_row(const SpmvArgs<T>& a, int r, int rs, int re, int lane,
                                      ColAt col_at, ValAt val_at)
{
  T xi = 0;
  if (SYM)
    xi = __ldg(a.x + r);
  T part = 0;
  for (int j = rs + lane; j < re; j += 32) {
    const int c = col_at(j);
    const T v = val_at(j);
    part = fma(v, __ldg(a.x + c), part);
    if (SYM)
      red_add(a.y + c, a.alpha * v * xi);
  }
  part = warp_sum(part);
  if (SYM)
    part += __ldg(a.diag + r) * xi;
  return part;
}

template <typename T, bool SYM>
__device__ __forceinline__ T store_row(const SpmvArgs<T>& a, int r, T sum)
{
  T out = mul_rn(a.alpha, sum);
  if (SYM) {
    red_add(a.y + r, out);
  } else {
    if (a.beta != T(0))
      out = add_rn(out, mul_rn(a.beta, a.y[r]));
    a.y[r] = out;
  }
  return out;
}

// One row's share of x . (A x) for the DOT epilogue.
//   general storage: x_r * out_r.
//   symmetric (lower + diagonal) storage: out_r is incomplete when the row is walked -- the transposed
//   entries of later rows are still to be scattered into it -- but the whole quadratic form is not:
//     x'Ax = sum_r x_r (d_r x_r + 2 sum_{c<r} a_rc x_c) = sum_r x_r (2 s_r - d_r x_r),  s_r = the row sum just formed
//   (alpha = 1).  dxr = d_r * x_r as the walk rounded it.  Differs from ddot(x, A x) by roundings only, which
//   symmetric storage (unordered RED adds) does not reproduce bit for bit in the first place.
template <typename T, bool SYM>
__device__ __forceinline__ void dot_row(dd& dot, const SpmvArgs<T>& a, int r, T out, T xr, T dxr)
{
  if (SYM) {
    // plain fp64 per thread (one FMA per row; a thread sees ~100 rows), twice the working precision only from
    // the warp sums on: two error-free products per row cost the 7-point kernel 11 % (fp64 issue, 56-register cap)
    dot.hi = fma((double)xr, 2.0 * (double)out - (double)dxr, dot.hi);
  } else {
    dd_add_prod(dot, (double)__ldg(a.x + r), (double)out);
  }
}
template <typename T, bool SYM>
__device__ __forceinline__ void dot_row(dd& dot, const SpmvArgs<T>& a, int r, T out)
{
  if (SYM) {
    const T xr = __ldg(a.x + r);
    dot_row<T, SYM>(dot, a, r, out, xr, mul_rn(__ldg(a.diag + r), xr));
  } else {
    dd_add_prod(dot, (double)__ldg(a.x + r), (double)out);
  }
}

// Walk a tile that sits in shared memory, one thread per row.  Returns this
// thread's share of sum_i x[i]*y[i] over the tile's rows (general storage only).
// CHUNK entries of a row are fetched together (16 for stencil-length rows, 8 / 4
// for short ones: fewer predicated-off slots).  LAY_DX: the tile's diagonal and x
// slices are staged too, and gathers that fall on the tile's own rows read the
// staged x.
template <typename T, bool SYM, bool LONG, bool DOT, int CHUNK, int LAY, bool SINGLE = false>
__device__ __forceinline__ dd walk_staged(const SpmvArgs<T>& a, const unsigned char* stage,
                                              StageMeta* m, int tid, int nthreads, int warp,
                                              int nwarps, int lane, int bar_id)
{
  constexpr bool DX = LAY >= LAY_DX;
  const int r0 = m->r0, nrows = m->nstaged, lo = m->lo, hi = m->hi;
  const TileLayout<T, LAY> L(r0, nrows, lo, hi);
  const T* sv = reinterpret_cast<const T*>(stage);
  const int* sc = reinterpret_cast<const int*>(stage + L.off_c());
  const int* sr = reinterpret_cast<const int*>(stage + L.off_r());
  const T* sd = reinterpret_cast<const T*>(stage + L.off_d()) + (r0 - L.lo_d); // diag[r0 + k]
  const T* sx = reinterpret_cast<const T*>(stage + L.off_x()) + (r0 - L.lo_d); // x[r0 + k]
  dd dot = {0.0, 0.0};

  for (int k = tid; k < nrows; k += nthreads) {
    const int r = r0 + k;
    const int rs = sr[r - L.lo_r];
    const int re = sr[r + 1 - L.lo_r];
    if (LONG && re - rs > SHORT_ROW) {
      const int q = atomicAdd(&m->nlong, 1);
      m->long_rows[q] = k;
      continue;
    }
    T xi = 0, sum = 0;
    if (SYM) {
      xi = DX ? sx[k] : __ldg(a.x + r);
      sum = mul_rn(DX ? sd[k] : __ldg(a.diag + r), xi);
    }
    // Chunks of CHUNK entries: all column indices, values and x gathers of a
    // chunk are issued before the first add, so one row keeps up to CHUNK
    // independent L2-latency gathers in flight; the adds then run in j order.
    for (int base = rs; base < re; base += CHUNK) {
      if (SINGLE && base != rs)
        break; // (never taken: the plan saw no row longer than CHUNK; lets the compiler drop the loop)
      int c[CHUNK];
      T v[CHUNK], xv[CHUNK];
#pragma unroll
      for (int u = 0; u < CHUNK; ++u) {
        const bool ok = base + u < re;
        c[u] = ok ? sc[base + u - L.lo_c] : (DX ? r0 : 0);
        v[u] = ok ? sv[base + u - L.lo_v] : T(0);
      }
#pragma unroll
      for (int u = 0; u < CHUNK; ++u) {
        if (DX && (unsigned)(c[u] - r0) < (unsigned)nrows)
          xv[u] = sx[c[u] - r0];
        else
          xv[u] = (base + u < re) ? __ldg(a.x + c[u]) : T(0);
      }
#pragma unroll
      for (int u = 0; u < CHUNK; ++u) {
        if (base + u < re) {
          sum = add_rn(sum, mul_rn(v[u], xv[u]));
          if (SYM)
            red_add(a.y + c[u], mul_rn(mul_rn(a.alpha, v[u]), xi));
        }
      }
    }
    const T out = store_row<T, SYM>(a, r, sum);
    if (DOT) {
      if (SYM)
        dot_row<T, SYM>(dot, a, r, out, xi, mul_rn(DX ? sd[k] : __ldg(a.diag + r), xi));
      else
        dd_add_prod(dot, (double)__ldg(a.x + r), (double)out);
    }
  }

  if (LONG) {
    named_bar_sync(bar_id, nthreads);
    const int nlong = m->nlong;
    for (int q = warp; q < nlong; q += nwarps) {
      const int r = r0 + m->long_rows[q];
      const int rs = sr[r - L.lo_r], re = sr[r + 1 - L.lo_r];
      const T sum = warp_row<T, SYM>(
          a, r, rs, re, lane, [&](int j) { return sc[j - L.lo_c]; },
          [&](int j) { return sv[j - L.lo_v]; });
      if (lane == 0) {
        const T out = store_row<T, SYM>(a, r, sum);
        if (DOT)
          dot_row<T, SYM>(dot, a, r, out);
      }
    }
  }
  return dot;
}

// Two-phase "stream" walk of a staged tile (short or irregular rows).
//   phase A, entry-parallel: products[j] = val[j] * x[col[j]] for every staged
//     entry, thread t taking entries t, t+T, ...: all gathers of the tile are
//     independent, so the memory pipeline stays full whatever the row lengths.
//     General storage overwrites val[j] in place; the symmetric (EXT) layout
//     keeps val for the transpose scatter and uses its own products region.
//   phase B, row-parallel: one thread per row adds its products in j order --
//     the same sequence of roundings as csr_kernels.cpp:45-47 -- and, symmetric
//     storage, scatters alpha*val*x[row] into y[col] (csr_kernels.cpp:33-37).
template <typename T, bool SYM, bool LONG, bool DOT>
__device__ __forceinline__ dd walk_stream(const SpmvArgs<T>& a, unsigned char* stage,
                                              StageMeta* m, int tid, int nthreads, int warp,
                                              int nwarps, int lane, int bar_id)
{
  constexpr int LAY = SYM ? LAY_DXP : LAY_PLAIN;
  constexpr bool EXT = SYM;
  constexpr int U = 4;
  const int r0 = m->r0, nrows = m->nstaged, lo = m->lo, hi = m->hi;
  const TileLayout<T, LAY> L(r0, nrows, lo, hi);
  T* sv = reinterpret_cast<T*>(stage) + (lo - L.lo_v);                  // sv[j - lo]
  const int* sc = reinterpret_cast<const int*>(stage + L.off_c()) + (lo - L.lo_c);
  const int* sr = reinterpret_cast<const int*>(stage + L.off_r()) + (r0 - L.lo_r); // sr[k]
  const T* sd = reinterpret_cast<const T*>(stage + L.off_d()) + (r0 - L.lo_d);     // diag[r0 + k]
  const T* sx = reinterpret_cast<const T*>(stage + L.off_x()) + (r0 - L.lo_d);     // x[r0 + k]
  T* sp = EXT ? reinterpret_cast<T*>(stage + L.off_p()) : sv;
  const int n = hi - lo;
  dd dot = {0.0, 0.0};

  for (int j0 = tid; j0 < n; j0 += nthreads * U) {
    int c[U];
    T v[U], xv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = j0 + u * nthreads;
      const bool ok = j < n;
      c[u] = ok ? sc[j] : (EXT ? r0 : 0);
      v[u] = ok ? sv[j] : T(0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (EXT && (unsigned)(c[u] - r0) < (unsigned)nrows)
        xv[u] = sx[c[u] - r0];
      else // padded slots gather nothing: a block may have more rows than columns, x[r0] need not exist
        xv[u] = (j0 + u * nthreads < n) ? __ldg(a.x + c[u]) : T(0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = j0 + u * nthreads;
      if (j < n)
        sp[j] = mul_rn(v[u], xv[u]);
    }
  }
  named_bar_sync(bar_id, nthreads);

  for (int k = tid; k < nrows; k += nthreads) {
    const int r = r0 + k;
    const int rs = sr[k] - lo, re = sr[k + 1] - lo;
    if (LONG && re - rs > SHORT_ROW) {
      const int q = atomicAdd(&m->nlong, 1);
      m->long_rows[q] = k;
      continue;
    }
    T xi = 0, sum = 0;
    if (SYM) {
      xi = sx[k];
      sum = mul_rn(sd[k], xi);
    }
    for (int j = rs; j < re; ++j) {
      sum = add_rn(sum, sp[j]);
      if (SYM)
        red_add(a.y + sc[j], mul_rn(mul_rn(a.alpha, sv[j]), xi));
    }
    const T out = store_row<T, SYM>(a, r, sum);
    if (DOT) {
      if (SYM)
        dot_row<T, SYM>(dot, a, r, out, xi, mul_rn(sd[k], xi));
      else
        dd_add_prod(dot, (double)__ldg(a.x + r), (double)out);
    }
  }

  if (LONG) {
    named_bar_sync(bar_id, nthreads);
    const int nlong = m->nlong;
    for (int q = warp; q < nlong; q += nwarps) {
      const int k = m->long_rows[q];
      const int r = r0 + k;
      const int rs = sr[k] - lo, re = sr[k + 1] - lo;
      T xi = 0;
      if (SYM)
        xi = sx[k];
      T part = 0;
      for (int j = rs + lane; j < re; j += 32) {
        part += sp[j];
        if (SYM)
          red_add(a.y + sc[j], a.alpha * sv[j] * xi);
      }
      part = warp_sum(part);
      if (lane == 0) {
        if (SYM)
          part += sd[k] * xi;
        const T out = store_row<T, SYM>(a, r, part);
        if (DOT)
          dot_row<T, SYM>(dot, a, r, out);
      }
    }
  }
  return dot;
}

// Walk a tile straight from global memory, one warp per row.
template <typename T, bool SYM, bool DOT>
__device__ __forceinline__ dd walk_direct(const SpmvArgs<T>& a, int r0, int nrows, int warp,
                                              int nwarps, int lane)
{
  dd dot = {0.0, 0.0};
  for (int k = warp; k < nrows; k += nwarps) {
    const int r = r0 + k;
    const int rs = __ldg(a.rowptr + r), re = __ldg(a.rowptr + r + 1);
    const T sum = warp_row<T, SYM>(
        a, r, rs, re, lane, [&](int j) { return __ldg(a.colind + j); },
        [&](int j) { return __ldg(a.val + j); });
    if (lane == 0) {
      const T out = store_row<T, SYM>(a, r, sum);
      if (DOT)
        dot_row<T, SYM>(dot, a, r, out);
    }
  }
  return dot;
}

// Rows too long for a stage, straight from global memory, by ALL consumer warps of the
// CTA: every thread takes entries rs + tid + k*nthreads, four gathers in flight each, then
// warp sums -> shared -> one thread adds them in warp order.  (One warp walking such a row
// alone is a chain of ~125 dependent global round trips for 4000 entries, ~0.15 ms: longer
// than the rest of the kernel, and the CTA that drew the row held everyone up.)
template <typename T, bool SYM, bool DOT>
__device__ __forceinline__ dd walk_direct_cta(const SpmvArgs<T>& a, int r0, int nrows, int tid, int nthreads,
                                              int warp, int nwarps, int lane, int bar_id, double* s_part)
{
  constexpr int U = 4;
  dd dot = {0.0, 0.0};
  for (int k = 0; k < nrows; ++k) {
    const int r = r0 + k;
    const int rs = __ldg(a.rowptr + r), re = __ldg(a.rowptr + r + 1);
    T xi = 0;
    if (SYM)
      xi = __ldg(a.x + r);
    T part = 0;
    for (int j0 = rs + tid; j0 < re; j0 += nthreads * U) {
      int c[U];
      T v[U], xv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int j = j0 + u * nthreads;
        const bool ok = j < re;
        c[u] = ok ? __ldg(a.colind + j) : 0;
        v[u] = ok ? __ldg(a.val + j) : T(0);
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        xv[u] = j0 + u * nthreads < re ? __ldg(a.x + c[u]) : T(0);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        part = fma(v[u], xv[u], part);
        if (SYM && j0 + u * nthreads < re)
          red_add(a.y + c[u], a.alpha * v[u] * xi);
      }
    }
    part = warp_sum(part);
    if (lane == 0)
      s_part[warp] = (double)part;
    named_bar_sync(bar_id, nthreads);
    if (tid == 0) {
      double sum = 0.0;
      for (int w = 0; w < nwarps; ++w)
        sum += s_part[w];
      T tot = (T)sum;
      if (SYM)
        tot += __ldg(a.diag + r) * xi;
      const T out = store_row<T, SYM>(a, r, tot);
      if (DOT)
        dot_row<T, SYM>(dot, a, r, out);
    }
    named_bar_sync(bar_id, nthreads); // s_part is free again
  }
  return dot;
}

// ---------------------------------------------------------------------------
// Warp-private, software-pipelined stream walk (WALK_SEG / WALK_SEGNA; general storage,
// irregular row lengths).  The stream walk's two phases -- products by entry, sums by
// row -- but each consumer warp owns a SEGMENT of every tile: the rows that start in its
// equal share of the tile's entries, cut at plan time (plan_warp_rows_kernel; the
// producer stages the tile's cw + 1 row offsets next to the rowptr slice).  No warp ever
// waits for another one: the only synchronisation is the full / empty mbarrier pair of
// the stage, on which every warp arrives for itself.  (The two CTA-wide barriers per tile
// of walk_stream were 11 of 30 stall cycles per issue on the power-law matrix,
// profiles/r01_powerlaw_stream_ncu_summary.txt.)
//   open   : wait for the tile, read this warp's segment, ISSUE the x gathers of its first
//            32*SEG_U entries into registers;
//   finish : products val*x written back over the staged values (rest of an over-long
//            segment gathered here, four in flight per lane);
//   sums   : lanes over the segment's rows, left-to-right sums of the products in chunks of
//            eight loads followed by eight dependent adds (the reference's order,
//            csr_kernels.cpp:45-47: rows <= 32 entries stay bit-identical); longer rows by
//            the whole warp.
// The loop runs `open` of tile i+1 BEFORE `sums` of tile i whenever that tile has already
// landed (mbarrier.test_wait), so the L2 latency of the gathers hides behind the
// shared-memory sums instead of adding to them.
// ---------------------------------------------------------------------------
template <typename T>
struct SegU {
  static constexpr int value = sizeof(T) == 8 ? 4 : 8; // gathers per lane carried across the previous tile's sums
};

// Shared-memory byte offsets (32-bit: two tiles' worth of these live in registers per thread)
template <typename T>
struct SegTile {
  int sv; // byte offset of the staged value of entry `lo` (overwritten by the products)
  int sc; // ... of its column index
  int sr; // ... of rowptr[r0]
  int r0, lo, ns, nrows;
  int ka, kb; // this warp's rows of the tile
  int ea, eb; // ... and their entries, relative to lo
};

template <typename T, bool NA>
__device__ __forceinline__ T seg_gather(const T* p)
{
  if constexpr (NA)
    return ldg_na(p);
  else
    return __ldg(p);
}

// This warp's view of the tile in stage s (the stage must have landed).
template <typename T>
__device__ __forceinline__ SegTile<T> seg_tile(const SpmvArgs<T>& a, const unsigned char* smem, const StageMeta* meta,
                                               int s, int warp)
{
  const StageMeta* m = meta + s;
  const int base = s * a.stage_bytes;
  SegTile<T> t;
  const int r0 = m->r0, ns = m->nstaged, lo = m->lo, hi = m->hi;
  t.r0 = r0;
  t.lo = lo;
  t.ns = ns;
  t.nrows = m->nrows;
  t.sv = t.sc = t.sr = base;
  t.ka = t.kb = t.ea = t.eb = 0;
  if (ns > 0) {
    const TileLayout<T, LAY_PLAIN> L(r0, ns, lo, hi);
    t.sv = base + (lo - L.lo_v) * (int)sizeof(T);
    t.sc = base + L.off_c() + (lo - L.lo_c) * 4;
    t.sr = base + L.off_r() + (r0 - L.lo_r) * 4;
    const int* wr = reinterpret_cast<const int*>(smem + base + L.off_d());
    const int* sr = reinterpret_cast<const int*>(smem + t.sr);
    t.ka = min(wr[warp], ns); // the producer may have staged a prefix of the tile's rows only
    t.kb = min(wr[warp + 1], ns);
    if (t.ka < t.kb) {
      t.ea = sr[t.ka] - lo;
      t.eb = sr[t.kb] - lo;
    }
  }
  return t;
}

template <typename T, bool NA>
__device__ __forceinline__ void seg_issue(const SpmvArgs<T>& a, const unsigned char* smem, const SegTile<T>& t, int lane,
                                          T (&xv)[SegU<T>::value])
{
  const int* sc = reinterpret_cast<const int*>(smem + t.sc);
#pragma unroll
  for (int u = 0; u < SegU<T>::value; ++u) {
    const int j = t.ea + lane + 32 * u;
    xv[u] = j < t.eb ? seg_gather<T, NA>(a.x + sc[j]) : T(0);
  }
}

template <typename T, bool NA>
__device__ __forceinline__ void seg_finish(const SpmvArgs<T>& a, unsigned char* smem, const SegTile<T>& t,
                                           const T (&xv)[SegU<T>::value], int lane)
{
  T* sv = reinterpret_cast<T*>(smem + t.sv);
  const int* sc = reinterpret_cast<const int*>(smem + t.sc);
#pragma unroll
  for (int u = 0; u < SegU<T>::value; ++u) {
    const int j = t.ea + lane + 32 * u;
    if (j < t.eb)
      sv[j] = mul_rn(sv[j], xv[u]);
  }
  constexpr int U = 4;
  for (int j0 = t.ea + 32 * SegU<T>::value + lane; j0 < t.eb; j0 += 32 * U) {
    T w[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      w[u] = j0 + 32 * u < t.eb ? seg_gather<T, NA>(a.x + sc[j0 + 32 * u]) : T(0);
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (j0 + 32 * u < t.eb)
        sv[j0 + 32 * u] = mul_rn(sv[j0 + 32 * u], w[u]);
  }
  __syncwarp();
}

template <typename T, bool DOT>
__device__ __forceinline__ void seg_sums(const SpmvArgs<T>& a, const unsigned char* smem, const SegTile<T>& t, int lane,
                                         dd& dot)
{
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int C = 8;
  const T* sv = reinterpret_cast<const T*>(smem + t.sv);
  const int* sr = reinterpret_cast<const int*>(smem + t.sr);
  for (int k0 = t.ka; k0 < t.kb; k0 += 32) {
    const int k = k0 + lane;
    int rs = 0, re = 0;
    if (k < t.kb) {
      rs = sr[k] - t.lo;
      re = sr[k + 1] - t.lo;
    }
    const bool is_long = re - rs > SHORT_ROW;
    if (k < t.kb && !is_long) {
      T sum = T(0);
      for (int b = rs; b < re; b += C) {
        T pv[C];
#pragma unroll
        for (int u = 0; u < C; ++u)
          pv[u] = b + u < re ? sv[b + u] : T(0);
#pragma unroll
        for (int u = 0; u < C; ++u)
          if (b + u < re)
            sum = add_rn(sum, pv[u]);
      }
      const T out = store_row<T, false>(a, t.r0 + k, sum);
      if (DOT)
        dd_add_prod(dot, (double)__ldg(a.x + t.r0 + k), (double)out);
    }
    unsigned todo = __ballot_sync(FULL, is_long);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const int qs = __shfl_sync(FULL, rs, src), qe = __shfl_sync(FULL, re, src);
      T part = T(0);
      for (int j = qs + lane; j < qe; j += 32)
        part += sv[j];
      part = warp_sum(part);
      if (lane == 0) {
        const T out = store_row<T, false>(a, t.r0 + k0 + src, part);
        if (DOT)
          dd_add_prod(dot, (double)__ldg(a.x + t.r0 + k0 + src), (double)out);
      }
    }
  }
}

__device__ __forceinline__ void fence_proxy_async_smem()
{
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

template <typename T, bool DOT, int CW, bool NA>
__device__ __forceinline__ dd seg_consumer(const SpmvArgs<T>& a, unsigned char* smem, StageMeta* meta, uint64_t* full,
                                           uint64_t* empty, int my_tiles, int warp, int lane, double* s_warp)
{
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int U = SegU<T>::value;
  const int stages = a.stages;
  dd dot = {0.0, 0.0};
  if (my_tiles <= 0)
    return dot;
  T xv[U];
  int s = 0, use = 0;
  mbar_wait(full, 0);
  SegTile<T> cur = seg_tile<T>(a, smem, meta, 0, warp);
  seg_issue<T, NA>(a, smem, cur, lane, xv);
  for (int it = 0; it < my_tiles; ++it) {
    seg_finish<T, NA>(a, smem, cur, xv, lane);
    const int s1 = s + 1 == stages ? 0 : s + 1;
    const int use1 = use + (s1 == 0);
    const bool more = it + 1 < my_tiles;
    bool early = false;
    if (more) {
      // has the next tile landed already?  then its gathers go out now and fly during the sums
      early = __shfl_sync(FULL, (int)mbar_test(full + s1, use1 & 1), 0) != 0;
      if (early) {
        const SegTile<T> nxt = seg_tile<T>(a, smem, meta, s1, warp);
        seg_issue<T, NA>(a, smem, nxt, lane, xv);
      }
    }
    seg_sums<T, DOT>(a, smem, cur, lane, dot);
    if (cur.ns < cur.nrows)
      dot = dd_merge(dot, walk_direct_cta<T, false, DOT>(a, cur.r0 + cur.ns, cur.nrows - cur.ns, threadIdx.x, CW * 32,
                                                         warp, CW, lane, 1, s_warp));
    fence_proxy_async_smem(); // our generic-proxy writes precede the next bulk copy into this stage
    __syncwarp();
    if (lane == 0)
      mbar_arrive(empty + s);
    if (more) {
      if (!early)
        mbar_wait(full + s1, use1 & 1);
      cur = seg_tile<T>(a, smem, meta, s1, warp); // (recomputed rather than carried: registers)
      if (!early)
        seg_issue<T, NA>(a, smem, cur, lane, xv);
    }
    s = s1;
    use = use1;
  }
  return dot;
}

// ---------------------------------------------------------------------------
// Split stream walk (WALK_SPLIT; general storage, irregular row lengths).  The stream walk's two
// phases run on DIFFERENT warps, one tile apart:
//   A warps (first half of the consumers): products.  Entry-parallel over the whole tile, every lane
//     busy whatever the row lengths, eight independent x gathers in flight per lane; the product
//     overwrites the staged value.  They arrive on prod[s] and go straight on to the next tile.
//   B warps (second half): sums.  One lane per row over the WHOLE tile (a 24 KB tile of the
//     power-law matrix has ~235 rows for 256 lanes; the warp-private walk had 15 rows per warp and
//     ran its sum loop with 7 lanes active, profiles/r02_powerlaw_walks.md), left-to-right in chunks
//     of eight loads then eight adds -- the reference's order (csr_kernels.cpp:45-47), rows <= 32
//     entries bit-identical; longer rows by the warp that found them.  They arrive on empty[s].
// Hand-over between the groups is the prod[s] mbarrier: no bar.sync anywhere, a waiting warp sleeps
// in mbarrier.try_wait instead of holding an issue slot, and the gather latency of tile i+1 hides
// behind the shared-memory sums of tile i by construction.
// Padding adds of +0.0 are exact: a left-to-right sum that starts at +0.0 is never -0.0.
// ---------------------------------------------------------------------------
template <typename T, int NA>
__device__ __forceinline__ void split_products(const SpmvArgs<T>& a, unsigned char* smem, StageMeta* meta, uint64_t* full,
                                               uint64_t* prod, int my_tiles, int warp, int lane)
{
  constexpr int U = 8;
  constexpr int NT = NA * 32;
  const int stages = a.stages, sb = a.stage_bytes;
  const int tid = warp * 32 + lane;
  for (int it = 0, s = 0, use = 0; it < my_tiles; ++it) {
    mbar_wait(full + s, use & 1);
    const StageMeta* m = meta + s;
    const int ns = m->nstaged;
    if (ns > 0) {
      const int r0 = m->r0, lo = m->lo, hi = m->hi;
      const TileLayout<T, LAY_PLAIN> L(r0, ns, lo, hi);
      unsigned char* stage = smem + (size_t)s * sb;
      T* sv = reinterpret_cast<T*>(stage) + (lo - L.lo_v);
      const int* sc = reinterpret_cast<const int*>(stage + L.off_c()) + (lo - L.lo_c);
      const int n = hi - lo;
      for (int j0 = tid; j0 < n; j0 += NT * U) {
        T xv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int j = j0 + u * NT;
          xv[u] = j < n ? __ldg(a.x + sc[j]) : T(0);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int j = j0 + u * NT;
          if (j < n)
            sv[j] = mul_rn(sv[j], xv[u]);
        }
      }
      fence_proxy_async_smem(); // these generic-proxy writes precede the bulk copy that will refill the stage
    }
    __syncwarp();
    if (lane == 0)
      mbar_arrive(prod + s);
    s = s + 1 == stages ? 0 : s + 1;
    use += s == 0;
  }
}

template <typename T, bool DOT, int NB>
__device__ __forceinline__ dd split_sums(const SpmvArgs<T>& a, unsigned char* smem, StageMeta* meta, uint64_t* prod,
                                         uint64_t* empty, int my_tiles, int warp, int lane, double* s_warp)
{
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int C = 8;
  constexpr int NT = NB * 32;
  const int stages = a.stages, sb = a.stage_bytes;
  const int tid = warp * 32 + lane;
  dd dot = {0.0, 0.0};
  for (int it = 0, s = 0, use = 0; it < my_tiles; ++it) {
    mbar_wait(prod + s, use & 1);
    const StageMeta* m = meta + s;
    const int ns = m->nstaged, nrows = m->nrows, r0 = m->r0;
    if (ns > 0) {
      const int lo = m->lo, hi = m->hi;
      const TileLayout<T, LAY_PLAIN> L(r0, ns, lo, hi);
      const unsigned char* stage = smem + (size_t)s * sb;
      const T* sv = reinterpret_cast<const T*>(stage) + (lo - L.lo_v);
      const int par = (lo - L.lo_v) & 1; // sv[j] is 16-byte aligned iff j + par is even
      (void)par;
      const int* sr = reinterpret_cast<const int*>(stage + L.off_r()) + (r0 - L.lo_r);
      for (int k0 = warp * 32; k0 < ns; k0 += NT) {
        const int k = k0 + lane;
        int rs = 0, re = 0;
        if (k < ns) {
          rs = sr[k] - lo;
          re = sr[k + 1] - lo;
        }
        const bool is_long = re - rs > SHORT_ROW;
        if (k < ns && !is_long) {
          T sum = T(0);
          if constexpr (sizeof(T) == 8) {
            // 16-byte loads: the shared-memory data pipe is what this kernel runs out of (80 % busy, profiles/
            // r02_powerlaw_split_ncu_summary.txt) and a quarter-warp of random 16-byte reads costs fewer
            // wavefronts per value than a half-warp of 8-byte ones.  One scalar load brings the row to alignment.
            int b = rs;
            if (((b + par) & 1) && b < re) {
              sum = add_rn(sum, sv[b]);
              ++b;
            }
            for (; b < re; b += C) {
              double2 q[C / 2];
#pragma unroll
              for (int u = 0; u < C / 2; ++u)
                q[u] = b + 2 * u < re ? *reinterpret_cast<const double2*>(sv + b + 2 * u) : make_double2(0.0, 0.0);
#pragma unroll
              for (int u = 0; u < C / 2; ++u) {
                sum = add_rn(sum, q[u].x);
                sum = add_rn(sum, b + 2 * u + 1 < re ? q[u].y : 0.0); // the pair's second half may be the next row's
              }
            }
          } else {
            for (int b = rs; b < re; b += C) {
              T pv[C];
#pragma unroll
              for (int u = 0; u < C; ++u)
                pv[u] = b + u < re ? sv[b + u] : T(0);
#pragma unroll
              for (int u = 0; u < C; ++u)
                sum = add_rn(sum, pv[u]);
            }
          }
          const T out = store_row<T, false>(a, r0 + k, sum);
          if (DOT)
            dd_add_prod(dot, (double)__ldg(a.x + r0 + k), (double)out);
        }
        unsigned todo = __ballot_sync(FULL, is_long);
        while (todo) {
          const int src = __ffs(todo) - 1;
          todo &= todo - 1;
          const int qs = __shfl_sync(FULL, rs, src), qe = __shfl_sync(FULL, re, src);
          T part = T(0);
          for (int j = qs + lane; j < qe; j += 32)
            part += sv[j];
          part = warp_sum(part);
          if (lane == 0) {
            const T out = store_row<T, false>(a, r0 + k0 + src, part);
            if (DOT)
              dd_add_prod(dot, (double)__ldg(a.x + r0 + k0 + src), (double)out);
          }
        }
      }
    }
    if (ns < nrows) // a row too long for a stage: straight from global memory, by the B group
      dot = dd_merge(dot, walk_direct_cta<T, false, DOT>(a, r0 + ns, nrows - ns, tid, NT, warp, NB, lane, 1, s_warp));
    __syncwarp();
    if (lane == 0)
      mbar_arrive(empty + s);
    s = s + 1 == stages ? 0 : s + 1;
    use += s == 0;
  }
  return dot;
}

// ---------------------------------------------------------------------------
// Variant 0: persistent CTAs, bulk-async staged ring, warp-specialised.
//   warps [0, CW): consumers;  warp CW: producer (lane 0 only).
//   WALK_ROW16/8/4: one thread per row (chunked gathers);  WALK_STREAM: two-phase stream walk.
// ---------------------------------------------------------------------------
// ---------------------------------------------------------------------------
// Two kernels share the pipeline design.  spmv_pipe_kernel carries the row walks and the CTA-wide
// stream walk and is kept EXACTLY as tuned in round 1: ptxas' register allocation for it is fragile
// (adding the split walk's third barrier array and tile-order variables to the same function took the
// 27-point kernel from 96 to 120 registers and the symmetric 7-point one from 56 to 64 -- one resident
// CTA per SM instead of two, 0.85 -> 1.42 ms and 0.215 -> 0.365 ms).  spmv_irr_kernel below is the same
// pipeline generalised for the irregular-row walks of general storage (WALK_SEG, WALK_SEGNA, WALK_SPLIT).
// ---------------------------------------------------------------------------
template <typename T, bool SYM, bool DOT, bool LONG, int CW, int WALK>
__global__ void __launch_bounds__((CW + 1) * 32) spmv_pipe_kernel(const SpmvArgs<T> a)
{
  constexpr int LAY = SYM ? (WALK == WALK_STREAM ? LAY_DXP : LAY_DX) : LAY_PLAIN;
  constexpr int CHUNK = (WALK == WALK_ROW4 || WALK == WALK_ROW4S) ? 4 : ((WALK == WALK_ROW8 || WALK == WALK_ROW8S) ? 8 : 16);
  constexpr bool SINGLE = WALK == WALK_ROW4S || WALK == WALK_ROW8S;
  extern __shared__ __align__(128) unsigned char smem[];
  const int stages = a.stages;
  const int sb = a.stage_bytes;
  StageMeta* meta = reinterpret_cast<StageMeta*>(smem + (size_t)stages * sb);
  uint64_t* full = reinterpret_cast<uint64_t*>(meta + stages);
  uint64_t* empty = full + stages;
  double* s_warp = reinterpret_cast<double*>(empty + stages);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, CW);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const int span = a.num_tiles - a.tile0;
  const int my_tiles = ((int)blockIdx.x < span) ? (span - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == CW) {
    if (lane == 0) {
      // general storage: evict_first only on request (flags bit 2; run_spmv sets it for blocks small enough that
      // the caller's vectors can live in L2 across SpMVs if the matrix stream does not sweep them out)
      const bool hint = SYM || (a.flags & 4);
      const uint64_t pol_stream = hint ? l2_policy_evict_first() : 0ull;
      auto stream_g2s = [&](void* d, const void* g, uint32_t bytes, uint64_t* bar) {
        if (hint)
          bulk_g2s_hint(d, g, bytes, bar, pol_stream);
        else
          bulk_g2s(d, g, bytes, bar);
      };
      for (int it = 0, s = 0, use = 0; it < my_tiles; ++it, s = (s + 1 == stages ? 0 : s + 1), use += (s == 0)) {
        if (use > 0)
          mbar_wait(empty + s, (use - 1) & 1);
        const int tile = a.tile0 + blockIdx.x + it * gridDim.x;
        const int2 t0 = __ldg(reinterpret_cast<const int2*>(a.tile_meta) + tile);
        const int2 t1 = __ldg(reinterpret_cast<const int2*>(a.tile_meta) + tile + 1);
        const int r0 = t0.x, lo = t0.y, nrows = t1.x - t0.x;
        // Stage the longest prefix of rows that fits: a tile overshoots its byte
        // budget by at most its last row, so this drops one (long) row at most.
        int ns = nrows, hi = t1.y;
        while (ns > 0 && !tile_fits<T, LAY>(r0, ns, lo, hi, sb)) {
          --ns;
          hi = __ldg(a.rowptr + r0 + ns);
        }
        StageMeta* m = meta + s;
        m->r0 = r0;
        m->nrows = nrows;
        m->lo = lo;
        m->hi = hi;
        m->nstaged = ns;
        m->nlong = 0;
        if (ns > 0) {
          const TileLayout<T, LAY> L(r0, ns, lo, hi);
          unsigned char* dst = smem + (size_t)s * sb;
          mbar_arrive_expect_tx(full + s, (uint32_t)L.loaded());
          // Symmetric storage: the matrix is read once per SpMV, while the y lines are touched several
          // times (zero-fill, the row's own add, the adds from the rows that list it).  evict_first on
          // the matrix stream keeps it from displacing them in L2: sym 7-pt 0.220 -> 0.213 ms, sym 27-pt
          // 0.64 -> 0.62 (profiles/r01_exp_evict_first_probe.txt).  General storage gains nothing
          // (7-pt 0.2526 vs 0.2515, 27-pt 0.864 vs 0.849) and keeps the default policy.
          if (L.bytes_v > 0)
            stream_g2s(dst, a.val + L.lo_v, L.bytes_v, full + s);
          if (L.bytes_c > 0)
            stream_g2s(dst + L.off_c(), a.colind + L.lo_c, L.bytes_c, full + s);
          stream_g2s(dst + L.off_r(), a.rowptr + L.lo_r, L.bytes_r, full + s);
          if (LAY >= LAY_DX) {
            stream_g2s(dst + L.off_d(), a.diag + L.lo_d, L.bytes_d, full + s);
            bulk_g2s(dst + L.off_x(), a.x + L.lo_d, L.bytes_d, full + s);
          }
        } else {
          mbar_arrive(full + s);
        }
      }
    }
    return;
  }

  dd dot = {0.0, 0.0};
  // Two spellings of the same loop.  ptxas schedules the 16-entry row walk ~20% worse (27-point:
  // 0.85 -> 1.04 ms) with the incremental stage counters and the CTA-wide direct walk in the loop
  // body, and the 8-entry general walk ~2% better (7-point: 0.257 -> 0.251 ms), A/B-measured in
  // profiles/r01_powerlaw_walks.md; the irregular walks need the CTA-wide direct walk.
  constexpr bool CTA_DIRECT = WALK == WALK_STREAM || (!SYM && CHUNK == 8);
  for (int it = 0, cs = 0, cu = 0; it < my_tiles; ++it) {
    int s, use;
    if constexpr (CTA_DIRECT) {
      s = cs, use = cu;
      cs = cs + 1 == stages ? 0 : cs + 1;
      cu += cs == 0;
    } else {
      s = it % stages, use = it / stages;
    }
    mbar_wait(full + s, use & 1);
    StageMeta* m = meta + s;
    const int ns = m->nstaged, nrows = m->nrows, r0 = m->r0;
    if (ns > 0) {
      if (WALK == WALK_STREAM) {
        dot = dd_merge(dot, walk_stream<T, SYM, LONG, DOT>(a, smem + (size_t)s * sb, m, threadIdx.x, CW * 32,
                                                           warp, CW, lane, 1));
        fence_proxy_async_smem(); // our generic-proxy writes precede the next bulk copy into this stage
      } else {
        dot = dd_merge(dot, walk_staged<T, SYM, LONG, DOT, CHUNK, LAY, SINGLE>(a, smem + (size_t)s * sb, m,
                                                                               threadIdx.x, CW * 32, warp, CW, lane, 1));
      }
    }
    if (ns < nrows) {
      if constexpr (CTA_DIRECT)
        dot = dd_merge(dot, walk_direct_cta<T, SYM, DOT>(a, r0 + ns, nrows - ns, threadIdx.x, CW * 32, warp, CW,
                                                         lane, 1, s_warp));
      else
        dot = dd_merge(dot, walk_direct<T, SYM, DOT>(a, r0 + ns, nrows - ns, warp, CW, lane));
    }
    __syncwarp();
    if (lane == 0)
      mbar_arrive(empty + s);
  }
  if (DOT)
    grid_dd_reduce(dot, threadIdx.x, CW * 32, 2, s_warp, a.partials, a.partials_len, a.counter, a.dot_out);
}


// resident CTAs per SM the register allocation must allow (only the warp-private walk asks: its two
// tiles in flight per warp would otherwise take ~94 registers and a single CTA per SM)
__host__ __device__ constexpr int pipe_min_ctas(bool sym, int walk, int cw)
{
  return (!sym && (walk_is_seg(walk) || walk == WALK_SPLIT)) ? (cw == 16 ? 2 : 3) : 1;
}

template <typename T, bool SYM, bool DOT, bool LONG, int CW, int WALK>
__global__ void __launch_bounds__((CW + 1) * 32, pipe_min_ctas(SYM, WALK, CW)) spmv_irr_kernel(const SpmvArgs<T> a)
{
  constexpr bool SEGW = !SYM && walk_is_seg(WALK); // warp-private pipelined walk (general storage only)
  constexpr bool SPLIT = !SYM && WALK == WALK_SPLIT; // products and sums on different warps
  constexpr int NA = CW / 2, NB = CW - NA;
  constexpr int WR_BYTES = SEGW ? seg_stride(CW) * 4 : 0;
  constexpr int LAY = SYM ? ((WALK == WALK_STREAM || walk_is_seg(WALK) || WALK == WALK_SPLIT) ? LAY_DXP : LAY_DX) : LAY_PLAIN;
  constexpr int CHUNK = (WALK == WALK_ROW4 || WALK == WALK_ROW4S) ? 4 : ((WALK == WALK_ROW8 || WALK == WALK_ROW8S) ? 8 : 16);
  constexpr bool SINGLE = WALK == WALK_ROW4S || WALK == WALK_ROW8S;
  extern __shared__ __align__(128) unsigned char smem[];
  const int stages = a.stages;
  const int sb = a.stage_bytes;
  StageMeta* meta = reinterpret_cast<StageMeta*>(smem + (size_t)stages * sb);
  uint64_t* full = reinterpret_cast<uint64_t*>(meta + stages);
  uint64_t* empty = full + stages;
  uint64_t* prod = empty + stages; // SPLIT: products of the stage are in place
  double* s_warp = reinterpret_cast<double*>(prod + stages);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, SPLIT ? NB : CW);
      mbar_init(prod + s, NA);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const int span = a.num_tiles - a.tile0;
  int my_tiles = ((int)blockIdx.x < span) ? (span - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  // tiles of this CTA: b, b + grid, ... (default) or, SPLIT with flag bit 1, one contiguous run of tiles
  int tile_first = a.tile0 + blockIdx.x, tile_step = gridDim.x;
  if (SPLIT && (a.flags & 2)) {
    const int base = span / (int)gridDim.x, rem = span % (int)gridDim.x;
    my_tiles = base + ((int)blockIdx.x < rem ? 1 : 0);
    tile_first = a.tile0 + (int)blockIdx.x * base + min((int)blockIdx.x, rem);
    tile_step = 1;
  }

  if (warp == CW) {
    if (lane == 0) {
      const uint64_t pol_stream = (SYM || SEGW || SPLIT) ? l2_policy_evict_first() : 0ull;
      const bool hint = SYM || ((SEGW || SPLIT) && (a.flags & 1));
      auto stream_g2s = [&](void* d, const void* g, uint32_t bytes, uint64_t* bar) {
        if constexpr (SYM || SEGW || SPLIT) {
          if (hint)
            bulk_g2s_hint(d, g, bytes, bar, pol_stream);
          else
            bulk_g2s(d, g, bytes, bar);
        } else {
          bulk_g2s(d, g, bytes, bar);
        }
      };
      const int2* tmeta = reinterpret_cast<const int2*>(a.tile_meta);
      // SEGW: the tile table entry of the NEXT tile is fetched before the wait for a free stage, so the
      // producer's chain per tile is not table load -> wait -> issue (~1.4 us per tile and CTA,
      // profiles/r01_powerlaw_walks.md) but max(table load, wait) -> issue
      int2 n0 = make_int2(0, 0), n1 = make_int2(0, 0);
      constexpr bool PREFETCH = SEGW || SPLIT;
      if (PREFETCH && my_tiles > 0) {
        n0 = __ldg(tmeta + tile_first);
        n1 = __ldg(tmeta + tile_first + 1);
      }
      for (int it = 0, s = 0, use = 0; it < my_tiles; ++it, s = (s + 1 == stages ? 0 : s + 1), use += (s == 0)) {
        const int tile = tile_first + it * tile_step;
        int2 t0, t1;
        if constexpr (SEGW || SPLIT) {
          t0 = n0, t1 = n1;
          if (it + 1 < my_tiles) {
            n0 = __ldg(tmeta + tile + tile_step);
            n1 = __ldg(tmeta + tile + tile_step + 1);
          }
        }
        if (use > 0)
          mbar_wait(empty + s, (use - 1) & 1);
        if constexpr (!(SEGW || SPLIT)) {
          t0 = __ldg(tmeta + tile);
          t1 = __ldg(tmeta + tile + 1);
        }
        const int r0 = t0.x, lo = t0.y, nrows = t1.x - t0.x;
        // Stage the longest prefix of rows that fits: a tile overshoots its byte
        // budget by at most its last row, so this drops one (long) row at most.
        int ns = nrows, hi = t1.y;
        while (ns > 0 && !tile_fits<T, LAY>(r0, ns, lo, hi, sb - WR_BYTES)) {
          --ns;
          hi = __ldg(a.rowptr + r0 + ns);
        }
        StageMeta* m = meta + s;
        m->r0 = r0;
        m->nrows = nrows;
        m->lo = lo;
        m->hi = hi;
        m->nstaged = ns;
        m->nlong = 0;
        if (ns > 0) {
          const TileLayout<T, LAY> L(r0, ns, lo, hi);
          unsigned char* dst = smem + (size_t)s * sb;
          mbar_arrive_expect_tx(full + s, (uint32_t)(L.loaded() + WR_BYTES));
          // Symmetric storage: the matrix is read once per SpMV, while the y lines are touched several
          // times (zero-fill, the row's own add, the adds from the rows that list it).  evict_first on
          // the matrix stream keeps it from displacing them in L2: sym 7-pt 0.220 -> 0.213 ms, sym 27-pt
          // 0.64 -> 0.62 (profiles/r01_exp_evict_first_probe.txt).  General storage gains nothing
          // (7-pt 0.2526 vs 0.2515, 27-pt 0.864 vs 0.849) and keeps the default policy.
          if (L.bytes_v > 0)
            stream_g2s(dst, a.val + L.lo_v, L.bytes_v, full + s);
          if (L.bytes_c > 0)
            stream_g2s(dst + L.off_c(), a.colind + L.lo_c, L.bytes_c, full + s);
          stream_g2s(dst + L.off_r(), a.rowptr + L.lo_r, L.bytes_r, full + s);
          if (LAY >= LAY_DX) {
            stream_g2s(dst + L.off_d(), a.diag + L.lo_d, L.bytes_d, full + s);
            bulk_g2s(dst + L.off_x(), a.x + L.lo_d, L.bytes_d, full + s);
          }
          if constexpr (SEGW) // the consumer warps' row segments of this tile, next to the rowptr slice
            bulk_g2s(dst + L.off_d(), a.warp_rows + (size_t)tile * seg_stride(CW), WR_BYTES, full + s);
        } else {
          mbar_arrive(full + s);
        }
      }
    }
    return;
  }

  dd dot = {0.0, 0.0};
  // Two spellings of the same loop.  ptxas schedules the 16-entry row walk ~20% worse (27-point:
  // 0.85 -> 1.04 ms) with the incremental stage counters and the CTA-wide direct walk in the loop
  // body, and the 8-entry general walk ~2% better (7-point: 0.257 -> 0.251 ms), A/B-measured in
  // profiles/r01_powerlaw_walks.md; the irregular walks need the CTA-wide direct walk.
  constexpr bool CTA_DIRECT = WALK == WALK_STREAM || walk_is_seg(WALK) || WALK == WALK_SPLIT || (!SYM && CHUNK == 8);
  if constexpr (SPLIT) {
    if (warp < NA) {
      split_products<T, NA>(a, smem, meta, full, prod, my_tiles, warp, lane);
      if (DOT) // the reduction below counts CW warps: A warps bring zeros
        grid_dd_reduce(dot, threadIdx.x, CW * 32, 2, s_warp, a.partials, a.partials_len, a.counter, a.dot_out);
      return;
    }
    // (scratch of the direct walk behind the 2 * CW doubles the dot reduction uses: the A warps may be in it already)
    dot = split_sums<T, DOT, NB>(a, smem, meta, prod, empty, my_tiles, warp - NA, lane, s_warp + 32);
  } else if constexpr (SEGW) {
    dot = seg_consumer<T, DOT, CW, WALK == WALK_SEGNA>(a, smem, meta, full, empty, my_tiles, warp, lane, s_warp);
  } else
  for (int it = 0, cs = 0, cu = 0; it < my_tiles; ++it) {
    int s, use;
    if constexpr (CTA_DIRECT) {
      s = cs, use = cu;
      cs = cs + 1 == stages ? 0 : cs + 1;
      cu += cs == 0;
    } else {
      s = it % stages, use = it / stages;
    }
    mbar_wait(full + s, use & 1);
    StageMeta* m = meta + s;
    const int ns = m->nstaged, nrows = m->nrows, r0 = m->r0;
    if (ns > 0) {
      if (WALK == WALK_STREAM || walk_is_seg(WALK) || WALK == WALK_SPLIT) {
        dot = dd_merge(dot, walk_stream<T, SYM, LONG, DOT>(a, smem + (size_t)s * sb, m, threadIdx.x, CW * 32,
                                                           warp, CW, lane, 1));
        fence_proxy_async_smem(); // our generic-proxy writes precede the next bulk copy into this stage
      } else {
        dot = dd_merge(dot, walk_staged<T, SYM, LONG, DOT, CHUNK, LAY, SINGLE>(a, smem + (size_t)s * sb, m,
                                                                               threadIdx.x, CW * 32, warp, CW, lane, 1));
      }
    }
    if (ns < nrows) {
      if constexpr (CTA_DIRECT)
        dot = dd_merge(dot, walk_direct_cta<T, SYM, DOT>(a, r0 + ns, nrows - ns, threadIdx.x, CW * 32, warp, CW,
                                                         lane, 1, s_warp));
      else
        dot = dd_merge(dot, walk_direct<T, SYM, DOT>(a, r0 + ns, nrows - ns, warp, CW, lane));
    }
    __syncwarp();
    if (lane == 0)
      mbar_arrive(empty + s);
  }
  if (DOT)
    grid_dd_reduce(dot, threadIdx.x, CW * 32, 2, s_warp, a.partials, a.partials_len, a.counter, a.dot_out);
}

// ---------------------------------------------------------------------------
// Variant 1: one tile per CTA, staged with plain vector loads, row walk.  Used
// when a base pointer is not 16-byte aligned, and as the comparison point for
// variant 0.
// ---------------------------------------------------------------------------
template <typename T, bool SYM, bool LONG, int NT>
__global__ void __launch_bounds__(NT) spmv_tile_kernel(const SpmvArgs<T> a)
{
  extern __shared__ __align__(128) unsigned char smem[];
  const int sb = a.stage_bytes;
  StageMeta* m = reinterpret_cast<StageMeta*>(smem + sb);
  const int tile = a.tile0 + blockIdx.x;
  const int r0 = __ldg(a.tile_meta + 2 * tile), lo = __ldg(a.tile_meta + 2 * tile + 1);
  const int r1 = __ldg(a.tile_meta + 2 * tile + 2);
  const int nrows = r1 - r0;
  if (nrows == 0)
    return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int ns = nrows, hi = __ldg(a.tile_meta + 2 * tile + 3);
  while (ns > 0 && !tile_fits<T, LAY_PLAIN>(r0, ns, lo, hi, sb)) {
    --ns;
    hi = __ldg(a.rowptr + r0 + ns);
  }
  if (ns > 0) {
    const TileLayout<T, LAY_PLAIN> L(r0, ns, lo, hi);
    if (threadIdx.x == 0) {
      m->r0 = r0;
      m->nrows = nrows;
      m->lo = lo;
      m->hi = hi;
      m->nstaged = ns;
      m->nlong = 0;
    }
    T* sv = reinterpret_cast<T*>(smem);
    int* sc = reinterpret_cast<int*>(smem + L.off_c());
    int* sr = reinterpret_cast<int*>(smem + L.off_r());
    for (int j = lo + threadIdx.x; j < hi; j += NT)
      sv[j - L.lo_v] = __ldg(a.val + j);
    for (int j = lo + threadIdx.x; j < hi; j += NT)
      sc[j - L.lo_c] = __ldg(a.colind + j);
    for (int r = r0 + threadIdx.x; r <= r0 + ns; r += NT)
      sr[r - L.lo_r] = __ldg(a.rowptr + r);
    __syncthreads();
    walk_staged<T, SYM, LONG, false, 16, LAY_PLAIN>(a, smem, m, threadIdx.x, NT, warp, NT / 32, lane, 1);
  }
  if (ns < nrows)
    walk_direct<T, SYM, false>(a, r0 + ns, nrows - ns, warp, NT / 32, lane);
}

// ---------------------------------------------------------------------------
// Sparse-row blocks: the remote block of a slab-partitioned operator has entries
// in the few rows that touch a ghost plane (2 planes of 256 for the Poisson slabs)
// and is empty elsewhere.  With beta == 1 an empty row leaves y untouched, so only
// the listed (non-empty) rows are visited: no pass over the whole rowptr and y
// (335 MB of traffic for 1.2 MB of entries).  One thread per listed row in j order
// (same roundings as csr_kernels.cpp:45-47), or one warp per row when rows are long.
// ---------------------------------------------------------------------------
template <typename T>
__global__ void spmv_rowlist_kernel(int n, const int32_t* __restrict__ rowlist,
                                    const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colind,
                                    const T* __restrict__ val, const T* __restrict__ x, T alpha, T beta,
                                    T* __restrict__ y)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n)
    return;
  const int r = __ldg(rowlist + k);
  const int rs = __ldg(rowptr + r), re = __ldg(rowptr + r + 1);
  T sum = 0;
  for (int j = rs; j < re; ++j)
    sum = add_rn(sum, mul_rn(__ldg(val + j), __ldg(x + __ldg(colind + j))));
  y[r] = add_rn(mul_rn(alpha, sum), mul_rn(beta, y[r]));
}

// Mixed row lengths (the remote block of an irregular matrix: a mean of 1-2 entries per listed
// row, a few rows of hundreds): lanes = consecutive listed rows as above; rows longer than
// SHORT_ROW are handed to the whole warp one after the other (lane-strided + xor-tree).  One
// warp per row for every row cost 430 us on config 5's remote block at 2 GPUs (1.8 M listed
// rows, 1.3 entries each: three dependent round trips per warp for one product).
template <typename T>
__global__ void spmv_rowlist_mixed_kernel(int n, const int32_t* __restrict__ rowlist,
                                          const int32_t* __restrict__ rowptr,
                                          const int32_t* __restrict__ colind, const T* __restrict__ val,
                                          const T* __restrict__ x, T alpha, T beta, T* __restrict__ y)
{
  constexpr unsigned FULL = 0xffffffffu;
  const int k = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
  int r = 0, rs = 0, re = 0;
  if (k < n) {
    r = __ldg(rowlist + k);
    rs = __ldg(rowptr + r);
    re = __ldg(rowptr + r + 1);
  }
  const bool is_long = re - rs > SHORT_ROW;
  if (k < n && !is_long) {
    T sum = 0;
    for (int j = rs; j < re; ++j)
      sum = add_rn(sum, mul_rn(__ldg(val + j), __ldg(x + __ldg(colind + j))));
    y[r] = add_rn(mul_rn(alpha, sum), mul_rn(beta, y[r]));
  }
  unsigned todo = __ballot_sync(FULL, is_long);
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const int qs = __shfl_sync(FULL, rs, src), qe = __shfl_sync(FULL, re, src), qr = __shfl_sync(FULL, r, src);
    T part = 0;
    for (int j = qs + lane; j < qe; j += 32)
      part = fma(__ldg(val + j), __ldg(x + __ldg(colind + j)), part);
    part = warp_sum(part);
    if (lane == 0)
      y[qr] = add_rn(mul_rn(alpha, part), mul_rn(beta, y[qr]));
  }
}

// The remote block of a two-block operator with the x.(A x) epilogue COMPLETED (fp64, alpha = beta = 1):
// y[r] += sum_j val[j] x[col[j]] over the listed rows as spmv_rowlist_kernel (same roundings), and
//   dot_out[0] = round( B + sum_listed c_r ),
//   B   = the double-double sum the LOCAL block's dot epilogue left as per-CTA partials in its plan
//         (nbase pairs: hi at base[i], lo at base[base_plen + i]);
//   c_r = x_r * y_new[r] - x_r * y_old[r]   general storage: the local epilogue counted x_r * y_old[r], and with
//         error-free products summed in twice the working precision the total is sum_r x_r * y_final[r] to
//         ~1e-32 -- the same double a single dot product over the final y rounds to;
//   c_r = x_r * (row sum)                   symmetric local block (its epilogue holds x'A_local x already).
// halo != nullptr: the kernel is also the WAIT half of the kernel-driven halo exchange (halo.cu): every CTA polls
//   this rank's flag words until each neighbour has published the epoch, and the ghost entries of x (columns >=
//   ghost_base) are read straight from the staging slot the neighbours wrote -- no unpack launch, no ghost-tail
//   round trip (the tail of x is left as it was).
// peer != nullptr: the last CTA also exchanges the total with the other ranks (the MPI_Allreduce of cg.cpp:65).
template <bool SYMLOCAL>
__global__ void __launch_bounds__(128)
    spmv_rowlist_dot_kernel(int n, const int32_t* __restrict__ rowlist, const int32_t* __restrict__ rowptr,
                            const int32_t* __restrict__ colind, const double* __restrict__ val,
                            const double* __restrict__ x, double* __restrict__ y, const double* __restrict__ base,
                            int base_plen, int nbase, double* __restrict__ part, int plen, unsigned int* counter,
                            double* __restrict__ dot_out, PeerDesc* peer, const HaloDesc* halo, int ghost_base)
{
  constexpr int NT = 128, NW = NT / 32;
  __shared__ double s_warp[2 * NW];
  __shared__ bool s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double* xg = x; // xg[c] = x[c] for ghost columns
  if (halo != nullptr) {
    const unsigned int epoch = halo->epoch; // the push of this exchange precedes us on the stream
    if (tid < halo->nnbr) {
      unsigned int seen;
      do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(halo->my_flag[tid]) : "memory");
      } while ((int)(seen - epoch) < 0);
    }
    __syncthreads();
    xg = reinterpret_cast<const double*>(halo->my_slot[epoch & 1u]) - ghost_base;
  }
  dd dot = {0.0, 0.0};
  for (int k = blockIdx.x * NT + tid; k < n; k += gridDim.x * NT) {
    const int r = __ldg(rowlist + k);
    const int rs = __ldg(rowptr + r), re = __ldg(rowptr + r + 1);
    double sum = 0;
    for (int j = rs; j < re; ++j) {
      const int c = __ldg(colind + j);
      const double xv = c >= ghost_base ? __ldcg(xg + c) : __ldg(x + c);
      sum = add_rn(sum, mul_rn(__ldg(val + j), xv));
    }
    const double yo = y[r];
    const double yn = add_rn(sum, yo);
    y[r] = yn;
    const double xr = __ldg(x + r);
    if (SYMLOCAL) {
      dd_add_prod(dot, xr, sum);
    } else {
      dd_add_prod(dot, xr, yn);
      dd_add_prod(dot, -xr, yo);
    }
  }
  dot = dd_warp_sum(dot);
  if (lane == 0) {
    s_warp[2 * warp] = dot.hi;
    s_warp[2 * warp + 1] = dot.lo;
  }
  __syncthreads();
  if (warp == 0) {
    dd t;
    t.hi = lane < NW ? s_warp[2 * lane] : 0.0;
    t.lo = lane < NW ? s_warp[2 * lane + 1] : 0.0;
    t = dd_warp_sum(t);
    if (lane == 0) {
      part[blockIdx.x] = t.hi;
      part[plen + blockIdx.x] = t.lo;
      __threadfence();
      s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
  }
  __syncthreads();
  if (!s_last)
    return;
  // the last CTA folds the local block's partials and this grid's with all its warps
  __threadfence();
  dd s = {0.0, 0.0};
  for (int i = tid; i < nbase; i += NT) {
    dd w;
    w.hi = __ldcg(base + i);
    w.lo = __ldcg(base + base_plen + i);
    s = dd_merge(s, w);
  }
  for (int i = tid; i < (int)gridDim.x; i += NT) {
    dd w;
    w.hi = __ldcg(part + i);
    w.lo = __ldcg(part + plen + i);
    s = dd_merge(s, w);
  }
  s = dd_warp_sum(s);
  if (lane == 0) {
    s_warp[2 * warp] = s.hi;
    s_warp[2 * warp + 1] = s.lo;
  }
  __syncthreads();
  if (warp == 0) {
    dd t;
    t.hi = lane < NW ? s_warp[2 * lane] : 0.0;
    t.lo = lane < NW ? s_warp[2 * lane + 1] : 0.0;
    t = dd_warp_sum(t);
    double mine[B200_PEER_REDUCE_MAX_COUNT] = {__dadd_rn(t.hi, t.lo), 0.0, 0.0, 0.0};
    double total[B200_PEER_REDUCE_MAX_COUNT] = {mine[0], 0.0, 0.0, 0.0};
    if (peer != nullptr)
      peer_exchange_warp(peer, lane, 1, mine, total);
    if (lane == 0) {
      dot_out[0] = total[0];
      *counter = 0u;
    }
  }
}

struct RowNonEmpty {
  const int32_t* rowptr;
  __host__ __device__ bool operator()(int r) const { return rowptr[r + 1] > rowptr[r]; }
};

// Diagonal-only symmetric block (nnz == 0, rowptr == NULL): csr_matrix.cpp:85,
// openmp/csr_kernels.openmp.cpp:222-225.
template <typename T>
__global__ void diag_only_kernel(int n, const T* __restrict__ diag, const T* __restrict__ x, T alpha,
                                 T beta, T* __restrict__ y)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    T out = mul_rn(alpha, mul_rn(diag[i], x[i]));
    if (beta != T(0))
      out = add_rn(out, mul_rn(beta, y[i]));
    y[i] = out;
  }
}

// Zero-fill of y in front of the symmetric kernel (beta == 0), instead of cudaMemsetAsync: from
// the LAST row down to the first and with an evict_last L2 policy, so that what is still in L2
// when the walk starts at row 0 is the low end of y, and the matrix stream (evict_first) does
// not push it out before the adds arrive.  n16 = number of 16-byte units.
__global__ void zero_fill_desc_kernel(size_t n16, float4* __restrict__ y)
{
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n16; k += stride)
    asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %1, %1, %1}, %2;" ::"l"(y + (n16 - 1 - k)), "f"(0.f),
                 "l"(pol)
                 : "memory");
}

template <typename T>
__global__ void scale_kernel(int n, T beta, T* __restrict__ y)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    y[i] = mul_rn(beta, y[i]);
}

// ---------------------------------------------------------------------------
// Plan kernels
// ---------------------------------------------------------------------------
__global__ void plan_tiles_kernel(const int32_t* __restrict__ rowptr, int rows, long long nnz,
                                  int num_tiles, long long tile_bytes, long long w_entry,
                                  long long w_row, int32_t* __restrict__ meta)
{
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > num_tiles)
    return;
  if (b == num_tiles) {
    meta[2 * b] = rows;
    meta[2 * b + 1] = (int32_t)nnz;
    return;
  }
  const long long target = (long long)b * tile_bytes;
  int lo = 0, hi = rows; // smallest r with w_entry*rowptr[r] + w_row*r >= target
  while (lo < hi) {
    const int mid = lo + ((hi - lo) >> 1);
    const long long g = w_entry * rowptr[mid] + w_row * mid;
    if (g >= target)
      hi = mid;
    else
      lo = mid + 1;
  }
  meta[2 * b] = lo;
  meta[2 * b + 1] = rowptr[lo];
}

// WALK_SEG: cut every tile into `cw` segments of (nearly) equal entry counts at row boundaries.
// out[t*(cw+1) + w] = first row of tile t (as an offset from the tile's first row) that starts at
// or after entry lo + n*w/cw; segment w = rows [out[w], out[w+1]).
__global__ void plan_warp_rows_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ meta,
                                      int num_tiles, int cw, int stride, int32_t* __restrict__ out)
{
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)num_tiles * stride)
    return;
  const int t = (int)(i / stride), w = (int)(i % stride);
  if (w > cw) { // padding up to the 16-byte multiple the bulk copy moves
    out[i] = 0;
    return;
  }
  const int r0 = meta[2 * t], lo = meta[2 * t + 1], r1 = meta[2 * t + 2], hi = meta[2 * t + 3];
  const int target = lo + (int)((long long)(hi - lo) * w / cw);
  int x0 = r0, x1 = r1; // first row in [r0, r1] with rowptr[row] >= target
  if (w == cw)
    x0 = r1;
  while (x0 < x1) {
    const int mid = x0 + ((x1 - x0) >> 1);
    if (rowptr[mid] >= target)
      x1 = mid;
    else
      x0 = mid + 1;
  }
  out[i] = x0 - r0;
}

__global__ void plan_hist_kernel(const int32_t* __restrict__ rowptr, int rows,
                                 unsigned long long* __restrict__ hist, int* __restrict__ max_len)
{
  __shared__ unsigned int s_hist[32];
  __shared__ int s_max;
  if (threadIdx.x < 32)
    s_hist[threadIdx.x] = 0;
  if (threadIdx.x == 0)
    s_max = 0;
  __syncthreads();
  int local_max = 0;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += gridDim.x * blockDim.x) {
    const int len = rowptr[r + 1] - rowptr[r];
    const int bucket = len <= 0 ? 0 : 32 - __clz(len);
    atomicAdd(&s_hist[bucket > 31 ? 31 : bucket], 1u);
    local_max = max(local_max, len);
  }
  atomicMax(&s_max, local_max);
  __syncthreads();
  if (threadIdx.x < 32 && s_hist[threadIdx.x] > 0)
    atomicAdd(&hist[threadIdx.x], (unsigned long long)s_hist[threadIdx.x]);
  if (threadIdx.x == 0)
    atomicMax(max_len, s_max);
}

// Column extent of the entries of each chunk of tiles (b200_csr_plan_chunking).
__global__ void chunk_extent_kernel(const int32_t* __restrict__ colind, const int32_t* __restrict__ ent_bounds,
                                    int32_t* __restrict__ cmin, int32_t* __restrict__ cmax)
{
  const int k = blockIdx.y;
  const long long lo = ent_bounds[k], hi = ent_bounds[k + 1];
  int mn = 0x7fffffff, mx = -1;
  for (long long j = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; j < hi; j += (long long)gridDim.x * blockDim.x) {
    const int c = colind[j];
    mn = min(mn, c);
    mx = max(mx, c);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if ((threadIdx.x & 31) == 0 && mx >= 0) {
    atomicMin(cmin + k, mn);
    atomicMax(cmax + k, mx);
  }
}

static int env_int(const char* name, int dflt)
{
  const char* v = std::getenv(name);
  return (v && *v) ? std::atoi(v) : dflt;
}

static bool plan_walk_is_irr(const Plan& p) { return !p.symmetric && (walk_is_seg(p.walk) || p.walk == WALK_SPLIT); }
static size_t pipe_smem_bytes(const Plan& p)
{
  if (plan_walk_is_irr(p)) // spmv_irr_kernel: a third barrier per stage, more reduction scratch
    return (size_t)p.stages * (p.stage_bytes + sizeof(StageMeta)) + 24 * p.stages + 8 * 64;
  return (size_t)p.stages * (p.stage_bytes + sizeof(StageMeta)) + 16 * p.stages + 8 * 32;
}

static int plan_partition(Plan* p, const int32_t* d_rowptr, cudaStream_t st)
{
  if (p->d_tile_meta) {
    B200_CUDA(cudaFree(p->d_tile_meta));
    p->d_tile_meta = nullptr;
  }
  if (p->d_warp_rows) {
    B200_CUDA(cudaFree(p->d_warp_rows));
    p->d_warp_rows = nullptr;
  }
  p->num_tiles = 0;
  if (p->nnz == 0 || p->rows == 0 || d_rowptr == nullptr)
    return B200_OK;
  // walk + storage decide what a row and an entry cost in the stage
  p->w_entry = 12;
  p->w_row = 4;
  if (p->symmetric && p->variant == 0) {
    p->w_row = 20; // rowptr + diagonal + x
    if (p->walk == WALK_STREAM || walk_is_seg(p->walk) || p->walk == WALK_SPLIT)
      p->w_entry = 20; // value + column + product
  }
  const long long total = (long long)p->w_entry * p->nnz + (long long)p->w_row * p->rows;
  const long long nt = (total + p->tile_bytes - 1) / p->tile_bytes;
  B200_REQUIRE(nt < (1ll << 30), "too many tiles");
  p->num_tiles = (int32_t)nt;
  B200_CUDA(cudaMalloc(&p->d_tile_meta, sizeof(int32_t) * 2 * (size_t)(nt + 1)));
  const int threads = 256;
  const int blocks = (int)((nt + 1 + threads - 1) / threads);
  plan_tiles_kernel<<<blocks, threads, 0, st>>>(d_rowptr, p->rows, p->nnz, p->num_tiles,
                                                p->tile_bytes, p->w_entry, p->w_row, p->d_tile_meta);
  B200_LAUNCH_CHECK();
  if (walk_is_seg(p->walk) && !p->symmetric && p->variant == 0) {
    const int stride = seg_stride(p->consumer_warps);
    const long long n = nt * stride;
    B200_CUDA(cudaMalloc(&p->d_warp_rows, sizeof(int32_t) * (size_t)n));
    plan_warp_rows_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, st>>>(
        d_rowptr, p->d_tile_meta, p->num_tiles, p->consumer_warps, stride, p->d_warp_rows);
    B200_LAUNCH_CHECK();
  }
  return B200_OK;
}

// The row-length histogram picks the walk (sweeps in profiles/r01_walk_sweeps.md):
//   general storage: 16-entry chunks for stencil-length rows (27-point: mean 26.8),
//     8-entry chunks when the mean is <= 10 (7-point: 0.306 -> 0.292 ms);
//   symmetric storage (lower-triangle rows are short): 4-entry chunks up to a mean of
//     4 (7-point lower: mean 3; 0.87 -> 0.23 ms), else 8-entry chunks (27-point lower:
//     mean 12.9; 1.49 -> 0.69 ms);
//   irregular rows (longest row > 16x the mean, e.g. power-law): the stream walk,
//     whose entry-parallel phase does not care how the entries fall into rows.
static void plan_choose_walk(Plan* p)
{
  if (p->walk_forced >= 0) {
    p->walk = p->walk_forced;
    return;
  }
  const double mean = p->rows > 0 ? (double)p->nnz / (double)p->rows : 0.0;
  if (p->max_row_len > SHORT_ROW && (double)p->max_row_len > 16.0 * mean)
    p->walk = p->symmetric ? WALK_STREAM : WALK_SPLIT; // general: 0.165 -> 0.139 ms on the 3 M-row power-law matrix
  else if (p->symmetric)
    p->walk = mean <= 4.0 ? WALK_ROW4 : WALK_ROW8;
  else
    p->walk = mean <= 10.0 ? WALK_ROW8 : WALK_ROW16;
}

// Launch shape that measured best for each walk (tile bytes, pipeline stages, consumer warps,
// CTAs/SM).  More than two stages only pays for the symmetric 4-entry walk, whose gathers are
// mostly served from the staged x slice: everywhere else the shared memory taken from L1 costs
// more than the deeper pipeline gives (profiles/r01_exp_za_probe3_stage_sweep.txt,
// r01_stage_sweep_head.txt: 27-point general 0.85 -> 1.05 ms with 3 stages).
static void plan_default_shape(const Plan* p, int* tile_bytes, int* stages, int* cw, int* ctas)
{
  *stages = 2;
  switch (p->walk) {
  case WALK_ROW4:
    if (p->symmetric)
      *tile_bytes = 20 * 1024, *stages = 3, *cw = 16, *ctas = 2; // sym 7-pt 0.222 -> 0.215 ms
    else
      *tile_bytes = 24 * 1024, *cw = 16, *ctas = 2;
    break;
  case WALK_ROW8:
    if (p->symmetric)
      *tile_bytes = 32 * 1024, *cw = 8, *ctas = 3; // sym 27-pt 0.702 -> 0.634 ms
    else
      *tile_bytes = 24 * 1024, *cw = 8, *ctas = 4;
    break;
  case WALK_STREAM:
    *tile_bytes = 24 * 1024, *cw = 16, *ctas = 2;
    break;
  case WALK_SEG:
  case WALK_SEGNA:
    *tile_bytes = 16 * 1024, *stages = 3, *cw = 16, *ctas = 2;
    break;
  case WALK_SPLIT:
    *tile_bytes = 16 * 1024, *stages = 3, *cw = 16, *ctas = 2;
    break;
  default:
    *tile_bytes = 48 * 1024, *cw = 8, *ctas = 2;
  }
}

// b200_csr_plan_tune's variant -> forced walk (-1: chosen from the histogram)
static int walk_of_variant(int variant)
{
  return variant == 8 ? WALK_SPLIT : variant == 7 ? WALK_SEGNA : variant == 6 ? WALK_SEG : (variant >= 2 && variant <= 5 ? variant - 2 : -1);
}

static int plan_apply_tuning(Plan* p, int variant, int tile_bytes, int stages, int cw, int ctas)
{
  B200_REQUIRE(variant >= 0 && variant <= 8, "variant must be 0..8");
  // 0: pipeline, walk chosen from the histogram; 1: one tile per CTA (row walk);
  // 2..5: pipeline with the walk forced to WALK_ROW16 / WALK_STREAM / WALK_ROW8 / WALK_ROW4;
  // 6: WALK_SEG (general storage; symmetric plans fall back to WALK_STREAM); 7: WALK_SEGNA; 8: WALK_SPLIT
  p->walk_forced = walk_of_variant(variant);
  if (variant >= 2)
    variant = 0;
  B200_REQUIRE(tile_bytes >= 1024 && tile_bytes + 2048 <= MAX_STAGE_BYTES, "tile_bytes out of range");
  B200_REQUIRE(stages >= 2 && stages <= 8, "stages out of range");
  B200_REQUIRE(cw == 8 || cw == 16, "consumer_warps must be 8 or 16");
  B200_REQUIRE(ctas >= 1 && ctas <= 8, "ctas_per_sm out of range");
  p->variant = variant;
  p->tile_bytes = tile_bytes;
  p->stage_bytes = tile_bytes + 2048; // slack: one row of up to ~160 entries
  p->stages = stages;
  p->consumer_warps = cw;
  p->ctas_per_sm = ctas;
  B200_REQUIRE(pipe_smem_bytes(*p) <= 227 * 1024, "stages * stage_bytes exceeds shared memory");
  return B200_OK;
}

// ---------------------------------------------------------------------------
// Launch
// ---------------------------------------------------------------------------
// cudaFuncAttributeMaxDynamicSharedMemorySize is per device: cache per device.
struct SmemCache {
  size_t configured[64] = {0};
};

template <typename Kernel>
static int set_smem(Kernel k, size_t bytes, SmemCache& cache)
{
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || cache.configured[dev] < bytes) {
    B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    if (dev >= 0 && dev < 64)
      cache.configured[dev] = bytes;
  }
  return B200_OK;
}

template <typename T, bool SYM, bool DOT, bool LONG, int CW, int WALK>
static int launch_pipe(const Plan& p, const SpmvArgs<T>& a, cudaStream_t st)
{
  const size_t smem = pipe_smem_bytes(p);
  static SmemCache cache;
  const int grid = std::min<long long>(a.num_tiles - a.tile0, (long long)num_sms() * p.ctas_per_sm);
  if constexpr (!SYM && (walk_is_seg(WALK) || WALK == WALK_SPLIT)) {
    if (int rc = set_smem(spmv_irr_kernel<T, SYM, DOT, LONG, CW, WALK>, smem, cache))
      return rc;
    spmv_irr_kernel<T, SYM, DOT, LONG, CW, WALK><<<grid, (CW + 1) * 32, smem, st>>>(a);
  } else {
    if (int rc = set_smem(spmv_pipe_kernel<T, SYM, DOT, LONG, CW, WALK>, smem, cache))
      return rc;
    spmv_pipe_kernel<T, SYM, DOT, LONG, CW, WALK><<<grid, (CW + 1) * 32, smem, st>>>(a);
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename T, bool SYM, bool DOT, bool LONG, int WALK>
static int launch_pipe_cw(const Plan& p, const SpmvArgs<T>& a, cudaStream_t st)
{
  switch (p.consumer_warps) {
  case 16:
    return launch_pipe<T, SYM, DOT, LONG, 16, WALK>(p, a, st);
  default:
    return launch_pipe<T, SYM, DOT, LONG, 8, WALK>(p, a, st);
  }
}

template <typename T, bool SYM, bool DOT, bool LONG>
static int launch_pipe_walk(const Plan& p, const SpmvArgs<T>& a, cudaStream_t st)
{
  switch (p.walk) {
  case WALK_STREAM:
    return launch_pipe_cw<T, SYM, DOT, LONG, WALK_STREAM>(p, a, st);
  case WALK_SEG:
    if constexpr (SYM) // the warp-private walk is general-storage only
      return launch_pipe_cw<T, SYM, DOT, LONG, WALK_STREAM>(p, a, st);
    else
      return launch_pipe_cw<T, SYM, DOT, false, WALK_SEG>(p, a, st);
  case WALK_SEGNA:
    if constexpr (SYM)
      return launch_pipe_cw<T, SYM, DOT, LONG, WALK_STREAM>(p, a, st);
    else
      return launch_pipe_cw<T, SYM, DOT, false, WALK_SEGNA>(p, a, st);
  case WALK_SPLIT:
    if constexpr (SYM)
      return launch_pipe_cw<T, SYM, DOT, LONG, WALK_STREAM>(p, a, st);
    else
      return launch_pipe_cw<T, SYM, DOT, false, WALK_SPLIT>(p, a, st);
  case WALK_ROW8:
    if (p.max_row_len <= 8)
      return launch_pipe_cw<T, SYM, DOT, false, WALK_ROW8S>(p, a, st);
    return launch_pipe_cw<T, SYM, DOT, LONG, WALK_ROW8>(p, a, st);
  case WALK_ROW4:
    if (p.max_row_len <= 4)
      return launch_pipe_cw<T, SYM, DOT, false, WALK_ROW4S>(p, a, st);
    return launch_pipe_cw<T, SYM, DOT, LONG, WALK_ROW4>(p, a, st);
  default:
    return launch_pipe_cw<T, SYM, DOT, LONG, WALK_ROW16>(p, a, st);
  }
}

template <typename T, bool SYM, bool LONG>
static int launch_tile(const Plan& p, const SpmvArgs<T>& a, cudaStream_t st)
{
  constexpr int NT = 256;
  const size_t smem = p.stage_bytes + sizeof(StageMeta);
  static SmemCache cache;
  if (int rc = set_smem(spmv_tile_kernel<T, SYM, LONG, NT>, smem, cache))
    return rc;
  spmv_tile_kernel<T, SYM, LONG, NT><<<a.num_tiles - a.tile0, NT, smem, st>>>(a);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename T, bool SYM>
static int run_spmv(void* plan_v, const int32_t* d_rowptr, const int32_t* d_colind, const T* d_val,
                    const T* d_diag, T alpha, const T* d_x, T beta, T* d_y, double* d_dot,
                    void* stream, int tile_begin = 0, int tile_end = -1)
{
  B200_REQUIRE(plan_v != nullptr, "plan is NULL");
  const Plan& p = *static_cast<const Plan*>(plan_v);
  B200_REQUIRE(SYM == (p.symmetric != 0), "plan / call symmetric mismatch");
  B200_REQUIRE(p.rows == 0 || (d_x != nullptr && d_y != nullptr), "x or y is NULL");
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  cudaStream_t st = as_stream(stream);
  if (p.rows == 0)
    return B200_OK;
  const int tb = 256, gb = (p.rows + tb - 1) / tb;
  // tile sub-range (b200_csr_*_tiles_f64): only the rows of tiles [tile_begin, tile_end);
  // symmetric storage then expects y already prepared (beta*y) by the caller
  const bool ranged = tile_end >= 0;
  if (ranged) {
    B200_REQUIRE(tile_begin >= 0 && tile_begin <= tile_end && tile_end <= p.num_tiles, "bad tile range");
    B200_REQUIRE(d_dot == nullptr, "tile ranges have no dot epilogue");
    if (tile_begin == tile_end)
      return B200_OK;
  }

  if (SYM && ranged) {
    B200_REQUIRE(d_diag != nullptr, "symmetric SpMV needs the diagonal");
  } else if (SYM) {
    B200_REQUIRE(d_diag != nullptr, "symmetric SpMV needs the diagonal");
    if (p.num_tiles == 0) {
      B200_REQUIRE(d_dot == nullptr, "symmetric dot epilogue needs a block with stored entries");
      diag_only_kernel<T><<<gb, tb, 0, st>>>(p.rows, d_diag, d_x, alpha, beta, d_y);
      B200_LAUNCH_CHECK();
      return B200_OK;
    }
    // every write below is an add: bring y to beta*y first
    if (beta == T(0)) {
      static const int zfill = env_int("B200_SPMV_ZFILL", 1);
      const size_t bytes = sizeof(T) * (size_t)p.rows;
      if (zfill && aligned16(d_y) && bytes >= (1u << 16)) {
        const size_t n16 = bytes / 16;
        zero_fill_desc_kernel<<<num_sms() * 8, 256, 0, st>>>(n16, reinterpret_cast<float4*>(d_y));
        B200_LAUNCH_CHECK();
        if (bytes > n16 * 16)
          B200_CUDA(cudaMemsetAsync(reinterpret_cast<char*>(d_y) + n16 * 16, 0, bytes - n16 * 16, st));
      } else {
        B200_CUDA(cudaMemsetAsync(d_y, 0, bytes, st));
      }
    } else if (beta != T(1)) {
      scale_kernel<T><<<gb, tb, 0, st>>>(p.rows, beta, d_y);
      B200_LAUNCH_CHECK();
    }
  } else if (p.num_tiles == 0) {
    // empty block: y = beta*y
    if (beta == T(0))
      B200_CUDA(cudaMemsetAsync(d_y, 0, sizeof(T) * (size_t)p.rows, st));
    else if (beta != T(1)) {
      scale_kernel<T><<<gb, tb, 0, st>>>(p.rows, beta, d_y);
      B200_LAUNCH_CHECK();
    }
    if (d_dot)
      B200_CUDA(cudaMemsetAsync(d_dot, 0, sizeof(double), st));
    return B200_OK;
  }
  B200_REQUIRE(d_rowptr && d_colind && d_val, "matrix arrays are NULL but nnz > 0");

  if (!SYM && !ranged && p.d_rowlist != nullptr && beta == T(1) && d_dot == nullptr) {
    const int n = p.n_rowlist;
    if (p.max_row_len <= SHORT_ROW)
      spmv_rowlist_kernel<T><<<(n + 127) / 128, 128, 0, st>>>(n, p.d_rowlist, d_rowptr, d_colind, d_val, d_x,
                                                              alpha, beta, d_y);
    else
      spmv_rowlist_mixed_kernel<T><<<(n + 127) / 128, 128, 0, st>>>(n, p.d_rowlist, d_rowptr, d_colind, d_val, d_x,
                                                                    alpha, beta, d_y);
    B200_LAUNCH_CHECK();
    return B200_OK;
  }

  SpmvArgs<T> a;
  a.rowptr = d_rowptr;
  a.colind = d_colind;
  a.val = d_val;
  a.diag = d_diag;
  a.x = d_x;
  a.y = d_y;
  a.alpha = alpha;
  a.beta = beta;
  a.tile_meta = p.d_tile_meta;
  a.warp_rows = p.d_warp_rows;
  {
    static const int seg_ef = env_int("B200_SPMV_SEG_EVICT_FIRST", 0);
    static const int contig = env_int("B200_SPMV_TILE_CONTIG", 0);
    // B200_SPMV_STREAM_EVICT_FIRST: -1 (default) by size -- general blocks of at most B200_SPMV_L2_ROWS rows
    // (default 6 M); 0 / 1 force.  CG on the 27-point operator, one B200, us per iteration without / with:
    // 2.1 M rows 158.7 / 147.0, 4.2 M 294.6 / 283.6, 8.4 M 572.2 / 574.3, 16.8 M 1114 / 1125 (the five CG vectors
    // of 2.1 M rows are 84 MB of the 126 MB L2; at 4.2 M part of them still survives the matrix stream)
    static const int gen_ef = env_int("B200_SPMV_STREAM_EVICT_FIRST", -1);
    static const int l2_rows = env_int("B200_SPMV_L2_ROWS", 6000000);
    const bool ef = gen_ef >= 0 ? gen_ef != 0 : (!SYM && p.rows <= l2_rows);
    a.flags = (seg_ef ? 1 : 0) | (contig ? 2 : 0) | (ef ? 4 : 0);
  }
  a.tile0 = ranged ? tile_begin : 0;
  a.num_tiles = ranged ? tile_end : p.num_tiles;
  a.stage_bytes = p.stage_bytes;
  a.stages = p.stages;
  a.partials = p.d_partials;
  a.partials_len = p.partials_len;
  a.counter = p.d_counter;
  a.dot_out = d_dot;

  const bool has_long = p.max_row_len > SHORT_ROW;
  const bool bulk_ok = aligned16(d_rowptr) && aligned16(d_colind) && aligned16(d_val)
                       && (!SYM || (aligned16(d_diag) && aligned16(d_x)));
  if (p.variant == 0 && bulk_ok) {
    if (d_dot != nullptr) {
      B200_REQUIRE((std::is_same<T, double>::value), "dot epilogue: fp64 only");
      // symmetric storage: the quadratic-form identity of dot_row() holds for y = A x only
      B200_REQUIRE(!SYM || (alpha == T(1) && beta == T(0) && !ranged), "symmetric dot epilogue: alpha = 1, beta = 0");
      B200_REQUIRE((long long)num_sms() * p.ctas_per_sm <= p.partials_len, "partials too small");
      constexpr bool D = !SYM || std::is_same<T, double>::value; // (no fp32 symmetric instantiation)
      if (has_long)
        return launch_pipe_walk<T, SYM, D, true>(p, a, st);
      return launch_pipe_walk<T, SYM, D, false>(p, a, st);
    }
    if (has_long)
      return launch_pipe_walk<T, SYM, false, true>(p, a, st);
    return launch_pipe_walk<T, SYM, false, false>(p, a, st);
  }
  B200_REQUIRE(d_dot == nullptr, "dot epilogue needs variant 0 and 16-byte aligned arrays");
  if (has_long)
    return launch_tile<T, SYM, true>(p, a, st);
  return launch_tile<T, SYM, false>(p, a, st);
}

} // namespace b200

using namespace b200;

extern "C" {

int b200_csr_plan_create(void** plan, int32_t rows, int32_t cols, int64_t nnz,
                         const int32_t* d_rowptr, int symmetric, void* stream)
{
  B200_REQUIRE(plan != nullptr, "plan is NULL");
  *plan = nullptr;
  B200_REQUIRE(rows >= 0 && cols >= 0 && nnz >= 0, "negative size");
  B200_REQUIRE(nnz < (1ll << 31), "nnz does not fit the int32 rowptr");
  B200_REQUIRE(nnz == 0 || d_rowptr != nullptr, "rowptr is NULL but nnz > 0");
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  Plan* p = new (std::nothrow) Plan();
  if (!p)
    return B200_ENOMEM;
  p->rows = rows;
  p->cols = cols;
  p->nnz = nnz;
  p->symmetric = symmetric ? 1 : 0;
  int rc = B200_OK;
  cudaStream_t st = as_stream(stream);
  p->partials_len = num_sms() * 8;
  unsigned long long* d_hist = nullptr;
  int* d_max = nullptr;
  auto fail = [&](int code) {
    cudaFree(d_hist);
    cudaFree(d_max);
    b200_csr_plan_destroy(p);
    return code;
  };
  if (cudaMalloc(&p->d_partials, sizeof(double) * 2 * p->partials_len) != cudaSuccess
      || cudaMalloc(&p->d_counter, sizeof(unsigned int)) != cudaSuccess
      || cudaMalloc(&d_hist, sizeof(unsigned long long) * 32) != cudaSuccess
      || cudaMalloc(&d_max, sizeof(int)) != cudaSuccess) {
    set_error("cudaMalloc failed in plan_create");
    return fail(B200_ENOMEM);
  }
  cudaMemsetAsync(p->d_counter, 0, sizeof(unsigned int), st);
  cudaMemsetAsync(d_hist, 0, sizeof(unsigned long long) * 32, st);
  cudaMemsetAsync(d_max, 0, sizeof(int), st);
  if (nnz > 0 && rows > 0) {
    const int blocks = std::min((rows + 255) / 256, num_sms() * 8);
    plan_hist_kernel<<<blocks, 256, 0, st>>>(d_rowptr, rows, d_hist, d_max);
  }
  unsigned long long h[32];
  int mx = 0;
  if (cudaMemcpyAsync(h, d_hist, sizeof(h), cudaMemcpyDeviceToHost, st) != cudaSuccess
      || cudaMemcpyAsync(&mx, d_max, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess
      || cudaStreamSynchronize(st) != cudaSuccess) {
    cuda_fail(cudaGetLastError(), "plan_create readback", __FILE__, __LINE__);
    return fail(B200_ECUDA);
  }
  for (int i = 0; i < 32; ++i)
    p->hist[i] = (int64_t)h[i];
  p->max_row_len = mx;
  cudaFree(d_hist);
  cudaFree(d_max);
  d_hist = nullptr;
  d_max = nullptr;
  // sparse-row block (more than half the rows empty): list the non-empty rows
  if (!p->symmetric && nnz > 0 && p->hist[0] * 2 > (int64_t)rows && env_int("B200_SPMV_ROWLIST", 1)) {
    const int n_list = (int)((int64_t)rows - p->hist[0]);
    int* d_num = nullptr;
    void* d_tmp = nullptr;
    size_t tmp_bytes = 0;
    cub::CountingInputIterator<int> ids(0);
    RowNonEmpty pred{d_rowptr};
    bool ok = cudaMalloc(&p->d_rowlist, sizeof(int32_t) * (size_t)std::max(n_list, 1)) == cudaSuccess
              && cudaMalloc(&d_num, sizeof(int)) == cudaSuccess
              && cub::DeviceSelect::If(nullptr, tmp_bytes, ids, p->d_rowlist, d_num, rows, pred, st) == cudaSuccess
              && cudaMalloc(&d_tmp, std::max<size_t>(tmp_bytes, 1)) == cudaSuccess
              && cub::DeviceSelect::If(d_tmp, tmp_bytes, ids, p->d_rowlist, d_num, rows, pred, st) == cudaSuccess;
    int got = -1;
    ok = ok && cudaMemcpyAsync(&got, d_num, sizeof(int), cudaMemcpyDeviceToHost, st) == cudaSuccess
         && cudaStreamSynchronize(st) == cudaSuccess && got == n_list;
    cudaFree(d_num);
    cudaFree(d_tmp);
    if (!ok) {
      set_error("plan_create: listing the non-empty rows failed (%s)", cudaGetErrorString(cudaGetLastError()));
      return fail(B200_ECUDA);
    }
    p->n_rowlist = n_list;
  }
  // the histogram picks the walk and its launch shape (environment overrides for
  // sweeps), the walk fixes the stage weights, the weights cut the tiles
  const int variant = env_int("B200_SPMV_VARIANT", 0);
  p->walk_forced = walk_of_variant(variant);
  plan_choose_walk(p);
  int d_tile, d_stages, d_cw, d_ctas;
  plan_default_shape(p, &d_tile, &d_stages, &d_cw, &d_ctas);
  rc = plan_apply_tuning(p, variant, env_int("B200_SPMV_TILE_BYTES", d_tile),
                         env_int("B200_SPMV_STAGES", d_stages), env_int("B200_SPMV_CWARPS", d_cw),
                         env_int("B200_SPMV_CTAS", d_ctas));
  if (rc)
    return fail(rc);
  rc = plan_partition(p, d_rowptr, st);
  if (rc || cudaStreamSynchronize(st) != cudaSuccess)
    return fail(rc ? rc : B200_ECUDA);
  *plan = p;
  return B200_OK;
}

int b200_csr_plan_destroy(void* plan)
{
  Plan* p = static_cast<Plan*>(plan);
  if (!p)
    return B200_OK;
  cudaFree(p->d_tile_meta);
  cudaFree(p->d_warp_rows);
  cudaFree(p->d_rowlist);
  cudaFree(p->d_partials);
  cudaFree(p->d_counter);
  delete p;
  return B200_OK;
}

int b200_csr_plan_info(void* plan, int32_t* num_tiles, int32_t* max_row_len, int64_t hist[32])
{
  B200_REQUIRE(plan != nullptr, "plan is NULL");
  const Plan* p = static_cast<const Plan*>(plan);
  if (num_tiles)
    *num_tiles = p->num_tiles;
  if (max_row_len)
    *max_row_len = p->max_row_len;
  if (hist)
    for (int i = 0; i < 32; ++i)
      hist[i] = p->hist[i];
  return B200_OK;
}

int b200_csr_plan_tune(void* plan, const int32_t* d_rowptr, int variant, int tile_bytes,
                       int stages, int consumer_warps, int ctas_per_sm, void* stream)
{
  B200_REQUIRE(plan != nullptr, "plan is NULL");
  Plan* p = static_cast<Plan*>(plan);
  int rc = plan_apply_tuning(p, variant, tile_bytes, stages, consumer_warps, ctas_per_sm);
  if (rc)
    return rc;
  plan_choose_walk(p);
  rc = plan_partition(p, d_rowptr, as_stream(stream));
  if (rc)
    return rc;
  B200_CUDA(cudaStreamSynchronize(as_stream(stream)));
  return B200_OK;
}

int b200_csr_plan_chunking(void* plan, const int32_t* d_colind, int32_t nchunks, int32_t* tile_bounds,
                           int32_t* row_bounds, int32_t* col_min, int32_t* col_max, void* stream)
{
  B200_REQUIRE(plan != nullptr && nchunks >= 1 && tile_bounds && row_bounds && col_min && col_max, "bad argument");
  const Plan* p = static_cast<const Plan*>(plan);
  cudaStream_t st = as_stream(stream);
  std::vector<int32_t> meta(2 * ((size_t)p->num_tiles + 1), 0), ent(nchunks + 1, 0);
  if (p->num_tiles > 0)
    B200_CUDA(cudaMemcpyAsync(meta.data(), p->d_tile_meta, sizeof(int32_t) * meta.size(), cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  for (int k = 0; k <= nchunks; ++k) {
    const long long t = (long long)p->num_tiles * k / nchunks;
    tile_bounds[k] = (int32_t)t;
    row_bounds[k] = p->num_tiles > 0 ? meta[2 * t] : (k == nchunks ? p->rows : 0);
    ent[k] = p->num_tiles > 0 ? meta[2 * t + 1] : 0;
  }
  row_bounds[nchunks] = p->rows;
  for (int k = 0; k < nchunks; ++k) {
    col_min[k] = 0x7fffffff;
    col_max[k] = -1;
  }
  if (p->num_tiles == 0 || d_colind == nullptr)
    return B200_OK;
  int32_t *d_ent = nullptr, *d_mm = nullptr;
  B200_CUDA(cudaMalloc(&d_ent, sizeof(int32_t) * (nchunks + 1)));
  if (cudaMalloc(&d_mm, sizeof(int32_t) * 2 * nchunks) != cudaSuccess) {
    cudaFree(d_ent);
    set_error("cudaMalloc failed in plan_chunking");
    return B200_ENOMEM;
  }
  std::vector<int32_t> mm(2 * (size_t)nchunks);
  std::copy(col_min, col_min + nchunks, mm.begin());
  std::copy(col_max, col_max + nchunks, mm.begin() + nchunks);
  cudaMemcpyAsync(d_ent, ent.data(), sizeof(int32_t) * (nchunks + 1), cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(d_mm, mm.data(), sizeof(int32_t) * 2 * nchunks, cudaMemcpyHostToDevice, st);
  chunk_extent_kernel<<<dim3(num_sms() * 2, nchunks), 256, 0, st>>>(d_colind, d_ent, d_mm, d_mm + nchunks);
  cudaMemcpyAsync(mm.data(), d_mm, sizeof(int32_t) * 2 * nchunks, cudaMemcpyDeviceToHost, st);
  const cudaError_t e = cudaStreamSynchronize(st);
  cudaFree(d_ent);
  cudaFree(d_mm);
  if (e != cudaSuccess)
    return cuda_fail(e, "plan_chunking", __FILE__, __LINE__);
  std::copy(mm.begin(), mm.begin() + nchunks, col_min);
  std::copy(mm.begin() + nchunks, mm.end(), col_max);
  return B200_OK;
}

int b200_csr_spmv_tiles_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind, const double* d_val,
                            double alpha, const double* d_x, double beta, double* d_y, int32_t tile_begin,
                            int32_t tile_end, void* stream)
{
  B200_REQUIRE(tile_end >= 0, "bad tile range");
  return run_spmv<double, false>(plan, d_rowptr, d_colind, d_val, nullptr, alpha, d_x, beta, d_y, nullptr, stream,
                                 tile_begin, tile_end);
}

int b200_csr_symv_tiles_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind, const double* d_val,
                            const double* d_diag, double alpha, const double* d_x, double* d_y,
                            int32_t tile_begin, int32_t tile_end, void* stream)
{
  B200_REQUIRE(tile_end >= 0, "bad tile range");
  return run_spmv<double, true>(plan, d_rowptr, d_colind, d_val, d_diag, alpha, d_x, 1.0, d_y, nullptr, stream,
                                tile_begin, tile_end);
}

int b200_csr_spmv_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                      const double* d_val, double alpha, const double* d_x, double beta,
                      double* d_y, void* stream)
{
  return run_spmv<double, false>(plan, d_rowptr, d_colind, d_val, nullptr, alpha, d_x, beta, d_y,
                                 nullptr, stream);
}

int b200_csr_spmv_f32(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                      const float* d_val, float alpha, const float* d_x, float beta, float* d_y,
                      void* stream)
{
  return run_spmv<float, false>(plan, d_rowptr, d_colind, d_val, nullptr, alpha, d_x, beta, d_y,
                                nullptr, stream);
}

int b200_csr_spmv_dot_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                          const double* d_val, const double* d_x, double* d_y, double* d_dot,
                          void* stream)
{
  B200_REQUIRE(d_dot != nullptr, "d_dot is NULL");
  return run_spmv<double, false>(plan, d_rowptr, d_colind, d_val, nullptr, 1.0, d_x, 0.0, d_y,
                                 d_dot, stream);
}

int b200_csr_plan_rowlist_dot_ok(void* plan, int32_t* ok)
{
  B200_REQUIRE(plan != nullptr && ok != nullptr, "NULL argument");
  const Plan& p = *static_cast<const Plan*>(plan);
  *ok = (!p.symmetric && p.d_rowlist != nullptr && p.n_rowlist > 0 && p.max_row_len <= SHORT_ROW && p.d_partials != nullptr) ? 1 : 0;
  return B200_OK;
}

int b200_csr_spmv_rowlist_dot_f64(void* remote_plan, void* local_plan, const int32_t* d_rowptr, const int32_t* d_colind,
                                  const double* d_val, const double* d_x, double* d_y, double* d_dot, void* stream)
{
  return b200_csr_spmv_rowlist_dot_halo_f64(remote_plan, local_plan, d_rowptr, d_colind, d_val, d_x, d_y, d_dot, nullptr,
                                            nullptr, 0, stream);
}

int b200_csr_spmv_rowlist_dot_halo_f64(void* remote_plan, void* local_plan, const int32_t* d_rowptr,
                                       const int32_t* d_colind, const double* d_val, const double* d_x, double* d_y,
                                       double* d_dot, void* peer_ctx, void* halo_desc, int32_t ghost_base, void* stream)
{
  B200_REQUIRE(remote_plan && local_plan && d_dot, "NULL argument");
  const Plan& p = *static_cast<const Plan*>(remote_plan);
  const Plan& l = *static_cast<const Plan*>(local_plan);
  B200_REQUIRE(!p.symmetric && p.d_rowlist != nullptr && p.n_rowlist > 0 && p.max_row_len <= SHORT_ROW,
               "rowlist dot: the block needs a row list of short rows (b200_csr_plan_rowlist_dot_ok)");
  B200_REQUIRE(d_rowptr && d_colind && d_val && d_x && d_y, "NULL array");
  B200_REQUIRE(l.num_tiles > 0 && l.variant == 0 && l.d_partials != nullptr, "rowlist dot: the local plan ran no dot epilogue");
  B200_REQUIRE(halo_desc == nullptr || ghost_base >= 0, "rowlist dot: bad ghost base");
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  // the grid the local block's pipeline kernel was launched with (launch_pipe): that many partials are live
  const int nbase = (int)std::min<long long>(l.num_tiles, (long long)num_sms() * l.ctas_per_sm);
  B200_REQUIRE(nbase <= l.partials_len, "partials too small");
  const int n = p.n_rowlist;
  const int grid = std::max(1, std::min((n + 127) / 128, p.partials_len));
  const int gb = halo_desc ? ghost_base : 0x7fffffff; // no halo: every column reads x
  cudaStream_t st = as_stream(stream);
  PeerDesc* peer = static_cast<PeerDesc*>(peer_ctx);
  const HaloDesc* halo = static_cast<const HaloDesc*>(halo_desc);
  if (l.symmetric)
    spmv_rowlist_dot_kernel<true><<<grid, 128, 0, st>>>(n, p.d_rowlist, d_rowptr, d_colind, d_val, d_x, d_y, l.d_partials,
                                                        l.partials_len, nbase, p.d_partials, p.partials_len, p.d_counter, d_dot,
                                                        peer, halo, gb);
  else
    spmv_rowlist_dot_kernel<false><<<grid, 128, 0, st>>>(n, p.d_rowlist, d_rowptr, d_colind, d_val, d_x, d_y, l.d_partials,
                                                         l.partials_len, nbase, p.d_partials, p.partials_len, p.d_counter, d_dot,
                                                         peer, halo, gb);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_csr_symv_dot_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                          const double* d_val, const double* d_diag, const double* d_x, double* d_y,
                          double* d_dot, void* stream)
{
  B200_REQUIRE(d_dot != nullptr, "d_dot is NULL");
  return run_spmv<double, true>(plan, d_rowptr, d_colind, d_val, d_diag, 1.0, d_x, 0.0, d_y, d_dot, stream);
}

int b200_csr_symv_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                      const double* d_val, const double* d_diag, double alpha, const double* d_x,
                      double beta, double* d_y, void* stream)
{
  return run_spmv<double, true>(plan, d_rowptr, d_colind, d_val, d_diag, alpha, d_x, beta, d_y,
                                nullptr, stream);
}

int b200_csr_symv_f32(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                      const float* d_val, const float* d_diag, float alpha, const float* d_x,
                      float beta, float* d_y, void* stream)
{
  return run_spmv<float, true>(plan, d_rowptr, d_colind, d_val, d_diag, alpha, d_x, beta, d_y,
                               nullptr, stream);
}

} // extern "C"
