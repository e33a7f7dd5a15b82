// common.cuh -- shared helpers for the sm_100a kernels behind include/b200spmv.h
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "b200spmv.h"

namespace b200
{

// Thread-local error text behind b200_last_error().
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t err, const char* what, const char* file, int line);

#define B200_CUDA(call)                                                        \
  do {                                                                         \
    cudaError_t err__ = (call);                                                \
    if (err__ != cudaSuccess)                                                  \
      return ::b200::cuda_fail(err__, #call, __FILE__, __LINE__);              \
  } while (0)

#define B200_LAUNCH_CHECK()                                                    \
  do {                                                                         \
    cudaError_t err__ = cudaGetLastError();                                    \
    if (err__ != cudaSuccess)                                                  \
      return ::b200::cuda_fail(err__, "kernel launch", __FILE__, __LINE__);    \
  } while (0)

#define B200_REQUIRE(cond, msg)                                                \
  do {                                                                         \
    if (!(cond)) {                                                             \
      ::b200::set_error("%s (%s:%d)", msg, __FILE__, __LINE__);                \
      return B200_EINVAL;                                                      \
    }                                                                          \
  } while (0)

inline cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

int num_sms(); // of the current device, cached

// ---------------------------------------------------------------------------
// SpMV plan (the `_aux_data` of csr_kernels.h:82).
//
// Rows are cut into tiles of roughly `tile_bytes` of matrix stream
// (12 B per stored entry + 4 B per row).  tile_meta[2b] = first row of tile b,
// tile_meta[2b+1] = rowptr[first row]; entry num_tiles is the end sentinel
// {rows, nnz}.  A tile whose stream fits the stage buffer is staged in shared
// memory by bulk-async copies and walked one thread per row; longer rows in it
// are walked by a warp.  A tile that does not fit (one very long row) is read
// straight from global memory, one warp per row.
// ---------------------------------------------------------------------------
struct Plan {
  int32_t rows = 0;
  int32_t cols = 0;
  int64_t nnz = 0;
  int symmetric = 0;
  int variant = 0;          // 0 persistent bulk-async pipeline, 1 one tile per CTA
  int walk = 0;             // 0 one thread per row, 1 two-phase stream walk (short / irregular rows)
  int walk_forced = -1;     // -1: chosen from the row-length histogram
  int32_t w_entry = 12;     // shared-memory bytes per stored entry / per row in a stage:
  int32_t w_row = 4;        //   the weights the tile partition balances
  int32_t num_tiles = 0;
  int32_t tile_bytes = 0;   // target bytes of matrix stream per tile
  int32_t stage_bytes = 0;  // shared-memory stage size (tile_bytes + slack)
  int32_t stages = 0;       // pipeline depth
  int32_t consumer_warps = 0;
  int32_t ctas_per_sm = 0;
  int32_t* d_tile_meta = nullptr; // 2*(num_tiles+1) ints
  int32_t* d_warp_rows = nullptr; // WALK_SEG: num_tiles*(consumer_warps+1) row offsets
  int32_t max_row_len = 0;
  int64_t hist[32] = {0};
  int32_t* d_rowlist = nullptr;    // ids of the non-empty rows, ascending (sparse-row blocks only)
  int32_t n_rowlist = 0;
  double* d_partials = nullptr;    // dot epilogue: one partial per CTA
  unsigned int* d_counter = nullptr; // last-CTA-done ticket
  int32_t partials_len = 0;
};

// ---------------------------------------------------------------------------
// PTX wrappers: mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP).
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)
               : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
  asm volatile("{\n"
               ".reg .pred p;\n"
               "WAIT_%=:\n"
               "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
               "@p bra DONE_%=;\n"
               "bra WAIT_%=;\n"
               "DONE_%=:\n"
               "}" ::"r"(smem_u32(bar)),
               "r"(parity)
               : "memory");
}
// Non-blocking probe of the same phase (true: the phase with this parity has completed).
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity)
{
  uint32_t ok;
  asm volatile("{\n"
               ".reg .pred p;\n"
               "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
               "selp.u32 %0, 1, 0, p;\n"
               "}"
               : "=r"(ok)
               : "r"(smem_u32(bar)), "r"(parity)
               : "memory");
  return ok != 0;
}
// Read-only gather that does not take an L1 line (random x gathers with no reuse inside the SM).
__device__ __forceinline__ double ldg_na(const double* p)
{
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float ldg_na(const float* p)
{
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
// Global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], "
               "%2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// Same with an L2 cache policy (createpolicy): evict_first for streams that are read once.
__device__ __forceinline__ void bulk_g2s_hint(void* dst, const void* src, uint32_t bytes, uint64_t* bar,
                                              uint64_t policy)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], "
               "%2, [%3], %4;" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
               : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first()
{
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads)
{
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Arithmetic that is never contracted into FMA: the one-thread-per-row walk
// then reproduces csr_kernels.cpp:41-51 (built for baseline x86-64, no FMA) bit
// for bit.
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }

template <typename T>
__device__ __forceinline__ T warp_sum(T v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------------------
// Twice-working-precision dot products (Ogita, Rump & Oishi's Dot2): every product is
// split exactly with an FMA, every addition with TwoSum, the rounding errors are carried
// in `lo`.  The result hi+lo is what a sum in ~106-bit arithmetic would give, so two
// evaluations in DIFFERENT orders (GPU tree here, left-to-right loop in the CPU oracle,
// any grid size) round to the same double except with probability ~1e-14 per dot.
// That is what keeps a CG residual history reproducible across implementations: a
// 1e-16 difference in one dot product grows by ~2x per iteration otherwise.  The
// kernels are bandwidth-bound; the ~10 extra flops per element are free.
// ---------------------------------------------------------------------------
struct dd {
  double hi, lo;
};
__device__ __forceinline__ void dd_two_sum(double a, double b, double& s, double& e)
{
  s = __dadd_rn(a, b);
  const double bb = __dadd_rn(s, -a);
  e = __dadd_rn(__dadd_rn(a, -__dadd_rn(s, -bb)), __dadd_rn(b, -bb));
}
__device__ __forceinline__ void dd_add_prod(dd& acc, double a, double b)
{
  const double p = __dmul_rn(a, b);
  const double pe = __fma_rn(a, b, -p);
  double s, e;
  dd_two_sum(acc.hi, p, s, e);
  acc.hi = s;
  acc.lo = __dadd_rn(acc.lo, __dadd_rn(e, pe));
}
__device__ __forceinline__ dd dd_merge(dd a, dd b)
{
  double s, e;
  dd_two_sum(a.hi, b.hi, s, e);
  dd r;
  r.hi = s;
  r.lo = __dadd_rn(__dadd_rn(a.lo, b.lo), e);
  return r;
}
__device__ __forceinline__ dd dd_warp_sum(dd v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    dd w;
    w.hi = __shfl_xor_sync(0xffffffffu, v.hi, o);
    w.lo = __shfl_xor_sync(0xffffffffu, v.lo, o);
    v = dd_merge(v, w);
  }
  return v;
}
// Grid-wide dd sum of one value per thread of the first `nthreads` threads of every CTA:
// warp tree -> per-CTA partial (hi at partials[cta], lo at partials[plen + cta]) -> the
// last CTA folds the partials.  s_warp: 2 * (nthreads/32) doubles of shared memory.
// Returns true in the threads of the folding warp of the last CTA.
__device__ __forceinline__ bool grid_dd_reduce(dd v, int tid, int nthreads, int bar_id, double* s_warp,
                                               double* partials, int plen, unsigned int* counter,
                                               double* out)
{
  const int lane = tid & 31, warp = tid >> 5, nwarps = nthreads >> 5;
  v = dd_warp_sum(v);
  if (lane == 0) {
    s_warp[2 * warp] = v.hi;
    s_warp[2 * warp + 1] = v.lo;
  }
  named_bar_sync(bar_id, nthreads);
  bool last = false;
  if (warp == 0) {
    dd t;
    t.hi = lane < nwarps ? s_warp[2 * lane] : 0.0;
    t.lo = lane < nwarps ? s_warp[2 * lane + 1] : 0.0;
    t = dd_warp_sum(t);
    unsigned int ticket = 0;
    if (lane == 0) {
      partials[blockIdx.x] = t.hi;
      partials[plen + blockIdx.x] = t.lo;
      __threadfence();
      ticket = atomicAdd(counter, 1u);
    }
    ticket = __shfl_sync(0xffffffffu, ticket, 0);
    if (ticket == gridDim.x - 1) {
      __threadfence();
      dd s;
      s.hi = 0.0;
      s.lo = 0.0;
      for (int i = lane; i < (int)gridDim.x; i += 32) {
        dd w;
        w.hi = __ldcg(partials + i);
        w.lo = __ldcg(partials + plen + i);
        s = dd_merge(s, w);
      }
      s = dd_warp_sum(s);
      if (lane == 0) {
        out[0] = __dadd_rn(s.hi, s.lo);
        *counter = 0u;
      }
      last = true;
    }
  }
  return last;
}

// Fire-and-forget global add (SASS RED.E.ADD.F64 / .F32).
// (B200_EXPERIMENT_NO_RED: timing experiments only -- wrong results -- to read the
// in-situ cost of the scatter off a kernel that otherwise does the same work.)
#ifdef B200_EXPERIMENT_NO_RED
__device__ __forceinline__ void red_add(double* addr, double v) { if (v == 1.2345e300) *addr = v; }
__device__ __forceinline__ void red_add(float* addr, float v) { if (v == 1.2345e30f) *addr = v; }
#else
// Explicit PTX: `atomicAdd` with an unused result is only turned into RED when the compiler
// feels free to; with a fence elsewhere in the kernel it emitted ATOMG (a round trip per
// entry: 1.6x slower symmetric SpMV).
__device__ __forceinline__ void red_add(double* addr, double v)
{
  asm volatile("red.relaxed.gpu.global.add.f64 [%0], %1;" ::"l"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ void red_add(float* addr, float v)
{
  asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
#endif

// Descriptor of the kernel-driven peer halo (halo.cu; b200_halo_desc_create), in device memory.
struct HaloDesc {
  int nnbr;
  unsigned int epoch;  // pushes completed
  unsigned int ticket; // last-CTA-done counter of the push
  int pad;
  long long nghosts;
  int send_off[B200_HALO_MAX_NEIGHBOURS + 1];       // my packed slices, neighbour by neighbour (elements)
  long long land[B200_HALO_MAX_NEIGHBOURS];         // element offset of my slice in the neighbour's slot
  char* peer_slot[2][B200_HALO_MAX_NEIGHBOURS];     // the neighbour's staging slots, mapped here
  unsigned int* peer_flag[B200_HALO_MAX_NEIGHBOURS]; // my flag word in the neighbour's buffer
  unsigned int* my_flag[B200_HALO_MAX_NEIGHBOURS];   // the neighbour's flag word in my buffer
  char* my_slot[2];
};

// ---------------------------------------------------------------------------
// Scalar allreduce over peer memory (the MPI_Allreduce of cg.cpp:49,65,75 on device
// scalars).  Every rank owns a mailbox in IPC-exported device memory: one slot per
// (parity, source rank).  One warp per rank: lane q stores this rank's values and then
// the sequence number (release, system scope) into rank q's mailbox over NVLink; lane q
// then spins (acquire) on the slot rank q writes here.  The sum runs over the ranks in
// rank order on every rank, so all ranks get the same bits.  Two parities: a rank can
// only be one allreduce ahead of the slowest one (it needs that rank's current value
// to finish), so the slot it overwrites has been consumed.
// ---------------------------------------------------------------------------
struct PeerSlot {
  double v[B200_PEER_REDUCE_MAX_COUNT];
  unsigned long long seq;
  unsigned long long pad[3];
};
static_assert(sizeof(PeerSlot) == 64, "one slot per 64-byte line");

struct PeerDesc {
  int rank, world;
  unsigned long long seq;
  PeerSlot* box[B200_PEER_REDUCE_MAX_RANKS]; // box[q] = rank q's mailbox (box[rank]: mine, local)
};

// One full warp.  mine[k], k < count: this rank's values; returns the sums over the ranks (all lanes).
__device__ __forceinline__ void peer_exchange_warp(PeerDesc* d, int lane, int count,
                                                   const double (&mine)[B200_PEER_REDUCE_MAX_COUNT],
                                                   double (&total)[B200_PEER_REDUCE_MAX_COUNT])
{
  const int rank = d->rank, world = d->world;
  const unsigned long long seq = d->seq + 1;
  const int par = (int)(seq & 1ull);
  __syncwarp();
  double got[B200_PEER_REDUCE_MAX_COUNT];
#pragma unroll
  for (int k = 0; k < B200_PEER_REDUCE_MAX_COUNT; ++k)
    got[k] = 0.0;
  if (lane < world) {
    PeerSlot* dst = d->box[lane] + par * world + rank;
#pragma unroll
    for (int k = 0; k < B200_PEER_REDUCE_MAX_COUNT; ++k)
      if (k < count)
        asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(&dst->v[k]), "d"(mine[k]) : "memory");
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(&dst->seq), "l"(seq) : "memory");
    const PeerSlot* src = d->box[rank] + par * world + lane;
    unsigned long long seen;
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(&src->seq) : "memory");
    } while (seen != seq);
#pragma unroll
    for (int k = 0; k < B200_PEER_REDUCE_MAX_COUNT; ++k)
      if (k < count)
        asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(got[k]) : "l"(&src->v[k]) : "memory");
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < B200_PEER_REDUCE_MAX_COUNT; ++k) {
    double t = 0.0;
    for (int q = 0; q < world; ++q)
      t = add_rn(t, __shfl_sync(0xffffffffu, got[k], q));
    total[k] = t;
  }
  if (lane == 0)
    d->seq = seq;
}


} // namespace b200
