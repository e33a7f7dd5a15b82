// cg.cu -- fused vector kernels of the Conjugate Gradient iteration.
//
// Replaces the BLAS-1 calls of cg.cpp:47-85 (cblas_ddot / daxpy / dscal; on the
// reference's CUDA backend cublasDdot x3, cublasDaxpy x3, cublasDscal x1 and the
// <<<1,1>>> scalar kernels compute_alpha / compute_beta / compute_sqrt,
// cuda/cg.cuda.cu:14-38,88-150) with three single-pass kernels:
//   dot        : s = a.b                                   (2 reads)
//   update_xr  : x += alpha p; r -= alpha Ap; rr = r.r      (4 reads, 2 writes)
//   update_p   : p = r + beta p  (+ scalar bookkeeping)     (2 reads, 1 write)
// All scalars stay in a device state block, so the loop never syncs with the
// host; once the `done` flag is set every kernel is a no-op, which lets a
// fixed-length CUDA graph reproduce the reference's early exit (cg.cpp:80-81).
// Reductions are deterministic: fixed grid, per-thread strided partial, warp
// xor-tree, per-CTA partial, last CTA folds the partials in index order.
// Element-wise arithmetic is un-contracted (mul then add) like the oracle's
// plain-loop restatement of daxpy/dscal.
#include "common.cuh"

#include <cstdlib>
#include <cstring>

#include <algorithm>

namespace b200
{

// state block layout (doubles), see b200spmv.h
enum { ST_RNORM_OLD = 0, ST_PAP = 1, ST_RR = 2, ST_RNORM0 = 3, ST_DONE = 4, ST_ITER = 5,
       ST_RTOL = 6, ST_KMAX = 7, ST_GEN = 8 /* grid-barrier generation of the fused step (u32) */,
       ST_COUNTER = 16, ST_PARTIALS = 32 };

constexpr int VEC_THREADS = 256;
// per-CTA partials: hi parts in [0, DOT_PARTIALS), lo parts in [DOT_PARTIALS, 2*DOT_PARTIALS)
constexpr int DOT_PARTIALS = 2000;
static_assert(ST_PARTIALS + 2 * DOT_PARTIALS + 1 <= B200_CG_STATE_DOUBLES, "state block too small");
static_assert(8 + 2 * DOT_PARTIALS <= B200_DOT_SCRATCH_DOUBLES, "dot scratch too small");

static int vec_grid(int n)
{
  const int want = (n + VEC_THREADS * 4 - 1) / (VEC_THREADS * 4);
  const int cap = std::min(num_sms() * 8, DOT_PARTIALS);
  return std::max(1, std::min(want, cap));
}

// (PeerSlot / PeerDesc / peer_exchange_warp: common.cuh -- the SpMV epilogues exchange through them too)
// Fold one dd value per thread over the whole grid into out[0] (deterministic, and in twice
// the working precision: see common.cuh).  partials: 2 * plen doubles.
// The last CTA to arrive folds the per-CTA partials with all its warps: one warp alone (grid_dd_reduce, which the
// SpMV epilogue keeps for its ~300 partials) walks the ~1200 of these grids in 37 dependent L2 round trips, a
// tail of several microseconds on kernels that take ten.
// peer != nullptr: the last CTA also exchanges the sum with the other ranks (peer_exchange_warp) and out[0] is the
// sum over all of them -- the MPI_Allreduce of cg.cpp:65,75 in the tail of the kernel that produced the value,
// instead of a launch of its own.
__device__ __forceinline__ void grid_sum(dd v, double* partials, int plen, unsigned int* counter, double* out,
                                         PeerDesc* peer = nullptr)
{
  constexpr int NW = VEC_THREADS / 32;
  __shared__ double s_warp[2 * NW];
  __shared__ bool s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  v = dd_warp_sum(v);
  if (lane == 0) {
    s_warp[2 * warp] = v.hi;
    s_warp[2 * warp + 1] = v.lo;
  }
  __syncthreads();
  if (warp == 0) {
    dd t;
    t.hi = lane < NW ? s_warp[2 * lane] : 0.0;
    t.lo = lane < NW ? s_warp[2 * lane + 1] : 0.0;
    t = dd_warp_sum(t);
    if (lane == 0) {
      partials[blockIdx.x] = t.hi;
      partials[plen + blockIdx.x] = t.lo;
      __threadfence();
      s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
  }
  __syncthreads();
  if (!s_last)
    return;
  __threadfence();
  dd s = {0.0, 0.0};
  for (int i = tid; i < (int)gridDim.x; i += VEC_THREADS) {
    dd w;
    w.hi = __ldcg(partials + i);
    w.lo = __ldcg(partials + plen + i);
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
      out[0] = total[0];
      *counter = 0u;
    }
  }
}

__global__ void __launch_bounds__(VEC_THREADS)
    dot_kernel(int n, const double* __restrict__ a, const double* __restrict__ b,
               double* __restrict__ out, double* __restrict__ scratch, const double* done_flag, PeerDesc* peer)
{
  if (done_flag != nullptr && *done_flag != 0.0)
    return;
  dd acc = {0.0, 0.0};
  const int stride = gridDim.x * blockDim.x;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int n2 = n >> 1;
  const double2* a2 = reinterpret_cast<const double2*>(a);
  const double2* b2 = reinterpret_cast<const double2*>(b);
  for (int i = tid; i < n2; i += stride) {
    const double2 u = __ldg(a2 + i), w = __ldg(b2 + i);
    dd_add_prod(acc, u.x, w.x);
    dd_add_prod(acc, u.y, w.y);
  }
  if ((n & 1) && tid == 0)
    dd_add_prod(acc, a[n - 1], b[n - 1]);
  grid_sum(acc, scratch + 8, DOT_PARTIALS, reinterpret_cast<unsigned int*>(scratch), out, peer);
}

// Unaligned fallback (pointer not 16-byte aligned): scalar loads.
__global__ void __launch_bounds__(VEC_THREADS)
    dot_kernel_scalar(int n, const double* __restrict__ a, const double* __restrict__ b,
                      double* __restrict__ out, double* __restrict__ scratch,
                      const double* done_flag, PeerDesc* peer)
{
  if (done_flag != nullptr && *done_flag != 0.0)
    return;
  dd acc = {0.0, 0.0};
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dd_add_prod(acc, __ldg(a + i), __ldg(b + i));
  grid_sum(acc, scratch + 8, DOT_PARTIALS, reinterpret_cast<unsigned int*>(scratch), out, peer);
}

__global__ void init_state_kernel(double* st, const double* rr, double rtol, int kmax)
{
  const double rnorm0 = sqrt(rr[0]); // cg.cpp:50
  st[ST_RNORM_OLD] = rnorm0;         // cg.cpp:53
  st[ST_RNORM0] = rnorm0;
  st[ST_PAP] = 0.0;
  st[ST_RR] = rr[0];
  st[ST_DONE] = (kmax <= 0) ? 1.0 : 0.0;
  st[ST_ITER] = 0.0;
  st[ST_RTOL] = rtol;
  st[ST_KMAX] = (double)kmax;
  *reinterpret_cast<unsigned int*>(st + ST_COUNTER) = 0u;
}

__global__ void __launch_bounds__(VEC_THREADS)
    update_xr_kernel(int n, double* __restrict__ st, const double* __restrict__ p,
                     const double* __restrict__ Ap, double* __restrict__ x, double* __restrict__ r, PeerDesc* peer)
{
  if (st[ST_DONE] != 0.0)
    return;
  const double rn = st[ST_RNORM_OLD];
  const double alpha = (rn * rn) / st[ST_PAP]; // cg.cpp:66
  const double nalpha = -alpha;
  dd acc = {0.0, 0.0};
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double pi = p[i], api = Ap[i];
    x[i] = __dadd_rn(x[i], __dmul_rn(alpha, pi));          // cg.cpp:69
    const double ri = __dadd_rn(r[i], __dmul_rn(nalpha, api)); // cg.cpp:70
    r[i] = ri;
    dd_add_prod(acc, ri, ri);                              // cg.cpp:73
  }
  grid_sum(acc, st + ST_PARTIALS, DOT_PARTIALS, reinterpret_cast<unsigned int*>(st + ST_COUNTER), st + ST_RR, peer);
}

__global__ void __launch_bounds__(VEC_THREADS)
    update_p_kernel(int n, double* __restrict__ st, const double* __restrict__ r,
                    double* __restrict__ p, double* __restrict__ hist, int hist_len)
{
  if (st[ST_DONE] != 0.0)
    return;
  const double rnorm_old = st[ST_RNORM_OLD];
  const double rnorm_new = sqrt(st[ST_RR]);                                  // cg.cpp:76
  const double beta = (rnorm_new * rnorm_new) / (rnorm_old * rnorm_old);     // cg.cpp:77
  const bool converged = (rnorm_new / st[ST_RNORM0]) < st[ST_RTOL];          // cg.cpp:80
  if (!converged) {
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
      p[i] = __dadd_rn(__dmul_rn(beta, p[i]), r[i]);                         // cg.cpp:84-85
  }
  // the last CTA to get here advances the scalars (nobody reads them after this)
  __shared__ bool s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned int* counter = reinterpret_cast<unsigned int*>(st + ST_COUNTER);
    const unsigned int ticket = atomicAdd(counter, 1u);
    s_last = (ticket == gridDim.x - 1);
    if (s_last) {
      *counter = 0u;
      const int iter = (int)st[ST_ITER] + 1;
      st[ST_ITER] = (double)iter;
      if (hist != nullptr && iter - 1 < hist_len)
        hist[iter - 1] = rnorm_new;
      st[ST_RNORM_OLD] = rnorm_new;                                          // cg.cpp:78
      if (converged || iter >= (int)st[ST_KMAX])
        st[ST_DONE] = 1.0;
    }
  }
}

__global__ void __launch_bounds__(32) peer_allreduce_kernel(PeerDesc* d, double* val, int count)
{
  const int lane = threadIdx.x;
  double mine[B200_PEER_REDUCE_MAX_COUNT], total[B200_PEER_REDUCE_MAX_COUNT];
#pragma unroll
  for (int k = 0; k < B200_PEER_REDUCE_MAX_COUNT; ++k)
    mine[k] = k < count ? val[k] : 0.0;
  peer_exchange_warp(d, lane, count, mine, total);
#pragma unroll
  for (int k = 0; k < B200_PEER_REDUCE_MAX_COUNT; ++k)
    if (lane == 0 && k < count)
      val[k] = total[k];
}

// ---------------------------------------------------------------------------
// One launch for the vector half of a CG iteration (cg.cpp:63-85):
//   [p.Ap]  ->  allreduce  ->  x += alpha p, r -= alpha Ap, r.r  ->  allreduce  ->  p = r + beta p
// Cooperative launch (every CTA resident), grid-wide hand-over through a generation word in the
// state block: each CTA leaves its partial, the LAST one to arrive folds them in index order,
// exchanges the sum with the other ranks over the peer mailboxes (one warp, as above), stores the
// global value and releases the generation; everybody else spins on it.  Replaces five launches
// (dot, allreduce, update_xr, allreduce, update_p) whose gaps are what strong scaling at 8 GPUs
// paid for (profiles/r01_bench_n8_strong_final.json: 0.29 ms per iteration for 0.19 ms of kernels).
// Same arithmetic as the separate kernels: un-contracted element updates, Dot2 sums.
// ---------------------------------------------------------------------------
__device__ __forceinline__ double fused_grid_allreduce(dd v, double* st, PeerDesc* peer, unsigned int gen_target,
                                                       double* s_warp, double* s_bcast)
{
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = VEC_THREADS / 32;
  __shared__ bool s_last;
  double* partials = st + ST_PARTIALS;
  unsigned int* counter = reinterpret_cast<unsigned int*>(st + ST_COUNTER);
  unsigned int* gen = reinterpret_cast<unsigned int*>(st + ST_GEN);
  v = dd_warp_sum(v);
  if (lane == 0) {
    s_warp[2 * warp] = v.hi;
    s_warp[2 * warp + 1] = v.lo;
  }
  __syncthreads();
  if (warp == 0) {
    dd t;
    t.hi = lane < NW ? s_warp[2 * lane] : 0.0;
    t.lo = lane < NW ? s_warp[2 * lane + 1] : 0.0;
    t = dd_warp_sum(t);
    if (lane == 0) {
      partials[blockIdx.x] = t.hi;
      partials[DOT_PARTIALS + blockIdx.x] = t.lo;
      __threadfence();
      s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
  }
  __syncthreads();
  if (s_last) {
    // The last CTA to arrive folds the partials with ALL its warps (one warp alone walked ~1200 partials in 37
    // dependent L2 round trips: 8-9 us of the hand-over, twice per iteration), exchanges the sum with the other
    // ranks, stores the global value and releases the generation.
    __threadfence();
    dd s = {0.0, 0.0};
    for (int i = tid; i < (int)gridDim.x; i += VEC_THREADS) {
      dd w;
      w.hi = __ldcg(partials + i);
      w.lo = __ldcg(partials + DOT_PARTIALS + i);
      s = dd_merge(s, w);
    }
    s = dd_warp_sum(s);
    __syncthreads(); // s_warp: everybody has read its warp's slot above
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
        st[ST_PARTIALS + 2 * DOT_PARTIALS] = total[0];
        *counter = 0u;
        __threadfence();
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(gen), "r"(gen_target) : "memory");
      }
    }
  }
  if (tid == 0) {
    unsigned int seen;
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(gen) : "memory");
      if ((int)(seen - gen_target) >= 0)
        break;
      __nanosleep(64); // ~1200 CTAs poll one word: leave the L2 slice some room for the last CTA's fold
    }
    *s_bcast = __ldcg(st + ST_PARTIALS + 2 * DOT_PARTIALS);
  }
  __syncthreads();
  const double out = *s_bcast;
  __syncthreads(); // s_warp / s_bcast are free again
  return out;
}

// The push half of the kernel-driven halo (halo.cu: halo_push_kernel) from inside the fused step: pack AND send
// x[idx[k]] into the neighbours' staging slots over NVLink; the last CTA to finish publishes the epoch.
// x was written by this grid: the loads go to L2 (ld.cg), never through a stale L1 line.
__device__ __forceinline__ void fused_halo_push(HaloDesc* d, int n, const int32_t* __restrict__ idx, const double* x)
{
  __shared__ bool s_last;
  const unsigned int epoch = d->epoch + 1; // advanced by the last CTA only after everybody's ticket: all read the same
  const int slot = (int)(epoch & 1u), nnbr = d->nnbr;
  const int stride = gridDim.x * blockDim.x;
  int i = 0;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
    while (k >= d->send_off[i + 1])
      ++i;
    double* dst = reinterpret_cast<double*>(d->peer_slot[slot][i]) + d->land[i] + (k - d->send_off[i]);
    *dst = __ldcg(x + __ldg(idx + k));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    s_last = atomicAdd(&d->ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last) {
    __threadfence_system();
    if ((int)threadIdx.x < nnbr)
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(d->peer_flag[threadIdx.x]), "r"(epoch) : "memory");
    if (threadIdx.x == 0) {
      d->ticket = 0u;
      d->epoch = epoch;
    }
  }
}

__global__ void __launch_bounds__(VEC_THREADS)
    cg_fused_step_kernel(int n, double* __restrict__ st, double* __restrict__ p, const double* __restrict__ Ap,
                         double* __restrict__ x, double* __restrict__ r, double* __restrict__ hist, int hist_len,
                         PeerDesc* peer, int need_dot, HaloDesc* halo, int n_idx, const int32_t* __restrict__ idx)
{
  __shared__ double s_warp[2 * (VEC_THREADS / 32)];
  __shared__ double s_bcast;
  if (st[ST_DONE] != 0.0) {
    // every CTA sees the same flag: it only changes at the very end of a launch.  The halo of the next
    // iteration still goes out (p unchanged): every iteration, live or not, is one push and one wait on every rank.
    if (halo != nullptr)
      fused_halo_push(halo, n_idx, idx, p);
    return;
  }
  // read before anybody can pass the first hand-over (which needs all CTAs to have arrived)
  const unsigned int gen0 = *reinterpret_cast<volatile unsigned int*>(st + ST_GEN);
  const double rnorm_old = st[ST_RNORM_OLD], rnorm0 = st[ST_RNORM0], rtol = st[ST_RTOL];
  const int stride = gridDim.x * blockDim.x;
  const int tid0 = blockIdx.x * blockDim.x + threadIdx.x;
  // B200_CG_TIMELINE (host/cg.cpp sets st[9]): CTA 0 leaves %globaltimer stamps of its phases in st[10..15]
  const bool timeline = st[9] != 0.0;
  auto stamp = [&](int k) {
    if (timeline && tid0 == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      st[10 + k] = (double)t;
    }
  };
  stamp(0);

  // 16-byte accesses when the vectors allow it (they come from cudaMalloc: always, unless a caller offsets them)
  const bool vec2 = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(Ap) | reinterpret_cast<uintptr_t>(x)
                      | reinterpret_cast<uintptr_t>(r)) & 15) == 0;
  const int n2 = vec2 ? n >> 1 : 0; // pairs handled as double2; [2*n2, n) one by one
  double2* p2 = reinterpret_cast<double2*>(p);
  const double2* Ap2 = reinterpret_cast<const double2*>(Ap);
  double2* x2 = reinterpret_cast<double2*>(x);
  double2* r2 = reinterpret_cast<double2*>(r);

  double pAp;
  if (need_dot) { // cg.cpp:63-65
    dd acc = {0.0, 0.0};
    for (int i = tid0; i < n2; i += stride) {
      const double2 u = p2[i], w = Ap2[i];
      dd_add_prod(acc, u.x, w.x);
      dd_add_prod(acc, u.y, w.y);
    }
    for (int i = 2 * n2 + tid0; i < n; i += stride)
      dd_add_prod(acc, p[i], Ap[i]);
    stamp(1);
    pAp = fused_grid_allreduce(acc, st, peer, gen0 + 1, s_warp, &s_bcast);
  } else if (peer != nullptr) { // the SpMV epilogue left this rank's p.Ap in st[ST_PAP]
    dd acc = {0.0, 0.0};
    if (tid0 == 0)
      acc.hi = st[ST_PAP];
    pAp = fused_grid_allreduce(acc, st, peer, gen0 + 1, s_warp, &s_bcast);
  } else {
    pAp = st[ST_PAP];
  }
  stamp(2);
  const double alpha = (rnorm_old * rnorm_old) / pAp; // cg.cpp:66
  const double nalpha = -alpha;
  dd acc = {0.0, 0.0};
  for (int i = tid0; i < n2; i += stride) {
    const double2 pi = p2[i], api = Ap2[i];
    double2 xi = x2[i], ri = r2[i];
    xi.x = __dadd_rn(xi.x, __dmul_rn(alpha, pi.x));  // cg.cpp:69
    xi.y = __dadd_rn(xi.y, __dmul_rn(alpha, pi.y));
    ri.x = __dadd_rn(ri.x, __dmul_rn(nalpha, api.x)); // cg.cpp:70
    ri.y = __dadd_rn(ri.y, __dmul_rn(nalpha, api.y));
    x2[i] = xi;
    r2[i] = ri;
    dd_add_prod(acc, ri.x, ri.x);                     // cg.cpp:73
    dd_add_prod(acc, ri.y, ri.y);
  }
  for (int i = 2 * n2 + tid0; i < n; i += stride) {
    const double pi = p[i], api = Ap[i];
    x[i] = __dadd_rn(x[i], __dmul_rn(alpha, pi));
    const double ri = __dadd_rn(r[i], __dmul_rn(nalpha, api));
    r[i] = ri;
    dd_add_prod(acc, ri, ri);
  }
  stamp(3);
  const double rr = fused_grid_allreduce(acc, st, peer, gen0 + 2, s_warp, &s_bcast);
  stamp(4);
  const double rnorm_new = sqrt(rr);                                      // cg.cpp:76
  const double beta = (rnorm_new * rnorm_new) / (rnorm_old * rnorm_old);  // cg.cpp:77
  const bool converged = (rnorm_new / rnorm0) < rtol;                     // cg.cpp:80
  if (!converged) {
    for (int i = tid0; i < n2; i += stride) {                             // cg.cpp:84-85
      double2 pi = p2[i];
      const double2 ri = r2[i];
      pi.x = __dadd_rn(__dmul_rn(beta, pi.x), ri.x);
      pi.y = __dadd_rn(__dmul_rn(beta, pi.y), ri.y);
      p2[i] = pi;
    }
    for (int i = 2 * n2 + tid0; i < n; i += stride)
      p[i] = __dadd_rn(__dmul_rn(beta, p[i]), r[i]);
  }
  stamp(5);
  if (tid0 == 0) { // nobody reads these again in this launch (everybody took its copies above)
    const int iter = (int)st[ST_ITER] + 1;
    st[ST_ITER] = (double)iter;
    if (hist != nullptr && iter - 1 < hist_len)
      hist[iter - 1] = rnorm_new;
    st[ST_PAP] = pAp;
    st[ST_RR] = rr;
    st[ST_RNORM_OLD] = rnorm_new;                                         // cg.cpp:78
    if (converged || iter >= (int)st[ST_KMAX])
      st[ST_DONE] = 1.0;
  }
  if (halo != nullptr) {
    // L2GMap::update(p) of the NEXT iteration (cg.cpp:59), here instead of in a launch of its own: one more
    // grid-wide hand-over (every CTA's share of the new p is out), then the boundary entries go to the neighbours.
    const dd zero = {0.0, 0.0};
    (void)fused_grid_allreduce(zero, st, nullptr, gen0 + 3, s_warp, &s_bcast);
    fused_halo_push(halo, n_idx, idx, p);
  }
}

static int need_gpu()
{
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  return B200_OK;
}

} // namespace b200

using namespace b200;


extern "C" {

int b200_cg_dot_f64(int32_t n, const double* d_a, const double* d_b, double* d_out,
                    double* d_scratch, const double* d_done_flag, void* stream)
{
  return b200_cg_dot_peer_f64(n, d_a, d_b, d_out, d_scratch, d_done_flag, nullptr, stream);
}

int b200_cg_dot_peer_f64(int32_t n, const double* d_a, const double* d_b, double* d_out, double* d_scratch,
                         const double* d_done_flag, void* peer_ctx, void* stream)
{
  B200_REQUIRE(n >= 0 && d_out && d_scratch, "bad argument");
  B200_REQUIRE(n == 0 || (d_a && d_b), "NULL vector");
  if (int rc = need_gpu())
    return rc;
  PeerDesc* peer = static_cast<PeerDesc*>(peer_ctx);
  const int grid = std::min(vec_grid(n), B200_DOT_SCRATCH_DOUBLES - 8);
  const bool al = ((reinterpret_cast<uintptr_t>(d_a) | reinterpret_cast<uintptr_t>(d_b)) & 15) == 0;
  if (al)
    dot_kernel<<<grid, VEC_THREADS, 0, as_stream(stream)>>>(n, d_a, d_b, d_out, d_scratch, d_done_flag, peer);
  else
    dot_kernel_scalar<<<grid, VEC_THREADS, 0, as_stream(stream)>>>(n, d_a, d_b, d_out, d_scratch, d_done_flag, peer);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_cg_init_state(double* d_state, const double* d_rr, double rtol, int32_t kmax, void* stream)
{
  B200_REQUIRE(d_state && d_rr, "NULL pointer");
  if (int rc = need_gpu())
    return rc;
  init_state_kernel<<<1, 1, 0, as_stream(stream)>>>(d_state, d_rr, rtol, kmax);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_cg_update_xr_f64(int32_t n, double* d_state, const double* d_p, const double* d_Ap,
                          double* d_x, double* d_r, void* stream)
{
  return b200_cg_update_xr_peer_f64(n, d_state, d_p, d_Ap, d_x, d_r, nullptr, stream);
}

int b200_cg_update_xr_peer_f64(int32_t n, double* d_state, const double* d_p, const double* d_Ap, double* d_x,
                               double* d_r, void* peer_ctx, void* stream)
{
  B200_REQUIRE(n >= 0 && d_state, "bad argument");
  B200_REQUIRE(n == 0 || (d_p && d_Ap && d_x && d_r), "NULL vector");
  if (int rc = need_gpu())
    return rc;
  update_xr_kernel<<<vec_grid(n), VEC_THREADS, 0, as_stream(stream)>>>(n, d_state, d_p, d_Ap, d_x, d_r,
                                                                       static_cast<PeerDesc*>(peer_ctx));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_cg_update_p_f64(int32_t n, double* d_state, const double* d_r, double* d_p, double* d_hist,
                         int32_t hist_len, void* stream)
{
  B200_REQUIRE(n >= 0 && d_state, "bad argument");
  B200_REQUIRE(n == 0 || (d_r && d_p), "NULL vector");
  if (int rc = need_gpu())
    return rc;
  update_p_kernel<<<vec_grid(n), VEC_THREADS, 0, as_stream(stream)>>>(n, d_state, d_r, d_p, d_hist,
                                                                      hist_len);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_cg_fused_step_f64(int32_t n, double* d_state, double* d_p, const double* d_Ap, double* d_x, double* d_r,
                           double* d_hist, int32_t hist_len, void* peer_ctx, int need_dot, void* stream)
{
  return b200_cg_fused_step_halo_f64(n, d_state, d_p, d_Ap, d_x, d_r, d_hist, hist_len, peer_ctx, need_dot, nullptr, 0,
                                     nullptr, stream);
}

int b200_cg_fused_step_halo_f64(int32_t n, double* d_state, double* d_p, const double* d_Ap, double* d_x, double* d_r,
                                double* d_hist, int32_t hist_len, void* peer_ctx, int need_dot, void* halo_desc,
                                int32_t n_idx, const int32_t* d_idx, void* stream)
{
  B200_REQUIRE(halo_desc == nullptr || (n_idx >= 0 && (n_idx == 0 || d_idx != nullptr)), "halo push: bad index list");
  B200_REQUIRE(n >= 0 && d_state, "bad argument");
  B200_REQUIRE(n == 0 || (d_p && d_Ap && d_x && d_r), "NULL vector");
  if (int rc = need_gpu())
    return rc;
  // every CTA must be resident (they hand over through a spin): occupancy bounds the grid
  static int per_sm[64] = {0};
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < 64, "device index out of range");
  if (per_sm[dev] == 0) {
    int b = 0;
    B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, cg_fused_step_kernel, VEC_THREADS, 0));
    B200_REQUIRE(b >= 1, "fused CG step does not fit an SM");
    per_sm[dev] = b;
  }
  // B200_CG_FUSED_CTAS caps the CTAs per SM (sweeps; profiles/r02_fused_probe.txt: the full co-resident grid is
  // best at 16.8 M rows and no worse at 2.1 M)
  static const int want_per_sm = [] {
    const char* e = getenv("B200_CG_FUSED_CTAS");
    return (e && *e) ? atoi(e) : 0;
  }();
  const int use_per_sm = want_per_sm > 0 ? std::min(per_sm[dev], want_per_sm) : per_sm[dev];
  const int grid = std::max(1, std::min(std::min(vec_grid(n), num_sms() * use_per_sm), DOT_PARTIALS));
  PeerDesc* peer = static_cast<PeerDesc*>(peer_ctx);
  HaloDesc* halo = static_cast<HaloDesc*>(halo_desc);
  void* args[] = {&n, &d_state, &d_p, &d_Ap, &d_x, &d_r, &d_hist, &hist_len, &peer, &need_dot, &halo, &n_idx, &d_idx};
  B200_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(cg_fused_step_kernel), dim3(grid), dim3(VEC_THREADS),
                                        args, 0, as_stream(stream)));
  return B200_OK;
}

int b200_peer_reduce_mailbox_bytes(int32_t world, size_t* num_bytes)
{
  B200_REQUIRE(num_bytes != nullptr && world >= 1 && world <= B200_PEER_REDUCE_MAX_RANKS, "bad argument");
  *num_bytes = sizeof(PeerSlot) * 2 * (size_t)world;
  return B200_OK;
}

int b200_peer_reduce_create(void** ctx, int32_t rank, int32_t world, void* const* d_mailboxes, void* stream)
{
  B200_REQUIRE(ctx != nullptr && d_mailboxes != nullptr, "bad argument");
  *ctx = nullptr;
  B200_REQUIRE(world >= 1 && world <= B200_PEER_REDUCE_MAX_RANKS && rank >= 0 && rank < world, "bad rank / world");
  if (int rc = need_gpu())
    return rc;
  PeerDesc h;
  std::memset(&h, 0, sizeof(h));
  h.rank = rank;
  h.world = world;
  for (int q = 0; q < world; ++q) {
    B200_REQUIRE(d_mailboxes[q] != nullptr, "mailbox is NULL");
    h.box[q] = static_cast<PeerSlot*>(d_mailboxes[q]);
  }
  PeerDesc* d = nullptr;
  B200_CUDA(cudaMalloc(&d, sizeof(PeerDesc)));
  cudaStream_t st = as_stream(stream);
  // my mailbox starts at sequence 0 everywhere; peers zero theirs the same way before anyone reduces
  // (the caller puts a barrier between create and the first allreduce)
  if (cudaMemsetAsync(h.box[rank], 0, sizeof(PeerSlot) * 2 * (size_t)world, st) != cudaSuccess
      || cudaMemcpyAsync(d, &h, sizeof(h), cudaMemcpyHostToDevice, st) != cudaSuccess
      || cudaStreamSynchronize(st) != cudaSuccess) {
    cudaFree(d);
    return cuda_fail(cudaGetLastError(), "peer_reduce_create", __FILE__, __LINE__);
  }
  *ctx = d;
  return B200_OK;
}

int b200_peer_reduce_destroy(void* ctx)
{
  if (ctx)
    B200_CUDA(cudaFree(ctx));
  return B200_OK;
}

int b200_peer_allreduce_sum_f64(void* ctx, double* d_inout, int32_t n, void* stream)
{
  B200_REQUIRE(n >= 0 && n <= B200_PEER_REDUCE_MAX_COUNT, "count out of range");
  if (n == 0 || ctx == nullptr) // single rank: the local value is the sum
    return B200_OK;
  B200_REQUIRE(d_inout != nullptr, "NULL value");
  if (int rc = need_gpu())
    return rc;
  peer_allreduce_kernel<<<1, 32, 0, as_stream(stream)>>>(static_cast<PeerDesc*>(ctx), d_inout, n);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------
// Allocation-free dot product (no counterpart in the reference; added
// for callers that may not call cudaMalloc: code under CUDA graph capture,
// tight serving loops, harnesses that forbid it).
//
// Same kernels, same grid and same partial layout as b200_cg_dot_f64 -- so the same bits in
// d_out -- but the reduction scratch (ticket counter + per-CTA hi/lo partials) is this
// file-scope device array instead of a caller buffer: B200_DOT_SCRATCH_DOUBLES doubles =
// 32 KiB, held by every GPU that loads the library until the process exits.
// In HOST code the array's name is only a stand-in for the symbol, not a device address: the
// address each device maps it at comes from cudaGetSymbolAddress, resolved once per device with
// that device current, and that pointer is what the kernel is given.
// ---------------------------------------------------------------------------
static __device__ __align__(256) double g_dot_static_scratch[B200_DOT_SCRATCH_DOUBLES];

namespace
{
constexpr int STATIC_MAX_DEVICES = 64;
double* g_dot_static_ptr[STATIC_MAX_DEVICES]; // per device, nullptr until resolved

// -> the current device's address of g_dot_static_scratch (no allocation: the array is part of the module image)
int dot_static_scratch(double** out)
{
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < STATIC_MAX_DEVICES, "device ordinal beyond the static-scratch table");
  double* p = __atomic_load_n(&g_dot_static_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) { // racing first calls resolve the same address twice: harmless
    B200_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_dot_static_scratch));
    B200_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned NULL for the static dot scratch");
    __atomic_store_n(&g_dot_static_ptr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return B200_OK;
}
} // namespace

extern "C" int b200_cg_dot_static_f64(int32_t n, const double* d_a, const double* d_b, double* d_out, void* stream)
{
  B200_REQUIRE(n >= 0 && d_out, "bad argument");
  if (n > B200_CG_DOT_STATIC_MAX_N) {
    set_error("b200_cg_dot_static_f64: n = %d exceeds B200_CG_DOT_STATIC_MAX_N = %d (use b200_cg_dot_f64 with a caller "
              "scratch for longer vectors)", (int)n, (int)B200_CG_DOT_STATIC_MAX_N);
    return B200_EINVAL;
  }
  B200_REQUIRE(n == 0 || (d_a && d_b), "NULL vector");
  if (int rc = need_gpu())
    return rc;
  double* scratch = nullptr;
  if (int rc = dot_static_scratch(&scratch))
    return rc;
  // The array keeps whatever the last call left.  The partials are rewritten by every CTA before the last CTA
  // reads them; the ticket counter must start at zero (a call that was aborted mid-kernel would leave it odd),
  // so clear the 64-byte header in stream order.  A memset node, not an allocation.
  B200_CUDA(cudaMemsetAsync(scratch, 0, 8 * sizeof(double), as_stream(stream)));
  const int grid = std::min(vec_grid(n), B200_DOT_SCRATCH_DOUBLES - 8);
  const bool al = ((reinterpret_cast<uintptr_t>(d_a) | reinterpret_cast<uintptr_t>(d_b)) & 15) == 0;
  if (al)
    dot_kernel<<<grid, VEC_THREADS, 0, as_stream(stream)>>>(n, d_a, d_b, d_out, scratch, nullptr, nullptr);
  else
    dot_kernel_scalar<<<grid, VEC_THREADS, 0, as_stream(stream)>>>(n, d_a, d_b, d_out, scratch, nullptr, nullptr);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
