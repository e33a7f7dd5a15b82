// halo.cu -- halo pack (gather) and reverse-update accumulate (scatter-add).
//
// Replaces gather_ghosts<T> + its per-call device sync
// (cuda/cuda_executor.cu:9-51; contract reference_executor.cpp:150-164) and
// the host-side accumulate loop of L2GMap::reverse_update
// (L2GMap.cpp:920-921, 948-949).  Grid-stride, two independent loads in flight
// per thread; the index stream is the only coalesced read, the gather itself is
// whatever the halo plan dictates (one contiguous plane for slab partitions).
#include "common.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace b200
{

template <typename T>
__global__ void gather_kernel(int n, const int32_t* __restrict__ idx, const T* __restrict__ in,
                              T* __restrict__ out)
{
  const int stride = gridDim.x * blockDim.x;
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + stride < n; i += 2 * stride) {
    const int a = __ldg(idx + i), b = __ldg(idx + i + stride);
    const T va = __ldg(in + a), vb = __ldg(in + b);
    out[i] = va;
    out[i + stride] = vb;
  }
  if (i < n)
    out[i] = __ldg(in + __ldg(idx + i));
}

// Indices may repeat across neighbours (one owned entry ghosted by several
// ranks), so the accumulate is an atomic add.
template <typename T>
__global__ void scatter_add_kernel(int n, const int32_t* __restrict__ idx,
                                   const T* __restrict__ in, T* __restrict__ inout)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    red_add(inout + __ldg(idx + i), __ldg(in + i));
}

template <typename T>
static int run_gather(int32_t n, const int32_t* d_idx, const T* d_in, T* d_out, void* stream)
{
  B200_REQUIRE(n >= 0, "negative count");
  if (n == 0)
    return B200_OK;
  B200_REQUIRE(d_idx && d_in && d_out, "NULL pointer");
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  const int threads = 256;
  const int blocks = std::min((n + 2 * threads - 1) / (2 * threads), num_sms() * 8);
  gather_kernel<T><<<std::max(blocks, 1), threads, 0, as_stream(stream)>>>(n, d_idx, d_in, d_out);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename T>
static int run_scatter_add(int32_t n, const int32_t* d_idx, const T* d_in, T* d_inout, void* stream)
{
  B200_REQUIRE(n >= 0, "negative count");
  if (n == 0)
    return B200_OK;
  B200_REQUIRE(d_idx && d_in && d_inout, "NULL pointer");
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  const int threads = 256;
  const int blocks = std::min((n + threads - 1) / threads, num_sms() * 8);
  scatter_add_kernel<T><<<blocks, threads, 0, as_stream(stream)>>>(n, d_idx, d_in, d_inout);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ---------------------------------------------------------------------------
// Kernel-driven peer halo (the forward update of L2GMap.cpp:502-526 in two launches on ONE
// stream, no copy engine, no stream memory operations, nothing the host has to count):
//   push : pack AND send -- every packed element x[indexbuf[k]] is stored straight into the
//          owning neighbour's IPC-mapped staging slot over NVLink; the last CTA to finish
//          publishes the epoch into each neighbour's flag word (release, system scope);
//   wait : every CTA polls this rank's flag words until each neighbour has published the
//          epoch, then the staging slot is copied into the ghost tail of x.
// The epoch lives in the descriptor in device memory and is advanced by the push kernel
// itself, so the pair is a fixed sequence of kernel nodes: a CUDA graph can replay it.
// Two slots by epoch parity are enough: a neighbour's push of epoch e+1 is stream-ordered
// after its wait for my push of epoch e, which follows my unpack of epoch e-1.
// ---------------------------------------------------------------------------
// (struct HaloDesc: common.cuh -- the fused CG step of cg.cu pushes through it too)

template <typename T>
__global__ void __launch_bounds__(256) halo_push_kernel(HaloDesc* d, int n, const int32_t* __restrict__ idx,
                                                        const T* __restrict__ x)
{
  __shared__ bool s_last;
  const unsigned int epoch = d->epoch + 1;
  const int slot = (int)(epoch & 1u), nnbr = d->nnbr;
  const int stride = gridDim.x * blockDim.x;
  int i = 0;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
    while (k >= d->send_off[i + 1])
      ++i;
    T* dst = reinterpret_cast<T*>(d->peer_slot[slot][i]) + d->land[i] + (k - d->send_off[i]);
    *dst = __ldg(x + __ldg(idx + k));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system(); // cumulative: the CTA's stores (ordered before me by the barrier) precede the ticket
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

template <typename T>
__global__ void __launch_bounds__(256) halo_wait_unpack_kernel(const HaloDesc* d, T* __restrict__ ghost_tail)
{
  const unsigned int epoch = d->epoch; // the push of this exchange precedes us on the stream
  const int nnbr = d->nnbr;
  if ((int)threadIdx.x < nnbr) {
    unsigned int seen;
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(d->my_flag[threadIdx.x]) : "memory");
    } while ((int)(seen - epoch) < 0);
  }
  __syncthreads();
  const T* slot = reinterpret_cast<const T*>(d->my_slot[epoch & 1u]);
  const long long n = d->nghosts;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
    T v;
    if constexpr (sizeof(T) == 8)
      asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(slot + k) : "memory"); // never a stale L1 line
    else
      asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(slot + k) : "memory");
    ghost_tail[k] = v;
  }
}

template <typename T>
static int run_halo_push(void* desc, int32_t n, const int32_t* d_idx, const T* d_x, void* stream)
{
  B200_REQUIRE(desc != nullptr && n >= 0, "bad argument");
  B200_REQUIRE(n == 0 || (d_idx && d_x), "NULL pointer");
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  const int threads = 256;
  static const int cap = [] {
    const char* e = std::getenv("B200_HALO_PUSH_CTAS");
    return (e && *e) ? std::atoi(e) : 0;
  }();
  // four elements per thread: every CTA ends with a system-scope fence and a ticket, and 512 of them for a
  // 131072-entry halo cost more than they moved (2 GPUs, CG iteration 175.0 us with 512 CTAs, 171.7 with 128)
  const int blocks = std::max(1, std::min((n + 4 * threads - 1) / (4 * threads), cap > 0 ? cap : num_sms() * 4));
  halo_push_kernel<T><<<blocks, threads, 0, as_stream(stream)>>>(static_cast<HaloDesc*>(desc), n, d_idx, d_x);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename T>
static int run_halo_wait(void* desc, int64_t nghosts, T* d_ghost_tail, void* stream)
{
  B200_REQUIRE(desc != nullptr && nghosts >= 0, "bad argument");
  B200_REQUIRE(nghosts == 0 || d_ghost_tail, "NULL pointer");
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  const int threads = 256;
  const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((nghosts + 4 * threads - 1) / (4 * threads), num_sms()));
  halo_wait_unpack_kernel<T><<<blocks, threads, 0, as_stream(stream)>>>(static_cast<const HaloDesc*>(desc), d_ghost_tail);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

} // namespace b200

extern "C" {

int b200_halo_desc_create(void** desc, int32_t nnbr, const int32_t* send_off, const int64_t* land,
                          void* const* d_peer_slot0, void* const* d_peer_slot1, void* const* d_peer_flag,
                          void* d_my_slot0, void* d_my_slot1, void* const* d_my_flag, int64_t nghosts, void* stream)
{
  using namespace b200;
  B200_REQUIRE(desc != nullptr, "desc is NULL");
  *desc = nullptr;
  B200_REQUIRE(nnbr >= 1 && nnbr <= B200_HALO_MAX_NEIGHBOURS, "neighbour count out of range");
  B200_REQUIRE(send_off && land && d_peer_slot0 && d_peer_slot1 && d_peer_flag && d_my_flag, "NULL array");
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  HaloDesc h;
  memset(&h, 0, sizeof(h));
  h.nnbr = nnbr;
  h.nghosts = nghosts;
  for (int i = 0; i <= nnbr; ++i)
    h.send_off[i] = send_off[i];
  for (int i = 0; i < nnbr; ++i) {
    B200_REQUIRE(d_peer_slot0[i] && d_peer_slot1[i] && d_peer_flag[i] && d_my_flag[i], "NULL peer pointer");
    h.land[i] = land[i];
    h.peer_slot[0][i] = static_cast<char*>(d_peer_slot0[i]);
    h.peer_slot[1][i] = static_cast<char*>(d_peer_slot1[i]);
    h.peer_flag[i] = static_cast<unsigned int*>(d_peer_flag[i]);
    h.my_flag[i] = static_cast<unsigned int*>(d_my_flag[i]);
  }
  h.my_slot[0] = static_cast<char*>(d_my_slot0);
  h.my_slot[1] = static_cast<char*>(d_my_slot1);
  HaloDesc* d = nullptr;
  B200_CUDA(cudaMalloc(&d, sizeof(HaloDesc)));
  if (cudaMemcpyAsync(d, &h, sizeof(h), cudaMemcpyHostToDevice, as_stream(stream)) != cudaSuccess
      || cudaStreamSynchronize(as_stream(stream)) != cudaSuccess) {
    cudaFree(d);
    return cuda_fail(cudaGetLastError(), "halo_desc_create", __FILE__, __LINE__);
  }
  *desc = d;
  return B200_OK;
}

int b200_halo_desc_destroy(void* desc)
{
  if (desc)
    B200_CUDA(cudaFree(desc));
  return B200_OK;
}

int b200_halo_push_f64(void* desc, int32_t n, const int32_t* d_idx, const double* d_x, void* stream)
{
  return b200::run_halo_push<double>(desc, n, d_idx, d_x, stream);
}
int b200_halo_push_f32(void* desc, int32_t n, const int32_t* d_idx, const float* d_x, void* stream)
{
  return b200::run_halo_push<float>(desc, n, d_idx, d_x, stream);
}
int b200_halo_wait_unpack_f64(void* desc, int64_t nghosts, double* d_ghost_tail, void* stream)
{
  return b200::run_halo_wait<double>(desc, nghosts, d_ghost_tail, stream);
}
int b200_halo_wait_unpack_f32(void* desc, int64_t nghosts, float* d_ghost_tail, void* stream)
{
  return b200::run_halo_wait<float>(desc, nghosts, d_ghost_tail, stream);
}

int b200_gather_f64(int32_t n, const int32_t* d_idx, const double* d_in, double* d_out,
                    void* stream)
{
  return b200::run_gather<double>(n, d_idx, d_in, d_out, stream);
}
int b200_gather_f32(int32_t n, const int32_t* d_idx, const float* d_in, float* d_out, void* stream)
{
  return b200::run_gather<float>(n, d_idx, d_in, d_out, stream);
}
int b200_scatter_add_f64(int32_t n, const int32_t* d_idx, const double* d_in, double* d_inout,
                         void* stream)
{
  return b200::run_scatter_add<double>(n, d_idx, d_in, d_inout, stream);
}
int b200_scatter_add_f32(int32_t n, const int32_t* d_idx, const float* d_in, float* d_inout,
                         void* stream)
{
  return b200::run_scatter_add<float>(n, d_idx, d_in, d_inout, stream);
}

} // extern "C"
