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

} // namespace b200

extern "C" {

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
