// ubench_atomics.cu -- micro-benchmarks that bound the symmetric-SpMV scatter on B200.
// Build+run on the GPU box: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/ub ubench_atomics.cu && /tmp/ub
// Pattern: N = 256^3 rows, each row scatters to y[i-1], y[i-256], y[i-65536] and adds to y[i]
// (the 7-point symmetric lower triangle).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int N = 256 * 256 * 256;

__global__ void t1_red(double* y, int n)
{
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double v = 1.0 + i * 1e-9;
    if (i >= 1) atomicAdd(y + i - 1, v);
    if (i >= 256) atomicAdd(y + i - 256, v);
    if (i >= 65536) atomicAdd(y + i - 65536, v);
    atomicAdd(y + i, v);
  }
}
// k reds per row to y[i - j*stride_k]
__global__ void t1b_red_k(double* y, int n, int k)
{
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double v = 1.0 + i * 1e-9;
    for (int j = 0; j < k; ++j) {
      int c = i - j * 65536;
      if (c >= 0) atomicAdd(y + c, v);
    }
  }
}
__global__ void t2_rmw(double* y, int n)
{
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double v = 1.0 + i * 1e-9;
    if (i >= 1) y[i - 1] += v;
    if (i >= 256) y[i - 256] += v;
    if (i >= 65536) y[i - 65536] += v;
    y[i] += v;
  }
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// T3: bulk reduce smem -> global, W doubles per op; each CTA handles windows b, b+grid...; 4 reduce ops per
// window (offsets 0, -1 (unaligned -> use -2), -256, -65536) so the element count equals t1's.
template <int W>
__global__ void t3_bulk(double* y, int n, int nops)
{
  extern __shared__ __align__(128) double win[];
  for (int i = threadIdx.x; i < W; i += blockDim.x) win[i] = 1.0 + i * 1e-9;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  const int nwin = n / W;
  if (threadIdx.x == 0) {
    for (int w = blockIdx.x; w < nwin; w += gridDim.x) {
      const int base = w * W;
      const int offs[4] = {0, 2, 256, 65536};
      for (int k = 0; k < nops; ++k) {
        int c = base - offs[k];
        if (c < 0) continue;
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(y + c),
                     "r"(smem_u32(win)), "r"(W * 8)
                     : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

// T4: shared-memory fp64 atomicAdd, 3 per row + own, consecutive addresses; window of 2048 doubles
__global__ void t4_smem_atomic(double* y, int n, int reps)
{
  __shared__ double win[4096];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) win[i] = 0;
  __syncthreads();
  for (int r = 0; r < reps; ++r) {
    int i = 300 + (threadIdx.x + r * blockDim.x) % 3000;
    double v = 1.0 + i * 1e-9;
    atomicAdd(&win[i - 1], v);
    atomicAdd(&win[i - 256], v);
    atomicAdd(&win[i - 37], v);
    atomicAdd(&win[i], v);
  }
  __syncthreads();
  if (threadIdx.x == 0) y[blockIdx.x] = win[5] + win[700];
}
__global__ void t5_smem_rmw(double* y, int n, int reps)
{
  __shared__ double win[4096];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) win[i] = 0;
  __syncthreads();
  for (int r = 0; r < reps; ++r) {
    int i = 300 + (threadIdx.x + r * blockDim.x) % 3000;
    double v = 1.0 + i * 1e-9;
    win[i - 1] += v;
    win[i - 256] += v;
    win[i - 37] += v;
    win[i] += v;
  }
  __syncthreads();
  if (threadIdx.x == 0) y[blockIdx.x] = win[5] + win[700];
}
// T6: plain coalesced write of y (store bandwidth reference) and read-modify-write streaming
__global__ void t6_axpy(double* y, int n)
{
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) y[i] = y[i] * 1.0000001 + 1.0;
}

template <typename F>
float timeit(F f, int reps = 10)
{
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  f();
  f();
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  for (int i = 0; i < reps; ++i) f();
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  return ms / reps;
}

int main()
{
  double* y;
  CK(cudaMalloc(&y, sizeof(double) * N));
  CK(cudaMemset(y, 0, sizeof(double) * N));
  int sms = 148;
  for (int bs : {256, 512}) {
    for (int per : {2, 4, 8}) {
      float ms = timeit([&] { t1_red<<<sms * per, bs>>>(y, N); });
      printf("t1_red grid=%d block=%d: %.4f ms  (%.1f G red/s)\n", sms * per, bs, ms, 4.0 * N / ms / 1e6);
    }
  }
  {
    float ms = timeit([&] { t1_red<<<N / 256, 256>>>(y, N); });
    printf("t1_red grid=N/256: %.4f ms (%.1f G red/s)\n", ms, 4.0 * N / ms / 1e6);
  }
  for (int k : {1, 2, 4, 8, 13}) {
    float ms = timeit([&] { t1b_red_k<<<N / 256, 256>>>(y, N, k); });
    printf("t1b_red k=%d: %.4f ms (%.1f G red/s)\n", k, ms, (double)k * N / ms / 1e6);
  }
  {
    float ms = timeit([&] { t2_rmw<<<N / 256, 256>>>(y, N); });
    printf("t2_rmw (racy ld+add+st) grid=N/256: %.4f ms (%.1f G rmw/s)\n", ms, 4.0 * N / ms / 1e6);
  }
  {
    float ms = timeit([&] { t6_axpy<<<sms * 8, 512>>>(y, N); });
    printf("t6 y=y*a+1 streaming: %.4f ms (%.1f GB/s)\n", ms, 16.0 * N / ms / 1e6);
  }
  for (int nops : {1, 2, 4}) {
    {
      cudaFuncSetAttribute(t3_bulk<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
      float ms = timeit([&] { t3_bulk<1024><<<sms * 4, 64, 1024 * 8>>>(y, N, nops); });
      printf("t3_bulk W=1024 nops=%d: %.4f ms (%.1f G elem/s)\n", nops, ms, (double)nops * N / ms / 1e6);
    }
    {
      cudaFuncSetAttribute(t3_bulk<2048>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
      float ms = timeit([&] { t3_bulk<2048><<<sms * 4, 64, 2048 * 8>>>(y, N, nops); });
      printf("t3_bulk W=2048 nops=%d: %.4f ms (%.1f G elem/s)\n", nops, ms, (double)nops * N / ms / 1e6);
    }
    {
      float ms = timeit([&] { t3_bulk<256><<<sms * 8, 64, 256 * 8>>>(y, N, nops); });
      printf("t3_bulk W=256 nops=%d: %.4f ms (%.1f G elem/s)\n", nops, ms, (double)nops * N / ms / 1e6);
    }
  }
  {
    int reps = 200;
    float ms = timeit([&] { t4_smem_atomic<<<sms * 2, 256>>>(y, N, reps); });
    printf("t4 smem atomicAdd f64: %.4f ms (%.2f adds/clk/SM @1.8GHz)\n", ms, 4.0 * reps * 256 * 2 / (ms * 1e-3 * 1.8e9));
    ms = timeit([&] { t5_smem_rmw<<<sms * 2, 256>>>(y, N, reps); });
    printf("t5 smem plain rmw f64: %.4f ms (%.2f adds/clk/SM @1.8GHz)\n", ms, 4.0 * reps * 256 * 2 / (ms * 1e-3 * 1.8e9));
  }
  CK(cudaDeviceSynchronize());
  CK(cudaGetLastError());
  printf("done\n");
  return 0;
}
