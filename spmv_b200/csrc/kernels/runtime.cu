// runtime.cu -- memory / stream / event entry points of include/b200spmv.h.
// Replaces the body of the reference's CudaExecutor (cuda/cuda_executor.cpp:
// 54-105): cudaMalloc / cudaFree / cudaMemset / cudaMemcpy on one settable
// stream, without the per-call device synchronisations.
#include "common.cuh"

#include <stdarg.h>
#include <string.h>

namespace b200
{

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t err, const char* what, const char* file, int line)
{
  set_error("CUDA error %d (%s) in %s at %s:%d", (int)err, cudaGetErrorString(err), what, file,
            line);
  return err == cudaErrorMemoryAllocation ? B200_ENOMEM : B200_ECUDA;
}

int num_sms()
{
  static thread_local int cached_dev = -1;
  static thread_local int cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess)
    return 0;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      return 0;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

} // namespace b200

using namespace b200;

extern "C" {

const char* b200_last_error(void) { return g_err; }
int b200_abi_version(void) { return 1; }

int b200_device_count(int* count)
{
  B200_REQUIRE(count != nullptr, "count is NULL");
  *count = 0;
  B200_CUDA(cudaGetDeviceCount(count));
  return B200_OK;
}

int b200_set_device(int device_id)
{
  B200_CUDA(cudaSetDevice(device_id));
  return B200_OK;
}

int b200_get_num_sms(int* n)
{
  B200_REQUIRE(n != nullptr, "num_sms is NULL");
  *n = num_sms();
  if (*n <= 0) {
    set_error("no CUDA device available");
    return B200_ECUDA;
  }
  return B200_OK;
}

int b200_malloc(void** d_ptr, size_t num_bytes)
{
  B200_REQUIRE(d_ptr != nullptr, "d_ptr is NULL");
  *d_ptr = nullptr;
  if (num_bytes == 0)
    return B200_OK;
  B200_CUDA(cudaMalloc(d_ptr, num_bytes));
  return B200_OK;
}

int b200_free(void* d_ptr)
{
  if (d_ptr != nullptr)
    B200_CUDA(cudaFree(d_ptr));
  return B200_OK;
}

int b200_memset_async(void* d_ptr, int value, size_t num_bytes, void* stream)
{
  if (num_bytes > 0)
    B200_CUDA(cudaMemsetAsync(d_ptr, value, num_bytes, as_stream(stream)));
  return B200_OK;
}

int b200_memcpy_d2d_async(void* d_dst, const void* d_src, size_t num_bytes, void* stream)
{
  if (num_bytes > 0 && d_dst != d_src)
    B200_CUDA(cudaMemcpyAsync(d_dst, d_src, num_bytes, cudaMemcpyDeviceToDevice,
                              as_stream(stream)));
  return B200_OK;
}

int b200_memcpy_h2d_async(void* d_dst, const void* h_src, size_t num_bytes, void* stream)
{
  if (num_bytes > 0)
    B200_CUDA(cudaMemcpyAsync(d_dst, h_src, num_bytes, cudaMemcpyHostToDevice,
                              as_stream(stream)));
  return B200_OK;
}

int b200_memcpy_d2h_async(void* h_dst, const void* d_src, size_t num_bytes, void* stream)
{
  if (num_bytes > 0)
    B200_CUDA(cudaMemcpyAsync(h_dst, d_src, num_bytes, cudaMemcpyDeviceToHost,
                              as_stream(stream)));
  return B200_OK;
}

int b200_host_alloc_pinned(void** h_ptr, size_t num_bytes)
{
  B200_REQUIRE(h_ptr != nullptr, "h_ptr is NULL");
  *h_ptr = nullptr;
  if (num_bytes == 0)
    return B200_OK;
  B200_CUDA(cudaMallocHost(h_ptr, num_bytes));
  return B200_OK;
}

int b200_host_free_pinned(void* h_ptr)
{
  if (h_ptr != nullptr)
    B200_CUDA(cudaFreeHost(h_ptr));
  return B200_OK;
}

int b200_stream_create(void** stream)
{
  B200_REQUIRE(stream != nullptr, "stream is NULL");
  cudaStream_t s;
  B200_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *stream = s;
  return B200_OK;
}

int b200_stream_create_priority(void** stream, int high)
{
  B200_REQUIRE(stream != nullptr, "stream is NULL");
  int least = 0, greatest = 0;
  B200_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
  cudaStream_t s;
  B200_CUDA(cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, high ? greatest : least));
  *stream = s;
  return B200_OK;
}

// ---- CUDA IPC + stream memory operations (direct peer halo) ------------------------------
int b200_ipc_alloc(void** d_ptr, size_t num_bytes, char handle[B200_IPC_HANDLE_BYTES])
{
  B200_REQUIRE(d_ptr != nullptr && handle != nullptr && num_bytes > 0, "bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == B200_IPC_HANDLE_BYTES, "IPC handle size");
  *d_ptr = nullptr;
  void* p = nullptr;
  B200_CUDA(cudaMalloc(&p, num_bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaMemset(p, 0, num_bytes);
  if (e == cudaSuccess)
    e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return cuda_fail(e, "b200_ipc_alloc", __FILE__, __LINE__);
  }
  memcpy(handle, &h, sizeof(h));
  *d_ptr = p;
  return B200_OK;
}

int b200_ipc_open(void** d_peer_ptr, const char handle[B200_IPC_HANDLE_BYTES])
{
  B200_REQUIRE(d_peer_ptr != nullptr && handle != nullptr, "bad argument");
  *d_peer_ptr = nullptr;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  cudaError_t e = cudaIpcOpenMemHandle(d_peer_ptr, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    cudaGetLastError(); // not sticky: the caller falls back to NCCL
    *d_peer_ptr = nullptr;
    return cuda_fail(e, "cudaIpcOpenMemHandle", __FILE__, __LINE__);
  }
  return B200_OK;
}

int b200_ipc_close(void* d_peer_ptr)
{
  if (d_peer_ptr)
    B200_CUDA(cudaIpcCloseMemHandle(d_peer_ptr));
  return B200_OK;
}

namespace
{
// Driver entry points fetched through the runtime (the library does not link libcuda).
using MemOp32 = int (*)(cudaStream_t, unsigned long long, unsigned int, unsigned int);
struct MemOps {
  MemOp32 write = nullptr, wait = nullptr;
  bool tried = false;
};
MemOps& memops()
{
  static MemOps m;
  if (!m.tried) {
    m.tried = true;
    void *w = nullptr, *a = nullptr;
    cudaDriverEntryPointQueryResult q1, q2;
    if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &w, cudaEnableDefault, &q1) == cudaSuccess
        && cudaGetDriverEntryPoint("cuStreamWaitValue32", &a, cudaEnableDefault, &q2) == cudaSuccess
        && q1 == cudaDriverEntryPointSuccess && q2 == cudaDriverEntryPointSuccess) {
      m.write = reinterpret_cast<MemOp32>(w);
      m.wait = reinterpret_cast<MemOp32>(a);
    } else {
      cudaGetLastError();
    }
  }
  return m;
}
} // namespace

int b200_stream_write_u32(void* stream, void* d_flag, uint32_t value)
{
  B200_REQUIRE(d_flag != nullptr, "flag is NULL");
  MemOps& m = memops();
  if (!m.write) {
    set_error("cuStreamWriteValue32 is not available from this driver");
    return B200_ECUDA;
  }
  const int rc = m.write(as_stream(stream), (unsigned long long)(uintptr_t)d_flag, value, 0u /* default */);
  if (rc != 0) {
    set_error("cuStreamWriteValue32 failed with CUresult %d", rc);
    return B200_ECUDA;
  }
  return B200_OK;
}

int b200_stream_wait_u32_geq(void* stream, void* d_flag, uint32_t value)
{
  B200_REQUIRE(d_flag != nullptr, "flag is NULL");
  MemOps& m = memops();
  if (!m.wait) {
    set_error("cuStreamWaitValue32 is not available from this driver");
    return B200_ECUDA;
  }
  const int rc = m.wait(as_stream(stream), (unsigned long long)(uintptr_t)d_flag, value, 0u /* CU_STREAM_WAIT_VALUE_GEQ */);
  if (rc != 0) {
    set_error("cuStreamWaitValue32 failed with CUresult %d", rc);
    return B200_ECUDA;
  }
  return B200_OK;
}

int b200_graph_capture_begin(void* stream)
{
  B200_REQUIRE(stream != nullptr, "the legacy default stream cannot be captured");
  B200_CUDA(cudaStreamBeginCapture(as_stream(stream), cudaStreamCaptureModeThreadLocal));
  return B200_OK;
}

int b200_graph_capture_end(void* stream, void** graph_exec)
{
  B200_REQUIRE(graph_exec != nullptr, "graph_exec is NULL");
  *graph_exec = nullptr;
  cudaGraph_t g = nullptr;
  {
    const cudaError_t err = cudaStreamEndCapture(as_stream(stream), &g);
    if (err != cudaSuccess) {
      cudaGetLastError(); // an invalidated capture must not leave its error for the next launch check to find
      if (g)
        cudaGraphDestroy(g);
      return cuda_fail(err, "cudaStreamEndCapture", __FILE__, __LINE__);
    }
  }
  cudaGraphExec_t e = nullptr;
  const cudaError_t err = cudaGraphInstantiate(&e, g, 0);
  cudaGraphDestroy(g);
  if (err != cudaSuccess)
    return cuda_fail(err, "cudaGraphInstantiate", __FILE__, __LINE__);
  *graph_exec = e;
  return B200_OK;
}

int b200_graph_launch(void* graph_exec, void* stream)
{
  B200_REQUIRE(graph_exec != nullptr, "graph_exec is NULL");
  B200_CUDA(cudaGraphLaunch(static_cast<cudaGraphExec_t>(graph_exec), as_stream(stream)));
  return B200_OK;
}

int b200_graph_destroy(void* graph_exec)
{
  if (graph_exec)
    B200_CUDA(cudaGraphExecDestroy(static_cast<cudaGraphExec_t>(graph_exec)));
  return B200_OK;
}

int b200_stream_destroy(void* stream)
{
  if (stream != nullptr)
    B200_CUDA(cudaStreamDestroy(as_stream(stream)));
  return B200_OK;
}

int b200_stream_synchronize(void* stream)
{
  B200_CUDA(cudaStreamSynchronize(as_stream(stream)));
  return B200_OK;
}

int b200_event_create(void** event, int timing)
{
  B200_REQUIRE(event != nullptr, "event is NULL");
  cudaEvent_t e;
  B200_CUDA(cudaEventCreateWithFlags(&e, timing ? cudaEventDefault : cudaEventDisableTiming));
  *event = e;
  return B200_OK;
}

int b200_event_destroy(void* event)
{
  if (event != nullptr)
    B200_CUDA(cudaEventDestroy(static_cast<cudaEvent_t>(event)));
  return B200_OK;
}

int b200_event_record(void* event, void* stream)
{
  B200_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(event), as_stream(stream)));
  return B200_OK;
}

int b200_event_synchronize(void* event)
{
  B200_CUDA(cudaEventSynchronize(static_cast<cudaEvent_t>(event)));
  return B200_OK;
}

int b200_stream_wait_event(void* stream, void* event)
{
  B200_CUDA(cudaStreamWaitEvent(as_stream(stream), static_cast<cudaEvent_t>(event), 0));
  return B200_OK;
}

int b200_event_elapsed_ms(void* start, void* stop, float* ms)
{
  B200_REQUIRE(ms != nullptr, "ms is NULL");
  B200_CUDA(cudaEventElapsedTime(ms, static_cast<cudaEvent_t>(start),
                                 static_cast<cudaEvent_t>(stop)));
  return B200_OK;
}

} // extern "C"
