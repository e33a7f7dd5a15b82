// b200_executor.h -- the B200 (sm_100a) backend of the DeviceExecutor interface.
//
// Takes the place of the reference's CudaExecutor (cuda/cuda_executor.h:17-101):
// same factory `create(device_id, host)`, same stream setters.  Differences, all
// behind the interface:
//   - kernels are this project's (include/b200spmv.h), not cuSPARSE / cuBLAS;
//   - no hot-path entry synchronises the host (the reference syncs after every pack
//     and every symmetric SpMV, cuda/cuda_executor.cu:28-32, csr_kernels.cuda.cu:105);
//   - the executor also owns the NCCL communicator and the side stream the halo
//     exchange runs on (the reference hands device pointers to CUDA-aware MPI).
#pragma once

#include "csr_kernels.h"
#include "device_executor.h"

#include <mpi.h>

struct CUstream_st;

namespace spmv
{

class SPMV_EXPORT B200Executor final : public DeviceExecutor
{
public:
  using stream_t = CUstream_st*; // == cudaStream_t

  ~B200Executor() override;

  template <typename... Ts>
  static std::unique_ptr<B200Executor> create(Ts&&... params)
  {
    return std::unique_ptr<B200Executor>(new B200Executor(std::forward<Ts>(params)...));
  }

  void synchronize() const override;
  const DeviceExecutor& get_host() const override { return *_host; }
  int get_num_devices() const override;
  int get_num_cus() const override; // SM count

  void spmv_init(CSRSpMV<float>& op, const CSRMatrix<float>& mat) const override;
  void spmv_init(CSRSpMV<double>& op, const CSRMatrix<double>& mat) const override;
  void spmv_run(const CSRSpMV<float>& op, const CSRMatrix<float>& mat, float alpha,
                float* __restrict__ in, float beta, float* __restrict__ out) const override;
  void spmv_run(const CSRSpMV<double>& op, const CSRMatrix<double>& mat, double alpha,
                double* __restrict__ in, double beta, double* __restrict__ out) const override;
  void spmv_finalize(CSRSpMV<float>& op) const override;
  void spmv_finalize(CSRSpMV<double>& op) const override;
  void gather_ghosts_run(int num_indices, const int32_t* indices, const float* in,
                         float* out) const override;
  void gather_ghosts_run(int num_indices, const int32_t* indices, const double* in,
                         double* out) const override;

  // ---- extended API (cuda/cuda_executor.h:71-76) ----
  void set_cuda_stream(stream_t stream) { _stream = stream; }
  void reset_cuda_stream() { _stream = nullptr; }
  stream_t get_cuda_stream() const { return _stream; }

  // ---- halo / reduction plumbing used by L2GMap and cg ----
  /// Collective over `comm`: make sure an NCCL communicator spanning comm's ranks
  /// exists (no-op for a single rank).  Called by L2GMap's constructor.
  void attach_comm(MPI_Comm comm) const;
  void* nccl_comm() const { return _nccl; }
  int comm_size() const { return _comm_size; }
  /// Side stream + events for the overlapped (non-blocking) exchange.
  void* comm_stream() const { return _comm_stream; }
  void* packed_event() const { return _ev_packed; }
  void* arrived_event() const { return _ev_arrived; }
  /// Slot for the CG solver's cached workspace (vectors, state block, private stream, captured graph):
  /// spmv::cg allocates it on first use and re-uses it across calls; freed with the executor.
  std::shared_ptr<void>& cg_workspace() const { return _cg_workspace; }
  /// scatter-add used by L2GMap::reverse_update
  void scatter_add_run(int num_indices, const int32_t* indices, const float* in, float* inout) const;
  void scatter_add_run(int num_indices, const int32_t* indices, const double* in, double* inout) const;

protected:
  void* _alloc(size_t num_bytes) const override;
  void _free(void* ptr) const override;
  void _memset(void* ptr, int value, size_t num_bytes) const override;
  void _copy(void* dst, const void* src, size_t num_bytes) const override;
  void _copy_async(void* dst, const void* src, size_t num_bytes, void* stream) const override;
  void _copy_from(void* dst, const DeviceExecutor& src_exec, const void* src,
                  size_t num_bytes) const override;
  void _copy_to(void* dst, const DeviceExecutor& dst_exec, const void* src,
                size_t num_bytes) const override;

private:
  B200Executor(int device_id, std::shared_ptr<DeviceExecutor> host);
  void bind() const; // cudaSetDevice(id) on the calling thread

  std::shared_ptr<DeviceExecutor> _host;
  stream_t _stream = nullptr;
  mutable void* _nccl = nullptr;
  mutable int _comm_size = 1;
  void* _comm_stream = nullptr;
  void* _ev_packed = nullptr;
  void* _ev_arrived = nullptr;
  mutable std::shared_ptr<void> _cg_workspace;
};

// Source compatibility with code written for the reference's CUDA backend.
using CudaExecutor = B200Executor;

} // namespace spmv
