// reference_executor.h -- the host-memory executor.
//
// In the reference, ReferenceExecutor is the sequential CPU backend
// (reference_executor.h:16-85, reference_executor.cpp:20-164).  This build has ONE
// compute backend, the B200; the class is kept because the API needs a host side:
// `get_host()`, the source of `copy_from`, the owner of host staging buffers, and the
// `host` argument of the GPU executor's factory (cuda/cuda_executor.h:24-30).
// Memory calls work (malloc / memcpy); the compute visitors throw
// std::runtime_error -- there is deliberately no CPU SpMV in this library.
#pragma once

#include "csr_kernels.h"
#include "device_executor.h"

namespace spmv
{

class SPMV_EXPORT ReferenceExecutor final : public DeviceExecutor
{
public:
  ~ReferenceExecutor() override = default;

  template <typename... Ts>
  static std::unique_ptr<ReferenceExecutor> create(Ts&&... params)
  {
    return std::unique_ptr<ReferenceExecutor>(new ReferenceExecutor(std::forward<Ts>(params)...));
  }

  void synchronize() const override {}
  const DeviceExecutor& get_host() const override { return *this; }
  int get_num_devices() const override { return 1; }
  int get_num_cus() const override { return 1; }

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
  ReferenceExecutor() { _dev_info.type = DeviceType::cpu; }
};

using HostExecutor = ReferenceExecutor;

} // namespace spmv
