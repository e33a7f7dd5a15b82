// csr_kernels.h -- the SpMV kernel object hung off every CSR block
// (reference: spmv/csr_kernels.h:22-83).  One init/run/finalize overload per
// executor type; `_aux_data` is the B200 plan handle (b200_csr_plan_*).
#pragma once

#include "config.h"
#include "spmv_export.h"

#include <cstdint>
#include <cstdlib>

namespace spmv
{

class ReferenceExecutor;
class B200Executor;

template <typename T>
class SPMV_EXPORT CSRSpMV
{
public:
  // Host executor: init records the flag; run throws (no CPU compute path).
  void init(int32_t num_rows, int32_t num_cols, int64_t num_non_zeros, const int32_t* rowptr,
            const int32_t* colind, const T* values, bool symmetric, const ReferenceExecutor& exec);
  void run(int32_t num_rows, int32_t num_cols, int64_t num_non_zeros, const int32_t* rowptr,
           const int32_t* colind, const T* values, const T* diagonal, T alpha, T* __restrict__ in,
           T beta, T* __restrict__ out, const ReferenceExecutor& exec) const;
  void finalize(const ReferenceExecutor& exec) const;

  // B200: init builds the tile plan from the device-resident rowptr; run launches
  // b200_csr_spmv_* / b200_csr_symv_* on the executor's stream.
  void init(int32_t num_rows, int32_t num_cols, int64_t num_non_zeros, const int32_t* rowptr,
            const int32_t* colind, const T* values, bool symmetric, const B200Executor& exec);
  void run(int32_t num_rows, int32_t num_cols, int64_t num_non_zeros, const int32_t* rowptr,
           const int32_t* colind, const T* values, const T* diagonal, T alpha, T* __restrict__ in,
           T beta, T* __restrict__ out, const B200Executor& exec) const;
  void finalize(const B200Executor& exec) const;

  /// The B200 plan (void* for b200_csr_plan_info / b200_csr_plan_tune), or nullptr.
  void* plan() const { return _aux_data; }

private:
  bool _symmetric = false;
  mutable void* _aux_data = nullptr;
};

} // namespace spmv
