// csr_matrix.h -- CSR block in executor memory (reference: spmv/csr_matrix.h:20-50,
// csr_matrix.cpp:23-87).  Host CSR in, executor-side copies owned by the object.
#pragma once

#include "csr_kernels.h"
#include "sub_matrix.h"

namespace spmv
{

template <typename T>
class SPMV_EXPORT CSRMatrix final : public SubMatrix<T>
{
public:
  /// rowptr/colind/values/diagonal are HOST arrays; they are copied to `exec`.
  /// rowptr may be anything when num_non_zeros == 0 (nothing is allocated then,
  /// csr_matrix.cpp:35); diagonal == nullptr means "no diagonal vector".
  CSRMatrix(std::shared_ptr<DeviceExecutor> exec, int32_t num_rows, int32_t num_cols,
            int64_t num_non_zeros, const int32_t* rowptr, const int32_t* colind, const T* values,
            const T* diagonal = nullptr, bool symmetric = false);
  /// Adopt arrays that already live in `exec` memory (allocated with exec->alloc; the
  /// block frees them).  rowptr/colind/values must be null when num_non_zeros == 0.
  struct adopt_t {};
  CSRMatrix(adopt_t, std::shared_ptr<DeviceExecutor> exec, int32_t num_rows, int32_t num_cols,
            int64_t num_non_zeros, int32_t* d_rowptr, int32_t* d_colind, T* d_values,
            T* d_diagonal = nullptr, bool symmetric = false);
  ~CSRMatrix() override;
  CSRMatrix(const CSRMatrix&) = delete;
  CSRMatrix& operator=(const CSRMatrix&) = delete;

  size_t format_size() const override;
  void mult(T alpha, T* __restrict__ in, T beta, T* __restrict__ out) const override;

  const int32_t* rowptr() const { return _rowptr; }
  const int32_t* colind() const { return _colind; }
  const T* values() const { return _values; }
  int32_t* rowptr() { return _rowptr; }
  int32_t* colind() { return _colind; }
  T* values() { return _values; }
  const CSRSpMV<T>& op() const { return _op; }

private:
  int32_t* _rowptr = nullptr;
  int32_t* _colind = nullptr;
  T* _values = nullptr;
  CSRSpMV<T> _op;
};

} // namespace spmv
