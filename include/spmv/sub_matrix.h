// sub_matrix.h -- one rank-local sparse block living in executor memory
// (reference: spmv/sub_matrix.h:26-122, minus the Eigen constructor: Eigen is not
// available here, blocks are built from raw CSR arrays).
#pragma once

#include "config.h"
#include "spmv_export.h"

#include <cstddef>
#include <cstdint>
#include <memory>

namespace spmv
{

class DeviceExecutor;

template <typename T>
class SubMatrix
{
public:
  SubMatrix() = default;
  virtual ~SubMatrix() = default;

  int32_t rows() const { return _num_rows; }
  int32_t cols() const { return _num_cols; }
  int64_t non_zeros() const { return _num_non_zeros; }
  bool symmetric() const { return _symmetric; }
  const T* diagonal() const { return _diagonal; }
  T* diagonal() { return _diagonal; }

  virtual size_t format_size() const = 0;
  /// out = alpha * A * in + beta * out
  virtual void mult(T alpha, T* __restrict__ in, T beta, T* __restrict__ out) const = 0;

protected:
  std::shared_ptr<DeviceExecutor> _exec = nullptr;
  int32_t _num_rows = 0;
  int32_t _num_cols = 0;
  int64_t _num_non_zeros = 0;
  bool _symmetric = false;
  T* _diagonal = nullptr;
};

} // namespace spmv
