// device_executor.h -- the plugin interface (reference: spmv/device_executor.h:29-151).
//
// A backend subclasses DeviceExecutor; Matrix, L2GMap, CSRMatrix and the builders
// reach memory and kernels only through it.  Names, argument meaning and const-ness
// follow the reference so code written against it compiles unchanged.
#pragma once

#include "config.h"
#include "spmv_export.h"

#include <cstddef>
#include <cstdint>
#include <memory>

namespace spmv
{

template <typename T>
class CSRMatrix;
template <typename T>
class CSRSpMV;

enum class DeviceType { undefined, cpu, gpu };

class SPMV_EXPORT DeviceExecutor
{
public:
  virtual ~DeviceExecutor() = default;

  // ---- typed memory helpers over the byte-level virtuals ----
  template <typename T>
  T* alloc(size_t num_elems) const
  {
    return static_cast<T*>(_alloc(num_elems * sizeof(T)));
  }
  void free(void* ptr) const { _free(ptr); }
  template <typename T>
  void memset(T* ptr, int value, size_t num_elems) const
  {
    _memset(ptr, value, num_elems * sizeof(T));
  }
  template <typename T>
  void copy(T* dst, const T* src, size_t num_elems) const
  {
    _copy(dst, src, num_elems * sizeof(T));
  }
  template <typename T>
  void copy_async(T* dst, const T* src, size_t num_elems, void* stream) const
  {
    _copy_async(dst, src, num_elems * sizeof(T), stream);
  }
  /// dst lives on *this, src on src_exec
  template <typename T>
  void copy_from(T* dst, const DeviceExecutor& src_exec, const T* src, size_t num_elems) const
  {
    _copy_from(dst, src_exec, src, num_elems * sizeof(T));
  }
  /// src lives on *this, dst on dst_exec
  template <typename T>
  void copy_to(T* dst, const DeviceExecutor& dst_exec, const T* src, size_t num_elems) const
  {
    _copy_to(dst, dst_exec, src, num_elems * sizeof(T));
  }

  virtual void synchronize() const = 0;
  virtual const DeviceExecutor& get_host() const = 0;
  virtual int get_num_devices() const = 0;
  virtual int get_num_cus() const = 0;

  // ---- CSR SpMV visitors (device_executor.h:88-99) ----
  virtual void spmv_init(CSRSpMV<float>& op, const CSRMatrix<float>& mat) const = 0;
  virtual void spmv_init(CSRSpMV<double>& op, const CSRMatrix<double>& mat) const = 0;
  virtual void spmv_run(const CSRSpMV<float>& op, const CSRMatrix<float>& mat, float alpha,
                        float* __restrict__ in, float beta, float* __restrict__ out) const = 0;
  virtual void spmv_run(const CSRSpMV<double>& op, const CSRMatrix<double>& mat, double alpha,
                        double* __restrict__ in, double beta, double* __restrict__ out) const = 0;
  virtual void spmv_finalize(CSRSpMV<float>& op) const = 0;
  virtual void spmv_finalize(CSRSpMV<double>& op) const = 0;

  // ---- halo pack: out[k] = in[indices[k]] (device_executor.h:123-126) ----
  virtual void gather_ghosts_run(int num_indices, const int32_t* indices, const float* in,
                                 float* out) const = 0;
  virtual void gather_ghosts_run(int num_indices, const int32_t* indices, const double* in,
                                 double* out) const = 0;

  struct DeviceInfo {
    DeviceType type = DeviceType::undefined;
    int id = -1;
  };
  const DeviceInfo& get_device_info() const { return _dev_info; }
  DeviceType get_device_type() const { return _dev_info.type; }
  int get_device_id() const { return _dev_info.id; }

protected:
  virtual void* _alloc(size_t num_bytes) const = 0;
  virtual void _free(void* ptr) const = 0;
  virtual void _memset(void* ptr, int value, size_t num_bytes) const = 0;
  virtual void _copy(void* dst, const void* src, size_t num_bytes) const = 0;
  virtual void _copy_async(void* dst, const void* src, size_t num_bytes, void* stream) const = 0;
  virtual void _copy_from(void* dst, const DeviceExecutor& src_exec, const void* src,
                          size_t num_bytes) const = 0;
  virtual void _copy_to(void* dst, const DeviceExecutor& dst_exec, const void* src,
                        size_t num_bytes) const = 0;

  DeviceInfo _dev_info;
};

} // namespace spmv
