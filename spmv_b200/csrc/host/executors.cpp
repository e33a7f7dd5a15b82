// executors.cpp -- ReferenceExecutor (host memory only), B200Executor, CSRSpMV<T>
// and CSRMatrix<T>.  Reference counterparts: reference_executor.cpp:20-164,
// cuda/cuda_executor.cpp:15-205, cuda/csr_kernels.cuda.cu:50-163, csr_matrix.cpp:23-87.
#include "b200_executor.h"
#include "csr_matrix.h"
#include "reference_executor.h"

#include "b200spmv.h"
#include "host_util.h"
#include "mpi_utils.h"

#include <cstring>
#include <stdexcept>
#include <string>
#include <typeinfo>

namespace spmv
{

// ---------------------------------------------------------------------------
// ReferenceExecutor: memory only
// ---------------------------------------------------------------------------
static const char* const kNoCpu
    = "spmv: the host executor has no compute path in this build (B200-only); use B200Executor";

void* ReferenceExecutor::_alloc(size_t num_bytes) const { return std::malloc(num_bytes ? num_bytes : 1); }
void ReferenceExecutor::_free(void* ptr) const { std::free(ptr); }
void ReferenceExecutor::_memset(void* ptr, int value, size_t num_bytes) const
{
  if (num_bytes > 0)
    std::memset(ptr, value, num_bytes);
}
void ReferenceExecutor::_copy(void* dst, const void* src, size_t num_bytes) const
{
  if (num_bytes > 0 && dst != src)
    std::memmove(dst, src, num_bytes);
}
void ReferenceExecutor::_copy_async(void* dst, const void* src, size_t num_bytes, void*) const
{
  _copy(dst, src, num_bytes);
}
void ReferenceExecutor::_copy_from(void* dst, const DeviceExecutor& src_exec, const void* src,
                                   size_t num_bytes) const
{
  if (src_exec.get_device_type() == DeviceType::cpu)
    _copy(dst, src, num_bytes);
  else
    src_exec.copy_to(static_cast<char*>(dst), *this, static_cast<const char*>(src), num_bytes);
}
void ReferenceExecutor::_copy_to(void* dst, const DeviceExecutor& dst_exec, const void* src,
                                 size_t num_bytes) const
{
  if (dst_exec.get_device_type() == DeviceType::cpu)
    _copy(dst, src, num_bytes);
  else
    dst_exec.copy_from(static_cast<char*>(dst), *this, static_cast<const char*>(src), num_bytes);
}
void ReferenceExecutor::spmv_init(CSRSpMV<float>& op, const CSRMatrix<float>& mat) const
{
  op.init(mat.rows(), mat.cols(), mat.non_zeros(), mat.rowptr(), mat.colind(), mat.values(),
          mat.symmetric(), *this);
}
void ReferenceExecutor::spmv_init(CSRSpMV<double>& op, const CSRMatrix<double>& mat) const
{
  op.init(mat.rows(), mat.cols(), mat.non_zeros(), mat.rowptr(), mat.colind(), mat.values(),
          mat.symmetric(), *this);
}
void ReferenceExecutor::spmv_run(const CSRSpMV<float>&, const CSRMatrix<float>&, float, float* __restrict__,
                                 float, float* __restrict__) const
{
  throw std::runtime_error(kNoCpu);
}
void ReferenceExecutor::spmv_run(const CSRSpMV<double>&, const CSRMatrix<double>&, double,
                                 double* __restrict__, double, double* __restrict__) const
{
  throw std::runtime_error(kNoCpu);
}
void ReferenceExecutor::spmv_finalize(CSRSpMV<float>& op) const { op.finalize(*this); }
void ReferenceExecutor::spmv_finalize(CSRSpMV<double>& op) const { op.finalize(*this); }
void ReferenceExecutor::gather_ghosts_run(int, const int32_t*, const float*, float*) const
{
  throw std::runtime_error(kNoCpu);
}
void ReferenceExecutor::gather_ghosts_run(int, const int32_t*, const double*, double*) const
{
  throw std::runtime_error(kNoCpu);
}

// ---------------------------------------------------------------------------
// B200Executor
// ---------------------------------------------------------------------------
B200Executor::B200Executor(int device_id, std::shared_ptr<DeviceExecutor> host) : _host(std::move(host))
{
  if (!_host)
    throw std::runtime_error("B200Executor: a host executor is required");
  _dev_info.type = DeviceType::gpu;
  _dev_info.id = device_id;
  detail::check(b200_set_device(device_id), "B200Executor: cannot select the device");
  detail::check(b200_stream_create_priority(&_comm_stream, 1), "B200Executor: comm stream");
  detail::check(b200_event_create(&_ev_packed, 0), "B200Executor: event");
  detail::check(b200_event_create(&_ev_arrived, 0), "B200Executor: event");
}

B200Executor::~B200Executor()
{
  b200_set_device(_dev_info.id);
  _cg_workspace.reset(); // device memory, stream, graph: before the communicator and the streams go
  if (_nccl)
    b200_nccl_comm_destroy(_nccl);
  b200_event_destroy(_ev_packed);
  b200_event_destroy(_ev_arrived);
  b200_stream_destroy(_comm_stream);
}

void B200Executor::bind() const { detail::check(b200_set_device(_dev_info.id), "cudaSetDevice"); }

void B200Executor::synchronize() const
{
  bind();
  detail::check(b200_stream_synchronize(_comm_stream), "synchronize (comm stream)");
  detail::check(b200_stream_synchronize(_stream), "synchronize");
}

int B200Executor::get_num_devices() const
{
  int n = 0;
  detail::check(b200_device_count(&n), "device count");
  return n;
}

int B200Executor::get_num_cus() const
{
  bind();
  int n = 0;
  detail::check(b200_get_num_sms(&n), "SM count");
  return n;
}

void* B200Executor::_alloc(size_t num_bytes) const
{
  bind();
  void* p = nullptr;
  detail::check(b200_malloc(&p, num_bytes ? num_bytes : 1), "device alloc");
  return p;
}
void B200Executor::_free(void* ptr) const
{
  if (ptr == nullptr)
    return;
  bind();
  detail::check(b200_free(ptr), "device free");
}
void B200Executor::_memset(void* ptr, int value, size_t num_bytes) const
{
  bind();
  detail::check(b200_memset_async(ptr, value, num_bytes, _stream), "memset");
}
void B200Executor::_copy(void* dst, const void* src, size_t num_bytes) const
{
  bind();
  detail::check(b200_memcpy_d2d_async(dst, src, num_bytes, _stream), "device copy");
}
void B200Executor::_copy_async(void* dst, const void* src, size_t num_bytes, void* stream) const
{
  bind();
  detail::check(b200_memcpy_d2d_async(dst, src, num_bytes, stream), "device copy");
}
void B200Executor::_copy_from(void* dst, const DeviceExecutor& src_exec, const void* src,
                              size_t num_bytes) const
{
  bind();
  if (src_exec.get_device_type() == DeviceType::cpu) {
    // the host buffer may be a temporary: complete before returning
    detail::check(b200_memcpy_h2d_async(dst, src, num_bytes, _stream), "host-to-device copy");
    detail::check(b200_stream_synchronize(_stream), "host-to-device copy");
  } else {
    detail::check(b200_memcpy_d2d_async(dst, src, num_bytes, _stream), "device-to-device copy");
  }
}
void B200Executor::_copy_to(void* dst, const DeviceExecutor& dst_exec, const void* src,
                            size_t num_bytes) const
{
  bind();
  if (dst_exec.get_device_type() == DeviceType::cpu) {
    detail::check(b200_memcpy_d2h_async(dst, src, num_bytes, _stream), "device-to-host copy");
    detail::check(b200_stream_synchronize(_stream), "device-to-host copy");
  } else {
    detail::check(b200_memcpy_d2d_async(dst, src, num_bytes, _stream), "device-to-device copy");
  }
}

void B200Executor::spmv_init(CSRSpMV<float>& op, const CSRMatrix<float>& mat) const
{
  op.init(mat.rows(), mat.cols(), mat.non_zeros(), mat.rowptr(), mat.colind(), mat.values(),
          mat.symmetric(), *this);
}
void B200Executor::spmv_init(CSRSpMV<double>& op, const CSRMatrix<double>& mat) const
{
  op.init(mat.rows(), mat.cols(), mat.non_zeros(), mat.rowptr(), mat.colind(), mat.values(),
          mat.symmetric(), *this);
}
void B200Executor::spmv_run(const CSRSpMV<float>& op, const CSRMatrix<float>& mat, float alpha,
                            float* __restrict__ in, float beta, float* __restrict__ out) const
{
  op.run(mat.rows(), mat.cols(), mat.non_zeros(), mat.rowptr(), mat.colind(), mat.values(),
         mat.diagonal(), alpha, in, beta, out, *this);
}
void B200Executor::spmv_run(const CSRSpMV<double>& op, const CSRMatrix<double>& mat, double alpha,
                            double* __restrict__ in, double beta, double* __restrict__ out) const
{
  op.run(mat.rows(), mat.cols(), mat.non_zeros(), mat.rowptr(), mat.colind(), mat.values(),
         mat.diagonal(), alpha, in, beta, out, *this);
}
void B200Executor::spmv_finalize(CSRSpMV<float>& op) const { op.finalize(*this); }
void B200Executor::spmv_finalize(CSRSpMV<double>& op) const { op.finalize(*this); }

void B200Executor::gather_ghosts_run(int n, const int32_t* indices, const float* in, float* out) const
{
  bind();
  detail::check(b200_gather_f32(n, indices, in, out, _stream), "gather_ghosts_run");
}
void B200Executor::gather_ghosts_run(int n, const int32_t* indices, const double* in, double* out) const
{
  bind();
  detail::check(b200_gather_f64(n, indices, in, out, _stream), "gather_ghosts_run");
}
void B200Executor::scatter_add_run(int n, const int32_t* indices, const float* in, float* inout) const
{
  bind();
  detail::check(b200_scatter_add_f32(n, indices, in, inout, _stream), "scatter_add_run");
}
void B200Executor::scatter_add_run(int n, const int32_t* indices, const double* in, double* inout) const
{
  bind();
  detail::check(b200_scatter_add_f64(n, indices, in, inout, _stream), "scatter_add_run");
}

void B200Executor::attach_comm(MPI_Comm comm) const
{
  int size = 1, rank = 0;
  CHECK_MPI(MPI_Comm_size(comm, &size));
  CHECK_MPI(MPI_Comm_rank(comm, &rank));
  if (size == 1 || _nccl != nullptr) {
    if (_nccl != nullptr && size != _comm_size)
      throw std::runtime_error("B200Executor: already attached to a communicator of another size");
    return;
  }
  bind();
  char id[B200_NCCL_UNIQUE_ID_BYTES] = {0};
  int rc = 0;
  if (rank == 0)
    rc = b200_nccl_unique_id(id);
  CHECK_MPI(MPI_Bcast(&rc, 1, MPI_INT, 0, comm));
  detail::check(rc, "ncclGetUniqueId");
  CHECK_MPI(MPI_Bcast(id, B200_NCCL_UNIQUE_ID_BYTES, MPI_BYTE, 0, comm));
  detail::check(b200_nccl_comm_create(&_nccl, size, rank, id), "ncclCommInitRank");
  _comm_size = size;
}

// ---------------------------------------------------------------------------
// CSRSpMV<T>
// ---------------------------------------------------------------------------
template <typename T>
void CSRSpMV<T>::init(int32_t, int32_t, int64_t, const int32_t*, const int32_t*, const T*, bool symmetric,
                      const ReferenceExecutor&)
{
  _symmetric = symmetric;
}
template <typename T>
void CSRSpMV<T>::run(int32_t, int32_t, int64_t, const int32_t*, const int32_t*, const T*, const T*, T,
                     T* __restrict__, T, T* __restrict__, const ReferenceExecutor&) const
{
  throw std::runtime_error(kNoCpu);
}
template <typename T>
void CSRSpMV<T>::finalize(const ReferenceExecutor&) const
{
}

template <typename T>
void CSRSpMV<T>::init(int32_t num_rows, int32_t num_cols, int64_t num_non_zeros, const int32_t* rowptr,
                      const int32_t*, const T*, bool symmetric, const B200Executor& exec)
{
  _symmetric = symmetric;
  detail::check(b200_set_device(exec.get_device_id()), "cudaSetDevice");
  detail::check(b200_csr_plan_create(&_aux_data, num_rows, num_cols, num_non_zeros,
                                     num_non_zeros > 0 ? rowptr : nullptr, symmetric ? 1 : 0,
                                     exec.get_cuda_stream()),
                "CSRSpMV::init");
}

static int run_general(void* plan, const int32_t* rp, const int32_t* ci, const double* v, double a,
                       const double* x, double b, double* y, void* s)
{
  return b200_csr_spmv_f64(plan, rp, ci, v, a, x, b, y, s);
}
static int run_general(void* plan, const int32_t* rp, const int32_t* ci, const float* v, float a,
                       const float* x, float b, float* y, void* s)
{
  return b200_csr_spmv_f32(plan, rp, ci, v, a, x, b, y, s);
}
static int run_symmetric(void* plan, const int32_t* rp, const int32_t* ci, const double* v, const double* d,
                         double a, const double* x, double b, double* y, void* s)
{
  return b200_csr_symv_f64(plan, rp, ci, v, d, a, x, b, y, s);
}
static int run_symmetric(void* plan, const int32_t* rp, const int32_t* ci, const float* v, const float* d,
                         float a, const float* x, float b, float* y, void* s)
{
  return b200_csr_symv_f32(plan, rp, ci, v, d, a, x, b, y, s);
}

template <typename T>
void CSRSpMV<T>::run(int32_t, int32_t, int64_t, const int32_t* rowptr, const int32_t* colind,
                     const T* values, const T* diagonal, T alpha, T* __restrict__ in, T beta,
                     T* __restrict__ out, const B200Executor& exec) const
{
  detail::check(b200_set_device(exec.get_device_id()), "cudaSetDevice");
  const int rc = _symmetric ? run_symmetric(_aux_data, rowptr, colind, values, diagonal, alpha, in, beta,
                                            out, exec.get_cuda_stream())
                            : run_general(_aux_data, rowptr, colind, values, alpha, in, beta, out,
                                          exec.get_cuda_stream());
  detail::check(rc, "CSRSpMV::run");
}

template <typename T>
void CSRSpMV<T>::finalize(const B200Executor& exec) const
{
  if (_aux_data) {
    b200_set_device(exec.get_device_id());
    b200_csr_plan_destroy(_aux_data);
    _aux_data = nullptr;
  }
}

template class CSRSpMV<float>;
template class CSRSpMV<double>;

// ---------------------------------------------------------------------------
// CSRMatrix<T>
// ---------------------------------------------------------------------------
template <typename T>
CSRMatrix<T>::CSRMatrix(std::shared_ptr<DeviceExecutor> exec, int32_t num_rows, int32_t num_cols,
                        int64_t num_non_zeros, const int32_t* rowptr, const int32_t* colind,
                        const T* values, const T* diagonal, bool symmetric)
{
  this->_exec = exec;
  this->_num_rows = num_rows;
  this->_num_cols = num_cols;
  this->_num_non_zeros = num_non_zeros;
  this->_symmetric = symmetric;
  const DeviceExecutor& host = exec->get_host();
  if (num_non_zeros > 0) {
    _rowptr = exec->template alloc<int32_t>(static_cast<size_t>(num_rows) + 1);
    _colind = exec->template alloc<int32_t>(num_non_zeros);
    _values = exec->template alloc<T>(num_non_zeros);
    exec->template copy_from<int32_t>(_rowptr, host, rowptr, static_cast<size_t>(num_rows) + 1);
    exec->template copy_from<int32_t>(_colind, host, colind, num_non_zeros);
    exec->template copy_from<T>(_values, host, values, num_non_zeros);
  }
  if (diagonal != nullptr) {
    this->_diagonal = exec->template alloc<T>(num_rows > 0 ? num_rows : 1);
    exec->template copy_from<T>(this->_diagonal, host, diagonal, num_rows);
  }
  exec->spmv_init(_op, *this);
}

template <typename T>
CSRMatrix<T>::CSRMatrix(adopt_t, std::shared_ptr<DeviceExecutor> exec, int32_t num_rows, int32_t num_cols,
                        int64_t num_non_zeros, int32_t* d_rowptr, int32_t* d_colind, T* d_values,
                        T* d_diagonal, bool symmetric)
{
  if (num_non_zeros == 0 && (d_rowptr || d_colind || d_values))
    throw std::runtime_error("CSRMatrix: an empty block carries no arrays (csr_matrix.cpp:35)");
  if (num_non_zeros > 0 && !(d_rowptr && d_colind && d_values))
    throw std::runtime_error("CSRMatrix: null array in a non-empty block");
  this->_exec = exec;
  this->_num_rows = num_rows;
  this->_num_cols = num_cols;
  this->_num_non_zeros = num_non_zeros;
  this->_symmetric = symmetric;
  _rowptr = d_rowptr;
  _colind = d_colind;
  _values = d_values;
  this->_diagonal = d_diagonal;
  exec->spmv_init(_op, *this);
}

template <typename T>
CSRMatrix<T>::~CSRMatrix()
{
  this->_exec->spmv_finalize(_op);
  this->_exec->free(_rowptr);
  this->_exec->free(_colind);
  this->_exec->free(_values);
  this->_exec->free(this->_diagonal);
}

template <typename T>
size_t CSRMatrix<T>::format_size() const
{
  return (static_cast<size_t>(this->_num_rows) + 1) * sizeof(int32_t)
         + static_cast<size_t>(this->_num_non_zeros) * (sizeof(int32_t) + sizeof(T));
}

template <typename T>
void CSRMatrix<T>::mult(T alpha, T* __restrict__ in, T beta, T* __restrict__ out) const
{
  if (this->_num_non_zeros > 0 || this->_diagonal != nullptr)
    this->_exec->spmv_run(_op, *this, alpha, in, beta, out);
}

template class CSRMatrix<float>;
template class CSRMatrix<double>;

} // namespace spmv
