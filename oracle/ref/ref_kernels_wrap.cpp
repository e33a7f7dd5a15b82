// oracle/ref/ref_kernels_wrap.cpp -- TEST INFRASTRUCTURE.
// C entry points around the reference's own CSRSpMV<T> objects, compiled from
// /root/reference/spmv/csr_kernels.cpp (sequential) and
// /root/reference/spmv/openmp/csr_kernels.openmp.cpp (threaded) where they lie.
// The executor argument of the sequential kernel is a reference to a type the
// kernel never touches (csr_kernels.cpp:11-57).
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <memory>

#include "csr_kernels.h"
#include "openmp/omp_executor.h"

namespace spmv
{
// Stand-in for the three OmpExecutor members the OpenMP kernel needs
// (openmp/omp_executor.cpp:22-34,41-52); the real file drags in Eigen.
// calloc instead of malloc: the kernel accumulates into its per-thread buffers
// without zeroing them (openmp/csr_kernels.openmp.cpp:94,204).
int OmpExecutor::get_num_cus() const
{
  const char* v = std::getenv("OMP_NUM_THREADS");
  return v ? std::atoi(v) : 1;
}
void* OmpExecutor::_alloc(size_t num_bytes) const { return std::calloc(1, num_bytes ? num_bytes : 1); }
void OmpExecutor::_free(void* ptr) const { std::free(ptr); }
} // namespace spmv

static const spmv::ReferenceExecutor& fake_ref_exec()
{
  static char storage[64];
  return *reinterpret_cast<const spmv::ReferenceExecutor*>(storage);
}

// The remaining OmpExecutor virtuals (openmp/omp_executor.cpp, not compiled here
// because it includes Eigen through csr_matrix.h) so the class can be
// instantiated through its own factory.  The kernel never calls them.
namespace spmv
{
void OmpExecutor::synchronize() const {}
const DeviceExecutor& OmpExecutor::get_host() const { return *this; }
int OmpExecutor::get_num_devices() const { return 1; }
void OmpExecutor::spmv_init(CSRSpMV<float>&, const CSRMatrix<float>&) const {}
void OmpExecutor::spmv_init(CSRSpMV<double>&, const CSRMatrix<double>&) const {}
void OmpExecutor::spmv_run(const CSRSpMV<float>&, const CSRMatrix<float>&, float, float* __restrict__,
                           float, float* __restrict__) const {}
void OmpExecutor::spmv_run(const CSRSpMV<double>&, const CSRMatrix<double>&, double,
                           double* __restrict__, double, double* __restrict__) const {}
void OmpExecutor::spmv_finalize(CSRSpMV<float>&) const {}
void OmpExecutor::spmv_finalize(CSRSpMV<double>&) const {}
void OmpExecutor::gather_ghosts_run(int n, const int32_t* idx, const float* in, float* out) const
{
  for (int i = 0; i < n; ++i) out[i] = in[idx[i]];
}
void OmpExecutor::gather_ghosts_run(int n, const int32_t* idx, const double* in, double* out) const
{
  for (int i = 0; i < n; ++i) out[i] = in[idx[i]];
}
void OmpExecutor::_memset(void* p, int v, size_t n) const { std::memset(p, v, n); }
void OmpExecutor::_copy(void* d, const void* s, size_t n) const { std::memcpy(d, s, n); }
void OmpExecutor::_copy_async(void* d, const void* s, size_t n, void*) const { std::memcpy(d, s, n); }
void OmpExecutor::_copy_from(void* d, const DeviceExecutor&, const void* s, size_t n) const { std::memcpy(d, s, n); }
void OmpExecutor::_copy_to(void* d, const DeviceExecutor&, const void* s, size_t n) const { std::memcpy(d, s, n); }
} // namespace spmv

#include <omp.h>

extern "C" {

// The reference sizes its row split from the OMP_NUM_THREADS environment
// variable (openmp/omp_executor.cpp:22-34) but runs `#pragma omp parallel` with
// whatever the runtime has; keep the two in step.
void ref_omp_set_threads(int n)
{
  char buf[32];
  snprintf(buf, sizeof(buf), "%d", n);
  setenv("OMP_NUM_THREADS", buf, 1);
  omp_set_num_threads(n);
}

// Sequential reference kernel, csr_kernels.cpp:20-52.
void ref_csr_run_f64(int32_t rows, int32_t cols, int64_t nnz, const int32_t* rowptr,
                     const int32_t* colind, const double* values, const double* diagonal,
                     int symmetric, double alpha, double* in, double beta, double* out)
{
  spmv::CSRSpMV<double> op;
  op.init(rows, cols, nnz, rowptr, colind, values, symmetric != 0, fake_ref_exec());
  op.run(rows, cols, nnz, rowptr, colind, values, diagonal, alpha, in, beta, out, fake_ref_exec());
  op.finalize(fake_ref_exec());
}

// OpenMP reference kernel, openmp/csr_kernels.openmp.cpp:27-262.  Handle-based so
// the analysis (row split, conflict map) is done once and `run` can be timed.
struct RefOmpOp {
  spmv::CSRSpMV<double> op;
  std::unique_ptr<spmv::OmpExecutor> exec;
};

void* ref_omp_create(int32_t rows, int32_t cols, int64_t nnz, const int32_t* rowptr,
                     const int32_t* colind, const double* values, int symmetric)
{
  RefOmpOp* h = new RefOmpOp;
  h->exec = spmv::OmpExecutor::create();
  h->op.init(rows, cols, nnz, rowptr, colind, values, symmetric != 0, *h->exec);
  return h;
}

void ref_omp_run(void* handle, int32_t rows, int32_t cols, int64_t nnz, const int32_t* rowptr,
                 const int32_t* colind, const double* values, const double* diagonal, double alpha,
                 double* in, double beta, double* out)
{
  RefOmpOp* h = static_cast<RefOmpOp*>(handle);
  h->op.run(rows, cols, nnz, rowptr, colind, values, diagonal, alpha, in, beta, out, *h->exec);
}

void ref_omp_destroy(void* handle)
{
  RefOmpOp* h = static_cast<RefOmpOp*>(handle);
  h->op.finalize(*h->exec);
  delete h;
}

} // extern "C"
