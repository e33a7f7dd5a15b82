// cg.h -- Conjugate Gradient on the distributed operator (reference: spmv/cg.h:33-35,
// cg.cpp:21-98; CUDA flavour cuda/cg_cuda.h:31-32, cuda/cg.cuda.cu:40-176).
#pragma once

#include "config.h"
#include "spmv_export.h"

#include <mpi.h>
#include <vector>

namespace spmv
{

template <typename T>
class Matrix;
class ReferenceExecutor;
class B200Executor;

/// Solve A x = b, x0 = 0.  b and x are executor (device) pointers of local_size
/// entries.  Stops when ||r||/||r0|| < rtol or after kmax iterations; returns the
/// number of iterations, exactly as cg.cpp:55-97 counts them.  The whole iteration runs
/// on the device (fused kernels, NCCL allreduce of the two scalars); the host only
/// polls a done flag a few iterations behind.
SPMV_EXPORT int cg(MPI_Comm comm, B200Executor& exec, const spmv::Matrix<double>& A, const double* b,
                   double* x, int kmax, double rtol);

/// Same, and also returns ||r_k|| for k = 1..iterations (the reference computes these
/// norms, cg.cpp:76, but never stores them).
SPMV_EXPORT int cg(MPI_Comm comm, B200Executor& exec, const spmv::Matrix<double>& A, const double* b,
                   double* x, int kmax, double rtol, std::vector<double>* residual_history);

/// Per-phase device time of one iteration, averaged over the iterations run (CUDA events, one iteration at
/// a time): halo update (pack + push), local-block SpMV, halo wait + remote-block SpMV, vector step with its
/// two scalar reductions.  What demos/cg.cpp's timers would show, split the way the kernels are.
struct CgPhaseTimes {
  int iterations = 0;
  double us_update = 0, us_local = 0, us_remote = 0, us_vector = 0;
};
/// Same solve in profile mode (no graph, a host synchronisation per iteration): fills `phases`.
SPMV_EXPORT int cg(MPI_Comm comm, B200Executor& exec, const spmv::Matrix<double>& A, const double* b,
                   double* x, int kmax, double rtol, std::vector<double>* residual_history, CgPhaseTimes* phases);

/// The host executor has no compute path: throws std::runtime_error.
SPMV_EXPORT int cg(MPI_Comm comm, ReferenceExecutor& exec, const spmv::Matrix<double>& A, const double* b,
                   double* x, int kmax, double rtol);

} // namespace spmv
