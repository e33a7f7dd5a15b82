// mpi_utils.h -- CHECK_MPI, the CommunicationModel enum and mpi_type<T>().
// API surface of the reference's spmv/mpi_utils.h:21-78.
#pragma once

#include <complex>
#include <cstdio>
#include <cstdlib>
#include <mpi.h>

namespace spmv
{

// Like the reference (mpi_utils.h:21-40): report and exit with the MPI status.
#define CHECK_MPI(func)                                                                            \
  do {                                                                                             \
    const int spmv_mpi_rc_ = (func);                                                               \
    if (spmv_mpi_rc_ != MPI_SUCCESS) {                                                             \
      char spmv_mpi_msg_[MPI_MAX_ERROR_STRING];                                                    \
      int spmv_mpi_len_ = 0;                                                                       \
      MPI_Error_string(spmv_mpi_rc_, spmv_mpi_msg_, &spmv_mpi_len_);                               \
      std::fprintf(stderr, "ERROR: MPI call \"%s\" at %s:%d: %s (%d)\n", #func, __FILE__, __LINE__, \
                   spmv_mpi_msg_, spmv_mpi_rc_);                                                   \
      std::exit(spmv_mpi_rc_);                                                                     \
    }                                                                                              \
  } while (0)

/// How L2GMap::update moves ghost values (mpi_utils.h:43-52).  On the B200 executor
/// the four in-scope models map onto one NCCL implementation: the *_blocking models
/// enqueue pack + exchange on the executor's stream, the *_nonblocking models run the
/// exchange on a side stream and update_finalise() joins it.  The one-sided and
/// shared-memory models are MPI-RMA specific and rejected.
enum class CommunicationModel {
  p2p_blocking,
  p2p_nonblocking,
  collective_blocking,
  collective_nonblocking,
  onesided_put_active,
  onesided_put_passive,
  shmem,
  shmem_nodup
};

template <typename T>
inline MPI_Datatype mpi_type();
template <>
inline MPI_Datatype mpi_type<float>()
{
  return MPI_FLOAT;
}
template <>
inline MPI_Datatype mpi_type<double>()
{
  return MPI_DOUBLE;
}
template <>
inline MPI_Datatype mpi_type<std::complex<float>>()
{
  return MPI_C_FLOAT_COMPLEX;
}
template <>
inline MPI_Datatype mpi_type<std::complex<double>>()
{
  return MPI_DOUBLE_COMPLEX;
}

} // namespace spmv
