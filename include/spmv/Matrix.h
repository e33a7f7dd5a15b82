// Matrix.h -- the distributed operator (reference: spmv/Matrix.h:33-140).
//
// Owns one or two CSR blocks in executor memory plus the column / row L2GMaps, and
// runs y = A x with one of four schedules (Matrix.cpp:131-141,483-552):
//   spmv              one block over [owned | ghost] columns
//   spmv_overlap      local block, update_finalise, remote block (beta = 1)
//   spmv_sym          symmetric local block (strictly lower + diagonal), remote block
//   spmv_sym_overlap  same with update_finalise in between
// The caller runs col_map()->update(x) first, as with the reference.
//
// Eigen is not available on the target, so host blocks travel as HostCSR<T> instead of
// Eigen::SparseMatrix<T, RowMajor>; the raw-pointer create_matrix overload is the
// reference's own (Matrix.h:106-111).
#pragma once

#include "config.h"
#include "spmv_export.h"

#include "mpi_utils.h"

#include <cstdint>
#include <memory>
#include <mpi.h>
#include <vector>

// Where Eigen is installed the reference's own signatures (Matrix.h:40,47-48,55-56,96: Eigen::SparseMatrix<T,
// RowMajor> blocks) are available too, so code written against the reference compiles unchanged; they convert to
// HostCSR and delegate.  SPMV_NO_EIGEN switches them off.
#if !defined(SPMV_NO_EIGEN) && defined(__has_include)
#if __has_include(<Eigen/Sparse>)
#include <Eigen/Sparse>
#define SPMV_HAVE_EIGEN 1
#endif
#endif

namespace spmv
{

template <typename T>
class SubMatrix;
class DeviceExecutor;
class L2GMap;

/// Host-side row-major compressed block: what Eigen::SparseMatrix<T, RowMajor> is to
/// the reference's constructors (outerIndexPtr / innerIndexPtr / valuePtr).
template <typename T>
struct HostCSR {
  std::int32_t rows = 0;
  std::int32_t cols = 0;
  std::vector<std::int32_t> rowptr; // rows+1 entries (or empty when rows == 0)
  std::vector<std::int32_t> colind;
  std::vector<T> values;
  std::int64_t non_zeros() const { return static_cast<std::int64_t>(values.size()); }
};

#ifdef SPMV_HAVE_EIGEN
/// Compressed row-major Eigen block -> HostCSR (outerIndexPtr / innerIndexPtr / valuePtr, csr_matrix.cpp:23-59).
template <typename T>
HostCSR<T> to_host_csr(const Eigen::SparseMatrix<T, Eigen::RowMajor>& m)
{
  HostCSR<T> h;
  h.rows = static_cast<std::int32_t>(m.rows());
  h.cols = static_cast<std::int32_t>(m.cols());
  const auto nnz = static_cast<std::size_t>(m.nonZeros());
  if (h.rows > 0)
    h.rowptr.assign(m.outerIndexPtr(), m.outerIndexPtr() + h.rows + 1);
  h.colind.assign(m.innerIndexPtr(), m.innerIndexPtr() + nnz);
  h.values.assign(m.valuePtr(), m.valuePtr() + nnz);
  return h;
}
#endif

template <typename T>
class SPMV_EXPORT Matrix
{
public:
#ifdef SPMV_HAVE_EIGEN
  using EigenBlock = Eigen::SparseMatrix<T, Eigen::RowMajor>;
  /// The reference's three constructors (Matrix.h:40,47-48,55-56); blocks must be compressed.
  Matrix(const EigenBlock& mat, std::shared_ptr<spmv::L2GMap> col_map, std::shared_ptr<spmv::L2GMap> row_map,
         std::shared_ptr<DeviceExecutor> exec)
      : Matrix(to_host_csr(mat), col_map, row_map, exec)
  {
  }
  Matrix(const EigenBlock& mat_local, const EigenBlock& mat_remote, std::shared_ptr<spmv::L2GMap> col_map,
         std::shared_ptr<spmv::L2GMap> row_map, std::shared_ptr<DeviceExecutor> exec)
      : Matrix(to_host_csr(mat_local), to_host_csr(mat_remote), col_map, row_map, exec)
  {
  }
  Matrix(const EigenBlock& mat_local, const EigenBlock& mat_remote,
         const Eigen::Matrix<T, Eigen::Dynamic, 1>& mat_diagonal, std::shared_ptr<spmv::L2GMap> col_map,
         std::shared_ptr<spmv::L2GMap> row_map, int nnz_full, std::shared_ptr<DeviceExecutor> exec)
      : Matrix(to_host_csr(mat_local), to_host_csr(mat_remote),
               std::vector<T>(mat_diagonal.data(), mat_diagonal.data() + mat_diagonal.size()), col_map, row_map, nnz_full,
               exec)
  {
  }
  /// Matrix.h:96: create_matrix from an Eigen block.
  static Matrix<T>* create_matrix(MPI_Comm comm, std::shared_ptr<DeviceExecutor> exec, const EigenBlock& mat,
                                  std::int64_t nrows_local, std::int64_t ncols_local,
                                  std::vector<std::int64_t> row_ghosts, std::vector<std::int64_t> col_ghosts,
                                  bool symmetric = false,
                                  CommunicationModel cm = CommunicationModel::collective_blocking)
  {
    return create_matrix(comm, exec, to_host_csr(mat), nrows_local, ncols_local, std::move(row_ghosts),
                         std::move(col_ghosts), symmetric, cm);
  }

#endif
  /// Single block; throws if the column map is overlapping (Matrix.cpp:28-29).
  Matrix(const HostCSR<T>& mat, std::shared_ptr<spmv::L2GMap> col_map,
         std::shared_ptr<spmv::L2GMap> row_map, std::shared_ptr<DeviceExecutor> exec);
  /// Local + remote blocks; throws unless the column map is overlapping (Matrix.cpp:44-45).
  Matrix(const HostCSR<T>& mat_local, const HostCSR<T>& mat_remote,
         std::shared_ptr<spmv::L2GMap> col_map, std::shared_ptr<spmv::L2GMap> row_map,
         std::shared_ptr<DeviceExecutor> exec);
  /// Symmetric: strictly-lower local block + diagonal + remote block (Matrix.cpp:53-72).
  Matrix(const HostCSR<T>& mat_local, const HostCSR<T>& mat_remote, const std::vector<T>& mat_diagonal,
         std::shared_ptr<spmv::L2GMap> col_map, std::shared_ptr<spmv::L2GMap> row_map, int nnz_full,
         std::shared_ptr<DeviceExecutor> exec);
  /// Device-side builders: wrap blocks that already live in executor
  /// memory (e.g. generated or split on the GPU) without a host round trip.  `remote`
  /// may be null for a single-block operator; `nnz` is what non_zeros() reports for
  /// symmetric storage (Matrix.cpp:56-59).  Same contract checks as the three
  /// host constructors.
  static Matrix<T> from_blocks(std::unique_ptr<SubMatrix<T>> local, std::unique_ptr<SubMatrix<T>> remote,
                               std::shared_ptr<spmv::L2GMap> col_map, std::shared_ptr<spmv::L2GMap> row_map,
                               std::int64_t nnz, bool symmetric, std::shared_ptr<DeviceExecutor> exec);
  ~Matrix();
  Matrix(Matrix&&) noexcept;
  Matrix& operator=(Matrix&&) noexcept;

  int rows() const;
  int cols() const;
  int non_zeros() const;
  bool symmetric() const { return _symmetric; }
  size_t format_size() const;

  /// y = A x; x has local_size()+num_ghosts() entries with ghosts updated.
  void mult(T* __restrict__ x, T* __restrict__ y) const;
  /// y_host = A x_host with both vectors in HOST memory (x_host: local_size() owned entries;
  /// pinned memory lets the copies overlap).  What the reference's user writes as
  /// copy_from + update + mult + copy_to (demos/spmv_cuda.cpp:71-110), pipelined: pieces of
  /// x go up, tile ranges run as their columns land, finished pieces of y come down, all
  /// at once (mult_host.cpp).  Blocks until y_host is complete.
  void mult_host(const T* x_host, T* y_host) const;
  /// A^T x: declared by the reference and unimplemented there (Matrix.cpp:143-148);
  /// throws std::runtime_error here too.
  void transpmult(const T* x, T* y) const;

  std::shared_ptr<L2GMap> row_map() const { return _row_map; }
  std::shared_ptr<const L2GMap> col_map() const { return _col_map; }
  std::shared_ptr<DeviceExecutor> executor() const { return _exec; }
  const SubMatrix<T>* local_block() const { return _mat_local.get(); }
  const SubMatrix<T>* remote_block() const { return _mat_remote.get(); }

  /// Build from a rank's CSR rows (local column ids; ghost columns numbered
  /// ncols_local + position in col_ghosts) such that the result has no row ghosts:
  /// ghost rows are shipped to their owners and merged (Matrix.cpp:164-480).
  /// Returns a new Matrix the caller deletes, like the reference.
  static Matrix<T>* create_matrix(MPI_Comm comm, std::shared_ptr<DeviceExecutor> exec,
                                  const std::int32_t* rowptr, const std::int32_t* colind,
                                  const T* values, std::int64_t nrows_local, std::int64_t ncols_local,
                                  std::vector<std::int64_t> row_ghosts, std::vector<std::int64_t> col_ghosts,
                                  bool symmetric = false,
                                  CommunicationModel cm = CommunicationModel::collective_blocking);
  static Matrix<T>* create_matrix(MPI_Comm comm, std::shared_ptr<DeviceExecutor> exec,
                                  const HostCSR<T>& mat, std::int64_t nrows_local,
                                  std::int64_t ncols_local, std::vector<std::int64_t> row_ghosts,
                                  std::vector<std::int64_t> col_ghosts, bool symmetric = false,
                                  CommunicationModel cm = CommunicationModel::collective_blocking);

private:
  Matrix() = default;
  void spmv(T* x, T* y) const;
  void spmv_overlap(T* x, T* y) const;
  void spmv_sym(T* x, T* y) const;
  void spmv_sym_overlap(T* x, T* y) const;

  std::shared_ptr<DeviceExecutor> _exec;
  std::unique_ptr<SubMatrix<T>> _mat_local;
  std::unique_ptr<SubMatrix<T>> _mat_remote;
  std::shared_ptr<spmv::L2GMap> _col_map;
  std::shared_ptr<spmv::L2GMap> _row_map;
  int _nnz = 0;
  bool _symmetric = false;
  mutable std::shared_ptr<void> _host_pipe; // device vectors, streams and chunk table of mult_host
};

} // namespace spmv
