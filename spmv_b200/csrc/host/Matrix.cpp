// Matrix.cpp -- the distributed operator: constructors, the four mult schedules.
// Reference: spmv/Matrix.cpp:21-148,483-552.  create_matrix lives in
// create_matrix.cpp.
#include "Matrix.h"

#include "L2GMap.h"
#include "csr_matrix.h"
#include "device_executor.h"
#include "host_util.h"

#include <atomic>
#include <stdexcept>

using namespace spmv;

namespace
{
template <typename T>
CSRMatrix<T>* make_block(std::shared_ptr<DeviceExecutor> exec, const HostCSR<T>& m, const T* diagonal = nullptr,
                         bool symmetric = false)
{
  if (m.non_zeros() > 0 && static_cast<std::int64_t>(m.rowptr.size()) != static_cast<std::int64_t>(m.rows) + 1)
    throw std::runtime_error("Matrix: rowptr must have rows+1 entries");
  return new CSRMatrix<T>(exec, m.rows, m.cols, m.non_zeros(), m.rowptr.data(), m.colind.data(),
                          m.values.data(), diagonal, symmetric);
}
} // namespace

//-----------------------------------------------------------------------------
template <typename T>
Matrix<T>::Matrix(const HostCSR<T>& mat, std::shared_ptr<spmv::L2GMap> col_map,
                  std::shared_ptr<spmv::L2GMap> row_map, std::shared_ptr<DeviceExecutor> exec)
    : _exec(exec), _col_map(col_map), _row_map(row_map), _nnz(static_cast<int>(mat.non_zeros()))
{
  if (col_map->overlapping())
    throw std::runtime_error("Ovelapping not supported in this format!");
  _mat_local.reset(make_block<T>(exec, mat));
}
//-----------------------------------------------------------------------------
template <typename T>
Matrix<T>::Matrix(const HostCSR<T>& mat_local, const HostCSR<T>& mat_remote,
                  std::shared_ptr<spmv::L2GMap> col_map, std::shared_ptr<spmv::L2GMap> row_map,
                  std::shared_ptr<DeviceExecutor> exec)
    : _exec(exec), _col_map(col_map), _row_map(row_map),
      _nnz(static_cast<int>(mat_local.non_zeros() + mat_remote.non_zeros()))
{
  if (!col_map->overlapping())
    throw std::runtime_error("Ovelapping not enabled in column mapping!");
  _mat_local.reset(make_block<T>(exec, mat_local));
  _mat_remote.reset(make_block<T>(exec, mat_remote));
}
//-----------------------------------------------------------------------------
template <typename T>
Matrix<T>::Matrix(const HostCSR<T>& mat_local, const HostCSR<T>& mat_remote,
                  const std::vector<T>& mat_diagonal, std::shared_ptr<spmv::L2GMap> col_map,
                  std::shared_ptr<spmv::L2GMap> row_map, int nnz_full, std::shared_ptr<DeviceExecutor> exec)
    : _exec(exec), _col_map(col_map), _row_map(row_map), _nnz(nnz_full), _symmetric(true)
{
  if (exec->get_device_type() == DeviceType::undefined)
    throw std::runtime_error("Device type not set!");
  if (static_cast<std::int64_t>(mat_diagonal.size()) != mat_local.rows)
    throw std::runtime_error("Matrix: diagonal must have one entry per local row");
  _mat_local.reset(make_block<T>(exec, mat_local, mat_diagonal.data(), true));
  _mat_remote.reset(make_block<T>(exec, mat_remote));
}
//-----------------------------------------------------------------------------
template <typename T>
Matrix<T> Matrix<T>::from_blocks(std::unique_ptr<SubMatrix<T>> local, std::unique_ptr<SubMatrix<T>> remote,
                                 std::shared_ptr<spmv::L2GMap> col_map, std::shared_ptr<spmv::L2GMap> row_map,
                                 std::int64_t nnz, bool symmetric, std::shared_ptr<DeviceExecutor> exec)
{
  if (!local)
    throw std::runtime_error("Matrix::from_blocks: the local block is required");
  if (symmetric && !remote)
    throw std::runtime_error("Matrix::from_blocks: symmetric storage carries a remote block (Matrix.cpp:63-71)");
  if (!symmetric && !remote && col_map->overlapping())
    throw std::runtime_error("Ovelapping not supported in this format!");
  if (!symmetric && remote && !col_map->overlapping())
    throw std::runtime_error("Ovelapping not enabled in column mapping!");
  Matrix<T> m;
  m._exec = exec;
  m._mat_local = std::move(local);
  m._mat_remote = std::move(remote);
  m._col_map = col_map;
  m._row_map = row_map;
  m._nnz = static_cast<int>(nnz);
  m._symmetric = symmetric;
  return m;
}
//-----------------------------------------------------------------------------
namespace spmv
{
namespace detail
{
static std::atomic<std::uint64_t> g_structure_epoch{1};
std::uint64_t structure_epoch() { return g_structure_epoch.load(); }
void bump_structure_epoch() { g_structure_epoch.fetch_add(1); }
} // namespace detail
} // namespace spmv
template <typename T>
Matrix<T>::~Matrix()
{
  detail::bump_structure_epoch(); // a graph captured over this operator must not be replayed
}
template <typename T>
Matrix<T>::Matrix(Matrix&& o) noexcept
    : _exec(std::move(o._exec)), _mat_local(std::move(o._mat_local)), _mat_remote(std::move(o._mat_remote)),
      _col_map(std::move(o._col_map)), _row_map(std::move(o._row_map)), _nnz(o._nnz), _symmetric(o._symmetric),
      _host_pipe(std::move(o._host_pipe))
{
  detail::bump_structure_epoch();
}
template <typename T>
Matrix<T>& Matrix<T>::operator=(Matrix&& o) noexcept
{
  if (this != &o) {
    _exec = std::move(o._exec);
    _mat_local = std::move(o._mat_local);
    _mat_remote = std::move(o._mat_remote);
    _col_map = std::move(o._col_map);
    _row_map = std::move(o._row_map);
    _nnz = o._nnz;
    _symmetric = o._symmetric;
    _host_pipe = std::move(o._host_pipe);
    detail::bump_structure_epoch();
  }
  return *this;
}
//-----------------------------------------------------------------------------
template <typename T>
int Matrix<T>::rows() const
{
  if (_mat_local)
    return _mat_local->rows();
  return _mat_remote ? _mat_remote->rows() : 0;
}
template <typename T>
int Matrix<T>::cols() const
{
  if (_mat_local)
    return _mat_local->cols();
  return _mat_remote ? _mat_remote->cols() : 0;
}
template <typename T>
int Matrix<T>::non_zeros() const
{
  if (_symmetric)
    return _nnz;
  if (_col_map->overlapping())
    return static_cast<int>(_mat_local->non_zeros() + _mat_remote->non_zeros());
  return static_cast<int>(_mat_local->non_zeros());
}
template <typename T>
size_t Matrix<T>::format_size() const
{
  // same accounting as Matrix.cpp:113-129
  size_t bytes = sizeof(int) * _mat_local->rows() + (sizeof(int) + sizeof(T)) * _mat_local->non_zeros();
  if (_symmetric || _col_map->overlapping())
    bytes += sizeof(int) * _mat_remote->rows() + (sizeof(int) + sizeof(T)) * _mat_remote->non_zeros();
  if (_symmetric)
    bytes += sizeof(T) * _mat_local->rows();
  return bytes;
}
//-----------------------------------------------------------------------------
template <typename T>
void Matrix<T>::mult(T* x, T* y) const
{
  if (_symmetric && _col_map->overlapping())
    spmv_sym_overlap(x, y);
  else if (_symmetric)
    spmv_sym(x, y);
  else if (_col_map->overlapping())
    spmv_overlap(x, y);
  else
    spmv(x, y);
}
template <typename T>
void Matrix<T>::transpmult(const T*, T*) const
{
  throw std::runtime_error("transpmult() operation not yet implemented");
}
//-----------------------------------------------------------------------------
template <typename T>
void Matrix<T>::spmv(T* x, T* y) const
{
  detail::NvtxRange r("spmv:single block");
  _mat_local->mult(1, x, 0, y);
}
template <typename T>
void Matrix<T>::spmv_overlap(T* x, T* y) const
{
  // local block runs while the halo values are in flight
  {
    detail::NvtxRange r("spmv:local block");
    _mat_local->mult(1, x, 0, y);
  }
  _col_map->update_finalise(x);
  detail::NvtxRange r("spmv:remote block");
  _mat_remote->mult(1, x, _mat_local->non_zeros() > 0 ? T(1) : T(0), y);
}
template <typename T>
void Matrix<T>::spmv_sym(T* x, T* y) const
{
  {
    detail::NvtxRange r("spmv:local block (symmetric)");
    _mat_local->mult(1, x, 0, y);
  }
  detail::NvtxRange r("spmv:remote block");
  _mat_remote->mult(1, x, 1, y);
}
template <typename T>
void Matrix<T>::spmv_sym_overlap(T* x, T* y) const
{
  {
    detail::NvtxRange r("spmv:local block (symmetric)");
    _mat_local->mult(1, x, 0, y);
  }
  _col_map->update_finalise(x);
  detail::NvtxRange r("spmv:remote block");
  _mat_remote->mult(1, x, 1, y);
}
//-----------------------------------------------------------------------------
template class spmv::Matrix<float>;
template class spmv::Matrix<double>;
