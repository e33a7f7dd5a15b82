// CreateLaplace3D.cpp -- see CreateLaplace3D.h.
#include "CreateLaplace3D.h"

#include <b200spmv.h>
#include <spmv/L2GMap.h>
#include <spmv/b200_executor.h>
#include <spmv/csr_matrix.h>

#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

namespace
{
void ck(int rc, const char* what)
{
  if (rc != B200_OK)
    throw std::runtime_error(std::string("create_laplace3d: ") + what + ": " + b200_last_error());
}

// read_petsc.cpp:20-36 applied to planes
std::int64_t plane_start(int rank, int size, std::int64_t n)
{
  const std::int64_t q = n / size, r = n % size;
  return rank < r ? rank * (q + 1) : rank * q + r;
}
} // namespace

spmv::Matrix<double> create_laplace3d(MPI_Comm comm, std::shared_ptr<spmv::DeviceExecutor> exec, int nx, int ny,
                                      std::int64_t nz_global, int stencil, bool symmetric,
                                      spmv::CommunicationModel cm)
{
  using namespace spmv;
  auto b200 = std::dynamic_pointer_cast<B200Executor>(exec);
  if (!b200)
    throw std::runtime_error("create_laplace3d: needs a B200Executor (no CPU generator path)");
  if (nx <= 0 || ny <= 0 || nz_global <= 0 || (stencil != 7 && stencil != 27))
    throw std::runtime_error("create_laplace3d: bad grid or stencil");
  int rank = 0, size = 1;
  MPI_Comm_rank(comm, &rank);
  MPI_Comm_size(comm, &size);
  const std::int64_t z0 = plane_start(rank, size, nz_global), z1 = plane_start(rank + 1, size, nz_global);
  const std::int64_t plane = static_cast<std::int64_t>(nx) * ny;
  const std::int64_t rows64 = plane * (z1 - z0);
  if (rows64 <= 0 || rows64 + 2 * plane > std::numeric_limits<std::int32_t>::max())
    throw std::runtime_error("create_laplace3d: every rank needs at least one plane and int32 local indices");
  const std::int32_t rows = static_cast<std::int32_t>(rows64);
  const bool overlap = cm == CommunicationModel::p2p_nonblocking || cm == CommunicationModel::collective_nonblocking;
  const int mode = symmetric ? 1 : (overlap ? 2 : 0);
  const bool two_blocks = mode != 0;

  ck(b200_set_device(b200->get_device_id()), "cudaSetDevice");
  void* stream = b200->get_cuda_stream();
  std::int32_t* rp_l = exec->alloc<std::int32_t>(static_cast<size_t>(rows) + 1);
  std::int32_t* rp_r = two_blocks ? exec->alloc<std::int32_t>(static_cast<size_t>(rows) + 1) : nullptr;
  std::int64_t nl = 0, nr = 0;
  ck(b200_gen_laplace3d_rowptr(nx, ny, nz_global, z0, z1, stencil, mode, rp_l, rp_r, &nl, &nr, stream), "row pointers");
  if (nl > std::numeric_limits<std::int32_t>::max() || nr > std::numeric_limits<std::int32_t>::max())
    throw std::runtime_error("create_laplace3d: block exceeds the int32 rowptr range");
  std::int32_t* ci_l = nl > 0 ? exec->alloc<std::int32_t>(nl) : nullptr;
  double* v_l = nl > 0 ? exec->alloc<double>(nl) : nullptr;
  std::int32_t* ci_r = nr > 0 ? exec->alloc<std::int32_t>(nr) : nullptr;
  double* v_r = nr > 0 ? exec->alloc<double>(nr) : nullptr;
  double* diag = symmetric ? exec->alloc<double>(rows) : nullptr;
  ck(b200_gen_laplace3d_fill(nx, ny, nz_global, z0, z1, stencil, mode, rp_l, ci_l, v_l, diag, rp_r, ci_r, v_r, stream),
     "fill");
  ck(b200_stream_synchronize(stream), "synchronize");
  // an empty block carries no arrays (csr_matrix.cpp:35)
  if (nl == 0) {
    exec->free(rp_l);
    rp_l = nullptr;
  }
  if (two_blocks && nr == 0) {
    exec->free(rp_r);
    rp_r = nullptr;
  }

  // ghost columns: the plane below and the plane above, ascending (Matrix.cpp:294-318)
  std::vector<std::int64_t> ghosts;
  if (z0 > 0)
    for (std::int64_t g = (z0 - 1) * plane; g < z0 * plane; ++g)
      ghosts.push_back(g);
  if (z1 < nz_global)
    for (std::int64_t g = z1 * plane; g < (z1 + 1) * plane; ++g)
      ghosts.push_back(g);
  const std::int32_t ncols = rows + static_cast<std::int32_t>(ghosts.size());
  // maps as create_matrix builds them (Matrix.cpp:440,464,476: the row map uses the default model)
  auto col_map = std::make_shared<L2GMap>(comm, rows, ghosts, exec, cm);
  auto row_map = std::make_shared<L2GMap>(comm, rows, std::vector<std::int64_t>(), exec);

  using Block = CSRMatrix<double>;
  std::unique_ptr<SubMatrix<double>> local(
      new Block(Block::adopt_t{}, exec, rows, (overlap && !symmetric) ? rows : ncols, nl, rp_l, ci_l, v_l, diag, symmetric));
  std::unique_ptr<SubMatrix<double>> remote;
  if (two_blocks)
    remote.reset(new Block(Block::adopt_t{}, exec, rows, ncols, nr, rp_r, ci_r, v_r, nullptr, false));
  // Matrix.cpp:443-446: symmetric non_zeros() = 2*lower + remote + diagonal rows
  const std::int64_t nnz = symmetric ? 2 * nl + nr + rows : nl + nr;
  return Matrix<double>::from_blocks(std::move(local), std::move(remote), col_map, row_map, nnz, symmetric, exec);
}
