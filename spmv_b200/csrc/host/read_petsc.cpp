// read_petsc.cpp -- PETSc binary readers / writers.
// Reference: spmv/read_petsc.cpp:39-303; format in SURVEY.md Appendix E (all fields
// big-endian; matrix: int32 classid 1211216, nrows, ncols, nnz, nnz_per_row[nrows],
// colidx[nnz], float64 values[nnz]; vector: int32 classid 1211214, n, values[n]).
//
// Same results as the reference (row split, ghost numbering, local / remote /
// diagonal rules, the maps), without its per-entry std::map lookups and Eigen
// inserts, and with 64-bit file offsets (the reference computes nnz_tot*4 in int,
// read_petsc.cpp:117-118).
#include "read_petsc.h"

#include "L2GMap.h"
#include "Matrix.h"
#include "block_split.h"
#include "device_executor.h"
#include "host_util.h"

#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <vector>

namespace
{

inline std::uint32_t bswap32(std::uint32_t v) { return __builtin_bswap32(v); }
inline std::uint64_t bswap64(std::uint64_t v) { return __builtin_bswap64(v); }

void read_exact(std::ifstream& f, void* dst, std::size_t bytes)
{
  f.read(static_cast<char*>(dst), static_cast<std::streamsize>(bytes));
  if (static_cast<std::size_t>(f.gcount()) != bytes)
    throw std::runtime_error("Unexpected end of PETSc file");
}

int g_threads = 1; // set per call from detail::host_threads

void read_i32_be(std::ifstream& f, std::int32_t* dst, std::size_t n)
{
  read_exact(f, dst, n * 4);
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (std::size_t i = 0; i < n; ++i) {
    std::uint32_t u;
    std::memcpy(&u, dst + i, 4);
    u = bswap32(u);
    std::memcpy(dst + i, &u, 4);
  }
}

void read_f64_be(std::ifstream& f, double* dst, std::size_t n)
{
  read_exact(f, dst, n * 8);
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (std::size_t i = 0; i < n; ++i) {
    std::uint64_t u;
    std::memcpy(&u, dst + i, 8);
    u = bswap64(u);
    std::memcpy(dst + i, &u, 8);
  }
}

} // namespace

namespace spmv
{

using namespace spmv::detail;

//-----------------------------------------------------------------------------
Matrix<double> read_petsc_binary_matrix(std::string filename, MPI_Comm comm, std::shared_ptr<DeviceExecutor> exec,
                                        bool symmetric, CommunicationModel cm)
{
  int mpi_rank = 0, mpi_size = 1;
  CHECK_MPI(MPI_Comm_rank(comm, &mpi_rank));
  CHECK_MPI(MPI_Comm_size(comm, &mpi_size));
  g_threads = host_threads(mpi_size);

  std::ifstream file(filename.c_str(), std::ios::in | std::ios::binary);
  if (!file.is_open())
    throw std::runtime_error("Could not open file");

  std::int32_t header[4];
  read_i32_be(file, header, 4);
  if (header[0] != 1211216)
    throw std::runtime_error("Bad signature in PETSc Matrix file");
  const std::int32_t nrows = header[1], ncols = header[2];
  const std::int64_t nnz_tot = header[3];
  const std::vector<std::int64_t> row_ranges = owner_ranges(mpi_size, nrows);
  const std::vector<std::int64_t> col_ranges = owner_ranges(mpi_size, ncols);
  if (mpi_rank == 0)
    std::cout << "Read matrix file: " << filename << ": " << nrows << "x" << ncols << " = " << nnz_tot << "\n";

  const std::int64_t row0 = row_ranges[mpi_rank], row1 = row_ranges[mpi_rank + 1];
  const std::int64_t col0 = col_ranges[mpi_rank], col1 = col_ranges[mpi_rank + 1];
  const std::int64_t nrows_local = row1 - row0, ncols_local = col1 - col0;

  // nnz per row for ALL rows (the reference reads the whole array too, :94-107)
  std::vector<std::int32_t> nnz(nrows);
  read_i32_be(file, nnz.data(), nrows);
  std::int64_t nnz_offset = 0, nnz_size = 0;
  for (std::int64_t i = 0; i < row0; ++i)
    nnz_offset += nnz[i];
  for (std::int64_t i = row0; i < row1; ++i)
    nnz_size += nnz[i];
  if (nnz_size >= (1ll << 31))
    throw std::runtime_error("read_petsc_binary_matrix: local block exceeds int32 nonzeros");

  const std::int64_t base = 16 + 4ll * nrows;
  std::vector<std::int32_t> gcol(nnz_size);
  std::vector<double> vals(nnz_size);
  file.seekg(base + 4 * nnz_offset, std::ios::beg);
  read_i32_be(file, gcol.data(), nnz_size);
  file.seekg(base + 4 * nnz_tot + 8 * nnz_offset, std::ios::beg);
  read_f64_be(file, vals.data(), nnz_size);
  file.close();

  // ghost columns, ascending by global index (read_petsc.cpp:126-148)
  std::vector<std::int64_t> ghosts;
  for (std::int64_t j = 0; j < nnz_size; ++j)
    if (gcol[j] < col0 || gcol[j] >= col1)
      ghosts.push_back(gcol[j]);
  std::sort(ghosts.begin(), ghosts.end());
  ghosts.erase(std::unique(ghosts.begin(), ghosts.end()), ghosts.end());
  const std::int64_t nghosts = static_cast<std::int64_t>(ghosts.size());

  // split with the rules of read_petsc.cpp:164-204 (== Matrix.cpp:337-358)
  const bool overlap
      = cm == CommunicationModel::p2p_nonblocking || cm == CommunicationModel::collective_nonblocking;
  const Splitter dest{symmetric, overlap, ncols_local, row0, col0};
  const std::int32_t n = static_cast<std::int32_t>(nrows_local);
  std::vector<std::int64_t> rowstart(n + 1, 0);
  for (std::int32_t r = 0; r < n; ++r)
    rowstart[r + 1] = rowstart[r] + nnz[row0 + r];

  HostCSR<double> blocks[2];
  for (HostCSR<double>& b : blocks) {
    b.rows = n;
    b.cols = static_cast<std::int32_t>(ncols_local + nghosts);
    b.rowptr.assign(static_cast<std::size_t>(n) + 1, 0);
  }
  std::vector<double> diagonal(symmetric ? n : 0, 0.0);
  std::vector<std::int32_t> lcol(nnz_size);
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (std::int32_t r = 0; r < n; ++r) {
    std::int32_t cnt[2] = {0, 0};
    for (std::int64_t j = rowstart[r]; j < rowstart[r + 1]; ++j) {
      const std::int64_t g = gcol[j];
      lcol[j] = (g >= col0 && g < col1)
                    ? static_cast<std::int32_t>(g - col0)
                    : static_cast<std::int32_t>(ncols_local
                                                + (std::lower_bound(ghosts.begin(), ghosts.end(), g) - ghosts.begin()));
      const Dest d = dest(r, lcol[j]);
      if (d <= REMOTE)
        ++cnt[d];
    }
    blocks[LOCAL].rowptr[r + 1] = cnt[LOCAL];
    blocks[REMOTE].rowptr[r + 1] = cnt[REMOTE];
  }
  for (HostCSR<double>& b : blocks) {
    for (std::int32_t r = 0; r < n; ++r)
      b.rowptr[r + 1] += b.rowptr[r];
    b.colind.resize(b.rowptr[n]);
    b.values.resize(b.rowptr[n]);
  }
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (std::int32_t r = 0; r < n; ++r) {
    std::int32_t w[2] = {blocks[LOCAL].rowptr[r], blocks[REMOTE].rowptr[r]};
    for (std::int64_t j = rowstart[r]; j < rowstart[r + 1]; ++j) {
      const Dest d = dest(r, lcol[j]);
      if (d <= REMOTE) {
        blocks[d].colind[w[d]] = lcol[j];
        blocks[d].values[w[d]] = vals[j];
        ++w[d];
      } else if (d == DIAG) {
        diagonal[r] = vals[j]; // assignment, as read_petsc.cpp:183-184
      }
    }
  }
  canonicalise(blocks[LOCAL], g_threads); // rows ordered by local column id, like insert + makeCompressed
  canonicalise(blocks[REMOTE], g_threads);

  auto col_map = std::make_shared<spmv::L2GMap>(comm, ncols_local, ghosts, exec, cm);
  auto row_map = std::make_shared<spmv::L2GMap>(comm, nrows_local, std::vector<std::int64_t>(), exec, cm);

  if (symmetric) // nnz_full is the GLOBAL count here, as in read_petsc.cpp:222-223
    return spmv::Matrix<double>(blocks[LOCAL], blocks[REMOTE], diagonal, col_map, row_map, static_cast<int>(nnz_tot),
                                exec);
  if (col_map->overlapping()) {
    blocks[LOCAL].cols = static_cast<std::int32_t>(ncols_local + nghosts); // read_petsc.cpp:150-153 sizes both alike
    return spmv::Matrix<double>(blocks[LOCAL], blocks[REMOTE], col_map, row_map, exec);
  }
  return spmv::Matrix<double>(blocks[LOCAL], col_map, row_map, exec);
}
//-----------------------------------------------------------------------------
double* read_petsc_binary_vector(MPI_Comm comm, const DeviceExecutor* exec, const std::string filename)
{
  int mpi_rank = 0, mpi_size = 1;
  CHECK_MPI(MPI_Comm_rank(comm, &mpi_rank));
  CHECK_MPI(MPI_Comm_size(comm, &mpi_size));
  std::ifstream file(filename.c_str(), std::ios::in | std::ios::binary);
  if (!file.is_open())
    throw std::runtime_error("Could not open file");
  std::int32_t header[2];
  read_i32_be(file, header, 2);
  if (header[0] != 1211214)
    throw std::runtime_error("Bad signature in PETSc Vector file");
  const std::int32_t nrows = header[1];
  const std::vector<std::int64_t> ranges = owner_ranges(mpi_size, nrows);
  if (mpi_rank == 0)
    std::cout << "Read vector file: " << filename << ": " << nrows << "\n";
  const std::int64_t n_local = ranges[mpi_rank + 1] - ranges[mpi_rank];
  std::vector<double> host(n_local);
  file.seekg(8 + 8 * ranges[mpi_rank], std::ios::beg);
  read_f64_be(file, host.data(), n_local);
  double* dev = exec->alloc<double>(n_local > 0 ? n_local : 1);
  exec->copy_from(dev, exec->get_host(), host.data(), n_local);
  return dev;
}
//-----------------------------------------------------------------------------
void write_petsc_binary_matrix(const std::string& filename, std::int32_t nrows, std::int32_t ncols,
                               const std::int32_t* rowptr, const std::int32_t* colind, const double* values)
{
  std::ofstream f(filename.c_str(), std::ios::out | std::ios::binary | std::ios::trunc);
  if (!f.is_open())
    throw std::runtime_error("Could not open file for writing");
  const std::int64_t nnz = rowptr[nrows];
  auto put32 = [&](std::uint32_t v) {
    v = bswap32(v);
    f.write(reinterpret_cast<const char*>(&v), 4);
  };
  put32(1211216);
  put32(static_cast<std::uint32_t>(nrows));
  put32(static_cast<std::uint32_t>(ncols));
  put32(static_cast<std::uint32_t>(nnz));
  std::vector<std::uint32_t> buf(std::max<std::int64_t>(nrows, nnz));
  for (std::int32_t r = 0; r < nrows; ++r)
    buf[r] = bswap32(static_cast<std::uint32_t>(rowptr[r + 1] - rowptr[r]));
  f.write(reinterpret_cast<const char*>(buf.data()), 4ll * nrows);
  for (std::int64_t j = 0; j < nnz; ++j)
    buf[j] = bswap32(static_cast<std::uint32_t>(colind[j]));
  f.write(reinterpret_cast<const char*>(buf.data()), 4 * nnz);
  std::vector<std::uint64_t> vbuf(nnz);
  for (std::int64_t j = 0; j < nnz; ++j) {
    std::uint64_t u;
    std::memcpy(&u, values + j, 8);
    vbuf[j] = bswap64(u);
  }
  f.write(reinterpret_cast<const char*>(vbuf.data()), 8 * nnz);
}

void write_petsc_binary_vector(const std::string& filename, std::int32_t n, const double* values)
{
  std::ofstream f(filename.c_str(), std::ios::out | std::ios::binary | std::ios::trunc);
  if (!f.is_open())
    throw std::runtime_error("Could not open file for writing");
  std::uint32_t h[2] = {bswap32(1211214u), bswap32(static_cast<std::uint32_t>(n))};
  f.write(reinterpret_cast<const char*>(h), 8);
  std::vector<std::uint64_t> vbuf(n);
  for (std::int32_t j = 0; j < n; ++j) {
    std::uint64_t u;
    std::memcpy(&u, values + j, 8);
    vbuf[j] = bswap64(u);
  }
  f.write(reinterpret_cast<const char*>(vbuf.data()), 8ll * n);
}

} // namespace spmv
