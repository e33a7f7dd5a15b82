// create_matrix.cpp -- Matrix<T>::create_matrix: ship ghost rows to their owners,
// renumber ghost columns, split into local / remote / diagonal blocks, build the maps.
// Reference: spmv/Matrix.cpp:150-480 (results must match it entry for entry).
//
// The reference goes through std::map lookups and Eigen triplets (16 B per entry +
// a sort inside setFromTriplets).  Here the split is two linear passes over the CSR
// (count, then fill), threaded over rows; a row is sorted only if the remap broke its
// column order, and duplicates (which only arise from merged ghost rows) are summed
// in input order, which is what setFromTriplets does.
#include "Matrix.h"

#include "L2GMap.h"
#include "block_split.h"
#include "device_executor.h"
#include "host_util.h"

#include <algorithm>
#include <numeric>
#include <set>
#include <stdexcept>

using namespace spmv;

using namespace spmv::detail;

//-----------------------------------------------------------------------------
template <typename T>
Matrix<T>* Matrix<T>::create_matrix(MPI_Comm comm, std::shared_ptr<DeviceExecutor> exec, const HostCSR<T>& mat,
                                    std::int64_t nrows_local, std::int64_t ncols_local,
                                    std::vector<std::int64_t> row_ghosts, std::vector<std::int64_t> col_ghosts,
                                    bool symmetric, CommunicationModel cm)
{
  return create_matrix(comm, exec, mat.rowptr.data(), mat.colind.data(), mat.values.data(), nrows_local,
                       ncols_local, std::move(row_ghosts), std::move(col_ghosts), symmetric, cm);
}
//-----------------------------------------------------------------------------
template <typename T>
Matrix<T>* Matrix<T>::create_matrix(MPI_Comm comm, std::shared_ptr<DeviceExecutor> exec,
                                    const std::int32_t* rowptr, const std::int32_t* colind, const T* values,
                                    std::int64_t nrows_local, std::int64_t ncols_local,
                                    std::vector<std::int64_t> row_ghosts, std::vector<std::int64_t> col_ghosts,
                                    bool symmetric, CommunicationModel cm)
{
  int mpi_size = 0, mpi_rank = 0;
  CHECK_MPI(MPI_Comm_size(comm, &mpi_size));
  CHECK_MPI(MPI_Comm_rank(comm, &mpi_rank));

  // ownership ranges of rows and columns (Matrix.cpp:174-186)
  std::vector<std::int64_t> row_ranges(mpi_size + 1, 0), col_ranges(mpi_size + 1, 0);
  CHECK_MPI(MPI_Allgather(&nrows_local, 1, MPI_INT64_T, row_ranges.data() + 1, 1, MPI_INT64_T, comm));
  CHECK_MPI(MPI_Allgather(&ncols_local, 1, MPI_INT64_T, col_ranges.data() + 1, 1, MPI_INT64_T, comm));
  for (int i = 0; i < mpi_size; ++i) {
    row_ranges[i + 1] += row_ranges[i];
    col_ranges[i + 1] += col_ranges[i];
  }
  const std::int64_t row0 = row_ranges[mpi_rank], col0 = col_ranges[mpi_rank], col1 = col_ranges[mpi_rank + 1];

  // ---- ghost rows go to their owners (Matrix.cpp:188-292) -------------------
  const std::size_t nrg = row_ghosts.size();
  std::vector<int> row_owner(nrg);
  for (std::size_t i = 0; i < nrg; ++i) {
    row_owner[i]
        = static_cast<int>(std::upper_bound(row_ranges.begin(), row_ranges.end(), row_ghosts[i]) - row_ranges.begin())
          - 1;
    if (row_owner[i] == mpi_rank || row_owner[i] < 0 || row_owner[i] >= mpi_size)
      throw std::runtime_error("create_matrix: a row ghost is owned locally or out of range");
  }
  std::set<int> dest_set(row_owner.begin(), row_owner.end());
  std::vector<int> dests(dest_set.begin(), dest_set.end());
  std::vector<char> is_dest(mpi_size, 0), is_source(mpi_size, 0);
  for (int d : dests)
    is_dest[d] = 1;
  CHECK_MPI(MPI_Alltoall(is_dest.data(), 1, MPI_CHAR, is_source.data(), 1, MPI_CHAR, comm));
  std::vector<int> sources;
  for (int i = 0; i < mpi_size; ++i)
    if (is_source[i])
      sources.push_back(i);

  MPI_Comm nbr_comm = MPI_COMM_NULL;
  CHECK_MPI(MPI_Dist_graph_create_adjacent(comm, static_cast<int>(sources.size()), sources.data(), MPI_UNWEIGHTED,
                                           static_cast<int>(dests.size()), dests.data(), MPI_UNWEIGHTED,
                                           MPI_INFO_NULL, false, &nbr_comm));

  // wire format per ghost row: (global_row, nnz, global cols...) with values (0, 0, vals...)
  std::vector<std::vector<std::int64_t>> to_index(dests.size());
  std::vector<std::vector<T>> to_val(dests.size());
  for (std::size_t i = 0; i < nrg; ++i) {
    const std::size_t p = std::lower_bound(dests.begin(), dests.end(), row_owner[i]) - dests.begin();
    const std::int32_t s = rowptr[nrows_local + i], e = rowptr[nrows_local + i + 1];
    to_index[p].push_back(row_ghosts[i]);
    to_val[p].push_back(T(0));
    to_index[p].push_back(e - s);
    to_val[p].push_back(T(0));
    for (std::int32_t j = s; j < e; ++j) {
      const std::int32_t c = colind[j];
      if (c >= ncols_local && static_cast<std::size_t>(c - ncols_local) >= col_ghosts.size())
        throw std::runtime_error("create_matrix: column index beyond the ghost list");
      to_index[p].push_back(c < ncols_local ? c + col0 : col_ghosts[c - ncols_local]);
      to_val[p].push_back(values[j]);
    }
  }
  std::vector<int> send_size(dests.size()), send_offset = {0};
  std::vector<std::int64_t> send_index;
  std::vector<T> send_val;
  for (std::size_t p = 0; p < dests.size(); ++p) {
    send_index.insert(send_index.end(), to_index[p].begin(), to_index[p].end());
    send_val.insert(send_val.end(), to_val[p].begin(), to_val[p].end());
    send_size[p] = static_cast<int>(to_index[p].size());
    send_offset.push_back(static_cast<int>(send_index.size()));
  }
  std::vector<int> recv_size(sources.size());
  if (sources.size() + dests.size() == 0) {
    send_size = {0};
    recv_size = {0};
  }
  CHECK_MPI(MPI_Neighbor_alltoall(send_size.data(), 1, MPI_INT, recv_size.data(), 1, MPI_INT, nbr_comm));
  std::vector<int> recv_offset = {0};
  for (int r : recv_size)
    recv_offset.push_back(recv_offset.back() + r);
  std::vector<std::int64_t> recv_index(recv_offset.back());
  std::vector<T> recv_val(recv_offset.back());
  CHECK_MPI(MPI_Neighbor_alltoallv(send_index.data(), send_size.data(), send_offset.data(), MPI_INT64_T,
                                   recv_index.data(), recv_size.data(), recv_offset.data(), MPI_INT64_T, nbr_comm));
  CHECK_MPI(MPI_Neighbor_alltoallv(send_val.data(), send_size.data(), send_offset.data(), mpi_type<T>(),
                                   recv_val.data(), recv_size.data(), recv_offset.data(), mpi_type<T>(), nbr_comm));
  CHECK_MPI(MPI_Comm_free(&nbr_comm));

  // ---- ghost columns: the given ones plus any that arrived with ghost rows,
  //      numbered ascending from ncols_local (Matrix.cpp:294-318) ---------------
  std::vector<std::int64_t> new_col_ghosts(col_ghosts);
  for (std::size_t pos = 0; pos < recv_index.size();) {
    const std::int64_t n = recv_index[pos + 1];
    pos += 2;
    for (std::int64_t k = 0; k < n; ++k, ++pos)
      if (recv_index[pos] >= col1 || recv_index[pos] < col0)
        new_col_ghosts.push_back(recv_index[pos]);
  }
  std::sort(new_col_ghosts.begin(), new_col_ghosts.end());
  new_col_ghosts.erase(std::unique(new_col_ghosts.begin(), new_col_ghosts.end()), new_col_ghosts.end());
  const std::int64_t nghosts = static_cast<std::int64_t>(new_col_ghosts.size());
  if (ncols_local + nghosts >= (1ll << 31))
    throw std::runtime_error("create_matrix: local column space exceeds int32");
  auto ghost_id = [&](std::int64_t g) {
    return static_cast<std::int32_t>(
        ncols_local + (std::lower_bound(new_col_ghosts.begin(), new_col_ghosts.end(), g) - new_col_ghosts.begin()));
  };
  std::vector<std::int32_t> old_to_new(col_ghosts.size());
  for (std::size_t k = 0; k < col_ghosts.size(); ++k)
    old_to_new[k] = ghost_id(col_ghosts[k]);
  auto remap = [&](std::int32_t c) { return c < ncols_local ? c : old_to_new[c - ncols_local]; };

  // received rows in local form: (row, col, val) in arrival order
  struct Recv {
    std::int32_t row, col;
    T val;
  };
  std::vector<Recv> received;
  for (std::size_t pos = 0; pos < recv_index.size();) {
    const std::int64_t grow = recv_index[pos];
    if (grow < row0 || grow >= row_ranges[mpi_rank + 1])
      throw std::runtime_error("create_matrix: received a row this rank does not own");
    const std::int64_t n = recv_index[pos + 1];
    pos += 2;
    for (std::int64_t k = 0; k < n; ++k, ++pos) {
      const std::int64_t gcol = recv_index[pos];
      const std::int32_t col
          = (gcol >= col1 || gcol < col0) ? ghost_id(gcol) : static_cast<std::int32_t>(gcol - col0);
      received.push_back({static_cast<std::int32_t>(grow - row0), col, recv_val[pos]});
    }
  }

  // ---- split (Matrix.cpp:320-408): count, scan, fill ------------------------
  const bool overlap
      = cm == CommunicationModel::p2p_nonblocking || cm == CommunicationModel::collective_nonblocking;
  const Splitter dest{symmetric, overlap, ncols_local, row0, col0};
  const std::int32_t n = static_cast<std::int32_t>(nrows_local);
  HostCSR<T> blocks[2];
  for (HostCSR<T>& b : blocks) {
    b.rows = n;
    b.cols = static_cast<std::int32_t>(ncols_local + nghosts);
    b.rowptr.assign(static_cast<std::size_t>(n) + 1, 0);
  }
  if (!symmetric && overlap)
    blocks[LOCAL].cols = static_cast<std::int32_t>(ncols_local);
  std::vector<T> diagonal;
  std::vector<char> has_diag;
  int mpi_size_for_threads = 1;
  MPI_Comm_size(comm, &mpi_size_for_threads);
  const int nthreads = detail::host_threads(mpi_size_for_threads);
  if (symmetric) {
    diagonal.assign(n, T(0));
    has_diag.assign(n, 0);
  }

#pragma omp parallel for schedule(static) num_threads(nthreads)
  for (std::int32_t r = 0; r < n; ++r) {
    std::int32_t cnt[2] = {0, 0};
    for (std::int32_t j = rowptr[r]; j < rowptr[r + 1]; ++j) {
      const Dest d = dest(r, remap(colind[j]));
      if (d <= REMOTE)
        ++cnt[d];
    }
    blocks[LOCAL].rowptr[r + 1] = cnt[LOCAL];
    blocks[REMOTE].rowptr[r + 1] = cnt[REMOTE];
  }
  for (const Recv& e : received) {
    const Dest d = dest(e.row, e.col);
    if (d <= REMOTE)
      ++blocks[d].rowptr[e.row + 1];
  }
  for (HostCSR<T>& b : blocks) {
    std::int64_t run = 0;
    for (std::int32_t r = 0; r < n; ++r) {
      run += b.rowptr[r + 1];
      if (run >= (1ll << 31))
        throw std::runtime_error("create_matrix: block exceeds int32 nonzeros");
      b.rowptr[r + 1] = static_cast<std::int32_t>(run);
    }
    b.colind.resize(run);
    b.values.resize(run);
  }
  std::vector<std::int32_t> filled[2];
  if (!received.empty()) {
    filled[LOCAL].assign(n, 0);
    filled[REMOTE].assign(n, 0);
  }
#pragma omp parallel for schedule(static) num_threads(nthreads)
  for (std::int32_t r = 0; r < n; ++r) {
    std::int32_t w[2] = {blocks[LOCAL].rowptr[r], blocks[REMOTE].rowptr[r]};
    for (std::int32_t j = rowptr[r]; j < rowptr[r + 1]; ++j) {
      const std::int32_t c = remap(colind[j]);
      const Dest d = dest(r, c);
      if (d <= REMOTE) {
        blocks[d].colind[w[d]] = c;
        blocks[d].values[w[d]] = values[j];
        ++w[d];
      } else if (d == DIAG) {
        diagonal[r] += values[j];
        has_diag[r] = 1;
      }
    }
    if (!received.empty()) {
      filled[LOCAL][r] = w[LOCAL] - blocks[LOCAL].rowptr[r];
      filled[REMOTE][r] = w[REMOTE] - blocks[REMOTE].rowptr[r];
    }
  }
  for (const Recv& e : received) {
    const Dest d = dest(e.row, e.col);
    if (d <= REMOTE) {
      const std::int32_t w = blocks[d].rowptr[e.row] + filled[d][e.row]++;
      blocks[d].colind[w] = e.col;
      blocks[d].values[w] = e.val;
    } else if (d == DIAG) {
      diagonal[e.row] += e.val;
      has_diag[e.row] = 1;
    }
  }
  canonicalise(blocks[LOCAL], nthreads);
  canonicalise(blocks[REMOTE], nthreads);

  // ---- maps and the Matrix (Matrix.cpp:410-479) ------------------------------
  auto col_map = std::make_shared<spmv::L2GMap>(comm, ncols_local, new_col_ghosts, exec, cm);
  auto row_map = std::make_shared<spmv::L2GMap>(comm, nrows_local, std::vector<std::int64_t>(), exec);
  if (symmetric) {
    const std::int64_t unique_rows = std::count(has_diag.begin(), has_diag.end(), char(1));
    const std::int64_t nnz_full = 2 * blocks[LOCAL].non_zeros() + blocks[REMOTE].non_zeros() + unique_rows;
    return new Matrix<T>(blocks[LOCAL], blocks[REMOTE], diagonal, col_map, row_map, static_cast<int>(nnz_full), exec);
  }
  if (overlap)
    return new Matrix<T>(blocks[LOCAL], blocks[REMOTE], col_map, row_map, exec);
  return new Matrix<T>(blocks[LOCAL], col_map, row_map, exec);
}
//-----------------------------------------------------------------------------
template Matrix<float>* Matrix<float>::create_matrix(MPI_Comm, std::shared_ptr<DeviceExecutor>, const std::int32_t*,
                                                     const std::int32_t*, const float*, std::int64_t, std::int64_t,
                                                     std::vector<std::int64_t>, std::vector<std::int64_t>, bool,
                                                     CommunicationModel);
template Matrix<double>* Matrix<double>::create_matrix(MPI_Comm, std::shared_ptr<DeviceExecutor>, const std::int32_t*,
                                                       const std::int32_t*, const double*, std::int64_t, std::int64_t,
                                                       std::vector<std::int64_t>, std::vector<std::int64_t>, bool,
                                                       CommunicationModel);
template Matrix<float>* Matrix<float>::create_matrix(MPI_Comm, std::shared_ptr<DeviceExecutor>, const HostCSR<float>&,
                                                     std::int64_t, std::int64_t, std::vector<std::int64_t>,
                                                     std::vector<std::int64_t>, bool, CommunicationModel);
template Matrix<double>* Matrix<double>::create_matrix(MPI_Comm, std::shared_ptr<DeviceExecutor>,
                                                       const HostCSR<double>&, std::int64_t, std::int64_t,
                                                       std::vector<std::int64_t>, std::vector<std::int64_t>, bool,
                                                       CommunicationModel);
