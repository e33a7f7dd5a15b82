// CreateA.cpp -- reference: demos/CreateA.cpp:31-103 (same matrix, same maps).
#include "CreateA.h"

#include <spmv/L2GMap.h>

#include <vector>

spmv::Matrix<double> create_A(MPI_Comm comm, int N, std::shared_ptr<spmv::DeviceExecutor> exec)
{
  int mpi_rank = 0, mpi_size = 1;
  MPI_Comm_rank(comm, &mpi_rank);
  MPI_Comm_size(comm, &mpi_size);

  // even row split, the first N % size ranks own one more (demos/CreateA.cpp:13-29)
  const std::int64_t q = N / mpi_size, rem = N % mpi_size;
  auto start = [&](int rank) { return rank < rem ? rank * (q + 1) : rank * q + rem; };
  const std::int64_t r0 = start(mpi_rank), r1 = start(mpi_rank + 1);
  const int M = static_cast<int>(r1 - r0);

  // ghosts: the global columns this rank touches but does not own, ascending
  std::vector<std::int64_t> ghosts;
  if (M > 0 && r0 > 0)
    ghosts.push_back(r0 - 1);
  if (M > 0 && r1 < N)
    ghosts.push_back(r1);

  auto col_l2g = std::make_shared<spmv::L2GMap>(comm, M, ghosts, exec);
  auto row_l2g = std::make_shared<spmv::L2GMap>(comm, M, std::vector<std::int64_t>(), exec);

  // rows in local numbering, sorted by local column: owned columns first, then ghosts
  const double gamma = 0.1;
  spmv::HostCSR<double> A;
  A.rows = M;
  A.cols = M + static_cast<int>(ghosts.size());
  A.rowptr.assign(M + 1, 0);
  for (int i = 0; i < M; ++i) {
    const std::int64_t c0 = r0 + i;
    struct E {
      std::int32_t col;
      double val;
    } e[3];
    int n = 0;
    auto add = [&](std::int64_t gcol, double v) { e[n++] = {col_l2g->global_to_local(gcol), v}; };
    if (c0 == 0) {
      add(c0, 1.0 - gamma);
      add(c0 + 1, gamma);
    } else if (c0 == N - 1) {
      add(c0 - 1, gamma);
      add(c0, 1.0 - gamma);
    } else {
      add(c0 - 1, gamma);
      add(c0, 1.0 - 2.0 * gamma);
      add(c0 + 1, gamma);
    }
    for (int a = 1; a < n; ++a) // insertion sort by local column (setFromTriplets order)
      for (int b = a; b > 0 && e[b].col < e[b - 1].col; --b)
        std::swap(e[b], e[b - 1]);
    for (int k = 0; k < n; ++k) {
      A.colind.push_back(e[k].col);
      A.values.push_back(e[k].val);
    }
    A.rowptr[i + 1] = static_cast<std::int32_t>(A.colind.size());
  }
  return spmv::Matrix<double>(A, col_l2g, row_l2g, exec);
}
