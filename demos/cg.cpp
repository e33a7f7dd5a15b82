// demos/cg.cpp -- Conjugate Gradient driver on the B200 executor.
//
// Protocol of the reference's driver (demos/cg.cpp:25-125; CUDA flavour demos/cg_cuda.cpp): read A
// and b, max_its = 100, rtol = 1e-10, solve, then recompute r = A x - b with one more SpMV and print
// ||r||, ||x|| and the iteration count.  b may also be synthesised as A*1 (`--ones`), and the matrix
// may be `--create-a N` (demos/CreateA.cpp:31-103).
//
//   demo_cg <matrix.petsc> <vector.petsc> [--symmetric] [--max-its K] [--rtol R]
//   demo_cg <matrix.petsc> --ones
//   demo_cg --create-a 100000 --ones
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include <mpi.h>
#include <b200spmv.h>
#include <spmv/spmv.h>

#include "CreateA.h"

int main(int argc, char** argv)
{
  MPI_Init(&argc, &argv);
  int rank = 0, size = 1;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  try {
    std::vector<std::string> files;
    long create_a = 0;
    int max_its = 100;
    double rtol = 1e-10;
    bool symmetric = false, ones = false;
    for (int i = 1; i < argc; ++i) {
      if (!std::strcmp(argv[i], "--create-a") && i + 1 < argc)
        create_a = std::atol(argv[++i]);
      else if (!std::strcmp(argv[i], "--max-its") && i + 1 < argc)
        max_its = std::atoi(argv[++i]);
      else if (!std::strcmp(argv[i], "--rtol") && i + 1 < argc)
        rtol = std::atof(argv[++i]);
      else if (!std::strcmp(argv[i], "--symmetric"))
        symmetric = true;
      else if (!std::strcmp(argv[i], "--ones"))
        ones = true;
      else
        files.push_back(argv[i]);
    }
    if ((files.empty() && create_a <= 0) || (!ones && files.size() < 2))
      throw std::runtime_error("Use: demo_cg <matrix_file> <vector_file> | <matrix_file> --ones | --create-a N --ones");

    int ndev = 0;
    if (b200_device_count(&ndev) != B200_OK)
      throw std::runtime_error(b200_last_error());
    std::shared_ptr<spmv::DeviceExecutor> host = spmv::ReferenceExecutor::create();
    std::shared_ptr<spmv::B200Executor> exec = spmv::B200Executor::create(rank % std::max(ndev, 1), host);
    void* stream = nullptr;
    if (b200_stream_create(&stream) != B200_OK)
      throw std::runtime_error(b200_last_error());
    exec->set_cuda_stream(static_cast<spmv::B200Executor::stream_t>(stream));

    const auto cm = spmv::CommunicationModel::p2p_nonblocking; // local block under the halo (demos/cg_cuda.cpp uses p2p)
    spmv::Matrix<double> A = create_a > 0 ? create_A(MPI_COMM_WORLD, static_cast<int>(create_a), exec)
                                          : spmv::read_petsc_binary_matrix(files[0], MPI_COMM_WORLD, exec, symmetric, cm);
    std::shared_ptr<const spmv::L2GMap> l2g = A.col_map();
    const std::int64_t M = l2g->local_size(), npad = M + l2g->num_ghosts();
    if (rank == 0)
      std::printf("Global vec size = %lld\n", (long long)l2g->global_size());

    double* b = nullptr;
    double* x_padded = exec->alloc<double>(npad);
    if (ones) { // b = A * 1
      std::vector<double> one(static_cast<size_t>(M), 1.0);
      exec->memset<double>(x_padded, 0, npad);
      exec->copy_from<double>(x_padded, *host, one.data(), M);
      b = exec->alloc<double>(M);
      l2g->update(x_padded);
      A.mult(x_padded, b);
    } else {
      b = spmv::read_petsc_binary_vector(MPI_COMM_WORLD, exec.get(), files[1]);
    }
    double* x = exec->alloc<double>(M);

    exec->synchronize();
    const auto t0 = std::chrono::steady_clock::now();
    const int num_its = spmv::cg(MPI_COMM_WORLD, *exec, A, b, x, max_its, rtol);
    const double solve_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    // r = A x - b with one more update + mult (demos/cg.cpp:75-83), norms on the host
    exec->memset<double>(x_padded, 0, npad);
    exec->copy<double>(x_padded, x, M);
    double* r = exec->alloc<double>(M);
    l2g->update(x_padded);
    A.mult(x_padded, r);
    std::vector<double> rh(static_cast<size_t>(M)), bh(static_cast<size_t>(M)), xh(static_cast<size_t>(M));
    exec->copy_to<double>(rh.data(), *host, r, M);
    exec->copy_to<double>(bh.data(), *host, b, M);
    exec->copy_to<double>(xh.data(), *host, x, M);
    double loc[2] = {0.0, 0.0}, glob[2] = {0.0, 0.0};
    for (std::int64_t i = 0; i < M; ++i) {
      loc[0] += (rh[i] - bh[i]) * (rh[i] - bh[i]);
      loc[1] += xh[i] * xh[i];
    }
    MPI_Allreduce(loc, glob, 2, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
    if (rank == 0) {
      std::printf("r.norm = %.15g\n", std::sqrt(glob[0]));
      std::printf("x.norm = %.15g in %d iterations\n", std::sqrt(glob[1]), num_its);
      std::printf("\nTimings (%d)\n----------------------------\n[1.Solve]\t%g s  (%g iterations/s)\n", size, solve_s,
                  num_its / solve_s);
    }
    exec->free(r);
    exec->free(x);
    exec->free(x_padded);
    exec->free(b);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "[rank %d] %s\n", rank, e.what());
    MPI_Finalize();
    return 1;
  }
  MPI_Finalize();
  return 0;
}
