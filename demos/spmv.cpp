// demos/spmv.cpp -- SpMV benchmark driver on the B200 executor.
//
// Protocol of the reference's driver (demos/spmv.cpp:25-140; CUDA flavour demos/spmv_cuda.cpp):
// build the operator, x_i = exp(-10 (5 (i/N - 0.5))^2) on the owned entries, one warm-up
// {update; mult}, then n_apply x {update; mult; x <- y}, report per-phase times and ||x||_2.
// Differences: the operator comes from a PETSc binary file as there OR from `--create-a N`
// (demos/CreateA.cpp:31-103, BASELINE configs[0]); phases are timed on the device with CUDA
// events recorded in stream order -- nothing here synchronises the host between phases, where
// the reference brackets every phase with MPI_Barrier and wall-clock reads (:84-103).
//
//   demo_spmv <matrix.petsc> [--symmetric] [--n-apply K] [--model 0..3]
//   demo_spmv --create-a 1000000
// One rank per process (MPI_COMM_WORLD of the bundled shim is this process unless a launcher
// installs a larger world).
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

namespace
{
void ck(int rc, const char* what)
{
  if (rc != B200_OK)
    throw std::runtime_error(std::string(what) + ": " + b200_last_error());
}
} // namespace

int main(int argc, char** argv)
{
  MPI_Init(&argc, &argv);
  int rank = 0, size = 1;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  try {
    std::string file;
    long create_a = 0;
    int n_apply = 100, model = 2; // collective_blocking: the library default (L2GMap.h:44)
    bool symmetric = false;
    for (int i = 1; i < argc; ++i) {
      if (!std::strcmp(argv[i], "--create-a") && i + 1 < argc)
        create_a = std::atol(argv[++i]);
      else if (!std::strcmp(argv[i], "--n-apply") && i + 1 < argc)
        n_apply = std::atoi(argv[++i]);
      else if (!std::strcmp(argv[i], "--model") && i + 1 < argc)
        model = std::atoi(argv[++i]);
      else if (!std::strcmp(argv[i], "--symmetric"))
        symmetric = true;
      else
        file = argv[i];
    }
    if (file.empty() && create_a <= 0)
      throw std::runtime_error("Use: demo_spmv <matrix_file> [--symmetric] | --create-a N   [--n-apply K] [--model 0..3]");

    int ndev = 0;
    ck(b200_device_count(&ndev), "device count");
    std::shared_ptr<spmv::DeviceExecutor> host = spmv::ReferenceExecutor::create();
    std::shared_ptr<spmv::B200Executor> exec = spmv::B200Executor::create(rank % std::max(ndev, 1), host);
    void* stream = nullptr;
    ck(b200_stream_create(&stream), "stream");
    exec->set_cuda_stream(static_cast<spmv::B200Executor::stream_t>(stream));

    const auto cm = static_cast<spmv::CommunicationModel>(model);
    spmv::Matrix<double> A = create_a > 0 ? create_A(MPI_COMM_WORLD, static_cast<int>(create_a), exec)
                                          : spmv::read_petsc_binary_matrix(file, MPI_COMM_WORLD, exec, symmetric, cm);
    std::shared_ptr<const spmv::L2GMap> l2g = A.col_map();
    const std::int64_t N = l2g->global_size(), M = A.rows(), r0 = l2g->global_offset();
    if (rank == 0)
      std::printf("Global vec size = %lld, %d rank(s), halo: %s\n", (long long)N, size, l2g->halo_mode());

    // x on the owned entries (demos/spmv.cpp:60-67); ghosts zero until the first update
    std::vector<double> xh(static_cast<size_t>(M));
    for (std::int64_t i = 0; i < M; ++i) {
      const double z = (double)(i + r0) / (double)N;
      xh[i] = std::exp(-10 * std::pow(5 * (z - 0.5), 2.0));
    }
    const std::int64_t npad = l2g->local_size() + l2g->num_ghosts();
    double* x = exec->alloc<double>(npad);
    double* y = exec->alloc<double>(M);
    exec->memset<double>(x, 0, npad);
    exec->copy_from<double>(x, *host, xh.data(), M);

    l2g->update(x); // warm-up (:81-82)
    A.mult(x, y);

    enum { UPDATE, SPMV, COPY, NPH };
    std::vector<void*> ev(static_cast<size_t>(n_apply) * (NPH + 1));
    for (void*& e : ev)
      ck(b200_event_create(&e, 1), "event");
    for (int i = 0; i < n_apply; ++i) {
      void** e = &ev[static_cast<size_t>(i) * (NPH + 1)];
      ck(b200_event_record(e[0], stream), "record");
      l2g->update(x);
      ck(b200_event_record(e[1], stream), "record");
      A.mult(x, y);
      ck(b200_event_record(e[2], stream), "record");
      exec->copy<double>(x, y, M);
      ck(b200_event_record(e[3], stream), "record");
    }
    exec->synchronize();
    double t[NPH] = {0, 0, 0};
    for (int i = 0; i < n_apply; ++i)
      for (int p = 0; p < NPH; ++p) {
        float ms = 0.f;
        ck(b200_event_elapsed_ms(ev[static_cast<size_t>(i) * (NPH + 1) + p], ev[static_cast<size_t>(i) * (NPH + 1) + p + 1], &ms),
           "elapsed");
        t[p] += ms * 1e-3;
      }
    for (void* e : ev)
      b200_event_destroy(e);

    // ||x||_2 (:105-108): the owned part comes home, the sum over ranks goes through the shim
    exec->copy_to<double>(xh.data(), *host, x, M);
    double local = 0.0, glob = 0.0;
    for (double v : xh)
      local += v * v;
    MPI_Allreduce(&local, &glob, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
    double tmin[NPH], tmax[NPH];
    MPI_Allreduce(t, tmin, NPH, MPI_DOUBLE, MPI_MIN, MPI_COMM_WORLD);
    MPI_Allreduce(t, tmax, NPH, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);
    if (rank == 0) {
      const char* names[NPH] = {"2.SpUpdate", "3.SpMV", "4.Copy"};
      std::printf("\nTimings (%d)\n----------------------------\n", size);
      for (int p = 0; p < NPH; ++p)
        std::printf("[%s]\t%g\t%g   (device seconds, min / max over ranks, %d applications)\n", names[p], tmin[p], tmax[p],
                    n_apply);
      std::printf("----------------------------\nnorm = %.15g\n", std::sqrt(glob));
      const double step = (tmax[UPDATE] + tmax[SPMV]) / n_apply;
      std::printf("update + mult: %.3f us per application, %lld stored entries on rank 0\n", step * 1e6,
                  (long long)A.local_block()->non_zeros());
    }
    exec->free(x);
    exec->free(y);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "[rank %d] %s\n", rank, e.what());
    MPI_Finalize();
    return 1;
  }
  MPI_Finalize();
  return 0;
}
