// oracle/ref/ref_driver.cpp -- TEST INFRASTRUCTURE.
// C entry points around the UNMODIFIED reference host code (L2GMap.cpp, Matrix.cpp,
// csr_matrix.cpp, reference_executor.cpp, cg.cpp, read_petsc.cpp, demos/CreateA.cpp),
// compiled where it lies under /root/reference against the header shims of
// oracle/shims (mini-Eigen, cblas) and the thread-rank MPI shim.  Built with
// -fno-access-control so the private halo plan can be read without touching the
// reference sources.  Ranks are threads: every call acts on the calling thread's
// MPI_COMM_WORLD (drive them with ref_run_ranks).
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <mpi.h>
#include <cblas.h>

#include "L2GMap.h"
#include "Matrix.h"
#include "cg.h"
#include "csr_matrix.h"
#include "read_petsc.h"
#include "reference_executor.h"

spmv::Matrix<double> create_A(MPI_Comm comm, int N); // demos/CreateA.cpp

namespace
{
thread_local std::string g_err;
template <typename F>
int guarded(F&& f)
{
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
  } catch (...) {
    g_err = "unknown exception";
  }
  return -1;
}
struct RefMat {
  std::shared_ptr<spmv::ReferenceExecutor> exec;
  std::unique_ptr<spmv::Matrix<double>> mat;
};
spmv::CommunicationModel as_cm(int cm) { return static_cast<spmv::CommunicationModel>(cm); }
template <typename V, typename O>
void copy_out(const V& v, void* out, int64_t cap, int64_t* n)
{
  *n = static_cast<int64_t>(v.size());
  O* o = static_cast<O*>(out);
  for (int64_t i = 0; i < std::min<int64_t>(cap, *n); ++i)
    o[i] = static_cast<O>(v[i]);
}
const spmv::CSRMatrix<double>* block_of(RefMat* m, int which)
{
  spmv::SubMatrix<double>* b = which == 0 ? m->mat->_mat_local.get() : m->mat->_mat_remote.get();
  return dynamic_cast<const spmv::CSRMatrix<double>*>(b);
}
} // namespace

extern "C" {

const char* ref_last_error(void) { return g_err.c_str(); }
int ref_run_ranks(int n, void (*fn)(void*), void* arg) { return spmv_shim_run_ranks(n, fn, arg); }
int ref_world_rank(void)
{
  int r = 0;
  MPI_Comm_rank(MPI_COMM_WORLD, &r);
  return r;
}
int ref_world_size(void)
{
  int s = 1;
  MPI_Comm_size(MPI_COMM_WORLD, &s);
  return s;
}
double ref_allreduce_sum(double v)
{
  double out = 0;
  MPI_Allreduce(&v, &out, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
  return out;
}

// Matrix<double>::create_matrix (Matrix.cpp:164-480) on a ReferenceExecutor
int ref_matrix_create(void** h, const int32_t* rowptr, const int32_t* colind, const double* values,
                      int64_t nrows_local, int64_t ncols_local, const int64_t* row_ghosts, int64_t nrg,
                      const int64_t* col_ghosts, int64_t ncg, int symmetric, int cm)
{
  return guarded([&] {
    auto m = new RefMat;
    m->exec = spmv::ReferenceExecutor::create();
    std::vector<int64_t> rg(row_ghosts, row_ghosts + nrg), cg(col_ghosts, col_ghosts + ncg);
    m->mat.reset(spmv::Matrix<double>::create_matrix(MPI_COMM_WORLD, m->exec, rowptr, colind, values, nrows_local,
                                                     ncols_local, rg, cg, symmetric != 0, as_cm(cm)));
    *h = m;
  });
}
// read_petsc_binary_matrix (read_petsc.cpp:39-228)
int ref_matrix_read_petsc(void** h, const char* filename, int symmetric, int cm)
{
  return guarded([&] {
    auto m = new RefMat;
    m->exec = spmv::ReferenceExecutor::create();
    m->mat.reset(new spmv::Matrix<double>(
        spmv::read_petsc_binary_matrix(filename, MPI_COMM_WORLD, m->exec, symmetric != 0, as_cm(cm))));
    *h = m;
  });
}
// create_A (demos/CreateA.cpp:31-103)
int ref_matrix_create_A(void** h, int N)
{
  return guarded([&] {
    auto m = new RefMat;
    m->exec = spmv::ReferenceExecutor::create();
    m->mat.reset(new spmv::Matrix<double>(create_A(MPI_COMM_WORLD, N)));
    *h = m;
  });
}
int ref_matrix_destroy(void* h)
{
  return guarded([&] { delete static_cast<RefMat*>(h); });
}
int ref_matrix_info(void* h, int64_t out[8])
{
  return guarded([&] {
    auto* m = static_cast<RefMat*>(h);
    out[0] = m->mat->rows();
    out[1] = m->mat->cols();
    out[2] = m->mat->non_zeros();
    out[3] = m->mat->symmetric();
    out[4] = static_cast<int64_t>(m->mat->format_size());
    out[5] = m->mat->_mat_remote ? 1 : 0;
    out[6] = m->mat->_mat_local ? m->mat->_mat_local->non_zeros() : 0;
    out[7] = m->mat->_mat_remote ? m->mat->_mat_remote->non_zeros() : 0;
  });
}
int ref_matrix_block_dims(void* h, int which, int64_t dims[4])
{
  return guarded([&] {
    const auto* b = block_of(static_cast<RefMat*>(h), which);
    dims[0] = b ? b->rows() : 0;
    dims[1] = b ? b->cols() : 0;
    dims[2] = b ? b->non_zeros() : -1;
    dims[3] = (b && b->diagonal()) ? 1 : 0;
  });
}
int ref_matrix_block_copy(void* h, int which, int32_t* rowptr, int32_t* colind, double* values, double* diagonal)
{
  return guarded([&] {
    const auto* b = block_of(static_cast<RefMat*>(h), which);
    if (!b)
      throw std::runtime_error("no such block");
    if (b->non_zeros() > 0) {
      std::memcpy(rowptr, b->rowptr(), sizeof(int32_t) * (b->rows() + 1));
      std::memcpy(colind, b->colind(), sizeof(int32_t) * b->non_zeros());
      std::memcpy(values, b->values(), sizeof(double) * b->non_zeros());
    } else {
      std::memset(rowptr, 0, sizeof(int32_t) * (b->rows() + 1));
    }
    if (diagonal && b->diagonal())
      std::memcpy(diagonal, b->diagonal(), sizeof(double) * b->rows());
  });
}
// The private plan of L2GMap (L2GMap.h:113-131,156).  which_map: 0 column map, 1 row map.
// what: 0 _ranges 1 _ghosts 2 _neighbours 3 _send_count 4 _recv_count 5 _send_offset
//       6 _recv_offset 7 _indexbuf[0.._num_indices)
int ref_l2g_get(void* h, int which_map, int what, void* out, int64_t cap, int64_t* n)
{
  return guarded([&] {
    auto* m = static_cast<RefMat*>(h);
    const spmv::L2GMap& g = which_map == 0 ? *m->mat->_col_map : *m->mat->_row_map;
    switch (what) {
    case 0:
      return copy_out<std::vector<int64_t>, int64_t>(g._ranges, out, cap, n);
    case 1:
      return copy_out<std::vector<int64_t>, int64_t>(g._ghosts, out, cap, n);
    case 2:
      return copy_out<std::vector<int>, int32_t>(g._neighbours, out, cap, n);
    case 3:
      return copy_out<std::vector<int32_t>, int32_t>(g._send_count, out, cap, n);
    case 4:
      return copy_out<std::vector<int32_t>, int32_t>(g._recv_count, out, cap, n);
    case 5:
      return copy_out<std::vector<int32_t>, int32_t>(g._send_offset, out, cap, n);
    case 6:
      return copy_out<std::vector<int32_t>, int32_t>(g._recv_offset, out, cap, n);
    case 7: {
      std::vector<int32_t> idx(g._indexbuf, g._indexbuf + g._num_indices);
      return copy_out<std::vector<int32_t>, int32_t>(idx, out, cap, n);
    }
    default:
      throw std::runtime_error("bad selector");
    }
  });
}
int ref_l2g_sizes(void* h, int which_map, int64_t out[8])
{
  return guarded([&] {
    auto* m = static_cast<RefMat*>(h);
    const spmv::L2GMap& g = which_map == 0 ? *m->mat->_col_map : *m->mat->_row_map;
    out[0] = g.local_size();
    out[1] = g.num_ghosts();
    out[2] = g.global_size();
    out[3] = g.global_offset();
    out[4] = static_cast<int64_t>(g._neighbours.size());
    out[5] = g._num_indices;
    out[6] = g.rank();
    out[7] = g.overlapping() ? 1 : 0;
  });
}
int ref_l2g_global_to_local(void* h, int64_t g, int32_t* out)
{
  return guarded([&] { *out = static_cast<RefMat*>(h)->mat->_col_map->global_to_local(g); });
}
// L2GMap::update (L2GMap.cpp:868-896) and Matrix::mult (Matrix.cpp:131-141)
int ref_update(void* h, double* x)
{
  return guarded([&] { static_cast<RefMat*>(h)->mat->_col_map->update(x); });
}
int ref_mult(void* h, double* x, double* y)
{
  return guarded([&] { static_cast<RefMat*>(h)->mat->mult(x, y); });
}
// L2GMap::reverse_update (L2GMap.cpp:908-959; only the collective_blocking variant is safe to run:
// the p2p one writes 2*nbrs requests into a _req sized nbrs)
int ref_reverse_update(void* h, double* x)
{
  return guarded([&] { static_cast<RefMat*>(h)->mat->_col_map->reverse_update(x); });
}
// The timed loop of demos/spmv.cpp:81-103 on this thread-rank: warm-up, then n_apply x {barrier; update;
// barrier; mult; barrier; x <- y}, wall-clock per phase accumulated like the demo's `timings` map
// (seconds: t[0] update, t[1] mult, t[2] copy), and the final ||x||_2 of demos/spmv.cpp:105-108.
int ref_demo_spmv_loop(void* h, double* x, double* y, int n_apply, double t[3], double* xnorm)
{
  return guarded([&] {
    auto* m = static_cast<RefMat*>(h);
    const int M = m->mat->rows();
    auto l2g = m->mat->col_map();
    using clk = std::chrono::steady_clock;
    t[0] = t[1] = t[2] = 0.0;
    l2g->update(x);
    m->mat->mult(x, y);
    for (int i = 0; i < n_apply; ++i) {
      MPI_Barrier(MPI_COMM_WORLD);
      auto a = clk::now();
      l2g->update(x);
      t[0] += std::chrono::duration<double>(clk::now() - a).count();
      MPI_Barrier(MPI_COMM_WORLD);
      a = clk::now();
      m->mat->mult(x, y);
      t[1] += std::chrono::duration<double>(clk::now() - a).count();
      MPI_Barrier(MPI_COMM_WORLD);
      a = clk::now();
      m->exec->copy<double>(x, y, M);
      t[2] += std::chrono::duration<double>(clk::now() - a).count();
    }
    double local = cblas_ddot(M, x, 1, x, 1), glob = 0.0;
    MPI_Allreduce(&local, &glob, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
    *xnorm = std::sqrt(glob);
  });
}
// spmv::cg (cg.cpp:21-98)
int ref_cg(void* h, const double* b, double* x, int kmax, double rtol, int* iterations)
{
  return guarded([&] {
    auto* m = static_cast<RefMat*>(h);
    *iterations = spmv::cg(MPI_COMM_WORLD, *m->exec, *m->mat, b, x, kmax, rtol);
  });
}

} // extern "C"
