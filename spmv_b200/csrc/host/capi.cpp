// capi.cpp -- include/spmv_capi.h over the C++ host layer.
#include "spmv_capi.h"

#include "../../../demos/CreateA.h"
#include "../../../demos/CreateLaplace3D.h"
#include "../../../demos/CreatePowerLaw.h"
#include "host_util.h"
#include <spmv/spmv.h>

#include <algorithm>
#include <cstring>
#include <string>

using namespace spmv;

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
    g_err = "unknown C++ exception";
  }
  return -1;
}

using ExecPtr = std::shared_ptr<DeviceExecutor>;
using MapPtr = std::shared_ptr<L2GMap>;

ExecPtr& as_exec(void* h)
{
  if (!h)
    throw std::runtime_error("executor handle is NULL");
  return *static_cast<ExecPtr*>(h);
}
L2GMap& as_map(void* h)
{
  if (!h)
    throw std::runtime_error("map handle is NULL");
  return **static_cast<MapPtr*>(h);
}
Matrix<double>& as_mat(void* h)
{
  if (!h)
    throw std::runtime_error("matrix handle is NULL");
  return *static_cast<Matrix<double>*>(h);
}
Matrix<float>& as_mat32(void* h)
{
  if (!h)
    throw std::runtime_error("matrix handle is NULL");
  return *static_cast<Matrix<float>*>(h);
}
CommunicationModel as_cm(int cm)
{
  if (cm < 0 || cm > 7)
    throw std::runtime_error("bad communication model");
  return static_cast<CommunicationModel>(cm);
}
const CSRMatrix<double>* block_of(Matrix<double>& m, int which)
{
  const SubMatrix<double>* b = which == 0 ? m.local_block() : m.remote_block();
  return dynamic_cast<const CSRMatrix<double>*>(b);
}
template <typename V, typename O>
void copy_out(const V& v, void* out, int64_t cap, int64_t* n)
{
  if (n)
    *n = static_cast<int64_t>(v.size());
  const int64_t m = std::min<int64_t>(cap, static_cast<int64_t>(v.size()));
  O* o = static_cast<O*>(out);
  for (int64_t i = 0; i < m; ++i)
    o[i] = static_cast<O>(v[i]);
}
} // namespace

extern "C" {

const char* spmvh_last_error(void) { return g_err.c_str(); }

int spmvh_exec_create_host(void** exec)
{
  return guarded([&] { *exec = new ExecPtr(ReferenceExecutor::create()); });
}
int spmvh_exec_create_b200(void** exec, int device_id)
{
  return guarded([&] {
    ExecPtr host(ReferenceExecutor::create());
    *exec = new ExecPtr(B200Executor::create(device_id, host));
  });
}
int spmvh_exec_destroy(void* exec)
{
  return guarded([&] { delete static_cast<ExecPtr*>(exec); });
}
int spmvh_exec_set_stream(void* exec, void* cuda_stream)
{
  return guarded([&] {
    auto* gpu = dynamic_cast<B200Executor*>(as_exec(exec).get());
    if (!gpu)
      throw std::runtime_error("not a B200 executor");
    gpu->set_cuda_stream(static_cast<B200Executor::stream_t>(cuda_stream));
  });
}
int spmvh_exec_synchronize(void* exec)
{
  return guarded([&] { as_exec(exec)->synchronize(); });
}
int spmvh_exec_alloc(void* exec, size_t num_bytes, void** d_ptr)
{
  return guarded([&] { *d_ptr = as_exec(exec)->alloc<char>(num_bytes); });
}
int spmvh_exec_free(void* exec, void* d_ptr)
{
  return guarded([&] { as_exec(exec)->free(d_ptr); });
}
int spmvh_exec_copy_from_host(void* exec, void* d_dst, const void* h_src, size_t num_bytes)
{
  return guarded([&] {
    ExecPtr& e = as_exec(exec);
    e->copy_from<char>(static_cast<char*>(d_dst), e->get_host(), static_cast<const char*>(h_src), num_bytes);
  });
}
int spmvh_exec_copy_to_host(void* exec, void* h_dst, const void* d_src, size_t num_bytes)
{
  return guarded([&] {
    ExecPtr& e = as_exec(exec);
    e->copy_to<char>(static_cast<char*>(h_dst), e->get_host(), static_cast<const char*>(d_src), num_bytes);
  });
}

int spmvh_l2g_create(void** map, void* exec, int64_t local_size, const int64_t* ghosts, int64_t num_ghosts, int cm)
{
  return guarded([&] {
    std::vector<std::int64_t> g(ghosts, ghosts + num_ghosts);
    *map = new MapPtr(std::make_shared<L2GMap>(MPI_COMM_WORLD, local_size, g, as_exec(exec), as_cm(cm)));
  });
}
int spmvh_l2g_destroy(void* map)
{
  return guarded([&] { delete static_cast<MapPtr*>(map); });
}
int spmvh_l2g_sizes(void* map, int64_t out[8])
{
  return guarded([&] {
    L2GMap& m = as_map(map);
    out[0] = m.local_size();
    out[1] = m.num_ghosts();
    out[2] = m.global_size();
    out[3] = m.global_offset();
    out[4] = static_cast<int64_t>(m.neighbours().size());
    out[5] = static_cast<int64_t>(m.indexbuf().size());
    out[6] = m.rank();
    out[7] = m.overlapping() ? 1 : 0;
  });
}
int spmvh_l2g_get(void* map, int what, void* out, int64_t cap, int64_t* n)
{
  return guarded([&] {
    L2GMap& m = as_map(map);
    switch (what) {
    case 0:
      return copy_out<std::vector<std::int64_t>, int64_t>(m.ranges(), out, cap, n);
    case 1:
      return copy_out<std::vector<std::int64_t>, int64_t>(m.ghosts(), out, cap, n);
    case 2:
      return copy_out<std::vector<int>, int32_t>(m.neighbours(), out, cap, n);
    case 3:
      return copy_out<std::vector<std::int32_t>, int32_t>(m.send_count(), out, cap, n);
    case 4:
      return copy_out<std::vector<std::int32_t>, int32_t>(m.recv_count(), out, cap, n);
    case 5:
      return copy_out<std::vector<std::int32_t>, int32_t>(m.send_offset(), out, cap, n);
    case 6:
      return copy_out<std::vector<std::int32_t>, int32_t>(m.recv_offset(), out, cap, n);
    case 7:
      return copy_out<std::vector<std::int32_t>, int32_t>(m.indexbuf(), out, cap, n);
    default:
      throw std::runtime_error("spmvh_l2g_get: bad selector");
    }
  });
}
int spmvh_l2g_global_to_local(void* map, int64_t global_index, int32_t* local_index)
{
  return guarded([&] { *local_index = as_map(map).global_to_local(global_index); });
}
int spmvh_l2g_update_f64(void* map, double* d_vec)
{
  return guarded([&] { as_map(map).update(d_vec); });
}
int spmvh_l2g_update_finalise_f64(void* map, double* d_vec)
{
  return guarded([&] { as_map(map).update_finalise(d_vec); });
}
int spmvh_l2g_reverse_update_f64(void* map, double* d_vec)
{
  return guarded([&] { as_map(map).reverse_update(d_vec); });
}
int spmvh_l2g_allreduce_f64(void* map, double* d_val, int32_t n, int use_nccl)
{
  return guarded([&] { as_map(map).allreduce_sum(d_val, n, use_nccl != 0); });
}

int spmvh_matrix_create(void** mat, void* exec, const int32_t* rowptr, const int32_t* colind, const double* values,
                        int64_t nrows_local, int64_t ncols_local, const int64_t* row_ghosts, int64_t num_row_ghosts,
                        const int64_t* col_ghosts, int64_t num_col_ghosts, int symmetric, int cm)
{
  return guarded([&] {
    std::vector<std::int64_t> rg(row_ghosts, row_ghosts + num_row_ghosts);
    std::vector<std::int64_t> cg(col_ghosts, col_ghosts + num_col_ghosts);
    *mat = Matrix<double>::create_matrix(MPI_COMM_WORLD, as_exec(exec), rowptr, colind, values, nrows_local,
                                         ncols_local, rg, cg, symmetric != 0, as_cm(cm));
  });
}
int spmvh_matrix_read_petsc(void** mat, void* exec, const char* filename, int symmetric, int cm)
{
  return guarded([&] {
    *mat = new Matrix<double>(
        read_petsc_binary_matrix(filename, MPI_COMM_WORLD, as_exec(exec), symmetric != 0, as_cm(cm)));
  });
}
int spmvh_matrix_create_A(void** mat, void* exec, int N)
{
  return guarded([&] { *mat = new Matrix<double>(create_A(MPI_COMM_WORLD, N, as_exec(exec))); });
}
int spmvh_matrix_create_laplace3d(void** mat, void* exec, int nx, int ny, int64_t nz_global, int stencil,
                                  int symmetric, int cm)
{
  return guarded([&] {
    *mat = new Matrix<double>(
        create_laplace3d(MPI_COMM_WORLD, as_exec(exec), nx, ny, nz_global, stencil, symmetric != 0, as_cm(cm)));
  });
}
int spmvh_matrix_destroy(void* mat)
{
  return guarded([&] { delete static_cast<Matrix<double>*>(mat); });
}
int spmvh_matrix_info(void* mat, int64_t out[8])
{
  return guarded([&] {
    Matrix<double>& m = as_mat(mat);
    out[0] = m.rows();
    out[1] = m.cols();
    out[2] = m.non_zeros();
    out[3] = m.symmetric() ? 1 : 0;
    out[4] = static_cast<int64_t>(m.format_size());
    out[5] = m.remote_block() ? 1 : 0;
    out[6] = m.local_block() ? m.local_block()->non_zeros() : 0;
    out[7] = m.remote_block() ? m.remote_block()->non_zeros() : 0;
  });
}
int spmvh_matrix_col_map(void* mat, void** map)
{
  return guarded([&] { *map = new MapPtr(std::const_pointer_cast<L2GMap>(as_mat(mat).col_map())); });
}
int spmvh_matrix_row_map(void* mat, void** map)
{
  return guarded([&] { *map = new MapPtr(as_mat(mat).row_map()); });
}
int spmvh_matrix_block_dims(void* mat, int which, int64_t dims[4])
{
  return guarded([&] {
    const CSRMatrix<double>* b = block_of(as_mat(mat), which);
    dims[0] = b ? b->rows() : 0;
    dims[1] = b ? b->cols() : 0;
    dims[2] = b ? b->non_zeros() : -1;
    dims[3] = (b && b->diagonal()) ? 1 : 0;
  });
}
int spmvh_matrix_block_copy(void* mat, int which, int32_t* rowptr, int32_t* colind, double* values, double* diagonal)
{
  return guarded([&] {
    Matrix<double>& m = as_mat(mat);
    const CSRMatrix<double>* b = block_of(m, which);
    if (!b)
      throw std::runtime_error("no such block");
    const std::shared_ptr<DeviceExecutor> exec = m.executor();
    const DeviceExecutor& host = exec->get_host();
    if (b->non_zeros() > 0) {
      if (rowptr)
        exec->copy_to<std::int32_t>(rowptr, host, b->rowptr(), static_cast<size_t>(b->rows()) + 1);
      if (colind)
        exec->copy_to<std::int32_t>(colind, host, b->colind(), b->non_zeros());
      if (values)
        exec->copy_to<double>(values, host, b->values(), b->non_zeros());
    } else if (rowptr) {
      std::memset(rowptr, 0, sizeof(std::int32_t) * (static_cast<size_t>(b->rows()) + 1));
    }
    if (diagonal && b->diagonal())
      exec->copy_to<double>(diagonal, host, b->diagonal(), b->rows());
  });
}
int spmvh_matrix_mult_f64(void* mat, double* d_x, double* d_y)
{
  return guarded([&] { as_mat(mat).mult(d_x, d_y); });
}
int spmvh_matrix_mult_host_f64(void* mat, const double* x_host, double* y_host)
{
  return guarded([&] { as_mat(mat).mult_host(x_host, y_host); });
}
int spmvh_matrix_tune(void* mat, int which, int variant, int tile_bytes, int stages, int consumer_warps,
                      int ctas_per_sm)
{
  return guarded([&] {
    const CSRMatrix<double>* b = block_of(as_mat(mat), which);
    if (!b || !b->op().plan())
      throw std::runtime_error("block has no B200 plan");
    detail::bump_structure_epoch(); // the tile table a captured CG graph points at is about to be replaced
    detail::check(b200_csr_plan_tune(b->op().plan(), b->rowptr(), variant, tile_bytes, stages, consumer_warps,
                                     ctas_per_sm, nullptr),
                  "plan tune");
  });
}

// ---- fp32: the float instantiations of the same classes (Matrix.cpp:565, L2GMap.cpp:1001,1007,1012,
// csr_kernels.cpp:62-63 in the reference) ----
int spmvh_l2g_update_f32(void* map, float* d_vec)
{
  return guarded([&] { as_map(map).update(d_vec); });
}
int spmvh_l2g_update_finalise_f32(void* map, float* d_vec)
{
  return guarded([&] { as_map(map).update_finalise(d_vec); });
}
int spmvh_l2g_reverse_update_f32(void* map, float* d_vec)
{
  return guarded([&] { as_map(map).reverse_update(d_vec); });
}
int spmvh_matrix_create_f32(void** mat, void* exec, const int32_t* rowptr, const int32_t* colind, const float* values,
                            int64_t nrows_local, int64_t ncols_local, const int64_t* row_ghosts,
                            int64_t num_row_ghosts, const int64_t* col_ghosts, int64_t num_col_ghosts, int symmetric,
                            int cm)
{
  return guarded([&] {
    std::vector<std::int64_t> rg(row_ghosts, row_ghosts + num_row_ghosts);
    std::vector<std::int64_t> cg(col_ghosts, col_ghosts + num_col_ghosts);
    *mat = Matrix<float>::create_matrix(MPI_COMM_WORLD, as_exec(exec), rowptr, colind, values, nrows_local,
                                        ncols_local, rg, cg, symmetric != 0, as_cm(cm));
  });
}
int spmvh_matrix_destroy_f32(void* mat)
{
  return guarded([&] { delete static_cast<Matrix<float>*>(mat); });
}
int spmvh_matrix_info_f32(void* mat, int64_t out[8])
{
  return guarded([&] {
    Matrix<float>& m = as_mat32(mat);
    out[0] = m.rows();
    out[1] = m.cols();
    out[2] = m.non_zeros();
    out[3] = m.symmetric() ? 1 : 0;
    out[4] = static_cast<int64_t>(m.format_size());
    out[5] = m.remote_block() ? 1 : 0;
    out[6] = m.local_block() ? m.local_block()->non_zeros() : 0;
    out[7] = m.remote_block() ? m.remote_block()->non_zeros() : 0;
  });
}
int spmvh_matrix_col_map_f32(void* mat, void** map)
{
  return guarded([&] { *map = new MapPtr(std::const_pointer_cast<L2GMap>(as_mat32(mat).col_map())); });
}
int spmvh_matrix_mult_f32(void* mat, float* d_x, float* d_y)
{
  return guarded([&] { as_mat32(mat).mult(d_x, d_y); });
}
int spmvh_matrix_mult_host_f32(void* mat, const float* x_host, float* y_host)
{
  return guarded([&] { as_mat32(mat).mult_host(x_host, y_host); });
}

int spmvh_cg(void* exec, void* mat, const double* d_b, double* d_x, int kmax, double rtol, double* hist,
             int hist_cap, int* iterations)
{
  return guarded([&] {
    auto* gpu = dynamic_cast<B200Executor*>(as_exec(exec).get());
    if (!gpu)
      throw std::runtime_error("spmv::cg needs the B200 executor (no CPU compute path)");
    std::vector<double> h;
    const int k = cg(MPI_COMM_WORLD, *gpu, as_mat(mat), d_b, d_x, kmax, rtol, hist ? &h : nullptr);
    if (iterations)
      *iterations = k;
    if (hist)
      for (int i = 0; i < std::min<int>(hist_cap, static_cast<int>(h.size())); ++i)
        hist[i] = h[i];
  });
}

int spmvh_cg_phases(void* exec, void* mat, const double* d_b, double* d_x, int kmax, double rtol, double out_us[4],
                    int* iterations)
{
  return guarded([&] {
    auto* gpu = dynamic_cast<B200Executor*>(as_exec(exec).get());
    if (!gpu)
      throw std::runtime_error("spmv::cg needs the B200 executor (no CPU compute path)");
    CgPhaseTimes t;
    const int k = cg(MPI_COMM_WORLD, *gpu, as_mat(mat), d_b, d_x, kmax, rtol, nullptr, &t);
    if (iterations)
      *iterations = k;
    out_us[0] = t.us_update;
    out_us[1] = t.us_local;
    out_us[2] = t.us_remote;
    out_us[3] = t.us_vector;
  });
}
int spmvh_l2g_halo_mode(void* map, char* out, int cap)
{
  return guarded([&] {
    const std::string both = std::string(as_map(map).halo_mode()) + "/" + as_map(map).halo_mode_cg();
    const char* m = both.c_str();
    std::strncpy(out, m, cap > 0 ? cap - 1 : 0);
    if (cap > 0)
      out[cap - 1] = 0;
  });
}

int spmvh_write_petsc_matrix(const char* filename, int32_t nrows, int32_t ncols, const int32_t* rowptr,
                             const int32_t* colind, const double* values)
{
  return guarded([&] { write_petsc_binary_matrix(filename, nrows, ncols, rowptr, colind, values); });
}
int spmvh_write_petsc_vector(const char* filename, int32_t n, const double* values)
{
  return guarded([&] { write_petsc_binary_vector(filename, n, values); });
}
int spmvh_read_petsc_vector(void* exec, const char* filename, double** d_vec, int64_t* n_local)
{
  return guarded([&] {
    int rank = 0, size = 1;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    *d_vec = read_petsc_binary_vector(MPI_COMM_WORLD, as_exec(exec).get(), filename);
    // local length: same split as the reader
    std::FILE* f = std::fopen(filename, "rb");
    unsigned char h[8] = {0};
    if (f) {
      if (std::fread(h, 1, 8, f) != 8)
        h[4] = h[5] = h[6] = h[7] = 0;
      std::fclose(f);
    }
    const std::int64_t n = (std::int64_t(h[4]) << 24) | (std::int64_t(h[5]) << 16) | (std::int64_t(h[6]) << 8) | h[7];
    const auto ranges = detail::owner_ranges(size, n);
    *n_local = ranges[rank + 1] - ranges[rank];
  });
}

namespace
{
PowerLawSpec as_spec(int64_t n, uint64_t seed, double alpha, double xmin, int64_t max_len, int64_t band, double p_global)
{
  PowerLawSpec s;
  s.n = n, s.seed = seed, s.alpha = alpha, s.xmin = xmin, s.max_len = max_len, s.band = band, s.p_global = p_global;
  return s;
}
} // namespace

int spmvh_powerlaw_counts(int64_t n, uint64_t seed, double alpha, double xmin, int64_t max_len, int64_t band,
                          double p_global, int64_t r0, int64_t r1, int32_t* counts)
{
  return guarded([&] { powerlaw_row_counts(as_spec(n, seed, alpha, xmin, max_len, band, p_global), r0, r1, counts); });
}
int spmvh_powerlaw_fill(int64_t n, uint64_t seed, double alpha, double xmin, int64_t max_len, int64_t band,
                        double p_global, int64_t r0, int64_t r1, const int64_t* rowstart, int64_t* col, double* val)
{
  return guarded(
      [&] { powerlaw_rows_fill(as_spec(n, seed, alpha, xmin, max_len, band, p_global), r0, r1, rowstart, col, val); });
}
int spmvh_write_powerlaw_petsc(const char* filename, int64_t n, uint64_t seed, double alpha, double xmin,
                               int64_t max_len, int64_t band, double p_global, int64_t* nnz)
{
  return guarded([&] {
    const std::int64_t m = write_powerlaw_petsc(filename, as_spec(n, seed, alpha, xmin, max_len, band, p_global));
    if (nnz)
      *nnz = m;
  });
}

} // extern "C"
