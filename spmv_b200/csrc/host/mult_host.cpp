// mult_host.cpp -- y_host = A * x_host with the vectors in HOST memory.
//
// The reference's user does this in three blocking steps: exec->copy_from(x), update +
// mult, exec->copy_to(y) (demos/spmv_cuda.cpp:71-110).  Over PCIe the two copies cost
// 20x the kernel, so here they are pipelined with it: the plan's tiles are consecutive
// row ranges, a chunk of rows only needs the columns up to its largest column index,
// and its y rows are final as soon as no later chunk can scatter into them.  x goes up
// in pieces on one stream, the tile ranges run on the executor's stream as their
// piece of x lands, finished pieces of y come down on a third stream -- both PCIe
// directions busy at once.  With neighbours and an overlapping communication model
// (local + remote block) the pages of x that the halo packs go up first, the exchange
// and the remote block run while the first chunk is still on its way, and the local
// chunks then add on top (beta = 1; same sum per row as Matrix::mult, operands swapped).
// Blocking models with ghosts (one merged block) fall back to the three blocking steps.
#include "Matrix.h"

#include "L2GMap.h"
#include "b200_executor.h"
#include "csr_matrix.h"
#include "host_util.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <vector>

namespace spmv
{

struct HostPipe {
  std::shared_ptr<DeviceExecutor> exec;
  double* x_dev = nullptr;
  double* y_dev = nullptr;
  void* h2d = nullptr;
  void* d2h = nullptr;
  int nchunks = 0;
  std::vector<std::int32_t> tile_bounds, row_bounds, x_need, y_ready;
  std::vector<void*> ev_x, ev_k, ev_y; // x piece landed, tile range done, y piece on the host (trace only)
  void* ev_start = nullptr;
  bool pipelined = false;
  bool halo = false; // neighbours: boundary pages first, exchange + remote block, then the chunks add
  std::vector<std::int32_t> edge; // [begin, end) pairs: the pages of x the halo packs
  void* ev_edge = nullptr;
  bool trace = false; // B200_HOST_TRACE=1: timed events, one timeline line per call on stderr
  ~HostPipe()
  {
    for (void* e : ev_x)
      b200_event_destroy(e);
    for (void* e : ev_k)
      b200_event_destroy(e);
    for (void* e : ev_y)
      b200_event_destroy(e);
    if (ev_start)
      b200_event_destroy(ev_start);
    if (ev_edge)
      b200_event_destroy(ev_edge);
    if (h2d)
      b200_stream_destroy(h2d);
    if (d2h)
      b200_stream_destroy(d2h);
    if (exec) {
      exec->free(x_dev);
      exec->free(y_dev);
    }
  }
};

namespace
{
constexpr std::int32_t PAGE = 4096 / sizeof(double); // vector elements per host page
void ck(int rc, const char* what) { detail::check(rc, what); }
} // namespace

template <typename T>
void Matrix<T>::mult_host(const T*, T*) const
{
  throw std::runtime_error("Matrix::mult_host is provided for double");
}

template <>
void Matrix<double>::mult_host(const double* x_host, double* y_host) const
{
  auto gpu = std::dynamic_pointer_cast<B200Executor>(_exec);
  if (!gpu)
    throw std::runtime_error("Matrix::mult_host needs the B200 executor (no CPU compute path)");
  ck(b200_set_device(gpu->get_device_id()), "cudaSetDevice");
  const std::int32_t nloc = _col_map->local_size(), nghost = _col_map->num_ghosts();
  const std::int32_t nrows = rows();
  void* S = gpu->get_cuda_stream();

  if (!_host_pipe) {
    auto hp = std::make_shared<HostPipe>();
    hp->exec = _exec;
    hp->x_dev = _exec->alloc<double>(static_cast<size_t>(nloc) + nghost + 1);
    hp->y_dev = _exec->alloc<double>(static_cast<size_t>(nrows) + 1);
    const auto* blk = dynamic_cast<const CSRMatrix<double>*>(_mat_local.get());
    const bool one_block = !_mat_remote || _mat_remote->non_zeros() == 0;
    const bool alone = nghost == 0 && one_block && _col_map->indexbuf().empty();
    hp->halo = !alone && _col_map->overlapping() && _mat_remote != nullptr;
    if (hp->halo) {
      // pages holding the entries the neighbours want; give up if that is a large part of x
      std::vector<std::int32_t> pages;
      for (std::int32_t i : _col_map->indexbuf())
        pages.push_back(i / PAGE);
      std::sort(pages.begin(), pages.end());
      pages.erase(std::unique(pages.begin(), pages.end()), pages.end());
      if (static_cast<std::int64_t>(pages.size()) * PAGE * 4 > nloc)
        hp->halo = false;
      for (size_t i = 0; hp->halo && i < pages.size();) {
        size_t j = i;
        while (j + 1 < pages.size() && pages[j + 1] == pages[j] + 1)
          ++j;
        hp->edge.push_back(pages[i] * PAGE);
        hp->edge.push_back(std::min<std::int32_t>((pages[j] + 1) * PAGE, nloc));
        i = j + 1;
      }
    }
    hp->pipelined = blk && blk->op().plan() && blk->non_zeros() > 0 && (alone || hp->halo);
    if (hp->pipelined) {
      // Equal chunks of tiles; 8 is the measured optimum on PCIe gen5 (profiles/r01_e2e_pipeline.md:
      // fewer leaves more of the first upload / last download un-overlapped, more adds per-chunk
      // latencies; a ramp of small first and last chunks measured slower than 8 equal ones).
      const char* env = std::getenv("B200_HOST_CHUNKS");
      hp->nchunks = env && *env ? std::max(1, std::atoi(env)) : 8;
      const int C = hp->nchunks;
      hp->tile_bounds.assign(C + 1, 0);
      hp->row_bounds.assign(C + 1, 0);
      std::vector<std::int32_t> cmin(C), cmax(C);
      ck(b200_csr_plan_chunking(blk->op().plan(), blk->colind(), C, hp->tile_bounds.data(), hp->row_bounds.data(),
                                cmin.data(), cmax.data(), S),
         "plan chunking");
      hp->x_need.assign(C, 0);
      hp->y_ready.assign(C, 0);
      std::int32_t run = 0;
      for (int k = 0; k < C; ++k) {
        run = std::max(run, cmax[k] + 1);
        if (_symmetric)
          run = std::max(run, hp->row_bounds[k + 1]); // x[row] and the diagonal term
        // whole 4 KB pages: a copy that starts or ends inside a page runs ~20% slower over PCIe
        hp->x_need[k] = std::min((run + PAGE - 1) / PAGE * PAGE, nloc);
      }
      hp->x_need[C - 1] = nloc; // nothing of x stays behind
      std::int32_t reach = nrows; // lowest row any later chunk still adds into
      for (int k = C - 1; k >= 0; --k) {
        hp->y_ready[k] = (_symmetric ? std::min(hp->row_bounds[k + 1], reach) : hp->row_bounds[k + 1]) / PAGE * PAGE;
        if (_symmetric)
          reach = std::min(reach, cmin[k]);
      }
      hp->y_ready[C - 1] = nrows;
      ck(b200_stream_create(&hp->h2d), "stream");
      ck(b200_stream_create(&hp->d2h), "stream");
      ck(b200_event_create(&hp->ev_edge, 0), "event");
      hp->ev_x.assign(C, nullptr);
      hp->ev_k.assign(C, nullptr);
      const char* tr = std::getenv("B200_HOST_TRACE");
      hp->trace = tr && *tr && *tr != '0';
      if (hp->trace) {
        hp->ev_y.assign(C, nullptr);
        ck(b200_event_create(&hp->ev_start, 1), "event");
      }
      for (int k = 0; k < C; ++k) {
        ck(b200_event_create(&hp->ev_x[k], hp->trace), "event");
        ck(b200_event_create(&hp->ev_k[k], hp->trace), "event");
        if (hp->trace)
          ck(b200_event_create(&hp->ev_y[k], 1), "event");
      }
    }
    _host_pipe = hp;
  }
  HostPipe& hp = *static_cast<HostPipe*>(_host_pipe.get());

  if (!hp.pipelined) {
    // the reference's three blocking steps
    _exec->copy_from<double>(hp.x_dev, _exec->get_host(), x_host, nloc);
    _col_map->update(hp.x_dev);
    mult(hp.x_dev, hp.y_dev);
    _exec->copy_to<double>(y_host, _exec->get_host(), hp.y_dev, nrows);
    return;
  }

  const auto* blk = static_cast<const CSRMatrix<double>*>(_mat_local.get());
  void* plan = blk->op().plan();
  const int C = hp.nchunks;
  if (_symmetric || hp.halo) // every write below is an add
    ck(b200_memset_async(hp.y_dev, 0, sizeof(double) * static_cast<size_t>(nrows), S), "memset");
  std::int32_t x_sent = 0, y_sent = 0;
  if (hp.trace)
    ck(b200_event_record(hp.ev_start, hp.h2d), "event");
  if (hp.halo) {
    // the chunks below bring the same pages again (identical bytes, ~1 MB for a slab)
    for (size_t i = 0; i + 1 < hp.edge.size(); i += 2)
      ck(b200_memcpy_h2d_async(hp.x_dev + hp.edge[i], x_host + hp.edge[i],
                               sizeof(double) * (hp.edge[i + 1] - hp.edge[i]), hp.h2d),
         "h2d");
    ck(b200_event_record(hp.ev_edge, hp.h2d), "event");
    ck(b200_stream_wait_event(S, hp.ev_edge), "wait");
    _col_map->update(hp.x_dev);
    _col_map->update_finalise(hp.x_dev);
    if (_mat_remote->non_zeros() > 0)
      _mat_remote->mult(1, hp.x_dev, 1, hp.y_dev);
  }
  const double beta = hp.halo ? 1.0 : 0.0;
  for (int k = 0; k < C; ++k) {
    if (hp.x_need[k] > x_sent) {
      ck(b200_memcpy_h2d_async(hp.x_dev + x_sent, x_host + x_sent, sizeof(double) * (hp.x_need[k] - x_sent), hp.h2d),
         "h2d");
      x_sent = hp.x_need[k];
    }
    ck(b200_event_record(hp.ev_x[k], hp.h2d), "event");
    ck(b200_stream_wait_event(S, hp.ev_x[k]), "wait");
    if (_symmetric)
      ck(b200_csr_symv_tiles_f64(plan, blk->rowptr(), blk->colind(), blk->values(), blk->diagonal(), 1.0, hp.x_dev,
                                 hp.y_dev, hp.tile_bounds[k], hp.tile_bounds[k + 1], S),
         "symv tiles");
    else
      ck(b200_csr_spmv_tiles_f64(plan, blk->rowptr(), blk->colind(), blk->values(), 1.0, hp.x_dev, beta, hp.y_dev,
                                 hp.tile_bounds[k], hp.tile_bounds[k + 1], S),
         "spmv tiles");
    ck(b200_event_record(hp.ev_k[k], S), "event");
    if (hp.y_ready[k] > y_sent) {
      ck(b200_stream_wait_event(hp.d2h, hp.ev_k[k]), "wait");
      ck(b200_memcpy_d2h_async(y_host + y_sent, hp.y_dev + y_sent, sizeof(double) * (hp.y_ready[k] - y_sent), hp.d2h),
         "d2h");
      y_sent = hp.y_ready[k];
    }
    if (hp.trace)
      ck(b200_event_record(hp.ev_y[k], hp.d2h), "event");
  }
  ck(b200_stream_synchronize(hp.d2h), "synchronize");
  ck(b200_stream_synchronize(S), "synchronize");
  if (hp.trace) {
    std::fprintf(stderr, "mult_host timeline (ms after the first copy was queued; x landed / tiles done / y home)\n");
    for (int k = 0; k < C; ++k) {
      float tx = 0, tk = 0, ty = 0;
      b200_event_elapsed_ms(hp.ev_start, hp.ev_x[k], &tx);
      b200_event_elapsed_ms(hp.ev_start, hp.ev_k[k], &tk);
      b200_event_elapsed_ms(hp.ev_start, hp.ev_y[k], &ty);
      std::fprintf(stderr, "  chunk %2d  x[..%9d] %7.3f   tiles [%6d,%6d) %7.3f   y[..%9d] %7.3f\n", k, hp.x_need[k], tx,
                   hp.tile_bounds[k], hp.tile_bounds[k + 1], tk, hp.y_ready[k], ty);
    }
  }
}

template void Matrix<float>::mult_host(const float*, float*) const;

} // namespace spmv
