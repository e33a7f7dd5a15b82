// cg.cpp -- Conjugate Gradient driver on the B200 executor.
// Reference: spmv/cg.cpp:21-98 (algorithm, iteration count semantics) and
// cuda/cg.cuda.cu:40-176 (what it replaces: 7 cuBLAS calls, 3 one-thread kernels and
// 3 host synchronisations per iteration).
//
// One iteration is a fixed sequence of kernels on ONE stream:
//   halo push (pack + NVLink stores + flag)           L2GMap::update(p)        cg.cpp:59
//   local-block SpMV [+ p.Ap in its epilogue]         Matrix::mult             cg.cpp:60
//   halo wait + unpack, remote-block SpMV             (update_finalise, two-block operators)
//   fused vector step: [p.Ap] -> sum over ranks -> x, r update -> r.r -> sum over ranks ->
//     p update -> bookkeeping, one cooperative launch (b200_cg_fused_step_f64)   cg.cpp:63-85
// Every per-iteration quantity -- scalars, convergence flag, iteration count, halo epochs,
// allreduce sequence numbers -- lives in device memory, so GRAPH_ITERS iterations are captured
// once into a CUDA graph and replayed; kernels are no-ops once the device-side `done` flag is
// set, and the host reads the flag two graphs behind.  The iteration count, x and the residual
// history are exactly those of the reference's loop.
// The workspace (vectors, state block, private stream, events, the graph) is cached on the
// executor: a second solve allocates nothing (the reference allocates per call, cg.cpp:39-42).
#include <cstdlib>
#include "cg.h"

#include "L2GMap.h"
#include "Matrix.h"
#include "b200_executor.h"
#include "csr_matrix.h"
#include "host_util.h"
#include "reference_executor.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <stdexcept>

namespace spmv
{

namespace
{
constexpr int LOOKAHEAD = 8;   // stream path: iterations enqueued ahead of the flag the host reads
constexpr int GRAPH_ITERS = 8; // iterations per captured graph

struct Flags {
  double done;
  double iter;
};

bool env_flag(const char* name, bool dflt)
{
  const char* e = std::getenv(name);
  return (e && *e) ? *e != '0' : dflt;
}

struct CgWorkspace {
  int device = -1;
  int M = -1, N_padded = -1, hist_cap = 0;
  double *r = nullptr, *Ap = nullptr, *x_padded = nullptr, *p = nullptr, *state = nullptr, *scratch = nullptr,
         *hist = nullptr;
  Flags* h_flags = nullptr; // pinned, LOOKAHEAD slots
  void* events[LOOKAHEAD] = {};
  void* stream = nullptr;   // private stream: the legacy default stream cannot be captured
  void* ev_in = nullptr;
  void* graph = nullptr;    // GRAPH_ITERS iterations, instantiated
  const void* graph_matrix = nullptr;
  std::uint64_t graph_epoch = 0; // detail::structure_epoch() at capture
  int graph_mode = -1;

  void drop_graph()
  {
    if (graph)
      b200_graph_destroy(graph);
    graph = nullptr;
    graph_matrix = nullptr;
    graph_mode = -1;
  }
  void release()
  {
    if (graph_mode == -2)
      graph_mode = -1;
    if (device >= 0)
      b200_set_device(device);
    drop_graph();
    for (double** q : {&r, &Ap, &x_padded, &p, &state, &scratch, &hist}) {
      b200_free(*q);
      *q = nullptr;
    }
    b200_host_free_pinned(h_flags);
    h_flags = nullptr;
    for (void*& e : events) {
      b200_event_destroy(e);
      e = nullptr;
    }
    b200_event_destroy(ev_in);
    ev_in = nullptr;
    if (stream)
      b200_stream_destroy(stream);
    stream = nullptr;
    M = N_padded = -1;
    hist_cap = 0;
  }
  ~CgWorkspace() { release(); }
};

double* dalloc(size_t n)
{
  void* q = nullptr;
  detail::check(b200_malloc(&q, sizeof(double) * std::max<size_t>(n, 1)), "cg workspace");
  return static_cast<double*>(q);
}

CgWorkspace& workspace(B200Executor& exec, int M, int N_padded, int hist_len)
{
  std::shared_ptr<void>& slot = exec.cg_workspace();
  if (!slot)
    slot = std::make_shared<CgWorkspace>();
  CgWorkspace& w = *static_cast<CgWorkspace*>(slot.get());
  if (w.device != exec.get_device_id() || w.M != M || w.N_padded != N_padded) {
    w.release();
    w.device = exec.get_device_id();
    w.M = M;
    w.N_padded = N_padded;
    w.r = dalloc(M);
    w.Ap = dalloc(M);
    w.x_padded = dalloc(N_padded);
    w.p = dalloc(N_padded);
    w.state = dalloc(B200_CG_STATE_DOUBLES);
    w.scratch = dalloc(B200_DOT_SCRATCH_DOUBLES);
    detail::check(b200_host_alloc_pinned(reinterpret_cast<void**>(&w.h_flags), sizeof(Flags) * LOOKAHEAD), "pinned alloc");
    for (void*& e : w.events)
      detail::check(b200_event_create(&e, 0), "event");
    detail::check(b200_event_create(&w.ev_in, 0), "event");
    detail::check(b200_stream_create(&w.stream), "stream");
    // the grid-barrier generation and the ticket counters start at zero, once
    detail::check(b200_memset_async(w.state, 0, sizeof(double) * B200_CG_STATE_DOUBLES, w.stream), "memset");
    detail::check(b200_memset_async(w.scratch, 0, sizeof(double) * B200_DOT_SCRATCH_DOUBLES, w.stream), "memset");
  }
  if (hist_len > w.hist_cap) { // the graph carries the history pointer: a new buffer needs a new capture
    w.drop_graph();
    b200_free(w.hist);
    w.hist_cap = std::max(hist_len, 1024);
    w.hist = dalloc(w.hist_cap);
  }
  return w;
}

// The solve runs on the workspace's private stream: point the executor(s) at it for the duration.
struct StreamScope {
  B200Executor *a, *b;
  B200Executor::stream_t sa, sb;
  StreamScope(B200Executor* a_, B200Executor* b_, void* s) : a(a_), b(b_ == a_ ? nullptr : b_), sa(a_->get_cuda_stream()), sb(nullptr)
  {
    a->set_cuda_stream(static_cast<B200Executor::stream_t>(s));
    if (b) {
      sb = b->get_cuda_stream();
      b->set_cuda_stream(static_cast<B200Executor::stream_t>(s));
    }
  }
  ~StreamScope()
  {
    a->set_cuda_stream(sa);
    if (b)
      b->set_cuda_stream(sb);
  }
};
} // namespace

int cg(MPI_Comm, ReferenceExecutor&, const Matrix<double>&, const double*, double*, int, double)
{
  throw std::runtime_error("spmv::cg: the host executor has no compute path in this build; use B200Executor");
}

int cg(MPI_Comm comm, B200Executor& exec, const Matrix<double>& A, const double* b, double* x, int kmax,
       double rtol)
{
  return cg(comm, exec, A, b, x, kmax, rtol, nullptr);
}

int cg(MPI_Comm comm, B200Executor& exec, const Matrix<double>& A, const double* b, double* x, int kmax,
       double rtol, std::vector<double>* residual_history)
{
  return cg(comm, exec, A, b, x, kmax, rtol, residual_history, nullptr);
}

int cg(MPI_Comm comm, B200Executor& exec, const Matrix<double>& A, const double* b, double* x, int kmax,
       double rtol, std::vector<double>* residual_history, CgPhaseTimes* phases)
{
  std::shared_ptr<const L2GMap> col_l2g = A.col_map();
  std::shared_ptr<const L2GMap> row_l2g = A.row_map();
  if (row_l2g->num_ghosts() > 0)
    throw std::runtime_error("spmv::cg - Error: A.row_map() has ghost entries");
  exec.attach_comm(comm);
  detail::check(b200_set_device(exec.get_device_id()), "cudaSetDevice");
  void* nccl = exec.nccl_comm();
  // scalar sums: over peer memory when the column map set it up, else NCCL
  void* peer = col_l2g->peer_reduce();

  const int M = row_l2g->local_size();
  const int N_padded = col_l2g->local_size() + col_l2g->num_ghosts();
  CgWorkspace& w = workspace(exec, M, N_padded, std::max(kmax, 0));
  double *r = w.r, *Ap = w.Ap, *x_padded = w.x_padded, *p = w.p, *state = w.state, *scratch = w.scratch, *hist = w.hist;
  const int hist_len = w.hist_cap;

  // the solve uses the kernel-driven halo exchange (graph-capturable); plain update() calls keep the map's default
  struct HaloScope {
    const L2GMap& m;
    bool was;
    explicit HaloScope(const L2GMap& m_) : m(m_), was(m_.set_kernel_exchange(true)) {}
    ~HaloScope() { m.set_kernel_exchange(was); }
  } halo_scope(*col_l2g);

  // the private stream starts after whatever the caller has queued on the executor's stream
  void* user_stream = exec.get_cuda_stream();
  void* stream = w.stream;
  detail::check(b200_event_record(w.ev_in, user_stream), "event record");
  detail::check(b200_stream_wait_event(stream, w.ev_in), "stream wait");
  auto* mat_exec = dynamic_cast<B200Executor*>(A.executor().get());
  StreamScope scope(&exec, mat_exec ? mat_exec : &exec, stream);

  auto allreduce = [&](double* d_val) {
    if (peer)
      detail::check(b200_peer_allreduce_sum_f64(peer, d_val, 1, stream), "cg allreduce (peer)");
    else
      detail::check(b200_allreduce_sum_f64(nccl, d_val, 1, stream), "cg allreduce");
  };

  // x0 = 0, explicitly (the reference relies on the allocator for this, cg.cpp:41)
  detail::check(b200_memset_async(x_padded, 0, sizeof(double) * N_padded, stream), "memset");
  detail::check(b200_memset_async(p, 0, sizeof(double) * N_padded, stream), "memset");
  detail::check(b200_memset_async(hist, 0, sizeof(double) * hist_len, stream), "memset");
  detail::check(b200_memcpy_d2d_async(r, b, sizeof(double) * M, stream), "copy"); // cg.cpp:44-45
  detail::check(b200_memcpy_d2d_async(p, b, sizeof(double) * M, stream), "copy");

  // rnorm0 = sqrt(allreduce(r.r))  (cg.cpp:47-50)
  detail::check(b200_cg_dot_f64(M, r, r, state + 2, scratch, nullptr, stream), "cg dot");
  allreduce(state + 2);
  detail::check(b200_cg_init_state(state, state + 2, rtol, kmax, stream), "cg init");
  const bool timeline = env_flag("B200_CG_TIMELINE", false); // debug: phase stamps of the fused step (st[9..15])
  {
    static const double on = 1.0, off = 0.0;
    detail::check(b200_memcpy_h2d_async(state + 9, timeline ? &on : &off, sizeof(double), stream), "timeline flag");
  }

  // fused p.Ap epilogue, general storage: it pays only where the SpMV instantiation keeps its
  // occupancy with the extra registers: stencil-length rows (16-entry walk; 27-point 256^3: 1.157 vs
  // 1.150 ms/iteration, one launch fewer).  With short rows the dot-carrying 8-entry walk loses a resident
  // CTA per SM (7-point general 256^3: 0.74 -> 0.56 ms/iteration, profiles/r01_cg_fused_dot_ab.txt).
  // B200_CG_FUSED_DOT=0/1 forces either.
  // Symmetric storage: x'Ax rides in the symmetric SpMV through the quadratic-form identity of
  // b200_csr_symv_dot_f64 (rounding differs from ddot(p, Ap); symmetric storage is not bit-reproducible anyway).
  // Two blocks (local, then remote with beta = 1): the remote product completes the sum
  // (b200_csr_spmv_rowlist_dot_f64; general storage: still the double ddot(p, Ap) rounds to).
  const CSRMatrix<double>* single = nullptr;     // general, one block
  const CSRMatrix<double>* sym_single = nullptr; // symmetric, no remote entries
  const CSRMatrix<double>* dot_local = nullptr;  // two blocks: local ...
  const CSRMatrix<double>* dot_remote = nullptr; //             ... and remote
  {
    const auto* loc = dynamic_cast<const CSRMatrix<double>*>(A.local_block());
    const auto* rem = A.remote_block() ? dynamic_cast<const CSRMatrix<double>*>(A.remote_block()) : nullptr;
    const bool rem_empty = A.remote_block() == nullptr || A.remote_block()->non_zeros() == 0;
    const bool loc_ok = loc && loc->non_zeros() > 0 && loc->op().plan() != nullptr && (!A.symmetric() || loc->diagonal() != nullptr);
    const double mean_row = loc ? (double)loc->non_zeros() / (double)std::max(M, 1) : 0.0;
    const bool want = env_flag("B200_CG_FUSED_DOT", A.symmetric() || mean_row > 10.0);
    if (loc_ok && want) {
      if (!A.symmetric() && A.remote_block() == nullptr) {
        single = loc;
      } else if (A.symmetric() && rem_empty) {
        sym_single = loc;
      } else if (!rem_empty && rem && rem->op().plan() != nullptr && (A.symmetric() || col_l2g->overlapping())
                 && env_flag("B200_CG_FUSED_DOT_REMOTE", true)) {
        // 2 GPUs, us per iteration without / with (separate vector kernels): 2.1 M rows each, 27-point 172.7 / 172.8,
        // symmetric 7-point 89.5 / 85.5; 16.8 M rows each: 1139.1 / 1127.3 and 475.9 / 446.2
        std::int32_t ok = 0;
        if (b200_csr_plan_rowlist_dot_ok(rem->op().plan(), &ok) == B200_OK && ok) {
          dot_local = loc;
          dot_remote = rem;
        }
      }
    }
  }
  const bool have_dot = single != nullptr || sym_single != nullptr || dot_local != nullptr;
  const bool wait_in_remote = env_flag("B200_CG_WAIT_IN_REMOTE", true);
  // Vector half of the iteration: three kernels (dot [unless it rode in the SpMV], x/r update + r.r, p update) with
  // the sums over the ranks in the tails of the first two (tail_ar), or ONE cooperative launch with grid-wide
  // hand-overs (B200_CG_FUSED_VEC=1; needs the peer mailboxes or one rank).  Inside a captured graph the launch gaps
  // the fused step was built to remove are gone, and its two in-kernel hand-overs (every CTA's ticket, the fold,
  // the release) cost more than two kernel boundaries: 2 GPUs, 27-point, 2.1 M rows per GPU, 180.9 us per
  // iteration fused against 173.5 with the three kernels; 16.8 M rows: 1142.9 against 1133.9.  Default: off.
  const bool tail_ar = peer != nullptr && env_flag("B200_CG_TAIL_ALLREDUCE", true);
  const bool can_fuse = peer != nullptr || exec.comm_size() == 1;
  const bool fused_vec = can_fuse && env_flag("B200_CG_FUSED_VEC", false);
  // CUDA graph of GRAPH_ITERS iterations: everything the iteration enqueues must be a kernel with
  // device-side state, i.e. the kernel-driven halo (or no halo) and no NCCL call
  const char* hm = col_l2g->halo_mode();
  const bool graphable = (std::strcmp(hm, "kernel") == 0 || std::strcmp(hm, "none") == 0) && (peer != nullptr || exec.comm_size() == 1);
  const bool use_graph = graphable && phases == nullptr && env_flag("B200_CG_GRAPH", true);

  // The fused step can also carry the halo push of the NEXT iteration's update(p) (B200_CG_FUSED_PUSH=1): one
  // launch less per iteration.  OFF by default: inside a captured graph the separate push launch costs less than
  // the extra grid-wide hand-over + per-CTA system fences it takes in the fused kernel (2 GPUs, 27-point, 2.1 M
  // rows per GPU: 196.4 us per iteration with the separate push, 200.8 fused; same bits).
  const L2GMap::KernelPush kpush = (fused_vec && env_flag("B200_CG_FUSED_PUSH", false)) ? col_l2g->kernel_push() : L2GMap::KernelPush();
  const bool fused_push = kpush.desc != nullptr;
  bool pushed_ahead = false; // the previous iteration's fused step already pushed p

  // ---- one iteration, enqueued on `stream` (cg.cpp:59-85) ----
  void* ev[6] = {};
  auto mark = [&](int i) {
    if (phases)
      detail::check(b200_event_record(ev[i], stream), "event record");
  };
  auto iteration = [&]() {
    mark(0);
    if (pushed_ahead)
      col_l2g->adopt_pushed(p);
    else
      col_l2g->update(p); // cg.cpp:59
    mark(1);
    if (single) { // cg.cpp:60-63
      detail::check(b200_csr_spmv_dot_f64(single->op().plan(), single->rowptr(), single->colind(), single->values(), p,
                                          Ap, state + 1, stream),
                    "cg spmv+dot");
      mark(2);
      mark(3);
    } else if (sym_single) {
      detail::check(b200_csr_symv_dot_f64(sym_single->op().plan(), sym_single->rowptr(), sym_single->colind(),
                                          sym_single->values(), sym_single->diagonal(), p, Ap, state + 1, stream),
                    "cg symv+dot");
      mark(2);
      mark(3);
    } else if (dot_local) {
      if (A.symmetric())
        detail::check(b200_csr_symv_dot_f64(dot_local->op().plan(), dot_local->rowptr(), dot_local->colind(),
                                            dot_local->values(), dot_local->diagonal(), p, Ap, state + 1, stream),
                      "cg symv+dot (local block)");
      else
        detail::check(b200_csr_spmv_dot_f64(dot_local->op().plan(), dot_local->rowptr(), dot_local->colind(),
                                            dot_local->values(), p, Ap, state + 1, stream),
                      "cg spmv+dot (local block)");
      mark(2);
      // the remote product is also the halo's wait (ghost values straight from the staging slot) and carries the
      // sum over the ranks in its tail
      void* hw = wait_in_remote ? col_l2g->kernel_wait_desc() : nullptr;
      if (hw == nullptr)
        col_l2g->update_finalise(p);
      detail::check(b200_csr_spmv_rowlist_dot_halo_f64(dot_remote->op().plan(), dot_local->op().plan(), dot_remote->rowptr(),
                                                       dot_remote->colind(), dot_remote->values(), p, Ap, state + 1,
                                                       (tail_ar && !fused_vec) ? peer : nullptr, hw, col_l2g->local_size(),
                                                       stream),
                    "cg remote block + dot");
      if (hw != nullptr)
        col_l2g->finalised_by_kernel();
      mark(3);
    } else if (phases && A.remote_block() != nullptr) { // Matrix::mult, opened up to time its parts
      A.local_block()->mult(1, p, 0, Ap);
      mark(2);
      col_l2g->update_finalise(p);
      if (A.symmetric() || col_l2g->overlapping())
        A.remote_block()->mult(1, p, (A.symmetric() || A.local_block()->non_zeros() > 0) ? 1.0 : 0.0, Ap);
      mark(3);
    } else {
      A.mult(p, Ap);
      mark(2);
      mark(3);
    }
    detail::NvtxRange nv("cg:vector step + reductions");
    if (fused_vec) {
      detail::check(b200_cg_fused_step_halo_f64(M, state, p, Ap, x_padded, r, hist, hist_len, peer, have_dot ? 0 : 1,
                                                kpush.desc, kpush.n_index, kpush.d_index, stream),
                    "cg fused step");
      pushed_ahead = fused_push;
    } else {
      // the sums over the ranks (cg.cpp:65,75) ride in the tail of the kernel that produces the value where the
      // peer mailboxes exist (tail_ar); a launch of their own otherwise
      if (!have_dot) {
        detail::check(b200_cg_dot_peer_f64(M, p, Ap, state + 1, scratch, state + 4, tail_ar ? peer : nullptr, stream), "cg dot");
        if (!tail_ar)
          allreduce(state + 1);
      } else if (!(dot_local && tail_ar)) { // (the two-block dot already summed over the ranks in its tail)
        allreduce(state + 1);
      }
      detail::check(b200_cg_update_xr_peer_f64(M, state, p, Ap, x_padded, r, tail_ar ? peer : nullptr, stream), "cg update x,r");
      if (!tail_ar)
        allreduce(state + 2);
      detail::check(b200_cg_update_p_f64(M, state, r, p, hist, hist_len, stream), "cg update p");
    }
    mark(4);
  };

  int iterations = 0;
  Flags last;
  if (phases) {
    // ---- profile mode: one iteration at a time, CUDA events between its parts ----
    for (void*& e : ev)
      detail::check(b200_event_create(&e, 1), "event");
    double acc[4] = {0, 0, 0, 0};
    int done_its = 0;
    for (int k = 0; k < kmax; ++k) {
      iteration();
      detail::check(b200_memcpy_d2h_async(&last, state + 4, sizeof(Flags), stream), "flag copy");
      detail::check(b200_stream_synchronize(stream), "cg sync");
      for (int i = 0; i < 4; ++i) {
        float ms = 0.f;
        detail::check(b200_event_elapsed_ms(ev[i], ev[i + 1], &ms), "event elapsed");
        acc[i] += ms;
      }
      ++done_its;
      if (last.done != 0.0)
        break;
    }
    for (void* e : ev)
      b200_event_destroy(e);
    const double s = done_its > 0 ? 1e3 / done_its : 0.0;
    phases->iterations = done_its;
    phases->us_update = acc[0] * s;
    phases->us_local = acc[1] * s;
    phases->us_remote = acc[2] * s;
    phases->us_vector = acc[3] * s;
  }
  // The first iteration always goes out un-captured: it runs whatever the path initialises lazily (kernel
  // attributes, occupancy queries, buffers), none of which may happen inside a capture.
  if (!phases && kmax > 0)
    iteration();
  bool graph_ready = false;
  if (!phases && use_graph && kmax > 1 && w.graph_mode != -2) {
    const int mode = (single ? 1 : 0) | (fused_vec ? 2 : 0) | (sym_single ? 4 : 0) | (fused_push ? 8 : 0) | (dot_local ? 16 : 0) | (tail_ar ? 32 : 0) | (wait_in_remote ? 64 : 0);
    graph_ready = w.graph != nullptr && w.graph_matrix == static_cast<const void*>(&A) && w.graph_mode == mode
                  && w.graph_epoch == detail::structure_epoch();
    if (!graph_ready) {
      w.drop_graph();
      detail::check(b200_stream_synchronize(stream), "cg sync");
      try { // capture enqueues nothing: if anything on the path cannot be captured, the stream path below runs instead
        detail::check(b200_graph_capture_begin(stream), "graph capture");
        try {
          for (int i = 0; i < GRAPH_ITERS; ++i)
            iteration();
        } catch (...) {
          void* dead = nullptr;
          b200_graph_capture_end(stream, &dead);
          b200_graph_destroy(dead);
          throw;
        }
        detail::check(b200_graph_capture_end(stream, &w.graph), "graph instantiate");
        w.graph_matrix = &A;
        w.graph_mode = mode;
        w.graph_epoch = detail::structure_epoch();
        graph_ready = true;
      } catch (const std::exception& e) {
        if (env_flag("B200_CG_GRAPH_STRICT", false))
          throw;
        if (env_flag("B200_CG_DEBUG", false))
          std::fprintf(stderr, "spmv::cg: graph capture failed (%s); using the stream path\n", e.what());
        w.drop_graph();
        w.graph_mode = -2; // do not try again with this workspace
      }
    }
  }
  if (phases) {
  } else if (graph_ready) {
    // ---- GRAPH_ITERS iterations per launch; the flag of graph g is read before graph g+2 goes out ----
    for (int g = 0;; ++g) {
      const int slot = g % LOOKAHEAD;
      if (g >= 2) {
        const int seen = (g - 2) % LOOKAHEAD;
        detail::check(b200_event_synchronize(w.events[seen]), "event sync");
        if (w.h_flags[seen].done != 0.0)
          break;
      }
      if (1 + g * GRAPH_ITERS >= kmax)
        break;
      detail::check(b200_graph_launch(w.graph, stream), "graph launch");
      detail::check(b200_memcpy_d2h_async(&w.h_flags[slot], state + 4, sizeof(Flags), stream), "flag copy");
      detail::check(b200_event_record(w.events[slot], stream), "event record");
    }
  } else {
    // ---- stream path: iterations enqueued LOOKAHEAD ahead of the flag the host reads ----
    for (int k = 1; k < kmax; ++k) {
      const int slot = k % LOOKAHEAD;
      if (k > LOOKAHEAD) {
        detail::check(b200_event_synchronize(w.events[slot]), "event sync");
        if (w.h_flags[slot].done != 0.0)
          break;
      }
      iteration();
      detail::check(b200_memcpy_d2h_async(&w.h_flags[slot], state + 4, sizeof(Flags), stream), "flag copy");
      detail::check(b200_event_record(w.events[slot], stream), "event record");
    }
  }
  if (pushed_ahead) { // the last fused step pushed a halo nobody multiplies with: drain it (every rank does)
    col_l2g->adopt_pushed(p);
    col_l2g->update_finalise(p);
    pushed_ahead = false;
  }
  detail::check(b200_memcpy_d2h_async(&last, state + 4, sizeof(Flags), stream), "flag copy");
  detail::check(b200_stream_synchronize(stream), "cg sync");
  iterations = static_cast<int>(last.iter);

  if (timeline) {
    double tl[6];
    int tl_rank = 0;
    MPI_Comm_rank(comm, &tl_rank);
    detail::check(b200_memcpy_d2h_async(tl, state + 10, sizeof(tl), stream), "timeline copy");
    detail::check(b200_stream_synchronize(stream), "cg sync");
    std::fprintf(stderr,
                 "[cg timeline rank %d] last live fused step, CTA 0, us: dot pass %.1f | sum over ranks %.1f | x,r update "
                 "%.1f | sum over ranks %.1f | p update %.1f | total %.1f\n",
                 tl_rank, (tl[1] - tl[0]) * 1e-3, (tl[2] - tl[1]) * 1e-3, (tl[3] - tl[2]) * 1e-3,
                 (tl[4] - tl[3]) * 1e-3, (tl[5] - tl[4]) * 1e-3, (tl[5] - tl[0]) * 1e-3);
  }
  detail::check(b200_memcpy_d2d_async(x, x_padded, sizeof(double) * M, stream), "copy"); // cg.cpp:89
  if (residual_history) {
    residual_history->assign(iterations, 0.0);
    if (iterations > 0)
      detail::check(b200_memcpy_d2h_async(residual_history->data(), hist, sizeof(double) * iterations, stream), "copy");
  }
  detail::check(b200_stream_synchronize(stream), "cg sync");
  return iterations;
}

} // namespace spmv
