// cg.cpp -- Conjugate Gradient driver on the B200 executor.
// Reference: spmv/cg.cpp:21-98 (algorithm, iteration count semantics) and
// cuda/cg.cuda.cu:40-176 (what it replaces: 7 cuBLAS calls, 3 one-thread kernels and
// 3 host synchronisations per iteration).
//
// Per iteration this enqueues: halo update of p, SpMV (with p.Ap folded into its
// epilogue when the operator is a single general block), [allreduce], x/r update +
// r.r, [allreduce], p update + scalar bookkeeping.  Nothing waits on the host: the
// convergence test runs on the device and sets a `done` flag that turns every later
// kernel into a no-op; the host reads the flag LOOKAHEAD iterations behind, so the
// iteration count and x are exactly those of the reference's loop.
#include <cstdlib>
#include "cg.h"

#include "L2GMap.h"
#include "Matrix.h"
#include "b200_executor.h"
#include "csr_matrix.h"
#include "host_util.h"
#include "reference_executor.h"

#include <algorithm>
#include <stdexcept>

namespace spmv
{

namespace
{
constexpr int LOOKAHEAD = 8;

struct Flags {
  double done;
  double iter;
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
  std::shared_ptr<const L2GMap> col_l2g = A.col_map();
  std::shared_ptr<const L2GMap> row_l2g = A.row_map();
  if (row_l2g->num_ghosts() > 0)
    throw std::runtime_error("spmv::cg - Error: A.row_map() has ghost entries");
  exec.attach_comm(comm);
  detail::check(b200_set_device(exec.get_device_id()), "cudaSetDevice");
  void* stream = exec.get_cuda_stream();
  void* nccl = exec.nccl_comm();
  // scalar sums: one small kernel over peer memory when the column map set it up, else NCCL
  void* peer = col_l2g->peer_reduce();
  auto allreduce = [&](double* d_val) {
    if (peer)
      detail::check(b200_peer_allreduce_sum_f64(peer, d_val, 1, stream), "cg allreduce (peer)");
    else
      detail::check(b200_allreduce_sum_f64(nccl, d_val, 1, stream), "cg allreduce");
  };

  const int M = row_l2g->local_size();
  const int N_padded = col_l2g->local_size() + col_l2g->num_ghosts();
  const int hist_len = residual_history ? std::max(kmax, 0) : 0;

  double* r = exec.alloc<double>(std::max(M, 1));
  double* Ap = exec.alloc<double>(std::max(M, 1));
  double* x_padded = exec.alloc<double>(std::max(N_padded, 1));
  double* p = exec.alloc<double>(std::max(N_padded, 1));
  double* state = exec.alloc<double>(B200_CG_STATE_DOUBLES);
  double* scratch = exec.alloc<double>(B200_DOT_SCRATCH_DOUBLES);
  double* hist = hist_len > 0 ? exec.alloc<double>(hist_len) : nullptr;
  Flags* h_flags = nullptr;
  detail::check(b200_host_alloc_pinned(reinterpret_cast<void**>(&h_flags), sizeof(Flags) * LOOKAHEAD),
                "pinned alloc");
  void* events[LOOKAHEAD];
  for (void*& e : events)
    detail::check(b200_event_create(&e, 0), "event");

  auto cleanup = [&]() {
    exec.free(r);
    exec.free(Ap);
    exec.free(x_padded);
    exec.free(p);
    exec.free(state);
    exec.free(scratch);
    exec.free(hist);
    b200_host_free_pinned(h_flags);
    for (void* e : events)
      b200_event_destroy(e);
  };

  int iterations = 0;
  try {
    // x0 = 0, explicitly (the reference relies on the allocator for this, cg.cpp:41)
    exec.memset<double>(x_padded, 0, N_padded);
    exec.memset<double>(p, 0, N_padded);
    exec.memset<double>(state, 0, B200_CG_STATE_DOUBLES);
    exec.memset<double>(scratch, 0, B200_DOT_SCRATCH_DOUBLES);
    if (hist)
      exec.memset<double>(hist, 0, hist_len);
    exec.copy<double>(r, b, M); // cg.cpp:44-45
    exec.copy<double>(p, b, M);

    // rnorm0 = sqrt(allreduce(r.r))  (cg.cpp:47-50)
    detail::check(b200_cg_dot_f64(M, r, r, state + 2, scratch, nullptr, stream), "cg dot");
    allreduce(state + 2);
    detail::check(b200_cg_init_state(state, state + 2, rtol, kmax, stream), "cg init");

    // fused p.Ap epilogue: single general block only
    const CSRMatrix<double>* single = nullptr;
    if (!A.symmetric() && A.remote_block() == nullptr)
      single = dynamic_cast<const CSRMatrix<double>*>(A.local_block());
    if (single && (single->non_zeros() == 0 || single->op().plan() == nullptr))
      single = nullptr;
    // The epilogue pays only where the SpMV instantiation keeps its occupancy with the extra
    // registers: stencil-length rows (16-entry walk; 27-point 256^3: 1.157 vs 1.150 ms/iteration,
    // one launch fewer).  With short rows the dot-carrying 8-entry walk loses a resident CTA per
    // SM and a separate 31 us dot kernel wins (7-point general 256^3: 0.74 -> 0.56 ms/iteration,
    // profiles/r01_cg_fused_dot_ab.txt).  B200_CG_FUSED_DOT=0/1 forces either.
    if (single) {
      const double mean_row = (double)single->non_zeros() / (double)std::max(M, 1);
      bool fused = mean_row > 10.0;
      if (const char* e = std::getenv("B200_CG_FUSED_DOT"))
        fused = *e != '0';
      if (!fused)
        single = nullptr;
    }

    bool stop = false;
    for (int k = 0; k < kmax && !stop; ++k) {
      const int slot = k % LOOKAHEAD;
      if (k >= LOOKAHEAD) {
        detail::check(b200_event_synchronize(events[slot]), "event sync");
        if (h_flags[slot].done != 0.0)
          break;
      }
      col_l2g->update(p); // cg.cpp:59
      if (single) {       // cg.cpp:60-63
        detail::check(b200_csr_spmv_dot_f64(single->op().plan(), single->rowptr(), single->colind(),
                                            single->values(), p, Ap, state + 1, stream),
                      "cg spmv+dot");
      } else {
        A.mult(p, Ap);
        detail::check(b200_cg_dot_f64(M, p, Ap, state + 1, scratch, state + 4, stream), "cg dot");
      }
      allreduce(state + 1); // cg.cpp:65
      detail::check(b200_cg_update_xr_f64(M, state, p, Ap, x_padded, r, stream), "cg update x,r");
      allreduce(state + 2); // cg.cpp:75
      detail::check(b200_cg_update_p_f64(M, state, r, p, hist, hist_len, stream), "cg update p");
      detail::check(b200_memcpy_d2h_async(&h_flags[slot], state + 4, sizeof(Flags), stream), "flag copy");
      detail::check(b200_event_record(events[slot], stream), "event record");
    }
    Flags last;
    detail::check(b200_memcpy_d2h_async(&last, state + 4, sizeof(Flags), stream), "flag copy");
    detail::check(b200_stream_synchronize(stream), "cg sync");
    iterations = static_cast<int>(last.iter);

    exec.copy<double>(x, x_padded, M); // cg.cpp:89
    if (residual_history) {
      residual_history->assign(iterations, 0.0);
      if (iterations > 0)
        exec.copy_to<double>(residual_history->data(), exec.get_host(), hist, iterations);
    }
    detail::check(b200_stream_synchronize(stream), "cg sync");
  } catch (...) {
    cleanup();
    throw;
  }
  cleanup();
  return iterations;
}

} // namespace spmv
