/* b200spmv.h -- the C ABI under the LIBSPMV plugin boundary, B200 (sm_100a).
 *
 * Every entry point here is what a backend of the reference has to provide
 * behind `spmv::DeviceExecutor` / `spmv::CSRSpMV<T>` / `spmv::L2GMap` /
 * `spmv::cg` (reference file:line cited per function).  Conventions:
 *   - plain C types only; `d_` pointers are device pointers on the current
 *     CUDA device; `stream` is a cudaStream_t passed as void* (NULL = legacy
 *     default stream); nothing here synchronises the host unless it says so;
 *   - every function returns 0 on success or a negative B200_E* code, and
 *     `b200_last_error()` gives a thread-local message.  The C++ host layer
 *     (include/spmv/) turns non-zero into std::runtime_error, the way the
 *     reference's constructors throw (Matrix.cpp:29, L2GMap.cpp:363);
 *   - there is no CPU fallback: without a usable GPU every compute call
 *     returns B200_ECUDA.
 */
#ifndef B200SPMV_H
#define B200SPMV_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

#define B200_OK 0
#define B200_EINVAL (-1) /* bad argument */
#define B200_ECUDA (-2)  /* CUDA runtime / launch failure, or no device */
#define B200_ENCCL (-3)  /* NCCL failure or NCCL not loadable */
#define B200_ENOMEM (-4)

B200_API const char* b200_last_error(void);
/* ABI version of this header: bump on any signature change. */
B200_API int b200_abi_version(void);

/* ---- device memory and streams: DeviceExecutor::_alloc/_free/_memset/_copy*
 *      (device_executor.h:129-139; cuda/cuda_executor.cpp:54-105) ---------- */
B200_API int b200_device_count(int* count);
B200_API int b200_set_device(int device_id);
B200_API int b200_get_num_sms(int* num_sms); /* DeviceExecutor::get_num_cus (device_executor.h:85) */
B200_API int b200_malloc(void** d_ptr, size_t num_bytes);
B200_API int b200_free(void* d_ptr);
B200_API int b200_memset_async(void* d_ptr, int value, size_t num_bytes, void* stream);
B200_API int b200_memcpy_d2d_async(void* d_dst, const void* d_src, size_t num_bytes, void* stream);
B200_API int b200_memcpy_h2d_async(void* d_dst, const void* h_src, size_t num_bytes, void* stream);
B200_API int b200_memcpy_d2h_async(void* h_dst, const void* d_src, size_t num_bytes, void* stream);
B200_API int b200_host_alloc_pinned(void** h_ptr, size_t num_bytes);
B200_API int b200_host_free_pinned(void* h_ptr);
B200_API int b200_stream_create(void** stream);
/* high != 0: greatest priority the device offers (the halo stream: its small copies and flag
 * operations should never queue behind a resident SpMV kernel's work). */
B200_API int b200_stream_create_priority(void** stream, int high);
B200_API int b200_stream_destroy(void* stream);
B200_API int b200_stream_synchronize(void* stream); /* DeviceExecutor::synchronize (device_executor.h:82) */
B200_API int b200_event_create(void** event, int timing);
B200_API int b200_event_destroy(void* event);
B200_API int b200_event_record(void* event, void* stream);
B200_API int b200_event_synchronize(void* event);
B200_API int b200_stream_wait_event(void* stream, void* event);
B200_API int b200_event_elapsed_ms(void* start, void* stop, float* ms);

/* ---- CSR SpMV plan: CSRSpMV<T>::init / finalize (csr_kernels.h:71-78,
 *      csr_kernels.cpp:11-17,55-57; the reference hangs its analysis off
 *      `void* _aux_data`, csr_kernels.h:82).  The plan cuts the rows into
 *      tiles of bounded matrix-stream bytes and takes the row-length histogram
 *      that selects the thread-per-row / warp-per-row walks.  `d_rowptr` may be NULL iff
 *      nnz == 0 (csr_matrix.cpp:35).  Synchronises `stream` once. ---------- */
B200_API int b200_csr_plan_create(void** plan, int32_t rows, int32_t cols, int64_t nnz,
                         const int32_t* d_rowptr, int symmetric, void* stream);
B200_API int b200_csr_plan_destroy(void* plan);
/* Introspection (tests, bench): number of tiles and the row-length histogram
 * (bucket k counts rows with length in [2^(k-1), 2^k), bucket 0 = empty rows;
 * 32 buckets). */
B200_API int b200_csr_plan_info(void* plan, int32_t* num_tiles, int32_t* max_row_len,
                       int64_t hist[32]);
/* Tuning hook (bench sweeps, tests).  variant 0 = persistent CTAs with a bulk-async staged
 * ring, walk chosen from the row-length histogram (default); 1 = one tile per CTA staged
 * with plain loads; 2..5 = variant 0 with the walk forced: 2 one thread per row in chunks
 * of 16 entries, 3 two-phase stream walk, 4 chunks of 8, 5 chunks of 4, 6 warp-private
 * software-pipelined stream walk (general storage; each consumer warp owns a plan-time segment of
 * the tile, no CTA barrier), 7 the same with L1::no_allocate gathers, 8 split stream walk (half of
 * the consumer warps make the products of tile i+1 while the other half sums the rows of tile i;
 * the default for irregular rows in general storage); symmetric plans fall back from 6-8 to 3.
 * consumer_warps is 8 or 16.  Re-partitions
 * the tiles (needs the same d_rowptr the plan was created with).
 * Defaults also read B200_SPMV_{VARIANT,TILE_BYTES,STAGES,CWARPS,CTAS} from the
 * environment at plan creation. */
B200_API int b200_csr_plan_tune(void* plan, const int32_t* d_rowptr, int variant, int tile_bytes,
                       int stages, int consumer_warps, int ctas_per_sm, void* stream);

/* Memory contract of the SpMV entries below.  The pipeline moves the matrix with 16-byte bulk copies, each
 * slice starting at the 16-byte aligned element at or below its first element and rounded up to a 16-byte
 * multiple: up to 12 bytes BEHIND the last element of d_val / d_colind / d_rowptr (and, symmetric storage,
 * d_diag and d_x) can be read and are never used.  The read stays inside the last element's 16-byte unit, so
 * any allocation whose size is rounded to 16 bytes (cudaMalloc rounds to 256) covers it; a caller that
 * sub-allocates exact-size arrays at the very end of a mapping must pad them to a 16-byte multiple (and
 * compute-sanitizer memcheck reports the read otherwise).  Base pointers that are not 16-byte aligned are
 * accepted and routed to the plain-load kernel.  With an overlapping schedule the over-read of x may touch the
 * first ghost entry while the exchange writes it: harmless, the value is not used. */

/* ---- general CSR SpMV: CSRSpMV<T>::run, general branch
 *      (csr_kernels.cpp:41-51; replaces cusparseSpMV at
 *      cuda/csr_kernels.cuda.cu:106-151).
 *      y[i] = alpha * sum_j val[j]*x[col[j]] + beta*y[i], j ascending.
 *      beta == 0 means overwrite (y is not read). ------------------------- */
B200_API int b200_csr_spmv_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                      const double* d_val, double alpha, const double* d_x, double beta,
                      double* d_y, void* stream);
B200_API int b200_csr_spmv_f32(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                      const float* d_val, float alpha, const float* d_x, float beta,
                      float* d_y, void* stream);
/* Same, and additionally d_dot[0] = sum_i x[i]*y[i] over the rows (the p.Ap of
 * cg.cpp:63 riding in the SpMV epilogue); deterministic two-stage reduction. */
B200_API int b200_csr_spmv_dot_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                          const double* d_val, const double* d_x, double* d_y,
                          double* d_dot, void* stream);
/* Two-block operators (local block, then remote block with beta = 1; Matrix::spmv_overlap / spmv_sym): the
 * remote product that COMPLETES the x.(A x) epilogue.  Call b200_csr_spmv_dot_f64 (or b200_csr_symv_dot_f64) on
 * the local block, then this on the remote block with the same x, y and d_dot on the same stream:
 * y[r] += (remote row sum) over the block's non-empty rows, and d_dot[0] = x.(A x) over both blocks -- for a
 * general local block the same double ddot(x, y_final) rounds to (error-free products, sums in twice the working
 * precision, carried between the two kernels in the local plan's per-CTA partials).  The remote plan must hold a
 * list of short non-empty rows (b200_csr_plan_rowlist_dot_ok: *ok = 1). */
B200_API int b200_csr_plan_rowlist_dot_ok(void* plan, int32_t* ok);
B200_API int b200_csr_spmv_rowlist_dot_f64(void* remote_plan, void* local_plan, const int32_t* d_rowptr,
                                  const int32_t* d_colind, const double* d_val, const double* d_x, double* d_y,
                                  double* d_dot, void* stream);
/* Same, optionally also (peer_ctx != NULL) summing d_dot[0] over the ranks in the kernel's tail (the MPI_Allreduce
 * of cg.cpp:65; b200_peer_reduce_create) and (halo_desc != NULL) acting as the WAIT half of the kernel-driven halo
 * exchange in place of b200_halo_wait_unpack_f64: the kernel polls the neighbours' flags itself and reads the
 * ghost entries of x (columns >= ghost_base = the map's local size) straight from the staging slot; the ghost
 * tail of d_x is then left untouched. */
B200_API int b200_csr_spmv_rowlist_dot_halo_f64(void* remote_plan, void* local_plan, const int32_t* d_rowptr,
                                       const int32_t* d_colind, const double* d_val, const double* d_x, double* d_y,
                                       double* d_dot, void* peer_ctx, void* halo_desc, int32_t ghost_base, void* stream);
/* Symmetric (lower + diagonal) storage, y = A x and d_dot[0] = x'Ax in the same pass.  The rows of y are
 * incomplete while the kernel runs (transposed entries still to be scattered), the quadratic form is not:
 * x'Ax = sum_r x_r (2 s_r - d_r x_r) with s_r = d_r x_r + sum_{c<r} a_rc x_c, the row sum the walk forms
 * anyway.  Equals ddot(x, A x) up to roundings (symmetric storage is not bit-reproducible in the first
 * place: csr_kernels.cpp:33-37 scatter order).  Square local block (every stored column < rows). */
B200_API int b200_csr_symv_dot_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                          const double* d_val, const double* d_diag, const double* d_x, double* d_y,
                          double* d_dot, void* stream);

/* ---- symmetric CSR SpMV: CSRSpMV<T>::run, symmetric branch
 *      (csr_kernels.cpp:26-40; replaces csr_sym_gmem_atomics at
 *      cuda/csr_kernels.cuda.cu:22-47,87-105).  Stored: strictly lower triangle
 *      of the local diagonal block (col < row) + `d_diag[rows]`.
 *      y = alpha*(L + D + L^T) x + beta*y.  d_rowptr/d_colind/d_val may be NULL
 *      iff the plan's nnz == 0 (diagonal-only block, csr_matrix.cpp:85). --- */
B200_API int b200_csr_symv_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                      const double* d_val, const double* d_diag, double alpha,
                      const double* d_x, double beta, double* d_y, void* stream);
B200_API int b200_csr_symv_f32(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                      const float* d_val, const float* d_diag, float alpha,
                      const float* d_x, float beta, float* d_y, void* stream);

/* ---- tile sub-ranges: the plan's tiles are consecutive row ranges of equal stream bytes, so a
 *      prefix of the rows can be multiplied while the rest of x is still on its way from the
 *      host (Matrix::mult_host pipelines H2D of x, these launches and D2H of y).
 *      b200_csr_plan_chunking cuts the tiles into nchunks equal groups and reports, per chunk k,
 *      tile_bounds[k..k+1], row_bounds[k..k+1] and the smallest / largest column its entries
 *      touch (col_min INT32_MAX / col_max -1 for an empty chunk).  Host arrays; synchronises.
 *      The *_tiles_* calls process only the rows of tiles [tile_begin, tile_end).  The
 *      symmetric one ADDS into y (rows of the range and the columns they scatter to): the
 *      caller zeroes / scales y once before the first range. ---- */
B200_API int b200_csr_plan_chunking(void* plan, const int32_t* d_colind, int32_t nchunks, int32_t* tile_bounds,
                           int32_t* row_bounds, int32_t* col_min, int32_t* col_max, void* stream);
B200_API int b200_csr_spmv_tiles_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                            const double* d_val, double alpha, const double* d_x, double beta, double* d_y,
                            int32_t tile_begin, int32_t tile_end, void* stream);
B200_API int b200_csr_symv_tiles_f64(void* plan, const int32_t* d_rowptr, const int32_t* d_colind,
                            const double* d_val, const double* d_diag, double alpha, const double* d_x,
                            double* d_y, int32_t tile_begin, int32_t tile_end, void* stream);

/* ---- halo pack / unpack: DeviceExecutor::gather_ghosts_run
 *      (device_executor.h:123-126; cuda/cuda_executor.cu:9-51) and the
 *      accumulate loop of L2GMap::reverse_update (L2GMap.cpp:920-921). ----- */
B200_API int b200_gather_f64(int32_t n, const int32_t* d_idx, const double* d_in, double* d_out, void* stream);
B200_API int b200_gather_f32(int32_t n, const int32_t* d_idx, const float* d_in, float* d_out, void* stream);
B200_API int b200_scatter_add_f64(int32_t n, const int32_t* d_idx, const double* d_in, double* d_inout, void* stream);
B200_API int b200_scatter_add_f32(int32_t n, const int32_t* d_idx, const float* d_in, float* d_inout, void* stream);

/* ---- NCCL communicator and halo exchange: the message part of
 *      L2GMap::update_p2p / update_p2p_start (L2GMap.cpp:564-635) and of
 *      reverse_update_p2p (:924-950).  One rank per GPU.
 *      b200_halo_exchange posts, in one NCCL group on `comm_stream`, for each
 *      neighbour i: send d_sendbuf[send_off[i] .. +send_cnt[i]) and receive
 *      recv_cnt[i] values into d_recvbuf[recv_off[i] ..).  Counts are elements
 *      of `elem_bytes` (4 or 8). ------------------------------------------- */
#define B200_NCCL_UNIQUE_ID_BYTES 128
B200_API int b200_nccl_unique_id(char id[B200_NCCL_UNIQUE_ID_BYTES]);
B200_API int b200_nccl_comm_create(void** comm, int nranks, int rank, const char id[B200_NCCL_UNIQUE_ID_BYTES]);
B200_API int b200_nccl_comm_destroy(void* comm);
B200_API int b200_halo_exchange(void* comm, int nnbr, const int32_t* nbr_rank, const void* d_sendbuf,
                       const int32_t* send_cnt, const int32_t* send_off, void* d_recvbuf,
                       const int32_t* recv_cnt, const int32_t* recv_off, int elem_bytes,
                       void* comm_stream);
/* In-stream sum-allreduce of n doubles (MPI_Allreduce at cg.cpp:49,65,75). */
B200_API int b200_allreduce_sum_f64(void* comm, double* d_inout, int32_t n, void* stream);

/* ---- direct peer halo over NVLink (the reference's one-sided models, L2GMap.cpp:645-721,
 *      re-thought for one process per GPU): a rank's receive staging buffer is exported with
 *      CUDA IPC, neighbours map it and push their packed slices into it with copy-engine
 *      copies; arrival is signalled by a stream-ordered 32-bit flag write into the peer and
 *      awaited with a stream-ordered wait -- no SM is used on either side and the host never
 *      blocks.  (An NCCL send/recv kernel needs SM resources that a persistent SpMV kernel
 *      holds for its whole run, so it cannot overlap with it reliably.)
 *      b200_ipc_alloc: cudaMalloc + zero + export.  b200_ipc_open maps a peer's buffer (fails
 *      for a handle created by this same process).  Flags are 4-byte aligned device words. ---- */
#define B200_IPC_HANDLE_BYTES 64
B200_API int b200_ipc_alloc(void** d_ptr, size_t num_bytes, char handle[B200_IPC_HANDLE_BYTES]);
B200_API int b200_ipc_open(void** d_peer_ptr, const char handle[B200_IPC_HANDLE_BYTES]);
B200_API int b200_ipc_close(void* d_peer_ptr);
/* *d_flag = value, in stream order after everything enqueued before it. */
B200_API int b200_stream_write_u32(void* stream, void* d_flag, uint32_t value);
/* The stream does not proceed until *d_flag >= value (wrap-around compare, driver semantics). */
B200_API int b200_stream_wait_u32_geq(void* stream, void* d_flag, uint32_t value);

/* ---- kernel-driven peer halo: the same exchange as two kernels on one stream.  b200_halo_push
 *      packs x[d_idx[k]] straight into the neighbours' staging slots (NVLink stores) and its last
 *      CTA publishes the epoch into each neighbour's flag word; b200_halo_wait_unpack polls this
 *      rank's flag words and copies the slot of the epoch into the ghost tail.  The epoch counter
 *      lives in the descriptor (device memory), so a captured CUDA graph replays the pair as is.
 *      Descriptor arrays are per neighbour, ascending rank: send_off[nnbr+1] = bounds of my packed
 *      slices (the map's _recv_offset), land[i] = element offset of my slice in neighbour i's
 *      slot, d_peer_slot0/1[i] and d_peer_flag[i] = its two slots and my flag word in its buffer
 *      (mapped with b200_ipc_open), d_my_slot0/1 and d_my_flag[i] = my slots and neighbour i's
 *      flag word in my buffer.  A pair of calls must not be mixed with the stream-memop exchange
 *      on the same flag words. ---- */
#define B200_HALO_MAX_NEIGHBOURS 32
B200_API int b200_halo_desc_create(void** desc, int32_t nnbr, const int32_t* send_off, const int64_t* land,
                          void* const* d_peer_slot0, void* const* d_peer_slot1, void* const* d_peer_flag,
                          void* d_my_slot0, void* d_my_slot1, void* const* d_my_flag, int64_t nghosts, void* stream);
B200_API int b200_halo_desc_destroy(void* desc);
B200_API int b200_halo_push_f64(void* desc, int32_t n, const int32_t* d_idx, const double* d_x, void* stream);
B200_API int b200_halo_push_f32(void* desc, int32_t n, const int32_t* d_idx, const float* d_x, void* stream);
B200_API int b200_halo_wait_unpack_f64(void* desc, int64_t nghosts, double* d_ghost_tail, void* stream);
B200_API int b200_halo_wait_unpack_f32(void* desc, int64_t nghosts, float* d_ghost_tail, void* stream);

/* ---- CG building blocks: cg.cpp:47-85 (cblas_ddot/daxpy/dscal) and the
 *      scalar kernels of cuda/cg.cuda.cu:14-38, fused.  All scalars live in a
 *      device "state" block of B200_CG_STATE_DOUBLES doubles so the loop never
 *      needs the host:
 *        st[0] rnorm_old   st[1] p.Ap (global)   st[2] r.r (global, raw sum)
 *        st[3] rnorm0      st[4] done flag (0/1) st[5] iterations completed
 *        st[6] rtol        st[7] kmax            st[16..] reduction scratch
 *      Every kernel is a no-op once st[4] != 0, so a fixed-length CUDA graph of
 *      iterations reproduces the reference's early exit (cg.cpp:80-81).
 *      Reductions are deterministic (fixed grid, fixed-order tree) and carried in twice
 *      the working precision (error-free products and sums, Dot2): a dot product rounds
 *      to the same double whatever the summation order, which keeps CG trajectories
 *      reproducible across implementations. -------- */
#define B200_CG_STATE_DOUBLES 4096
#define B200_DOT_SCRATCH_DOUBLES 4096
/* d_out[0] = sum a[i]*b[i].  d_scratch: B200_DOT_SCRATCH_DOUBLES doubles, zeroed
 * once by the caller before first use.  If d_done_flag != NULL and
 * *d_done_flag != 0 the kernel does nothing. */
B200_API int b200_cg_dot_f64(int32_t n, const double* d_a, const double* d_b, double* d_out,
                    double* d_scratch, const double* d_done_flag, void* stream);
/* Allocation-free variant of b200_cg_dot_f64.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  d_out[0] = sum a[i]*b[i],
 * same kernel, grid and summation order -- the same bits -- as b200_cg_dot_f64, but the
 * reduction scratch is a file-scope `static __device__` array of B200_DOT_SCRATCH_DOUBLES
 * doubles (32 KiB) that every GPU loading the library holds until exit; the call itself
 * performs no cudaMalloc / cudaFree / allocating library call (one 64-byte cudaMemsetAsync of
 * the scratch header + one kernel launch, both in stream order).
 * Limit: 0 <= n <= B200_CG_DOT_STATIC_MAX_N, else B200_EINVAL with a message naming the limit.
 * All calls on a device share the one array: calls that can OVERLAP on the GPU (different
 * streams without an ordering between them) race on it -- order them, or use b200_cg_dot_f64
 * with one scratch per stream. */
#define B200_CG_DOT_STATIC_MAX_N (1 << 28)
B200_API int b200_cg_dot_static_f64(int32_t n, const double* d_a, const double* d_b, double* d_out, void* stream);
/* b200_cg_dot_f64 / b200_cg_update_xr_f64 with the sum over the ranks (the MPI_Allreduce of cg.cpp:65,75) in the
 * tail of the same kernel: the last CTA to finish exchanges the folded sum over the peer mailboxes
 * (peer_ctx: b200_peer_reduce_create; NULL: the plain kernel) and stores the global value, every rank the same bits.
 * One launch instead of two per reduction.  Every rank of the communicator must make the call. */
B200_API int b200_cg_dot_peer_f64(int32_t n, const double* d_a, const double* d_b, double* d_out, double* d_scratch,
                         const double* d_done_flag, void* peer_ctx, void* stream);
B200_API int b200_cg_update_xr_peer_f64(int32_t n, double* d_state, const double* d_p, const double* d_Ap,
                               double* d_x, double* d_r, void* peer_ctx, void* stream);
/* st[0] = st[3] = sqrt(d_rr[0]) (d_rr = allreduced r.r, cg.cpp:47-53); clears
 * done / iteration count; stores rtol and kmax. */
B200_API int b200_cg_init_state(double* d_state, const double* d_rr, double rtol, int32_t kmax, void* stream);
/* alpha = st[0]^2/st[1];  x += alpha p;  r -= alpha Ap;  st[2] = r.r (local
 * part; the caller allreduces st[2] in place when there is more than one rank).
 * cg.cpp:66-73. */
B200_API int b200_cg_update_xr_f64(int32_t n, double* d_state, const double* d_p, const double* d_Ap,
                          double* d_x, double* d_r, void* stream);
/* After the allreduce of st[2]: rnorm_new = sqrt(st[2]); beta = rnorm_new^2 /
 * st[0]^2; unless rnorm_new/st[3] < rtol: p = beta p + r (cg.cpp:76-85).  Then
 * iterations += 1, d_hist[iterations-1] = rnorm_new (if d_hist != NULL and in
 * range), st[0] = rnorm_new, done = converged || iterations == kmax. */
B200_API int b200_cg_update_p_f64(int32_t n, double* d_state, const double* d_r, double* d_p,
                         double* d_hist, int32_t hist_len, void* stream);

/* The whole vector half of an iteration in ONE cooperative launch (cg.cpp:63-85): [p.Ap if need_dot, else
 * st[1] as left by b200_csr_spmv_dot_f64], sum over the ranks, x/r update, r.r, sum over the ranks, p
 * update, bookkeeping as b200_cg_update_p_f64.  peer_ctx: b200_peer_reduce_create context (NULL: one
 * rank).  Grid-wide hand-overs inside the kernel; no-op once st[4] != 0.  Call after b200_cg_init_state. */
B200_API int b200_cg_fused_step_f64(int32_t n, double* d_state, double* d_p, const double* d_Ap, double* d_x,
                           double* d_r, double* d_hist, int32_t hist_len, void* peer_ctx, int need_dot,
                           void* stream);
/* Same, and the launch also performs the PUSH half of the kernel-driven halo exchange of d_p for the next
 * iteration (L2GMap::update(p), cg.cpp:59; b200_halo_push_f64 with halo_desc / n_idx / d_idx) once the new p is
 * out grid-wide -- also when st[4] != 0 (p unchanged), so that every call is exactly one push.  The caller pairs
 * it with b200_halo_wait_unpack_f64 before the next remote-block product.  halo_desc == NULL: plain fused step. */
B200_API int b200_cg_fused_step_halo_f64(int32_t n, double* d_state, double* d_p, const double* d_Ap, double* d_x,
                                double* d_r, double* d_hist, int32_t hist_len, void* peer_ctx, int need_dot,
                                void* halo_desc, int32_t n_idx, const int32_t* d_idx, void* stream);

/* ---- CUDA graphs: a captured sequence of the calls above replayed with one launch (the CG loop:
 *      all per-iteration state, halo epochs included, lives in device memory). ---- */
B200_API int b200_graph_capture_begin(void* stream);
B200_API int b200_graph_capture_end(void* stream, void** graph_exec);
B200_API int b200_graph_launch(void* graph_exec, void* stream);
B200_API int b200_graph_destroy(void* graph_exec);

/* ---- synthetic operators generated on the device (bench / tests only; the
 *      host builders create_A, create_matrix, read_petsc are in include/spmv/).
 *      3-D Laplacian on an nx*ny*nzg grid, index i = x + nx*(y + ny*z), rows
 *      z in [z0, z1) owned; stencil 7 (diag 6, nbrs -1) or 27 (diag 26, nbrs
 *      -1), Dirichlet truncation.  Column numbering is the local one of
 *      Matrix.cpp:294-318: owned g - nx*ny*z0, ghosts (plane z0-1 then plane
 *      z1) ascending after the owned block; rows sorted by local column.
 *      mode 0: one general block;  mode 1: symmetric split (strictly lower
 *      local block + diagonal + remote block, Matrix.cpp:337-349);  mode 2:
 *      overlap split (local block + remote block, Matrix.cpp:350-355).
 *      Step 1 fills the row pointers (rows+1 ints each) and returns the entry
 *      counts (synchronises); step 2 fills the arrays the caller then sized. -- */
B200_API int b200_gen_laplace3d_rowptr(int nx, int ny, int64_t nzg, int64_t z0, int64_t z1, int stencil,
                              int mode, int32_t* d_rowptr_local, int32_t* d_rowptr_remote,
                              int64_t* nnz_local, int64_t* nnz_remote, void* stream);
B200_API int b200_gen_laplace3d_fill(int nx, int ny, int64_t nzg, int64_t z0, int64_t z1, int stencil,
                            int mode, const int32_t* d_rowptr_local, int32_t* d_colind_local,
                            double* d_val_local, double* d_diag, const int32_t* d_rowptr_remote,
                            int32_t* d_colind_remote, double* d_val_remote, void* stream);

/* ---- scalar allreduce over peer memory: the MPI_Allreduce on device scalars of the CG loop
 *      (cg.cpp:49,65,75; cuda/cg.cuda.cu:90,112,133), as one 32-thread kernel that stores into
 *      every rank's IPC-mapped mailbox over NVLink and spins on its own -- no NCCL kernel, same
 *      bits on every rank (sum in rank order).  Setup: each rank b200_ipc_alloc's a mailbox of
 *      b200_peer_reduce_mailbox_bytes(world), the handles are exchanged by the host, every rank
 *      opens the others' (b200_ipc_open) and passes the `world` device pointers (own one at
 *      [rank]) to b200_peer_reduce_create, then a host barrier.  All ranks must call the
 *      allreduce in the same order.  ctx == NULL: single rank, no-op. ------------------------ */
#define B200_PEER_REDUCE_MAX_RANKS 16
#define B200_PEER_REDUCE_MAX_COUNT 4
B200_API int b200_peer_reduce_mailbox_bytes(int32_t world, size_t* num_bytes);
B200_API int b200_peer_reduce_create(void** ctx, int32_t rank, int32_t world, void* const* d_mailboxes, void* stream);
B200_API int b200_peer_reduce_destroy(void* ctx);
B200_API int b200_peer_allreduce_sum_f64(void* ctx, double* d_inout, int32_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SPMV_H */
