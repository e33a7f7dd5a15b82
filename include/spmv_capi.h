/* spmv_capi.h -- flat C view of the C++ host layer (include/spmv/), for FFI.
 *
 * The C++ classes mirror the reference's API (spmv::Matrix, spmv::L2GMap, spmv::cg,
 * read_petsc_binary_matrix, create_A, DeviceExecutor); this header exposes them as
 * opaque handles so a host language without C++ interop (Python ctypes here; cgo / JNI
 * elsewhere) can drive the same objects.  Every call returns 0 or -1; on -1
 * spmvh_last_error() holds the text of the C++ exception that was caught.
 * Pointers named d_* are executor (device) memory; everything else is host memory.
 */
#ifndef SPMV_CAPI_H
#define SPMV_CAPI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SPMVH_API __attribute__((visibility("default")))
#else
#define SPMVH_API
#endif

SPMVH_API const char* spmvh_last_error(void);

/* ---- executors: ReferenceExecutor::create() / B200Executor::create(id, host) ---- */
SPMVH_API int spmvh_exec_create_host(void** exec);
SPMVH_API int spmvh_exec_create_b200(void** exec, int device_id);
SPMVH_API int spmvh_exec_destroy(void* exec);
SPMVH_API int spmvh_exec_set_stream(void* exec, void* cuda_stream); /* set_cuda_stream */
SPMVH_API int spmvh_exec_synchronize(void* exec);
SPMVH_API int spmvh_exec_alloc(void* exec, size_t num_bytes, void** d_ptr);
SPMVH_API int spmvh_exec_free(void* exec, void* d_ptr);
SPMVH_API int spmvh_exec_copy_from_host(void* exec, void* d_dst, const void* h_src, size_t num_bytes);
SPMVH_API int spmvh_exec_copy_to_host(void* exec, void* h_dst, const void* d_src, size_t num_bytes);

/* ---- L2GMap (spmv/L2GMap.h).  cm: 0 p2p_blocking, 1 p2p_nonblocking,
 *      2 collective_blocking, 3 collective_nonblocking (mpi_utils.h enum order).
 *      All constructors are collective over the calling thread's MPI_COMM_WORLD. ---- */
SPMVH_API int spmvh_l2g_create(void** map, void* exec, int64_t local_size, const int64_t* ghosts,
                               int64_t num_ghosts, int cm);
SPMVH_API int spmvh_l2g_destroy(void* map);
/* out: local_size, num_ghosts, global_size, global_offset, num_neighbours,
 *      num_indices, rank, overlapping */
SPMVH_API int spmvh_l2g_sizes(void* map, int64_t out[8]);
/* what: 0 ranges(int64) 1 ghosts(int64) 2 neighbours(int32) 3 send_count 4 recv_count
 *       5 send_offset 6 recv_offset 7 indexbuf (all int32).  Copies min(n, cap)
 *       elements into out and stores n. */
SPMVH_API int spmvh_l2g_get(void* map, int what, void* out, int64_t cap, int64_t* n);
SPMVH_API int spmvh_l2g_global_to_local(void* map, int64_t global_index, int32_t* local_index);
SPMVH_API int spmvh_l2g_update_f64(void* map, double* d_vec);
SPMVH_API int spmvh_l2g_update_finalise_f64(void* map, double* d_vec);
SPMVH_API int spmvh_l2g_reverse_update_f64(void* map, double* d_vec);
/* Sum d_val[0..n) (n <= 4) over the map's communicator on the executor's stream, as the CG loop
 * does: use_nccl == 0 -> peer-memory kernel when the map has one (else NCCL), != 0 -> NCCL. */
SPMVH_API int spmvh_l2g_allreduce_f64(void* map, double* d_val, int32_t n, int use_nccl);

/* ---- Matrix<double> builders ---- */
SPMVH_API int spmvh_matrix_create(void** mat, void* exec, const int32_t* rowptr, const int32_t* colind,
                                  const double* values, int64_t nrows_local, int64_t ncols_local,
                                  const int64_t* row_ghosts, int64_t num_row_ghosts, const int64_t* col_ghosts,
                                  int64_t num_col_ghosts, int symmetric, int cm);
SPMVH_API int spmvh_matrix_read_petsc(void** mat, void* exec, const char* filename, int symmetric, int cm);
SPMVH_API int spmvh_matrix_create_A(void** mat, void* exec, int N);
/* Device-side builder of the 7-/27-point Laplacian (demos/CreateLaplace3D.h); exec must be B200. */
SPMVH_API int spmvh_matrix_create_laplace3d(void** mat, void* exec, int nx, int ny, int64_t nz_global, int stencil,
                                            int symmetric, int cm);
SPMVH_API int spmvh_matrix_destroy(void* mat);
/* out: rows, cols, non_zeros, symmetric, format_size, has_remote_block, local nnz, remote nnz */
SPMVH_API int spmvh_matrix_info(void* mat, int64_t out[8]);
/* New handles (destroy with spmvh_l2g_destroy); the maps stay alive while either lives. */
SPMVH_API int spmvh_matrix_col_map(void* mat, void** map);
SPMVH_API int spmvh_matrix_row_map(void* mat, void** map);
/* which: 0 local block, 1 remote block.  dims: rows, cols, nnz, has_diagonal */
SPMVH_API int spmvh_matrix_block_dims(void* mat, int which, int64_t dims[4]);
/* Copies the block back to host arrays sized from block_dims (NULL = skip). */
SPMVH_API int spmvh_matrix_block_copy(void* mat, int which, int32_t* rowptr, int32_t* colind, double* values,
                                      double* diagonal);
SPMVH_API int spmvh_matrix_mult_f64(void* mat, double* d_x, double* d_y); /* Matrix::mult */
/* Matrix::mult_host: both vectors in host memory (pinned for overlap), copies pipelined with the kernel. */
SPMVH_API int spmvh_matrix_mult_host_f64(void* mat, const double* x_host, double* y_host);
/* b200_csr_plan_tune on one block's plan */
SPMVH_API int spmvh_matrix_tune(void* mat, int which, int variant, int tile_bytes, int stages, int consumer_warps,
                                int ctas_per_sm);

/* ---- fp32 twins: Matrix<float>, L2GMap::update<float> (the reference instantiates both: Matrix.cpp:565,
 *      L2GMap.cpp:1001,1007,1012).  A float matrix handle is only valid with the *_f32 entries; maps are
 *      untyped and take either update. ---- */
SPMVH_API int spmvh_l2g_update_f32(void* map, float* d_vec);
SPMVH_API int spmvh_l2g_update_finalise_f32(void* map, float* d_vec);
SPMVH_API int spmvh_l2g_reverse_update_f32(void* map, float* d_vec);
SPMVH_API int spmvh_matrix_create_f32(void** mat, void* exec, const int32_t* rowptr, const int32_t* colind,
                                      const float* values, int64_t nrows_local, int64_t ncols_local,
                                      const int64_t* row_ghosts, int64_t num_row_ghosts, const int64_t* col_ghosts,
                                      int64_t num_col_ghosts, int symmetric, int cm);
SPMVH_API int spmvh_matrix_destroy_f32(void* mat);
SPMVH_API int spmvh_matrix_info_f32(void* mat, int64_t out[8]);
SPMVH_API int spmvh_matrix_col_map_f32(void* mat, void** map);
SPMVH_API int spmvh_matrix_mult_f32(void* mat, float* d_x, float* d_y);
SPMVH_API int spmvh_matrix_mult_host_f32(void* mat, const float* x_host, float* y_host);

/* ---- spmv::cg.  hist (may be NULL) receives ||r_k||, k = 1..iterations. ---- */
SPMVH_API int spmvh_cg(void* exec, void* mat, const double* d_b, double* d_x, int kmax, double rtol, double* hist,
                       int hist_cap, int* iterations);

/* The same solve in profile mode (one iteration at a time, CUDA events between its parts): average device
 * microseconds per iteration of {halo update, local-block SpMV, halo wait + remote block, vector step}. */
SPMVH_API int spmvh_cg_phases(void* exec, void* mat, const double* d_b, double* d_x, int kmax, double rtol,
                              double out_us[4], int* iterations);
/* How this map moves its halo, "<update()>/<inside spmv::cg>", each of "kernel", "copy", "nccl", "none"
 * (NUL-terminated into out[cap]; e.g. "copy/kernel"). */
SPMVH_API int spmvh_l2g_halo_mode(void* map, char* out, int cap);

/* ---- PETSc binary writers (inputs for read_petsc) ---- */
SPMVH_API int spmvh_write_petsc_matrix(const char* filename, int32_t nrows, int32_t ncols, const int32_t* rowptr,
                                       const int32_t* colind, const double* values);
SPMVH_API int spmvh_write_petsc_vector(const char* filename, int32_t n, const double* values);
SPMVH_API int spmvh_read_petsc_vector(void* exec, const char* filename, double** d_vec, int64_t* n_local);

/* ---- config-5 input: power-law rows (demos/CreatePowerLaw.h), host cores only ----
 * counts: entries per row for rows [r0, r1); fill: the rows in CSR order at rowstart[]
 * (exclusive prefix sums of counts, r1 - r0 + 1 long), global column ids;
 * write: the whole n x n matrix as a PETSc binary file for spmvh_matrix_read_petsc. */
SPMVH_API int spmvh_powerlaw_counts(int64_t n, uint64_t seed, double alpha, double xmin, int64_t max_len,
                                    int64_t band, double p_global, int64_t r0, int64_t r1, int32_t* counts);
SPMVH_API int spmvh_powerlaw_fill(int64_t n, uint64_t seed, double alpha, double xmin, int64_t max_len, int64_t band,
                                  double p_global, int64_t r0, int64_t r1, const int64_t* rowstart, int64_t* col,
                                  double* val);
SPMVH_API int spmvh_write_powerlaw_petsc(const char* filename, int64_t n, uint64_t seed, double alpha, double xmin,
                                         int64_t max_len, int64_t band, double p_global, int64_t* nnz);

#ifdef __cplusplus
}
#endif
#endif /* SPMV_CAPI_H */
