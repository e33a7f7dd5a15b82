// gen.cu -- synthetic 3-D Laplacian blocks generated on the device.
//
// The reference has no stencil generator (its only synthetic operator is the
// 1-D create_A, demos/CreateA.cpp:31-103); BASELINE.json's configs 2-4 name
// 7- and 27-point Laplacians on 256^3 (per GPU).  This builds, per rank of a
// 1-D slab partition (whole z-planes, owner_ranges of read_petsc.cpp:20-36), the
// same blocks Matrix::create_matrix would produce from the host CSR
// (Matrix.cpp:320-408): local column numbering with ghosts ascending after the
// owned columns (Matrix.cpp:294-318), rows sorted by local column id
// (setFromTriplets), and the symmetric / overlap splits of Matrix.cpp:337-358.
// It exists so that full-size operators (449 M entries) appear in HBM in
// milliseconds; tests check it against the host path on small grids.
#include "common.cuh"

#include <cub/device/device_scan.cuh>

namespace b200
{

struct GridSpec {
  int nx, ny;
  long long nzg, z0, z1;
  int stencil; // 7 or 27
  int mode;    // 0 one block, 1 symmetric split, 2 overlap split
};

enum EntryClass { OWNED_LOWER, DIAGONAL, OWNED_UPPER, GHOST };

// Calls f(local_col, value, class) for the entries of global row (x, y, z) in
// ascending local-column order: owned columns, then lower-plane ghosts, then
// upper-plane ghosts.
template <typename F>
__device__ __forceinline__ void for_each_entry(const GridSpec& g, int x, int y, long long z, F f)
{
  const long long plane = (long long)g.nx * g.ny;
  const long long ncols_local = plane * (g.z1 - g.z0);
  const long long nlow = g.z0 > 0 ? plane : 0;
  const double diag = g.stencil == 7 ? 6.0 : 26.0;
  // segment 0: owned planes; 1: ghost plane z0-1; 2: ghost plane z1
  for (int seg = 0; seg < 3; ++seg) {
    for (int dz = -1; dz <= 1; ++dz) {
      const long long zz = z + dz;
      if (zz < 0 || zz >= g.nzg)
        continue;
      const int where = (zz >= g.z0 && zz < g.z1) ? 0 : (zz < g.z0 ? 1 : 2);
      if (where != seg)
        continue;
      for (int dy = -1; dy <= 1; ++dy) {
        const int yy = y + dy;
        if (yy < 0 || yy >= g.ny)
          continue;
        for (int dx = -1; dx <= 1; ++dx) {
          const int xx = x + dx;
          if (xx < 0 || xx >= g.nx)
            continue;
          const int nz_off = (dx != 0) + (dy != 0) + (dz != 0);
          if (g.stencil == 7 && nz_off > 1)
            continue;
          const long long in_plane = xx + (long long)g.nx * yy;
          long long col;
          EntryClass cls;
          if (where == 0) {
            col = in_plane + plane * (zz - g.z0);
            cls = nz_off == 0 ? DIAGONAL
                              : ((dz < 0 || (dz == 0 && (dy < 0 || (dy == 0 && dx < 0))))
                                     ? OWNED_LOWER
                                     : OWNED_UPPER);
          } else if (where == 1) {
            col = ncols_local + in_plane;
            cls = GHOST;
          } else {
            col = ncols_local + nlow + in_plane;
            cls = GHOST;
          }
          f((int)col, nz_off == 0 ? diag : -1.0, cls);
        }
      }
    }
  }
}

// destination of an entry under a mode: 0 local block, 1 remote block,
// 2 diagonal vector, -1 dropped
__device__ __forceinline__ int dest_of(int mode, EntryClass c)
{
  if (mode == 0)
    return 0;
  if (mode == 1)
    return c == OWNED_LOWER ? 0 : (c == DIAGONAL ? 2 : (c == GHOST ? 1 : -1));
  return c == GHOST ? 1 : 0;
}

__global__ void laplace_count_kernel(GridSpec g, int rows, int32_t* __restrict__ cnt_local,
                                     int32_t* __restrict__ cnt_remote)
{
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows)
    return;
  const int x = r % g.nx, y = (r / g.nx) % g.ny;
  const long long z = g.z0 + r / (g.nx * g.ny);
  int nl = 0, nr = 0;
  for_each_entry(g, x, y, z, [&](int, double, EntryClass c) {
    const int d = dest_of(g.mode, c);
    nl += d == 0;
    nr += d == 1;
  });
  cnt_local[r] = nl;
  if (cnt_remote)
    cnt_remote[r] = nr;
}

__global__ void laplace_fill_kernel(GridSpec g, int rows, const int32_t* __restrict__ rp_local,
                                    int32_t* __restrict__ ci_local, double* __restrict__ v_local,
                                    double* __restrict__ diag, const int32_t* __restrict__ rp_remote,
                                    int32_t* __restrict__ ci_remote, double* __restrict__ v_remote)
{
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows)
    return;
  const int x = r % g.nx, y = (r / g.nx) % g.ny;
  const long long z = g.z0 + r / (g.nx * g.ny);
  int pl = rp_local[r];
  int pr = rp_remote ? rp_remote[r] : 0;
  for_each_entry(g, x, y, z, [&](int col, double val, EntryClass c) {
    const int d = dest_of(g.mode, c);
    if (d == 0) {
      ci_local[pl] = col;
      v_local[pl] = val;
      ++pl;
    } else if (d == 1) {
      ci_remote[pr] = col;
      v_remote[pr] = val;
      ++pr;
    } else if (d == 2) {
      diag[r] = val;
    }
  });
}

static int check_spec(const GridSpec& g, int* rows_out)
{
  B200_REQUIRE(g.nx > 0 && g.ny > 0 && g.nzg > 0, "empty grid");
  B200_REQUIRE(g.z0 >= 0 && g.z0 <= g.z1 && g.z1 <= g.nzg, "bad slab");
  B200_REQUIRE(g.stencil == 7 || g.stencil == 27, "stencil must be 7 or 27");
  B200_REQUIRE(g.mode >= 0 && g.mode <= 2, "mode must be 0, 1 or 2");
  const long long rows = (long long)g.nx * g.ny * (g.z1 - g.z0);
  B200_REQUIRE(rows * 27 < (1ll << 31), "slab too large for int32 rowptr");
  *rows_out = (int)rows;
  if (num_sms() <= 0) {
    set_error("no CUDA device available: this build has no CPU path");
    return B200_ECUDA;
  }
  return B200_OK;
}

static int exclusive_scan(int32_t* d_data, int n, cudaStream_t st)
{
  void* tmp = nullptr;
  size_t tmp_bytes = 0;
  B200_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_data, d_data, n, st));
  B200_CUDA(cudaMalloc(&tmp, tmp_bytes));
  cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, d_data, d_data, n, st);
  if (e == cudaSuccess)
    e = cudaStreamSynchronize(st);
  cudaFree(tmp);
  B200_CUDA(e);
  return B200_OK;
}

} // namespace b200

using namespace b200;

extern "C" {

int b200_gen_laplace3d_rowptr(int nx, int ny, int64_t nzg, int64_t z0, int64_t z1, int stencil,
                              int mode, int32_t* d_rowptr_local, int32_t* d_rowptr_remote,
                              int64_t* nnz_local, int64_t* nnz_remote, void* stream)
{
  GridSpec g{nx, ny, (long long)nzg, (long long)z0, (long long)z1, stencil, mode};
  int rows = 0;
  if (int rc = check_spec(g, &rows))
    return rc;
  B200_REQUIRE(d_rowptr_local != nullptr, "rowptr_local is NULL");
  B200_REQUIRE(mode == 0 || d_rowptr_remote != nullptr, "rowptr_remote is NULL");
  cudaStream_t st = as_stream(stream);
  if (rows > 0) {
    laplace_count_kernel<<<(rows + 255) / 256, 256, 0, st>>>(g, rows, d_rowptr_local,
                                                            mode == 0 ? nullptr : d_rowptr_remote);
    B200_LAUNCH_CHECK();
  }
  // the scan runs over rows+1 entries so the last one becomes the total
  B200_CUDA(cudaMemsetAsync(d_rowptr_local + rows, 0, sizeof(int32_t), st));
  if (int rc = exclusive_scan(d_rowptr_local, rows + 1, st))
    return rc;
  int32_t tl = 0, tr = 0;
  B200_CUDA(cudaMemcpyAsync(&tl, d_rowptr_local + rows, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  if (mode != 0) {
    B200_CUDA(cudaMemsetAsync(d_rowptr_remote + rows, 0, sizeof(int32_t), st));
    if (int rc = exclusive_scan(d_rowptr_remote, rows + 1, st))
      return rc;
    B200_CUDA(
        cudaMemcpyAsync(&tr, d_rowptr_remote + rows, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  }
  B200_CUDA(cudaStreamSynchronize(st));
  if (nnz_local)
    *nnz_local = tl;
  if (nnz_remote)
    *nnz_remote = tr;
  return B200_OK;
}

int b200_gen_laplace3d_fill(int nx, int ny, int64_t nzg, int64_t z0, int64_t z1, int stencil,
                            int mode, const int32_t* d_rowptr_local, int32_t* d_colind_local,
                            double* d_val_local, double* d_diag, const int32_t* d_rowptr_remote,
                            int32_t* d_colind_remote, double* d_val_remote, void* stream)
{
  GridSpec g{nx, ny, (long long)nzg, (long long)z0, (long long)z1, stencil, mode};
  int rows = 0;
  if (int rc = check_spec(g, &rows))
    return rc;
  B200_REQUIRE(d_rowptr_local != nullptr, "rowptr_local is NULL");
  B200_REQUIRE(mode != 1 || d_diag != nullptr, "diag is NULL in symmetric mode");
  B200_REQUIRE(mode == 0 || d_rowptr_remote != nullptr, "rowptr_remote is NULL");
  if (rows == 0)
    return B200_OK;
  laplace_fill_kernel<<<(rows + 255) / 256, 256, 0, as_stream(stream)>>>(
      g, rows, d_rowptr_local, d_colind_local, d_val_local, d_diag,
      mode == 0 ? nullptr : d_rowptr_remote, d_colind_remote, d_val_remote);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

} // extern "C"
