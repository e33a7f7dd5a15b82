/* oracle/restated/kernels.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference's per-rank CPU arithmetic for the hot
 * path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg
 * may load this; the product (spmv_b200/, include/) never does.
 * Pinned against the reference's own compiled csr_kernels.cpp (oracle/_ref) and
 * the golden values of SURVEY.md Appendix D by tests/test_oracle.py.
 * Compile without FMA contraction (-ffp-contract=off) so that the result does
 * not depend on the host ISA.
 */
#include <stdint.h>
#include <stddef.h>

/* csr_kernels.cpp:41-51 -- general branch of CSRSpMV<T>::run */
void oracle_csr_spmv_f64(int32_t num_rows, const int32_t* rowptr, const int32_t* colind,
                         const double* values, double alpha, const double* in, double beta,
                         double* out)
{
  for (int32_t i = 0; i < num_rows; ++i) {
    double sum = 0.0;
    for (int32_t j = rowptr[i]; j < rowptr[i + 1]; ++j)
      sum += values[j] * in[colind[j]];
    out[i] = alpha * sum + beta * out[i];
  }
}

void oracle_csr_spmv_f32(int32_t num_rows, const int32_t* rowptr, const int32_t* colind,
                         const float* values, float alpha, const float* in, float beta, float* out)
{
  for (int32_t i = 0; i < num_rows; ++i) {
    float sum = 0.0f;
    for (int32_t j = rowptr[i]; j < rowptr[i + 1]; ++j)
      sum += values[j] * in[colind[j]];
    out[i] = alpha * sum + beta * out[i];
  }
}

/* csr_kernels.cpp:26-40 -- symmetric branch: strictly lower triangle + diagonal.
 * num_non_zeros == 0 (rowptr may be NULL) leaves only the diagonal term. */
void oracle_csr_symv_f64(int32_t num_rows, int64_t num_non_zeros, const int32_t* rowptr,
                         const int32_t* colind, const double* values, const double* diagonal,
                         double alpha, const double* in, double beta, double* out)
{
  for (int32_t i = 0; i < num_rows; ++i) {
    double sum = diagonal[i] * in[i];
    if (num_non_zeros > 0) {
      for (int32_t j = rowptr[i]; j < rowptr[i + 1]; ++j) {
        int32_t col = colind[j];
        double val = values[j];
        sum += val * in[col];
        out[col] += alpha * val * in[i];
      }
    }
    out[i] = alpha * sum + beta * out[i];
  }
}

void oracle_csr_symv_f32(int32_t num_rows, int64_t num_non_zeros, const int32_t* rowptr,
                         const int32_t* colind, const float* values, const float* diagonal,
                         float alpha, const float* in, float beta, float* out)
{
  for (int32_t i = 0; i < num_rows; ++i) {
    float sum = diagonal[i] * in[i];
    if (num_non_zeros > 0) {
      for (int32_t j = rowptr[i]; j < rowptr[i + 1]; ++j) {
        int32_t col = colind[j];
        float val = values[j];
        sum += val * in[col];
        out[col] += alpha * val * in[i];
      }
    }
    out[i] = alpha * sum + beta * out[i];
  }
}

/* reference_executor.cpp:158-164 -- gather_ghosts_run */
void oracle_gather_f64(int32_t n, const int32_t* indices, const double* in, double* out)
{
  for (int32_t i = 0; i < n; ++i)
    out[i] = in[indices[i]];
}

/* L2GMap.cpp:920-921 -- accumulate loop of reverse_update */
void oracle_scatter_add_f64(int32_t n, const int32_t* indices, const double* in, double* inout)
{
  for (int32_t i = 0; i < n; ++i)
    inout[indices[i]] += in[i];
}

/* cg.cpp:47,63,73 -- cblas_ddot with unit strides.  The reference's BLAS is an
 * un-vendored, un-pinned system library (CMakeLists.txt:71); its summation order and
 * accuracy are unknown, so the dot product is ours to define.  The definition: the
 * left-to-right sum evaluated in twice the working precision (Ogita, Rump & Oishi's
 * Dot2: products split exactly with fma, additions with TwoSum, errors carried along),
 * rounded once at the end.  Such a value does not depend on the summation order except
 * with probability ~1e-14, so an implementation that sums in another order (the GPU's
 * tree) follows the same CG trajectory instead of drifting apart by a factor ~2 per
 * iteration.  oracle_ddot_plain keeps the naive loop for comparison. */
#include <math.h>
double oracle_ddot(int32_t n, const double* x, const double* y)
{
  double hi = 0.0, lo = 0.0;
  for (int32_t i = 0; i < n; ++i) {
    const double p = x[i] * y[i];
    const double pe = fma(x[i], y[i], -p);
    const double s = hi + p;
    const double bb = s - hi;
    const double e = (hi - (s - bb)) + (p - bb);
    hi = s;
    lo = lo + (e + pe);
  }
  return hi + lo;
}

double oracle_ddot_plain(int32_t n, const double* x, const double* y)
{
  double s = 0.0;
  for (int32_t i = 0; i < n; ++i)
    s += x[i] * y[i];
  return s;
}

/* cg.cpp:69,70,85 -- cblas_daxpy: y += a*x */
void oracle_daxpy(int32_t n, double a, const double* x, double* y)
{
  for (int32_t i = 0; i < n; ++i)
    y[i] += a * x[i];
}

/* cg.cpp:84 -- cblas_dscal: x *= a */
void oracle_dscal(int32_t n, double a, double* x)
{
  for (int32_t i = 0; i < n; ++i)
    x[i] *= a;
}

/* ---- input generator for the CPU legs (bench.py cpu_baseline / --impl reference): rows of planes [z0, z1)
 * of the 7- or 27-point Laplacian on an nx*ny*nzg grid, the same definition as spmv_b200/synth.py::
 * laplace3d_rows (index i = x + nx*(y + ny*z), diagonal 6 / 26, neighbours -1, Dirichlet truncation),
 * GLOBAL column ids ascending within a row.  mode 0: every entry.  mode 1: the one-rank symmetric split of
 * Matrix.cpp:337-349 -- strictly lower triangle in (rowptr, colind, values), diagonal in diag[].
 * Call once with colind == NULL to fill rowptr (rows + 1 entries) and get the entry count, then again with
 * the arrays sized from it.  Plain C, one thread: 449 M entries in a couple of seconds. */
long long oracle_laplace3d(int nx, int ny, long long nzg, long long z0, long long z1, int stencil, int mode,
                           long long* rowptr, int* colind, double* values, double* diag)
{
  const double dval = stencil == 7 ? 6.0 : 26.0;
  long long n = 0, row = 0;
  for (long long z = z0; z < z1; ++z)
    for (int y = 0; y < ny; ++y)
      for (int x = 0; x < nx; ++x, ++row) {
        const long long self = x + (long long)nx * (y + (long long)ny * z);
        rowptr[row] = n;
        for (int dz = -1; dz <= 1; ++dz)
          for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) {
              const int off = (dx != 0) + (dy != 0) + (dz != 0);
              if (stencil == 7 && off > 1)
                continue;
              const long long zz = z + dz;
              const int yy = y + dy, xx = x + dx;
              if (xx < 0 || xx >= nx || yy < 0 || yy >= ny || zz < 0 || zz >= nzg)
                continue;
              const long long c = xx + (long long)nx * (yy + (long long)ny * zz);
              if (mode == 1) {
                if (c == self) {
                  if (diag)
                    diag[row] = dval;
                  continue;
                }
                if (c > self)
                  continue;
              }
              if (colind) {
                colind[n] = (int)c;
                values[n] = off == 0 ? dval : -1.0;
              }
              ++n;
            }
      }
  rowptr[row] = n;
  return n;
}
