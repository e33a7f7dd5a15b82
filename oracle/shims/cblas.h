/* oracle/shims/cblas.h -- TEST INFRASTRUCTURE.
 * The three BLAS-1 calls of cg.cpp:47-85.  The reference links an un-vendored,
 * un-pinned system BLAS (CMakeLists.txt:71), so its summation order is not recoverable;
 * ours is a documented choice ("parity unpinned" for CG): daxpy / dscal are
 * plain loops, ddot is the left-to-right sum in twice the working precision (Dot2), which
 * makes the value independent of the summation order.  Definitions live in
 * oracle/restated/kernels.c. */
#pragma once
#ifdef __cplusplus
extern "C" {
#endif
double oracle_ddot(int n, const double* x, const double* y);
void oracle_daxpy(int n, double a, const double* x, double* y);
void oracle_dscal(int n, double a, double* x);
#ifdef __cplusplus
}
#endif
static inline double cblas_ddot(const int n, const double* x, const int incx, const double* y, const int incy)
{
  (void)incx;
  (void)incy;
  return oracle_ddot(n, x, y);
}
static inline void cblas_daxpy(const int n, const double a, const double* x, const int incx, double* y, const int incy)
{
  (void)incx;
  (void)incy;
  oracle_daxpy(n, a, x, y);
}
static inline void cblas_dscal(const int n, const double a, double* x, const int incx)
{
  (void)incx;
  oracle_dscal(n, a, x);
}
