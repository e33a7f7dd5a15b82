// oracle/ref/config.h -- hand-written stand-in for the header CMake generates
// from /root/reference/spmv/config.h.in (all backends off except host OpenMP).
#pragma once
#ifndef _OPENMP_HOST
#define _OPENMP_HOST
#endif
#ifndef _BLAS_OPENBLAS
#define _BLAS_OPENBLAS
#endif
