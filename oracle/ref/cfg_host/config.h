// oracle/ref/cfg_host/config.h -- stand-in for the CMake-generated config.h when the
// reference's host layer is built for the oracle: sequential backend only, BLAS
// through oracle/shims/cblas.h.
#pragma once
#ifndef _BLAS_OPENBLAS
#define _BLAS_OPENBLAS
#endif
