// config.h -- build configuration of the B200 host layer (the reference generates
// its config.h from spmv/config.h.in with one macro per backend; here there is one
// backend, so the file is static).
#pragma once
#define SPMV_B200 1
// The reference's CUDA-backend macro: code written against `#ifdef _CUDA` blocks of
// spmv.h (spmv/spmv.h:22-25) finds the B200 executor under the same names.
#ifndef _CUDA
#define _CUDA
#endif
