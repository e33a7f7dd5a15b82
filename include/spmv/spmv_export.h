#pragma once
#if defined(__GNUC__)
#define SPMV_EXPORT __attribute__((visibility("default")))
#else
#define SPMV_EXPORT
#endif
