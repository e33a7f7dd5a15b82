// oracle/ref/spmv_export.h -- stand-in for CMake's generate_export_header output.
#pragma once
#define SPMV_EXPORT
#define SPMV_NO_EXPORT
