// read_petsc.h -- PETSc binary matrix / vector readers (reference: spmv/read_petsc.h:
// 29-46, read_petsc.cpp:39-303; file format in SURVEY.md Appendix E: big-endian,
// 32-bit indices).
#pragma once

#include "config.h"
#include "spmv_export.h"

#include "mpi_utils.h"

#include <cstdint>
#include <memory>
#include <mpi.h>
#include <string>

namespace spmv
{

class DeviceExecutor;
template <typename>
class Matrix;

/// Every rank reads its row slice (owner_ranges split, read_petsc.cpp:20-36), numbers
/// ghost columns ascending after the owned ones and builds the local / remote /
/// diagonal blocks with the same rules as Matrix::create_matrix.
SPMV_EXPORT Matrix<double>
read_petsc_binary_matrix(std::string filename, MPI_Comm comm, std::shared_ptr<DeviceExecutor> exec,
                         bool symmetric = false,
                         CommunicationModel cm = CommunicationModel::collective_blocking);

/// Returns this rank's slice in executor memory; free it with exec->free().
SPMV_EXPORT double* read_petsc_binary_vector(MPI_Comm comm, const DeviceExecutor* exec,
                                             const std::string filename);

/// Writers for the same formats (not in the reference; used to produce config-5 inputs).
SPMV_EXPORT void write_petsc_binary_matrix(const std::string& filename, std::int32_t nrows, std::int32_t ncols,
                                           const std::int32_t* rowptr, const std::int32_t* colind,
                                           const double* values);
SPMV_EXPORT void write_petsc_binary_vector(const std::string& filename, std::int32_t n, const double* values);

} // namespace spmv
