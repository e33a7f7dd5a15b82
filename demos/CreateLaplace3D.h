// CreateLaplace3D.h -- synthetic 3-D Laplacian operators of BASELINE.json's configs 2-4.
//
// The reference ships only the 1-D create_A (demos/CreateA.cpp:31-103); the 7- and
// 27-point operators are this project's inputs.  This is a device-side
// builder: the blocks are generated straight into executor memory in
// the layout Matrix::create_matrix would produce (Matrix.cpp:294-358), never
// materialising host CSR, so a 449 M-entry operator is ready in milliseconds.
#pragma once

#include <cstdint>
#include <memory>
#include <mpi.h>
#include <spmv/Matrix.h>
#include <spmv/device_executor.h>
#include <spmv/mpi_utils.h>

/// nx*ny*nz_global grid, index i = x + nx*(y + ny*z); stencil 7 (diagonal 6) or 27
/// (diagonal 26), neighbours -1, Dirichlet truncation.  Planes are split over the
/// ranks with the reference's owner_ranges rule (read_petsc.cpp:20-36).
///   symmetric          strictly-lower local block + diagonal + remote block
///   non-blocking `cm`  local + remote blocks (overlap schedule)
///   otherwise          one block over [owned | ghost] columns
/// `exec` must be a B200Executor (there is no CPU generator path).
spmv::Matrix<double> create_laplace3d(MPI_Comm comm, std::shared_ptr<spmv::DeviceExecutor> exec, int nx, int ny,
                                      std::int64_t nz_global, int stencil, bool symmetric,
                                      spmv::CommunicationModel cm);
