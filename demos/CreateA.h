// CreateA.h -- the reference's synthetic test operator (demos/CreateA.h:12).
#pragma once

#include <memory>
#include <mpi.h>
#include <spmv/Matrix.h>
#include <spmv/device_executor.h>

/// 1-D three-point operator: rows [gamma, 1-2 gamma, gamma], gamma = 0.1, first row
/// [1-gamma, gamma], last row [gamma, 1-gamma]; rows split evenly over the ranks.
/// The reference hard-wires a ReferenceExecutor (demos/CreateA.cpp:79-80); here the
/// executor is a parameter because the compute backend is the B200.
spmv::Matrix<double> create_A(MPI_Comm comm, int N, std::shared_ptr<spmv::DeviceExecutor> exec);
