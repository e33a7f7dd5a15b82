// spmv.h -- umbrella header (reference: spmv/spmv.h).
#pragma once

#include "config.h"

#include "L2GMap.h"
#include "Matrix.h"
#include "b200_executor.h"
#include "cg.h"
#include "csr_matrix.h"
#include "device_executor.h"
#include "read_petsc.h"
#include "reference_executor.h"
