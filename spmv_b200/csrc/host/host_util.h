// host_util.h -- internal helpers of the host layer.
#pragma once

#include "b200spmv.h"

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

namespace spmv
{
namespace detail
{

/// C-ABI status -> exception, the way the reference's constructors throw
/// std::runtime_error (Matrix.cpp:29, L2GMap.cpp:363).
inline void check(int rc, const char* what)
{
  if (rc != B200_OK)
    throw std::runtime_error(std::string("spmv: ") + what + ": " + b200_last_error());
}

/// read_petsc.cpp:20-36 / demos/CreateA.cpp:13-29: the first N % size ranks own one more.
inline std::vector<std::int64_t> owner_ranges(int size, std::int64_t N)
{
  const std::int64_t n = N / size, r = N % size;
  std::vector<std::int64_t> ranges(size + 1);
  for (int rank = 0; rank <= size; ++rank)
    ranges[rank] = rank < r ? rank * (n + 1) : rank * n + r;
  return ranges;
}

} // namespace detail
} // namespace spmv
