// host_util.h -- internal helpers of the host layer.
#pragma once

#include "b200spmv.h"

#include <cstdint>
#include <cstdlib>
#include <sched.h>
#include <stdexcept>
#include <string>
#include <vector>

// NVTX ranges around the phases of the path (pack / exchange / local block / remote block / vector step +
// reductions), so a timeline shows the same five bars the reference's demo timers print (demos/spmv.cpp:86-102).
// Header-only nvtx3: a no-op unless a profiler injects itself; compiled out where the header is missing.
#if defined(__has_include)
#if __has_include(<nvtx3/nvToolsExt.h>)
#include <nvtx3/nvToolsExt.h>
#define SPMV_HAVE_NVTX 1
#endif
#endif

namespace spmv
{
namespace detail
{
struct NvtxRange {
#ifdef SPMV_HAVE_NVTX
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
#else
  explicit NvtxRange(const char*) {}
#endif
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

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

/// Bumped whenever device arrays or launch plans a captured CUDA graph may have baked in go away or change
/// (a Matrix is destroyed or moved, a plan is re-tuned): cached graphs compare it with the value at capture.
std::uint64_t structure_epoch();
void bump_structure_epoch();

/// Threads for the host-side builders: SPMV_HOST_THREADS if set, else the CPUs this process may run
/// on divided by `sharers` (ranks doing the same work at the same time).  Launchers such as torchrun
/// export OMP_NUM_THREADS=1, which would serialise a 50 M-row setup; the builders pass this count to
/// their parallel regions explicitly instead of relying on the OpenMP default.
inline int host_threads(int sharers = 1)
{
  if (const char* e = std::getenv("SPMV_HOST_THREADS"))
    if (std::atoi(e) > 0)
      return std::atoi(e);
  cpu_set_t set;
  CPU_ZERO(&set);
  int n = sched_getaffinity(0, sizeof(set), &set) == 0 ? CPU_COUNT(&set) : 1;
  if (const char* e = std::getenv("LOCAL_WORLD_SIZE"))
    if (std::atoi(e) > 0)
      sharers = std::atoi(e);
  n /= sharers > 0 ? sharers : 1;
  return n > 0 ? n : 1;
}

} // namespace detail
} // namespace spmv
