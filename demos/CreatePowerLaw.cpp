// CreatePowerLaw.cpp -- see CreatePowerLaw.h.
#include "CreatePowerLaw.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <fcntl.h>
#include <cstdlib>
#include <limits>
#include <sched.h>
#include <stdexcept>
#include <unistd.h>
#include <vector>

namespace
{

inline std::uint64_t splitmix64(std::uint64_t x)
{
  std::uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
inline double u01(std::uint64_t key) { return static_cast<double>(splitmix64(key) >> 11) * (1.0 / 9007199254740992.0); }

struct Draw {
  std::int64_t col;
  std::int32_t k;
  double val;
};

// One row: the surviving off-diagonal draws, ascending by column, in `out`.
void row_draws(const PowerLawSpec& s, std::int64_t row, std::int64_t cap, bool want_values, std::vector<Draw>& out)
{
  out.clear();
  const std::uint64_t base = static_cast<std::uint64_t>(row) * 0x100000000ull + s.seed * 0x9E3779B1ull;
  const double u = u01(base);
  double len_f = std::floor(s.xmin / std::pow(1.0 - u, 1.0 / s.alpha));
  std::int64_t len = len_f >= 9.0e18 ? cap : static_cast<std::int64_t>(len_f);
  len = std::min(std::max<std::int64_t>(len, 1), cap);
  const std::int64_t lo = std::max<std::int64_t>(row - s.band, 0), hi = std::min(row + s.band + 1, s.n);
  const double width = static_cast<double>(hi - lo), nd = static_cast<double>(s.n);
  for (std::int64_t k = 0; k < len; ++k) {
    const std::uint64_t key = base + static_cast<std::uint64_t>(k + 1) * 3ull;
    const double sel = u01(key), pos = u01(key + 1);
    std::int64_t c = sel < s.p_global ? static_cast<std::int64_t>(std::floor(pos * nd))
                                      : lo + static_cast<std::int64_t>(std::floor(pos * width));
    c = std::min(std::max<std::int64_t>(c, 0), s.n - 1);
    if (c == row)
      continue;
    out.push_back(Draw{c, static_cast<std::int32_t>(k), want_values ? 2.0 * u01(key + 2) - 1.0 : 0.0});
  }
  std::sort(out.begin(), out.end(), [](const Draw& a, const Draw& b) { return a.col < b.col || (a.col == b.col && a.k < b.k); });
  out.erase(std::unique(out.begin(), out.end(), [](const Draw& a, const Draw& b) { return a.col == b.col; }), out.end());
}

// Writes row `row` (its sorted draws plus the diagonal in place) at col/val[w...].
inline void place_row(std::int64_t row, const std::vector<Draw>& d, std::int64_t w, std::int64_t* col, double* val)
{
  double absum = 0.0; // column order, like the numpy definition's bincount
  for (const Draw& e : d)
    absum += std::fabs(e.val);
  bool placed = false;
  for (const Draw& e : d) {
    if (!placed && e.col > row) {
      col[w] = row, val[w] = 1.0 + absum, ++w;
      placed = true;
    }
    col[w] = e.col, val[w] = e.val, ++w;
  }
  if (!placed)
    col[w] = row, val[w] = 1.0 + absum;
}

inline std::int64_t draw_cap(const PowerLawSpec& s) { return std::min(s.max_len, std::max<std::int64_t>(s.n - 1, 1)); }

void check(const PowerLawSpec& s, std::int64_t r0, std::int64_t r1)
{
  if (s.n <= 0 || r0 < 0 || r1 < r0 || r1 > s.n || s.band < 0 || s.max_len < 1 || !(s.alpha > 0) || !(s.xmin > 0))
    throw std::runtime_error("powerlaw: bad specification or row range");
}

// all the CPUs this process may use (SPMV_HOST_THREADS overrides): launchers export OMP_NUM_THREADS=1
int gen_threads()
{
  if (const char* e = std::getenv("SPMV_HOST_THREADS"))
    if (std::atoi(e) > 0)
      return std::atoi(e);
  cpu_set_t set;
  CPU_ZERO(&set);
  const int n = sched_getaffinity(0, sizeof(set), &set) == 0 ? CPU_COUNT(&set) : 1;
  return n > 0 ? n : 1;
}

void pwrite_all(int fd, const void* buf, std::size_t bytes, std::int64_t offset)
{
  const char* p = static_cast<const char*>(buf);
  while (bytes > 0) {
    const ssize_t w = ::pwrite(fd, p, bytes, offset);
    if (w <= 0)
      throw std::runtime_error("write_powerlaw_petsc: write failed");
    p += w, bytes -= static_cast<std::size_t>(w), offset += w;
  }
}

} // namespace

void powerlaw_row_counts(const PowerLawSpec& s, std::int64_t r0, std::int64_t r1, std::int32_t* counts)
{
  check(s, r0, r1);
  const std::int64_t cap = draw_cap(s);
#pragma omp parallel num_threads(gen_threads())
  {
    std::vector<Draw> d;
#pragma omp for schedule(dynamic, 4096)
    for (std::int64_t r = r0; r < r1; ++r) {
      row_draws(s, r, cap, false, d);
      counts[r - r0] = static_cast<std::int32_t>(d.size()) + 1; // + the diagonal
    }
  }
}

void powerlaw_rows_fill(const PowerLawSpec& s, std::int64_t r0, std::int64_t r1, const std::int64_t* rowstart,
                        std::int64_t* col, double* val)
{
  check(s, r0, r1);
  const std::int64_t cap = draw_cap(s);
#pragma omp parallel num_threads(gen_threads())
  {
    std::vector<Draw> d;
#pragma omp for schedule(dynamic, 4096)
    for (std::int64_t r = r0; r < r1; ++r) {
      row_draws(s, r, cap, true, d);
      place_row(r, d, rowstart[r - r0], col, val);
    }
  }
}

std::int64_t write_powerlaw_petsc(const std::string& filename, const PowerLawSpec& s)
{
  check(s, 0, s.n);
  if (s.n > std::numeric_limits<std::int32_t>::max())
    throw std::runtime_error("write_powerlaw_petsc: row count exceeds the format's int32 header");
  std::vector<std::int32_t> counts(s.n);
  powerlaw_row_counts(s, 0, s.n, counts.data());

  constexpr std::int64_t CHUNK = 65536;
  const std::int64_t nchunks = (s.n + CHUNK - 1) / CHUNK;
  std::vector<std::int64_t> chunk_start(nchunks + 1, 0);
  for (std::int64_t c = 0; c < nchunks; ++c) {
    std::int64_t t = 0;
    for (std::int64_t r = c * CHUNK; r < std::min(s.n, (c + 1) * CHUNK); ++r)
      t += counts[r];
    chunk_start[c + 1] = chunk_start[c] + t;
  }
  const std::int64_t nnz = chunk_start[nchunks];
  // the reference reader multiplies the int32 header field by 4 in int (read_petsc.cpp:80,117-118)
  if (nnz > std::numeric_limits<std::int32_t>::max())
    throw std::runtime_error("write_powerlaw_petsc: entry count exceeds the format's int32 header");

  const int fd = ::open(filename.c_str(), O_CREAT | O_TRUNC | O_WRONLY, 0644);
  if (fd < 0)
    throw std::runtime_error("write_powerlaw_petsc: could not open " + filename);
  bool failed = false;
  try {
    const std::uint32_t header[4] = {__builtin_bswap32(1211216u), __builtin_bswap32(static_cast<std::uint32_t>(s.n)),
                                     __builtin_bswap32(static_cast<std::uint32_t>(s.n)),
                                     __builtin_bswap32(static_cast<std::uint32_t>(nnz))};
    pwrite_all(fd, header, 16, 0);
    const std::int64_t col_base = 16 + 4 * s.n, val_base = col_base + 4 * nnz;
#pragma omp parallel num_threads(gen_threads())
    {
      std::vector<std::int64_t> rowstart, col;
      std::vector<double> val;
      std::vector<std::uint32_t> be32;
      std::vector<std::uint64_t> be64;
#pragma omp for schedule(dynamic, 1)
      for (std::int64_t c = 0; c < nchunks; ++c) {
        if (failed)
          continue;
        try {
          const std::int64_t r0 = c * CHUNK, r1 = std::min(s.n, (c + 1) * CHUNK), m = chunk_start[c + 1] - chunk_start[c];
          rowstart.assign(r1 - r0 + 1, 0);
          be32.resize(std::max<std::int64_t>(r1 - r0, m));
          for (std::int64_t r = r0; r < r1; ++r) {
            rowstart[r - r0 + 1] = rowstart[r - r0] + counts[r];
            be32[r - r0] = __builtin_bswap32(static_cast<std::uint32_t>(counts[r]));
          }
          pwrite_all(fd, be32.data(), 4 * (r1 - r0), 16 + 4 * r0);
          col.resize(m), val.resize(m), be64.resize(m);
          // serial inside the chunk: the enclosing loop already uses every thread
          const std::int64_t cap = draw_cap(s);
          std::vector<Draw> d;
          for (std::int64_t r = r0; r < r1; ++r) {
            row_draws(s, r, cap, true, d);
            place_row(r, d, rowstart[r - r0], col.data(), val.data());
          }
          for (std::int64_t j = 0; j < m; ++j) {
            be32[j] = __builtin_bswap32(static_cast<std::uint32_t>(col[j]));
            std::uint64_t bits;
            std::memcpy(&bits, &val[j], 8);
            be64[j] = __builtin_bswap64(bits);
          }
          pwrite_all(fd, be32.data(), 4 * m, col_base + 4 * chunk_start[c]);
          pwrite_all(fd, be64.data(), 8 * m, val_base + 8 * chunk_start[c]);
        } catch (...) {
          failed = true;
        }
      }
    }
  } catch (...) {
    failed = true;
  }
  ::close(fd);
  if (failed)
    throw std::runtime_error("write_powerlaw_petsc: writing " + filename + " failed");
  return nnz;
}
