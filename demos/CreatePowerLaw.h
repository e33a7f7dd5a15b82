// CreatePowerLaw.h -- the irregular power-law-row matrix of BASELINE.json's config 5.
//
// The reference ships no generator for this input (SURVEY.md 8d lists it as "synthetic");
// it is read with read_petsc_binary_matrix (read_petsc.cpp:39-228), so what this file
// produces is the PETSc binary file, written straight from a
// counter-based row generator: any thread can produce any row, nothing but one chunk of
// rows per thread is ever held in memory, and the 50 M-row / ~4e8-entry file is ready in
// seconds on the host cores.  spmv_b200/synth.py::powerlaw_rows is the same definition in
// numpy; tests/test_oracle.py holds the two bit-for-bit equal.
//
// Row i: length = clamp(floor(Pareto(alpha, xmin)), 1, min(max_len, n-1)) off-diagonal
// draws, each with probability p_global uniform over all columns and otherwise uniform in
// [i-band, i+band]; draws on the diagonal dropped, duplicates keep the first draw, columns
// ascending; off-diagonal values uniform in [-1, 1); diagonal = 1 + sum |off-diagonal|.
#pragma once

#include <cstdint>
#include <string>

struct PowerLawSpec {
  std::int64_t n = 0;             ///< global rows = columns
  std::uint64_t seed = 20261104;  ///< SURVEY.md 8d
  double alpha = 2.0, xmin = 4.0; ///< Pareto shape / scale of the row length
  std::int64_t max_len = 4096;    ///< clamp of the draw count
  std::int64_t band = 32768;      ///< half-width of the local band
  double p_global = 0.1;          ///< fraction of draws over all columns
};

/// Entries per row for rows [r0, r1): counts[r - r0].
void powerlaw_row_counts(const PowerLawSpec& spec, std::int64_t r0, std::int64_t r1, std::int32_t* counts);

/// Rows [r0, r1) in CSR order; `rowstart` (r1 - r0 + 1 exclusive prefix sums of the
/// counts) says where each row goes in `col` (global column ids) and `val`.
void powerlaw_rows_fill(const PowerLawSpec& spec, std::int64_t r0, std::int64_t r1, const std::int64_t* rowstart,
                        std::int64_t* col, double* val);

/// Writes the whole matrix as a PETSc binary matrix file; returns the entry count.
/// Throws std::runtime_error when n or the entry count leaves the format's int32 header.
std::int64_t write_powerlaw_petsc(const std::string& filename, const PowerLawSpec& spec);
