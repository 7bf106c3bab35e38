// sparse_matrix.h -- Matrix Market ingest and the operator the harness applies, with the reference's
// semantics (src/sparse_matrix.cpp:11-119, 409-456):
//   Q1  a file entry "I J v" is stored in ROW J, COLUMN I (the applied operator is the transpose of
//       the file's matrix)                                              src/sparse_matrix.cpp:86-87,105-109
//   Q2  values pass through `int` on their way into the rows (1474.779 -> 1474, |v|<1 -> 0)       :107
//   Q3  rows keep file order (the reference's per-row sort works on a copy)                   :113-118
//   Q4  pattern -> 1.0; symmetric -> mirror pushed right after its original when I != J;
//       skew / hermitian read as general; complex and array files are rejected with exit(-1) :29-31,65-69
// Where the reference lays the rows out as padded ELLPACK or "RSA" byte blobs for the Lift kernels
// (cl_encode, :122-399), this host emits CSR (csr_encode) for libshb; `truncate_values = false`
// gives the canonical operator (float values kept) used by the canonical PageRank / SSSP modes.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <limits>
#include <string>
#include <utility>
#include <vector>

#include "common.h"
#include "csds_timer.h"
#include "logger.h"
#include "mm_reader.h"

template <typename T> struct CsrArrays {
  int64_t n_rows = 0, n_cols = 0;
  std::vector<int64_t> row_ptr;  // n_rows + 1
  std::vector<int32_t> col_idx;
  std::vector<T> vals;
  int64_t nnz() const { return (int64_t)col_idx.size(); }
};

template <typename EType> class SparseMatrix {
public:
  template <typename T> using ellpack_row = std::vector<std::pair<int, T>>;
  template <typename T> using ellpack_matrix = std::vector<ellpack_row<T>>;

  explicit SparseMatrix(const std::string &filename) { load_from_file(filename); }
  // an operator that never touches a file (synthetic workloads are generated on the device)
  SparseMatrix(int n_rows, int n_cols) : rows(n_rows), cols(n_cols), nonz(0) {}

  int height() const { return rows; }
  int width() const { return cols; }
  int nonZeros() const { return nonz; }
  std::size_t storedEntries() const { return eI.size(); }

  // ragged rows exactly as the reference's calculate_ellpack builds them (Q1-Q3); used by the gold loop
  ellpack_matrix<EType> &ellpack_encode() {
    if (!ellpack_calculated) calculate_ellpack();
    return ellpackMatrix;
  }

  // CSR of the same operator; padded_height >= height() appends empty rows (the reference pads the
  // height for chunked kernels and the padded rows take part in the iterative apps' termination test, Q7)
  CsrArrays<EType> csr_encode(bool truncate_values, int padded_height = -1) const {
    start_timer(csr_encode, sparse_matrix);
    CsrArrays<EType> A;
    const int64_t H = std::max<int64_t>(rows, padded_height);
    A.n_rows = H;
    A.n_cols = std::max<int64_t>(cols, H);
    A.row_ptr.assign((std::size_t)H + 1, 0);
    for (std::size_t i = 0; i < eI.size(); ++i) A.row_ptr[(std::size_t)eJ[i] + 1]++;
    for (int64_t r = 0; r < H; ++r) A.row_ptr[(std::size_t)r + 1] += A.row_ptr[(std::size_t)r];
    A.col_idx.resize(eI.size());
    A.vals.resize(eI.size());
    std::vector<int64_t> cursor(A.row_ptr.begin(), A.row_ptr.end() - 1);
    for (std::size_t i = 0; i < eI.size(); ++i) {
      const int64_t p = cursor[(std::size_t)eJ[i]]++;
      A.col_idx[(std::size_t)p] = eI[i];
      A.vals[(std::size_t)p] = truncate_values ? through_int(eV[i]) : eV[i];
    }
    return A;
  }

  // v <- (|v| / colsum[I]) * d with the sums taken over the tuples' first index (src/sparse_matrix.cpp:409-431).
  // The sums accumulate in EType in tuple order; quotient and product are evaluated in double and
  // rounded once (the reference calls the double ::fabs on its float).
  void pagerank_normalise(float dampingFactor, EType zero) {
    start_timer(pagerank_normalise, sparse_matrix);
    std::vector<EType> column_sums((std::size_t)std::max(cols, 1), zero);
    for (std::size_t i = 0; i < eI.size(); ++i) column_sums[(std::size_t)eI[i]] = column_sums[(std::size_t)eI[i]] + eV[i];
    for (std::size_t i = 0; i < eI.size(); ++i) {
      const double q = (std::fabs((double)eV[i]) / (double)column_sums[(std::size_t)eI[i]]) * (double)dampingFactor;
      eV[i] = (EType)q;
    }
    ellpack_calculated = false;
  }

  // diagonal -> numeric_limits<EType>::min(), off-diagonal -> the tuple's second index (src/sparse_matrix.cpp:433-456)
  void scc_normalise() {
    start_timer(scc_normalise, sparse_matrix);
    for (std::size_t i = 0; i < eI.size(); ++i)
      eV[i] = eI[i] == eJ[i] ? std::numeric_limits<EType>::min() : (EType)(int)eJ[i];
    ellpack_calculated = false;
  }

  unsigned int maxRowLength() {
    if (!ellpack_calculated) calculate_ellpack();
    return max_width;
  }

private:
  static EType through_int(EType v) { return (EType)(int)v; }

  void load_from_file(const std::string &filename) {
    start_timer(load_from_file, SparseMatrix);
    std::FILE *f = std::fopen(filename.c_str(), "rb");
    if (!f) {
      std::cerr << "Failed to open matrix file " << filename << ENDL;
      std::exit(-1);
    }
    std::string buf;
    {
      char chunk[1 << 16];
      std::size_t got;
      while ((got = std::fread(chunk, 1, sizeof chunk, f)) > 0) buf.append(chunk, got);
    }
    std::fclose(f);
    MMHeader h;
    std::size_t pos = 0;
    if (mm_parse_banner(buf.data(), buf.size(), h, pos) != MM_OK) {
      std::cerr << "Could not read matrix market banner" << ENDL;
      std::exit(-1);
    }
    std::cerr << "Matcode: " << h.typecode() << ENDL;
    if (!(h.is_matrix && h.is_coordinate && (h.is_real || h.is_integer || h.is_pattern))) {
      std::cerr << "Cannot process this matrix type. Typecode: " << h.typecode() << ENDL;
      std::exit(-1);
    }
    if (mm_parse_crd_size(buf.data(), buf.size(), h, pos) != MM_OK) {
      std::cerr << "Cannot read matrix sizes and number of non-zeros" << ENDL;
      return;
    }
    rows = h.rows;
    cols = h.cols;
    nonz = h.nonz;
    std::cerr << "Rows " << rows << " cols " << cols << " non-zeros " << nonz << ENDL;
    const std::size_t cap = (std::size_t)std::max(nonz, 0) * (h.is_symmetric ? 2 : 1);
    eI.reserve(cap);
    eJ.reserve(cap);
    eV.reserve(cap);
    // entries are whitespace-separated tokens, line structure is irrelevant (fscanf semantics)
    const char *p = buf.c_str() + pos;
    for (int i = 0; i < nonz; ++i) {
      char *end = nullptr;
      const long I1 = std::strtol(p, &end, 10);
      if (end == p) break;
      p = end;
      const long J1 = std::strtol(p, &end, 10);
      if (end == p) break;
      p = end;
      double v = 1.0;
      if (!h.is_pattern) {
        v = std::strtod(p, &end);
        if (end == p) break;
        p = end;
      }
      const int32_t I = (int32_t)I1 - 1, J = (int32_t)J1 - 1;
      if (I < 0 || J < 0 || I >= cols_or_rows() || J >= rows) {
        std::cerr << "Matrix entry (" << I1 << ", " << J1 << ") lies outside the declared " << rows << " x " << cols
                  << " matrix" << ENDL;
        std::exit(-1);
      }
      const EType tv = static_cast<EType>(v);
      eI.push_back(I);
      eJ.push_back(J);
      eV.push_back(tv);
      if (h.is_symmetric && I != J) {
        eI.push_back(J);
        eJ.push_back(I);
        eV.push_back(tv);
      }
    }
  }

  int cols_or_rows() const { return std::max(cols, rows); }

  void calculate_ellpack() {
    start_timer(calculate_ellpack, sparse_matrix);
    ellpack_calculated = true;
    std::vector<unsigned int> lengths((std::size_t)std::max(rows, 0), 0);
    max_width = 0;
    for (std::size_t i = 0; i < eJ.size(); ++i) max_width = std::max(max_width, ++lengths[(std::size_t)eJ[i]]);
    LOG_DEBUG("max width: ", max_width);
    ellpackMatrix.assign((std::size_t)std::max(rows, 0), ellpack_row<EType>());
    for (int r = 0; r < rows; ++r) ellpackMatrix[(std::size_t)r].reserve(lengths[(std::size_t)r]);
    for (std::size_t i = 0; i < eI.size(); ++i)
      ellpackMatrix[(std::size_t)eJ[i]].push_back(std::pair<int, EType>(eI[i], through_int(eV[i])));
  }

  // tuples (I, J, value) in the loader's order, stored as three arrays
  std::vector<int32_t> eI, eJ;
  std::vector<EType> eV;
  int rows = 0, cols = 0, nonz = 0;

  bool ellpack_calculated = false;
  unsigned int max_width = 0;
  ellpack_matrix<EType> ellpackMatrix;
};
