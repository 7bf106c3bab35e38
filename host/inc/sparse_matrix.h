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
#include <cstring>
#include <functional>
#include <iostream>
#include <limits>
#include <string>
#include <utility>
#include <thread>
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
  // the tuple list itself (first index, second index = row, value), for the device-side encode (shb_csr_from_coo)
  const std::vector<int32_t> &tupleFirst() const { return eI; }
  const std::vector<int32_t> &tupleSecond() const { return eJ; }
  const std::vector<EType> &tupleValues() const { return eV; }

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
    A.col_idx.resize(eI.size());
    A.vals.resize(eI.size());
    // Stable counting sort of the tuples by row (the tuple's second index, Q1).  Large inputs: every thread owns a
    // contiguous range of ROWS, scans all tuples in file order and handles the ones that fall in its range -- the
    // reads are sequential, each thread's counters and output slab stay in its own cache, and the in-row order is the
    // file order whatever the thread count (same arrays as the sequential loop, bit for bit).
    unsigned n_thr = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    if (const char *e = std::getenv("SHB_MM_THREADS")) n_thr = (unsigned)std::max(1, std::atoi(e));
    std::size_t min_entries = (std::size_t)1 << 20;  // below this the threads do not pay
    if (const char *e = std::getenv("SHB_MM_PARALLEL_MIN_BYTES")) min_entries = (std::size_t)std::strtoull(e, nullptr, 10) / 8;
    if (eI.size() < min_entries || H < (int64_t)n_thr * (min_entries < ((std::size_t)1 << 20) ? 1 : 1024)) n_thr = 1;
    auto in_parallel = [&](const std::function<void(unsigned)> &fn) {
      std::vector<std::thread> th;
      for (unsigned t = 1; t < n_thr; ++t) th.emplace_back(fn, t);
      fn(0);
      for (auto &x : th) x.join();
    };
    auto row_lo = [&](unsigned t) { return (int64_t)((__int128)H * t / n_thr); };
    in_parallel([&](unsigned t) {
      const int64_t lo = row_lo(t), hi = row_lo(t + 1);
      for (std::size_t i = 0; i < eJ.size(); ++i) {
        const int64_t r = eJ[i];
        if (r >= lo && r < hi) A.row_ptr[(std::size_t)r + 1]++;
      }
    });
    for (int64_t r = 0; r < H; ++r) A.row_ptr[(std::size_t)r + 1] += A.row_ptr[(std::size_t)r];
    in_parallel([&](unsigned t) {
      const int64_t lo = row_lo(t), hi = row_lo(t + 1);
      std::vector<int64_t> cursor(A.row_ptr.begin() + lo, A.row_ptr.begin() + hi);
      for (std::size_t i = 0; i < eI.size(); ++i) {
        const int64_t r = eJ[i];
        if (r < lo || r >= hi) continue;
        const int64_t p = cursor[(std::size_t)(r - lo)]++;
        A.col_idx[(std::size_t)p] = eI[i];
        A.vals[(std::size_t)p] = truncate_values ? through_int(eV[i]) : eV[i];
      }
    });
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
    ellpack_calculated = false;  // values changed; the row lengths (max_width) did not
  }

  // diagonal -> numeric_limits<EType>::min(), off-diagonal -> the tuple's second index (src/sparse_matrix.cpp:433-456)
  void scc_normalise() {
    start_timer(scc_normalise, sparse_matrix);
    for (std::size_t i = 0; i < eI.size(); ++i)
      eV[i] = eI[i] == eJ[i] ? std::numeric_limits<EType>::min() : (EType)(int)eJ[i];
    ellpack_calculated = false;
  }

  // longest row (the ELL width the reference pads to); needs the row lengths only, not the ragged rows
  unsigned int maxRowLength() {
    if (!ellpack_calculated && !width_calculated) {
      start_timer(max_row_length, sparse_matrix);
      std::vector<unsigned int> lengths((std::size_t)std::max(rows, 0), 0);
      max_width = 0;
      for (std::size_t i = 0; i < eJ.size(); ++i) max_width = std::max(max_width, ++lengths[(std::size_t)eJ[i]]);
      width_calculated = true;
    }
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
      std::fseek(f, 0, SEEK_END);
      const long sz = std::ftell(f);
      std::rewind(f);
      if (sz > 0) {  // regular file: one read into a buffer of the right size
        buf.resize((std::size_t)sz);
        const std::size_t got = std::fread(&buf[0], 1, (std::size_t)sz, f);
        buf.resize(got);
      }
      char chunk[1 << 16];
      std::size_t got;
      while ((got = std::fread(chunk, 1, sizeof chunk, f)) > 0) buf.append(chunk, got);  // pipes, growing files
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
    const char *data = buf.c_str() + pos;
    const std::size_t data_len = buf.size() - pos;
    if (parse_entries_parallel(data, data_len, h.is_pattern, h.is_symmetric, nonz)) return;
    std::vector<long> rawI, rawJ;
    std::vector<double> rawV;
    parse_entries_sequential(data, h.is_pattern, nonz, rawI, rawJ, rawV);
    // range check, 1-based -> 0-based and symmetric expansion, in file order
    for (std::size_t k = 0; k < rawI.size(); ++k) {
      const long I1 = rawI[k], J1 = rawJ[k];
      const int32_t I = (int32_t)I1 - 1, J = (int32_t)J1 - 1;
      if (I < 0 || J < 0 || I >= cols_or_rows() || J >= rows) entry_out_of_range(I1, J1);
      const EType tv = static_cast<EType>(rawV[k]);
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

  void entry_out_of_range(long I1, long J1) const {
    std::cerr << "Matrix entry (" << I1 << ", " << J1 << ") lies outside the declared " << rows << " x " << cols
              << " matrix" << ENDL;
    std::exit(-1);
  }

  // Fast token parsers for the threaded path.  They accept exactly the common spellings and produce exactly what
  // strtol / strtod produce for them; anything else (hex, inf/nan, > 18 digits, exponents beyond +-22, ...) is left
  // to the C library, so the tuples are bit-identical to the sequential parser's by construction:
  //   integer: [+-]digits, at most 18 digits;
  //   decimal: [+-]digits[.digits][(e|E)[+-]digits] with at most 19 significant digits forming m <= 2^53 and a
  //            decimal exponent |e10| <= 22.  Then m and 10^|e10| are exact doubles and the single multiplication or
  //            division rounds once, i.e. correctly -- Clinger's fast path, the same value strtod returns.
  static inline long fast_long(const char *q, const char *eol, char **end) {
    const char *p = q;
    bool neg = false;
    if (p < eol && (*p == '+' || *p == '-')) neg = *p++ == '-';
    const char *d0 = p;
    unsigned long long v = 0;
    while (p < eol && *p >= '0' && *p <= '9') v = v * 10 + (unsigned)(*p++ - '0');
    if (p == d0 || p - d0 > 18) return std::strtol(q, end, 10);
    *end = const_cast<char *>(p);
    return neg ? -(long)v : (long)v;
  }

  static inline double fast_double(const char *q, const char *eol, char **end) {
    static const double P10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                   1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    const char *p = q;
    bool neg = false;
    if (p < eol && (*p == '+' || *p == '-')) neg = *p++ == '-';
    unsigned long long m = 0;
    int digits = 0, sig = 0, e10 = 0;
    while (p < eol && *p >= '0' && *p <= '9') {
      if (sig || *p != '0') { m = m * 10 + (unsigned)(*p - '0'); ++sig; }
      ++p; ++digits;
    }
    if (p < eol && *p == '.') {
      ++p;
      while (p < eol && *p >= '0' && *p <= '9') {
        if (sig || *p != '0') { m = m * 10 + (unsigned)(*p - '0'); ++sig; }
        ++p; ++digits; --e10;
      }
    }
    if (digits == 0 || sig > 19) return std::strtod(q, end);
    if (p < eol && (*p == 'e' || *p == 'E')) {
      const char *pe = p + 1;
      bool eneg = false;
      if (pe < eol && (*pe == '+' || *pe == '-')) eneg = *pe++ == '-';
      const char *e0 = pe;
      int ev = 0;
      while (pe < eol && *pe >= '0' && *pe <= '9' && pe - e0 < 6) ev = ev * 10 + (*pe++ - '0');
      if (pe == e0 || (pe < eol && *pe >= '0' && *pe <= '9')) return std::strtod(q, end);  // "1e" / huge exponent
      e10 += eneg ? -ev : ev;
      p = pe;
    }
    // a following character that strtod could still consume (hex "0x", "inf", "nan", a second '.') -> C library
    if (p < eol && ((*p >= '0' && *p <= '9') || *p == '.' || *p == 'x' || *p == 'X' || *p == 'n' || *p == 'N' || *p == 'i' ||
                    *p == 'I' || *p == 'p' || *p == 'P'))
      return std::strtod(q, end);
    if (m > (1ull << 53) || e10 > 22 || e10 < -22) return std::strtod(q, end);
    double v = (double)m;
    if (e10 > 0) v *= P10[e10];
    else if (e10 < 0) v /= P10[-e10];
    *end = const_cast<char *>(p);
    return neg ? -v : v;
  }

  // The reference reads entries with fscanf("%d %d %lg\n") (src/sparse_matrix.cpp:50-57): whitespace-separated
  // tokens, line structure irrelevant, stop at the first token that does not parse.
  static void parse_entries_sequential(const char *p, bool pattern, int nonz, std::vector<long> &oI, std::vector<long> &oJ,
                                       std::vector<double> &oV) {
    oI.clear(); oJ.clear(); oV.clear();
    for (int i = 0; i < nonz; ++i) {
      char *end = nullptr;
      const long I1 = std::strtol(p, &end, 10);
      if (end == p) break;
      p = end;
      const long J1 = std::strtol(p, &end, 10);
      if (end == p) break;
      p = end;
      double v = 1.0;
      if (!pattern) {
        v = std::strtod(p, &end);
        if (end == p) break;
        p = end;
      }
      oI.push_back(I1);
      oJ.push_back(J1);
      oV.push_back(v);
    }
  }

  // Same tokens, parsed by several threads: the data is cut at line ends, every thread parses its lines with the
  // same strtol / strtod calls (and applies the same range check and element conversion), and the pieces are
  // concatenated in file order -- by the same threads when no symmetric expansion is needed.  Only files with
  // exactly one entry per non-blank line qualify (every SuiteSparse file does); anything else returns false with
  // nothing stored and the sequential token parser runs, so the result is identical by construction.
  // SHB_MM_THREADS caps the thread count (1 = never parallel), SHB_MM_PARALLEL_MIN_BYTES (default 4 MiB) is the
  // size below which threads do not pay.
  bool parse_entries_parallel(const char *data, std::size_t len, bool pattern, bool symmetric, int nonz_decl) {
    std::size_t min_bytes = 4u << 20;
    if (const char *e = std::getenv("SHB_MM_PARALLEL_MIN_BYTES")) min_bytes = (std::size_t)std::strtoull(e, nullptr, 10);
    unsigned n_thr = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
    if (const char *e = std::getenv("SHB_MM_THREADS")) n_thr = (unsigned)std::max(1, std::atoi(e));
    if (n_thr < 2 || len < min_bytes || nonz_decl <= 0) return false;
    std::vector<std::size_t> cut(n_thr + 1, len);
    cut[0] = 0;
    for (unsigned t = 1; t < n_thr; ++t) {
      std::size_t c = std::max(cut[t - 1], len / n_thr * t);
      while (c < len && data[c] != '\n') ++c;
      cut[t] = std::min(len, c + (c < len ? 1 : 0));
    }
    struct Piece {
      std::vector<int32_t> I, J;
      std::vector<EType> V;
      bool regular = true;
      std::size_t bad_at = (std::size_t)-1;  // first entry of the piece outside the declared shape
      long badI = 0, badJ = 0;
    };
    std::vector<Piece> pieces(n_thr);
    const int lim_i = cols_or_rows(), lim_j = rows;
    auto work = [&](unsigned t) {
      Piece &pc = pieces[t];
      const char *p = data + cut[t], *stop = data + cut[t + 1];
      const std::size_t guess = (cut[t + 1] - cut[t]) / 10 + 16;
      pc.I.reserve(guess); pc.J.reserve(guess); pc.V.reserve(guess);
      while (p < stop) {
        const char *eol = (const char *)std::memchr(p, '\n', (std::size_t)(stop - p));
        if (!eol) eol = stop;
        const char *q = p;
        while (q < eol && (*q == ' ' || *q == '\t' || *q == '\r')) ++q;
        if (q < eol) {  // a non-blank line: exactly the entry's tokens, nothing else
          char *end = nullptr;
          const long I1 = fast_long(q, eol, &end);
          if (end == q || end > eol) { pc.regular = false; return; }
          q = end;
          while (q < eol && (*q == ' ' || *q == '\t')) ++q;  // strtol / strtod skip leading blanks themselves
          const long J1 = fast_long(q, eol, &end);
          if (end == q || end > eol) { pc.regular = false; return; }
          q = end;
          double v = 1.0;
          if (!pattern) {
            while (q < eol && (*q == ' ' || *q == '\t')) ++q;
            v = fast_double(q, eol, &end);
            if (end == q || end > eol) { pc.regular = false; return; }
            q = end;
          }
          while (q < eol && (*q == ' ' || *q == '\t' || *q == '\r')) ++q;
          if (q != eol) { pc.regular = false; return; }
          const int32_t I = (int32_t)I1 - 1, J = (int32_t)J1 - 1;
          if ((I < 0 || J < 0 || I >= lim_i || J >= lim_j) && pc.bad_at == (std::size_t)-1) {
            pc.bad_at = pc.I.size();
            pc.badI = I1;
            pc.badJ = J1;
          }
          pc.I.push_back(I); pc.J.push_back(J); pc.V.push_back(static_cast<EType>(v));
        }
        p = eol + 1;
      }
    };
    {
      std::vector<std::thread> threads;
      for (unsigned t = 1; t < n_thr; ++t) threads.emplace_back(work, t);
      work(0);
      for (auto &th : threads) th.join();
    }
    std::size_t total = 0;
    for (auto &pc : pieces) {
      if (!pc.regular) return false;
      total += pc.I.size();
    }
    const std::size_t keep = std::min<std::size_t>(total, (std::size_t)nonz_decl);
    std::vector<std::size_t> first(n_thr + 1, 0);  // position of each piece in the kept sequence
    for (unsigned t = 0; t < n_thr; ++t) first[t + 1] = std::min(keep, first[t] + pieces[t].I.size());
    for (unsigned t = 0; t < n_thr; ++t)
      if (pieces[t].bad_at != (std::size_t)-1 && first[t] + pieces[t].bad_at < first[t + 1])
        entry_out_of_range(pieces[t].badI, pieces[t].badJ);  // the first offender in file order
    if (!symmetric) {
      eI.resize(keep); eJ.resize(keep); eV.resize(keep);
      auto copy = [&](unsigned t) {
        const std::size_t n = first[t + 1] - first[t];
        if (!n) return;
        std::memcpy(&eI[first[t]], pieces[t].I.data(), n * sizeof(int32_t));
        std::memcpy(&eJ[first[t]], pieces[t].J.data(), n * sizeof(int32_t));
        std::copy(pieces[t].V.begin(), pieces[t].V.begin() + (std::ptrdiff_t)n, eV.begin() + (std::ptrdiff_t)first[t]);
      };
      std::vector<std::thread> threads;
      for (unsigned t = 1; t < n_thr; ++t) threads.emplace_back(copy, t);
      copy(0);
      for (auto &th : threads) th.join();
    } else {
      for (unsigned t = 0; t < n_thr; ++t)
        for (std::size_t k = 0; k < first[t + 1] - first[t]; ++k) {
          const int32_t I = pieces[t].I[k], J = pieces[t].J[k];
          eI.push_back(I); eJ.push_back(J); eV.push_back(pieces[t].V[k]);
          if (I != J) { eI.push_back(J); eJ.push_back(I); eV.push_back(pieces[t].V[k]); }
        }
    }
    return true;
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

  bool ellpack_calculated = false, width_calculated = false;
  unsigned int max_width = 0;
  ellpack_matrix<EType> ellpackMatrix;
};
