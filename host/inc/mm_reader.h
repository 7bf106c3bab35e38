// mm_reader.h -- Matrix Market banner + size line, with the acceptance rules of the NIST reader the
// reference links (src/mmio.cpp:92-199) as used by SparseMatrix::load_from_file
// (src/sparse_matrix.cpp:24-41):
//   line 1: "%%MatrixMarket <object> <format> <field> <symmetry>", the last four case-insensitive
//   object must be "matrix"; format coordinate|array; field real|complex|pattern|integer;
//   symmetry general|symmetric|hermitian|skew-symmetric
//   then any number of lines starting with '%', then "M N nz"
// This is an independent implementation working on an in-memory copy of the file; it also hands
// back the offset of the first entry so the entry parser can run over the same buffer.
#pragma once

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstring>
#include <string>

struct MMHeader {
  bool is_matrix = false, is_coordinate = false, is_array = false;
  bool is_real = false, is_complex = false, is_pattern = false, is_integer = false;
  bool is_general = false, is_symmetric = false, is_hermitian = false, is_skew = false;
  int rows = 0, cols = 0, nonz = 0;
  std::size_t data_offset = 0;  // first byte after the size line
  std::string typecode() const {  // the 4 letters NIST prints, e.g. "MCRS"
    std::string t = "    ";
    t[0] = is_matrix ? 'M' : ' ';
    t[1] = is_coordinate ? 'C' : (is_array ? 'A' : ' ');
    t[2] = is_real ? 'R' : is_complex ? 'C' : is_pattern ? 'P' : is_integer ? 'I' : ' ';
    t[3] = is_general ? 'G' : is_symmetric ? 'S' : is_hermitian ? 'H' : is_skew ? 'K' : ' ';
    return t;
  }
};

enum MMStatus { MM_OK = 0, MM_PREMATURE_EOF = 12, MM_NO_HEADER = 14, MM_UNSUPPORTED = 15 };

namespace mm_detail {
inline std::string lower(std::string s) {
  for (auto &c : s) c = (char)std::tolower((unsigned char)c);
  return s;
}
inline std::size_t line_end(const char *buf, std::size_t n, std::size_t pos) {
  while (pos < n && buf[pos] != '\n') ++pos;
  return pos;
}
}  // namespace mm_detail

inline int mm_parse_banner(const char *buf, std::size_t n, MMHeader &h, std::size_t &pos) {
  const std::size_t e = mm_detail::line_end(buf, n, 0);
  if (e == 0 && n == 0) return MM_PREMATURE_EOF;
  const std::string line(buf, e);
  pos = e < n ? e + 1 : e;
  char banner[65], mtx[65], crd[65], dtype[65], scheme[65];
  if (std::sscanf(line.c_str(), "%64s %64s %64s %64s %64s", banner, mtx, crd, dtype, scheme) != 5)
    return MM_PREMATURE_EOF;
  if (std::strncmp(banner, "%%MatrixMarket", 14) != 0) return MM_NO_HEADER;
  const std::string m = mm_detail::lower(mtx), c = mm_detail::lower(crd), d = mm_detail::lower(dtype),
                    s = mm_detail::lower(scheme);
  if (m != "matrix") return MM_UNSUPPORTED;
  h.is_matrix = true;
  if (c == "coordinate") h.is_coordinate = true;
  else if (c == "array") h.is_array = true;
  else return MM_UNSUPPORTED;
  if (d == "real") h.is_real = true;
  else if (d == "complex") h.is_complex = true;
  else if (d == "pattern") h.is_pattern = true;
  else if (d == "integer") h.is_integer = true;
  else return MM_UNSUPPORTED;
  if (s == "general") h.is_general = true;
  else if (s == "symmetric") h.is_symmetric = true;
  else if (s == "hermitian") h.is_hermitian = true;
  else if (s == "skew-symmetric") h.is_skew = true;
  else return MM_UNSUPPORTED;
  return MM_OK;
}

inline int mm_parse_crd_size(const char *buf, std::size_t n, MMHeader &h, std::size_t &pos) {
  // skip comment lines
  std::size_t e;
  std::string line;
  do {
    if (pos >= n) return MM_PREMATURE_EOF;
    e = mm_detail::line_end(buf, n, pos);
    line.assign(buf + pos, e - pos);
    pos = e < n ? e + 1 : e;
  } while (!line.empty() && line[0] == '%');
  if (std::sscanf(line.c_str(), "%d %d %d", &h.rows, &h.cols, &h.nonz) == 3) {
    h.data_offset = pos;
    return MM_OK;
  }
  // the NIST reader keeps scanning for three integers after a blank / malformed line
  for (;;) {
    if (pos >= n) return MM_PREMATURE_EOF;
    int used = 0;
    const std::string rest(buf + pos, std::min<std::size_t>(n - pos, 256));
    if (std::sscanf(rest.c_str(), "%d %d %d%n", &h.rows, &h.cols, &h.nonz, &used) == 3) {
      pos += (std::size_t)used;
      h.data_offset = pos;
      return MM_OK;
    }
    e = mm_detail::line_end(buf, n, pos);
    pos = e < n ? e + 1 : e;
  }
}
