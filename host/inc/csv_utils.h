// csv_utils.h -- runfile reader (reference: inc/csv_utils.h:14-49).
// A line is split on ','; a trailing comma therefore adds no token ("1280,1,1,1,1,1," -> 6 tokens).
// Reading stops at the first line without tokens (empty line or end of file).
#pragma once

#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "csds_timer.h"

namespace CSV {
using csv_token = std::string;
using csv_line = std::vector<csv_token>;

inline csv_line tokenise_line(const std::string &line) {
  csv_line tokens;
  std::stringstream ss(line);
  std::string cell;
  while (std::getline(ss, cell, ',')) tokens.push_back(cell);
  return tokens;
}

inline std::size_t read_size_t(const csv_token &str) {
  std::stringstream buffer(str);
  std::size_t value = 0;
  buffer >> value;
  return value;
}

inline std::vector<csv_line> load_csv(const std::string &filename) {
  start_timer(load_csv, csv_utils);
  std::vector<csv_line> lines;
  std::ifstream in(filename);
  std::string line;
  while (std::getline(in, line)) {
    csv_line tokens = tokenise_line(line);
    if (tokens.empty()) break;
    lines.push_back(tokens);
  }
  return lines;
}
}  // namespace CSV
