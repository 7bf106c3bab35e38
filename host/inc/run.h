// run.h -- one launch configuration from the runfile (reference: inc/run.h:9-36, src/run.cpp:4-23).
// The six fields are OpenCL NDRange sizes in the reference.  The CUDA build reads them as hints:
// local1 -> threads per CTA (64/128/256, anything else = library default), global1/local1 -> CTA cap
// of the thread-per-row schedule.  They are echoed unchanged into every result row.
#pragma once

#include <cassert>
#include <cstddef>
#include <ostream>

#include "csv_utils.h"

class Run {
public:
  Run(std::size_t g1, std::size_t g2, std::size_t g3, std::size_t l1, std::size_t l2, std::size_t l3)
      : global1(g1), global2(g2), global3(g3), local1(l1), local2(l2), local3(l3) {}

  explicit Run(const CSV::csv_line &line) {
    assert(line.size() == 6 && "Bad CSV format");
    global1 = CSV::read_size_t(line[0]);
    global2 = CSV::read_size_t(line[1]);
    global3 = CSV::read_size_t(line[2]);
    local1 = CSV::read_size_t(line[3]);
    local2 = CSV::read_size_t(line[4]);
    local3 = CSV::read_size_t(line[5]);
  }

  std::size_t num_work_items() const { return local1 * local2 * local3; }

  std::size_t global1 = 0, global2 = 0, global3 = 0;
  std::size_t local1 = 0, local2 = 0, local3 = 0;
};

inline std::ostream &operator<<(std::ostream &stream, const Run &r) {
  stream << "{" << r.global1 << " " << r.global2 << " " << r.global3 << " / " << r.local1 << " " << r.local2 << " "
         << r.local3 << "}";
  return stream;
}
