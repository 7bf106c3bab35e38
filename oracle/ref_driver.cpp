// ref_driver.cpp -- thin command-line driver around the UNMODIFIED reference CPU path.
//
// TEST INFRASTRUCTURE (oracle/): compiled by oracle/Makefile together with the reference's own
// sources, taken where they lie under /root/reference (never copied into this repo):
//   src/sparse_matrix.cpp  src/mmio.cpp  src/csds_timer.cpp  src/util/Logger.cpp  + inc/spmv_gold.h
// Output goes to oracle/_ref/ref_driver (git-ignored).  This file contains no reference code; it
// only calls the reference's public classes:
//   SparseMatrix<T>(filename)            src/sparse_matrix.cpp:5-70
//   SparseMatrix<T>::ellpack_encode()    src/sparse_matrix.cpp:401-407 (ragged rows after Q1-Q3)
//   SparseMatrix<T>::pagerank_normalise  src/sparse_matrix.cpp:409-431
//   SparseMatrix<T>::scc_normalise       src/sparse_matrix.cpp:433-456
//   Gold<T>::spmv                        inc/spmv_gold.h:9-28
//
// usage:
//   ref_driver rows <mtx> <float|int> <none|pr|scc>   dump the loader's ragged rows
//   ref_driver gold <mtx> <xmode>                      Gold<float>::spmv, alpha=1 beta=0 y=0
//                                                      xmode 0: x == 1 (app/spmv.cpp:117)
//                                                      xmode 1: x[i] = ((7 i + 3) mod 11) - 5
//   ref_driver time <mtx> <reps> [warmup]              wall-clock ns of Gold<float>::spmv per call
#include <chrono>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <limits>
#include <string>

#include "sparse_matrix.h"
#include "vector_generator.h"
#include "spmv_gold.h"

template <typename T> class PatternX : public XVectorGenerator<T> {
  int mode;
public:
  explicit PatternX(int m) : mode(m) {}
  virtual T get(int ix) { return mode == 0 ? static_cast<T>(1) : static_cast<T>(((7 * ix + 3) % 11) - 5); }
};

template <typename T> static void dump_rows(SparseMatrix<T> &m, const char *norm) {
  if (!strcmp(norm, "pr")) m.pagerank_normalise(0.85f, static_cast<T>(0));
  if (!strcmp(norm, "scc")) m.scc_normalise();
  auto &rows = m.ellpack_encode();
  std::printf("ROWS %d %d %d %zu\n", m.height(), m.width(), m.nonZeros(), rows.size());
  for (auto &row : rows) {
    std::printf("%zu", row.size());
    for (auto &e : row) std::printf(" %d %.17g", e.first, (double)e.second);
    std::printf("\n");
  }
}

int main(int argc, char **argv) {
  if (argc < 4) { std::fprintf(stderr, "usage: see header\n"); return 64; }
  const std::string cmd = argv[1], file = argv[2];
  if (cmd == "rows") {
    if (argc < 5) return 64;
    if (!strcmp(argv[3], "float")) { SparseMatrix<float> m(file); dump_rows(m, argv[4]); }
    else { SparseMatrix<int> m(file); dump_rows(m, argv[4]); }
    return 0;
  }
  if (cmd == "gold") {
    SparseMatrix<float> m(file);
    PatternX<float> x(std::atoi(argv[3]));
    ConstYVectorGenerator<float> y(0);
    auto g = Gold<float>::spmv(m, x, y, 1.0f, 0.0f, 0.0f);
    std::printf("GOLD %zu\n", g.size());
    for (float v : g) std::printf("%.17g\n", (double)v);
    return 0;
  }
  if (cmd == "time") {
    SparseMatrix<float> m(file);
    ConstXVectorGenerator<float> x(1.0f);
    ConstYVectorGenerator<float> y(0);
    const int reps = std::atoi(argv[3]);
    const int warmup = argc > 4 ? std::atoi(argv[4]) : 0;
    (void)m.ellpack_encode(); // build the ragged rows once, as the spmv app has by then
    for (int r = 0; r < warmup; ++r) (void)Gold<float>::spmv(m, x, y, 1.0f, 0.0f, 0.0f);
    double best = 1e300, sum = 0;
    float sink = 0;
    for (int r = 0; r < reps; ++r) {
      auto t0 = std::chrono::steady_clock::now();
      auto g = Gold<float>::spmv(m, x, y, 1.0f, 0.0f, 0.0f);
      auto t1 = std::chrono::steady_clock::now();
      double ns = (double)std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count();
      if (ns < best) best = ns;
      sum += ns;
      sink += g[0];
    }
    std::printf("TIME best_ns %.0f mean_ns %.0f reps %d sink %g\n", best, sum / reps, reps, (double)sink);
    return 0;
  }
  return 64;
}
