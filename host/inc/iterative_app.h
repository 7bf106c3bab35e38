// iterative_app.h -- the part the four iterative harnesses share.  In the reference
// app/sssp.cpp:44-177, app/bfs.cpp:43-174, app/pr.cpp:43-176 and app/scc.cpp:43-176 are copies of one
// class that differ only in the element type and the termination predicate; here that class is a
// template and each app instantiates it.
//   benchmark():  per trial: run to the fixed point, then append a MEDIAN_RESULT row (upper median of the
//                 per-iteration times) and a MULTI_ITERATION_SUM row -- whose sum, as in the reference,
//                 includes the median row that was pushed just before it (app/sssp.cpp:65-84) -- and
//                 resetInputs() for the next trial
//   termination:  ABSDIFF: all i, fabs(in[i] - out[i]) < delta   (app/sssp.cpp:157-176, app/pr.cpp:156-176)
//                 EQUAL:   all i, in[i] == out[i]                (app/bfs.cpp:154-174, app/scc.cpp:156-176)
#pragma once

#include <algorithm>
#include <cmath>
#include <numeric>
#include <vector>

#include "app_common.h"
#include "harness.h"

template <typename SemiRingType, int TERM_MODE>
class IterativeAppHarness : public IterativeHarness<std::vector<SqlStat>, SemiRingType> {
  typedef IterativeHarness<std::vector<SqlStat>, SemiRingType> Base;

public:
  IterativeAppHarness(std::string &kernel_source, unsigned int platform, unsigned int device,
                      ArgContainer<SemiRingType> args, unsigned int trials, std::chrono::milliseconds timeout,
                      double delta)
      : Base(kernel_source, platform, device, std::move(args), trials, timeout, delta) {
    this->allocateBuffers();
  }

  std::vector<std::vector<SqlStat>> benchmark(Run run, std::vector<SemiRingType> &gold) override {
    start_timer(benchmark, IterativeAppHarness);
    std::vector<std::vector<SqlStat>> runtimes;
    for (unsigned int t = 0; t < this->_trials; t++) {
      start_timer(benchmark_iteration, IterativeAppHarness);
      std::vector<SqlStat> run_runtimes = executeRun(run, t, gold);
      if (run_runtimes.size() > 0) {
        std::sort(run_runtimes.begin(), run_runtimes.end(), SqlStat::compare);
        const std::chrono::nanoseconds median_time = run_runtimes[run_runtimes.size() / 2].getTime();
        run_runtimes.push_back(SqlStat(median_time, NOT_CHECKED, run.global1, run.local1, MEDIAN_RESULT, t));
        const std::chrono::nanoseconds total_time =
            std::accumulate(run_runtimes.begin(), run_runtimes.end(), std::chrono::nanoseconds(0),
                            [](std::chrono::nanoseconds acc, const SqlStat &stat) { return acc + stat.getTime(); });
        run_runtimes.push_back(SqlStat(total_time, NOT_CHECKED, run.global1, run.local1, MULTI_ITERATION_SUM));
        runtimes.push_back(run_runtimes);
      }
      // keep the converged vector of the LAST trial on the host (for --dump) before the inputs are restored
      _final = this->outputVector(this->_mem_manager._output_host_buffer.size() / sizeof(SemiRingType));
      this->resetInputs();
    }
    return runtimes;
  }

  const std::vector<SemiRingType> &finalVector() const { return _final; }

protected:
  std::vector<SqlStat> executeRun(Run run, unsigned int trial, std::vector<SemiRingType> &) override {
    start_timer(executeRun, IterativeAppHarness);
    return this->iterate(run, trial, TERM_MODE);
  }

  bool should_terminate_iteration(std::vector<char> &input, std::vector<char> &output) override {
    start_timer(should_terminate_iteration, IterativeAppHarness);
    const SemiRingType *in = reinterpret_cast<const SemiRingType *>(input.data());
    const SemiRingType *out = reinterpret_cast<const SemiRingType *>(output.data());
    const std::size_t n = std::min(input.size(), output.size()) / sizeof(SemiRingType);
    bool equal = true;
    for (std::size_t i = 0; equal && i < n; i++)
      equal = TERM_MODE == SHB_TERM_EQUAL ? (in[i] == out[i]) : (std::fabs(in[i] - out[i]) < this->_delta);
    return equal;
  }

private:
  std::vector<SemiRingType> _final;
};

// the tail every iterative main() shares: per runfile row, benchmark and print one INSERT per trial
template <typename H, typename T>
void run_iterative_app(H &harness, HarnessSetup<T> &s, const std::string &kernel_name) {
  std::vector<T> gold;  // the reference has no CPU answer for these apps: rows say "notchecked"
  const std::string device_name = harness.getDeviceName();
  for (auto run : s.runs) {
    start_timer(run_iteration, main);
    std::cout << "Benchmarking run: " << run << ENDL;
    std::vector<std::vector<SqlStat>> runtimes = harness.benchmark(run, gold);
    for (auto &statList : runtimes)
      std::cout << SqlStat::makeSqlCommand(statList, kernel_name, s.hostname, device_name, s.matrix_name, s.experiment)
                << "\n";
    std::cout << "Iterations: " << harness.lastIterationCount() << (harness.lastRunConverged() ? "" : " (cap reached)")
              << ENDL;
  }
  std::vector<T> out = harness.finalVector();
  if (out.size() > (std::size_t)s.matrix->height()) out.resize((std::size_t)s.matrix->height());
  dump_vector(s.dump, out);
}

// ---- canonical modes (--mode canonical) ------------------------------------------------------------
// The reference's bfs and pr apps do not compute BFS levels / PageRank (TODO.md:3-9) and
// its sssp needs one dense sweep per relaxation round.  The canonical drivers run libshb's device-resident
// graph algorithms on the same operator: shb_bfs (integer levels), shb_sssp (exact min-plus fixed point),
// shb_pagerank (x <- A x + (1-d)/N with y == 1 never re-bound).  One RAW_RESULT row per trial carries the
// whole traversal's device time; rows are "notchecked" like every iterative row of the reference.
struct CanonicalResult {
  uint64_t ns = 0;
  uint32_t iterations = 0;
  int64_t edges = 0, reached = 0;
  bool converged = false;
};

template <typename T, typename RunOnce>
void run_canonical_app(HarnessSetup<T> &s, const std::string &kernel_name, const std::string &device_name,
                       const char *what, RunOnce run_once) {
  for (auto run : s.runs) {
    start_timer(run_iteration, main);
    std::cout << "Benchmarking run: " << run << ENDL;
    std::vector<SqlStat> rows;
    CanonicalResult last;
    for (unsigned int t = 0; t < s.trials; ++t) {
      last = run_once(run);
      rows.push_back(SqlStat(std::chrono::nanoseconds(last.ns), NOT_CHECKED, run.global1, run.local1, RAW_RESULT, t,
                             last.iterations));
    }
    std::vector<SqlStat> sorted(rows);
    std::sort(sorted.begin(), sorted.end(), SqlStat::compare);
    const std::chrono::nanoseconds median = sorted[sorted.size() / 2].getTime();
    rows.push_back(SqlStat(median, NOT_CHECKED, run.global1, run.local1, MEDIAN_RESULT));
    std::cout << SqlStat::makeSqlCommand(rows, kernel_name, s.hostname, device_name, s.matrix_name, s.experiment) << "\n";
    const double gteps = median.count() > 0 ? (double)last.edges / (double)median.count() : 0.0;
    std::cout << what << ": iterations " << last.iterations << ", reached " << last.reached << ", edges traversed "
              << last.edges << ", median " << ((double)median.count()) / 1e6 << " ms, " << gteps << " GTEPS"
              << (last.converged ? "" : " (cap reached)") << ENDL;
  }
}
