// spmv_harness -- single-shot plus-times SpMV, `trials` timed launches per runfile row, every result
// compared with the CPU gold loop (reference: app/spmv.cpp:43-169).
//   x = 1, y = 0, alpha = 1, beta = 0, zero = 0      (app/spmv.cpp:117-120)
// Output: "Benchmarking run: {..}", "runtimes: [ <row>... ]" and one INSERT statement per runfile row.
#include <algorithm>
#include <iostream>
#include <vector>

#include "app_common.h"
#include "harness.h"
#include "spmv_gold.h"

class HarnessSPMV : public Harness<SqlStat, float> {
public:
  HarnessSPMV(std::string &kernel_source, unsigned int platform, unsigned int device, ArgContainer<float> args,
              unsigned int trials, std::chrono::milliseconds timeout, double delta)
      : Harness(kernel_source, platform, device, std::move(args), trials, timeout, delta) {
    allocateBuffers();
  }

  std::vector<SqlStat> benchmark(Run run, std::vector<float> &gold) override {
    start_timer(benchmark, HarnessSPMV);
    std::vector<SqlStat> runtimes;
    for (unsigned int t = 0; t < _trials; t++) {
      start_timer(benchmark_iteration, HarnessSPMV);
      resetTempBuffers();
      SqlStat stat = executeRun(run, t, gold);
      runtimes.push_back(stat);
      if (stat.getTime() > _timeout) break;  // a trial slower than the timeout ends the trial loop
      lowerTimeout(stat.getTime());
      assertBuffersNotEqual(_mem_manager._output_host_buffer, _mem_manager._temp_out_buffer);
    }
    // upper median of the trials, appended as a MEDIAN_RESULT row
    std::sort(runtimes.begin(), runtimes.end(), SqlStat::compare);
    const std::chrono::nanoseconds median_time = runtimes[runtimes.size() / 2].getTime();
    runtimes.push_back(SqlStat(median_time, STATISTIC_VALUE, run.global1, run.local1, MEDIAN_RESULT));
    return runtimes;
  }

private:
  SqlStat executeRun(Run run, unsigned int, std::vector<float> &gold) override {
    const std::chrono::nanoseconds time = executeKernel(run);
    readFromGlobalArg(_mem_manager._output_host_buffer, _mem_manager._output);
    return SqlStat(time, check_result(gold), run.global1, run.local1, RAW_RESULT);
  }
};

int main(int argc, char *argv[]) {
  start_timer(main, global);
  HarnessSetup<float> s = harness_preamble<float>(argc, argv);
  SparseMatrix<float> &matrix = *s.matrix;
  KernelConfig<float> &kernel = *s.kernel;

  ConstXVectorGenerator<float> x(1.0f);
  ConstYVectorGenerator<float> y(0);
  const float alpha = 1.0f, beta = 0.0f;

  unsigned long max_alloc = deviceGetMaxAllocSize(s.platform, s.device);
  std::cout << "Got max alloc: " << max_alloc << "\n";

  ArgContainer<float> args;
  try {
    args = s.synthetic.empty() ? executorEncodeMatrix(max_alloc, kernel, matrix, 0.0f, x, y, alpha, beta, true, s.gpu_encode)
                               : executorSyntheticArgs(kernel, s.synthetic, matrix.height(), x, y, alpha, beta);
  } catch (unsigned long attempted_alloc_size) {
    LOG_ERROR("Attempted to allocate: ", attempted_alloc_size, " bytes, but this platform's max is ", max_alloc);
  }
  apply_variant_override(s, args);

  HarnessSPMV harness(kernel.getSource(), s.platform, s.device, std::move(args), s.trials,
                      std::chrono::milliseconds(s.timeout_ms), s.delta);

  // the gold values fill the "correct" column only; synthetic operators live on the device and are not checked here
  std::vector<float> gold;
  if (s.synthetic.empty()) gold = Gold<float>::spmv(matrix, x, y, alpha, beta, 0.0f);

  const std::string &kernel_name = kernel.getName();
  const std::string device_name = harness.getDeviceName();
  for (auto run : s.runs) {
    start_timer(run_iteration, main);
    std::cout << "Benchmarking run: " << run << ENDL;
    std::vector<SqlStat> runtimes = harness.benchmark(run, gold);
    std::cout << "runtimes: [";
    for (auto &time : runtimes)
      std::cout << "\n\t" << time.printStat(kernel_name, s.hostname, device_name, s.matrix_name, s.experiment);
    std::cout << "\n]" << ENDL;
    std::cout << SqlStat::makeSqlCommand(runtimes, kernel_name, s.hostname, device_name, s.matrix_name, s.experiment)
              << "\n";
  }
  dump_vector(s.dump, harness.outputVector((std::size_t)matrix.height()));
  return 0;
}
