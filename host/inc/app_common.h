// app_common.h -- what every harness main() starts with: the reference's COMMON_MAIN_PREAMBLE
// (inc/common.h:5-56) as a function.  Same flags, defaults, messages and exit codes:
//   -p/--platform (0) -d/--device (0) -i/--trials (10) -m/--matrix -f/--matrix_name -k/--kernel
//   -r/--runfile -n/--hostname -e/--experiment (all six required) -c/--delta (1e-4) -t/--timeout (100 ms)
//   non-square matrix -> message on stdout and stderr, exit(2)
// Additions of the CUDA build (long options only, all optional):
//   --variant json|auto|scalar|scalar-reversed|vector|stream   override the schedule the kernel JSON selects
//   --mode faithful|canonical   faithful = the reference's recurrences incl. its quirks (default);
//                               canonical = BFS levels / exact SSSP / true PageRank
//   --max_iterations N          cap for the iterative apps (the reference has none)
//   --host_loop                 step the iterative apps from the host like the reference does
//   --synthetic SPEC            build the operator on the device instead of reading -m:
//                               uniform:ROWS:K | uniform-int:ROWS:K | rmat:SCALE:EF | rmat-pr:SCALE:EF | grid:W:H
//   --dump FILE                 write the final output vector (raw little-endian elements) to FILE
//   --gpu_encode                COO tuples -> CSR on the device (shb_csr_from_coo) instead of the host counting sort
//   --gpus N                    canonical modes of pr / bfs / sssp: row-partition the operator over the first N
//                               devices and run the loop through shb_group_* (host/inc/group_app.h)
//   --symmetric                 with --gpus: the operator equals its transpose (no out-edge copy is built)
#pragma once

#include <algorithm>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

#include "common.h"
#include "csds_timer.h"
#include "csv_utils.h"
#include "kernel_config.h"
#include "options.h"
#include "run.h"
#include "sparse_matrix.h"

template <typename T> struct HarnessSetup {
  unsigned int platform = 0, device = 0, trials = 10, timeout_ms = 100, max_iterations = 100000, gpus = 1;
  bool symmetric = false, gpu_encode = false;
  double delta = 0.0001;
  std::string matrix_filename, matrix_name, kernel_filename, runs_filename, hostname, experiment;
  std::string variant = "json", mode = "faithful", synthetic, dump;
  bool host_loop = false;
  std::unique_ptr<SparseMatrix<T>> matrix;
  std::unique_ptr<KernelConfig<T>> kernel;
  std::vector<Run> runs;
  bool canonical() const { return mode == "canonical"; }
};

inline long long synthetic_rows(const std::string &spec) {
  std::string s = spec;
  for (auto &c : s)
    if (c == ':') c = ',';
  const CSV::csv_line f = CSV::tokenise_line(s);
  auto num = [&f](std::size_t i) { return i < f.size() ? std::atoll(f[i].c_str()) : 0ll; };
  if (f.empty()) return 0;
  if (f[0] == "uniform" || f[0] == "uniform-int") return num(1);
  if (f[0] == "rmat" || f[0] == "rmat-pr") return 1ll << num(1);
  if (f[0] == "grid") return num(1) * num(2);
  return 0;
}

template <typename T> HarnessSetup<T> harness_preamble(int argc, char *argv[]) {
  OptParser op("Harness for SPMV sparse matrix dense vector multiplication benchmarks");
  auto opt_platform = op.addOption<unsigned>({'p', "platform", "OpenCL platform index (default 0; ignored by the CUDA build).", 0});
  auto opt_device = op.addOption<unsigned>({'d', "device", "CUDA device index (default 0).", 0});
  auto opt_trials = op.addOption<unsigned>({'i', "trials", "Execute each kernel 'trials' times (default 10).", 10});
  auto opt_matrix_file = op.addOption<std::string>({'m', "matrix", "Input matrix"});
  auto opt_matrix_name = op.addOption<std::string>({'f', "matrix_name", "Input matrix name"});
  auto opt_kernel_file = op.addOption<std::string>({'k', "kernel", "Input kernel"});
  auto opt_run_file = op.addOption<std::string>({'r', "runfile", "Run configuration file"});
  auto opt_host_name = op.addOption<std::string>({'n', "hostname", "Host the harness is running on"});
  auto opt_experiment_id = op.addOption<std::string>({'e', "experiment", "An experiment ID for data reporting"});
  auto opt_float_delta = op.addOption<double>({'c', "delta", "Delta for floating point comparisons", 0.0001});
  auto opt_timeout = op.addOption<unsigned int>({'t', "timeout", "Timeout to avoid multiple executions (default 100ms).", 100});
  auto opt_variant = op.addOption<std::string>({0, "variant", "CUDA schedule: json|auto|scalar|scalar-reversed|vector|stream (default json).", "json"});
  auto opt_mode = op.addOption<std::string>({0, "mode", "faithful (reference recurrences, default) or canonical.", "faithful"});
  auto opt_max_iter = op.addOption<unsigned int>({0, "max_iterations", "Iteration cap of the iterative apps (default 100000).", 100000});
  auto opt_host_loop = op.addOption<bool>({0, "host_loop", "Step iterative apps from the host like the reference.", false});
  auto opt_synthetic = op.addOption<std::string>({0, "synthetic", "Device-generated operator instead of -m (uniform:R:K|rmat:S:EF|rmat-pr:S:EF|grid:W:H).", ""});
  auto opt_dump = op.addOption<std::string>({0, "dump", "Write the final output vector to this file (raw elements).", ""});
  auto opt_gpus = op.addOption<unsigned int>({0, "gpus", "Canonical pr/bfs/sssp: row-partition over the first N devices (default 1).", 1});
  auto opt_gpu_encode = op.addOption<bool>({0, "gpu_encode", "Encode the operator (COO -> CSR) on the device.", false});
  auto opt_symmetric = op.addOption<bool>({0, "symmetric", "With --gpus: the operator equals its transpose.", false});
  op.parse(argc, argv);

  HarnessSetup<T> s;
  s.synthetic = opt_synthetic->get();
  s.matrix_filename = s.synthetic.empty() ? opt_matrix_file->require() : opt_matrix_file->get();
  s.matrix_name = opt_matrix_name->require();
  s.kernel_filename = opt_kernel_file->require();
  s.runs_filename = opt_run_file->require();
  s.hostname = opt_host_name->require();
  s.experiment = opt_experiment_id->require();
  s.platform = opt_platform->get();
  s.device = opt_device->get();
  s.trials = opt_trials->get();
  s.timeout_ms = opt_timeout->get();
  s.delta = opt_float_delta->get();
  s.variant = opt_variant->get();
  s.mode = opt_mode->get();
  s.max_iterations = opt_max_iter->get();
  s.host_loop = opt_host_loop->get();
  s.dump = opt_dump->get();
  s.gpus = std::max(1u, opt_gpus->get());
  s.symmetric = opt_symmetric->get();
  s.gpu_encode = opt_gpu_encode->get();
  if (s.mode != "faithful" && s.mode != "canonical") {
    std::cout << "Error: invalid argument for option mode" << ENDL;
    op.usage(argv[0]);
    std::exit(0);
  }
  std::cerr << "matrix_filename " << s.matrix_filename << ENDL;
  std::cerr << "kernel_filename " << s.kernel_filename << ENDL;
  if (s.synthetic.empty()) {
    s.matrix.reset(new SparseMatrix<T>(s.matrix_filename));
  } else {
    const long long n = synthetic_rows(s.synthetic);
    if (n <= 0 || n >= 2147483647ll) {
      std::cout << "Error: invalid argument for option synthetic" << ENDL;
      std::exit(0);
    }
    s.matrix.reset(new SparseMatrix<T>((int)n, (int)n));
  }
  s.kernel.reset(new KernelConfig<T>(s.kernel_filename));
  for (auto &line : CSV::load_csv(s.runs_filename)) s.runs.push_back(Run(line));
  if (s.matrix->height() != s.matrix->width()) {
    std::cout << "Matrix is not square. Failing computation." << ENDL;
    std::cerr << "Matrix is not square. Failing computation." << ENDL;
    std::exit(2);
  } else {
    std::cout << " Matrix is square - width = " << s.matrix->width() << " and height = " << s.matrix->height() << "\n";
  }
  return s;
}

template <typename T> void dump_vector(const std::string &path, const std::vector<T> &v) {
  if (path.empty()) return;
  std::ofstream out(path, std::ios::binary);
  out.write(reinterpret_cast<const char *>(v.data()), (std::streamsize)(v.size() * sizeof(T)));
}

template <typename T, typename Args> void apply_variant_override(const HarnessSetup<T> &s, Args &args) {
  if (!KernelConfig<T>::applyVariantName(args.launch, s.variant)) {
    std::cout << "Error: invalid argument for option variant" << ENDL;
    std::exit(0);
  }
}

// The semiring is whatever the kernel JSON's source implements (as in the reference, where it lives
// only in the Lift-generated user functions).  Each app states the one it was written for: a different
// one of the same element type is honoured with a warning, a different element type is an error.
template <typename Args> void expect_semiring(Args &args, int expected, const char *app) {
  auto is_float = [](int sr) { return sr == SHB_PLUS_TIMES_F32 || sr == SHB_MIN_PLUS_F32; };
  if (args.semiring == expected) return;
  if (is_float(args.semiring) != is_float(expected)) {
    std::cerr << app << ": the kernel's semiring (" << args.semiring << ") has a different element type than this "
              << "harness computes in" << ENDL;
    std::exit(1);
  }
  std::cerr << app << ": warning: kernel JSON implements semiring " << args.semiring << ", this harness expects "
            << expected << "; running the kernel's" << ENDL;
}
