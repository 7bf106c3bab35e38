// harness.h -- Harness<TimingType, SemiRingType> / IterativeHarness: the seam between the algorithm
// drivers (app/*.cpp) and the device.  Same class shape, method names and argument ABI as the
// reference (inc/harness.h:11-501); everything that was OpenCL is now a call into libshb's C ABI
// (include/shb.h):
//   ctor: platform/device/context/JIT/queue (:13-82)       -> shb_ctx_create (kernels are precompiled sm_100a)
//   executeKernel: clEnqueueNDRangeKernel + event profiling -> shb_spmv, device time from a cudaEvent pair
//   createAndUploadGlobalArg / writeToGlobalArg / fillGlobalArg / readFromGlobalArg / createGlobalArg
//                                                           -> shb_buf_create / upload / fill_zero / download
//   setGlobalArg(2|3|6, &mem): the positional kernel-argument ABI (0 idx, 1 vals, 2 x, 3 y, 4 alpha, 5 beta,
//                              6 out; :199-249) is kept as a table of bound handles that executeKernel reads
//   IterativeHarness: + iterateOnDevice(), the device-resident replacement of the apps' host-driven
//                       do/while loops (app/sssp.cpp:112-153): vectors, ping-pong and the convergence flag stay
//                       in HBM, the host reads one 4-byte flag per batch of launches.  iterateOnHost() keeps the
//                       reference's step-by-step loop (one blocking read-back per iteration) for comparison.
// Error convention is the reference's: any failing device call logs and exit(1)s (device_utils.h).
#pragma once

#include <type_traits>

#include <chrono>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "device_memory_manager.h"
#include "device_utils.h"
#include "kernel_utils.h"
#include "run.h"
#include "sql_stat.h"

template <typename TimingType, typename SemiRingType> class Harness {
public:
  Harness(std::string &kernel_source, unsigned int platform, unsigned int device, ArgContainer<SemiRingType> args,
          unsigned int trials, std::chrono::milliseconds timeout, double delta)
      : _kernel_source(kernel_source), _device(device), _args(std::move(args)), _mem_manager(_args), _trials(trials),
        _timeout(timeout), _delta(delta) {
    (void)platform;  // OpenCL platform index: meaningless for CUDA, accepted for CLI compatibility
    int count = 0;
    checkShbError(shb_device_count(&count));
    if (count == 0) {
      LOG_ERROR("No CUDA devices found!");
      std::exit(1);
    }
    LOG_DEBUG_INFO("Found ", count, " devices");
    checkShbError(shb_ctx_create((int)device, &_ctx));
    LOG_INFO("Running on CUDA device: ", getDeviceName());
    for (auto &b : _bound) b = nullptr;
  }

  virtual ~Harness() {
    // the reference never releases anything (no clRelease* call exists); the CUDA library owns real
    // memory behind its handles, so give it back
    shb_csr_destroy(_mem_manager._matrix);
    shb_buf_destroy(_mem_manager._x_vect);
    shb_buf_destroy(_mem_manager._y_vect);
    shb_buf_destroy(_mem_manager._output);
    shb_ctx_destroy(_ctx);
  }
  Harness(const Harness &) = delete;
  Harness &operator=(const Harness &) = delete;

  virtual std::vector<TimingType> benchmark(Run run, std::vector<SemiRingType> &gold) = 0;

  // a trial faster than half the timeout lowers the timeout to twice that trial (:92-98)
  void lowerTimeout(std::chrono::nanoseconds measured_time) {
    auto ms_measured = std::chrono::duration_cast<std::chrono::milliseconds>(measured_time);
    if (ms_measured * 2 < _timeout) _timeout = ms_measured * 2;
  }

  std::string getDeviceName() {
    char name[256];
    checkShbError(shb_ctx_device_name(_ctx, name, sizeof name));
    return std::string(name);
  }

  // the last output as typed elements (first `count` of them); used by --dump and by the drivers' reports
  std::vector<SemiRingType> outputVector(std::size_t count) const {
    std::vector<SemiRingType> v(std::min(count, _mem_manager._output_host_buffer.size() / sizeof(SemiRingType)));
    if (!v.empty()) std::memcpy(v.data(), _mem_manager._output_host_buffer.data(), v.size() * sizeof(SemiRingType));
    return v;
  }

  shb_ctx *context() { return _ctx; }
  shb_csr *matrixHandle() { return _mem_manager._matrix; }

protected:
  virtual TimingType executeRun(Run run, unsigned int trial, std::vector<SemiRingType> &gold) = 0;

  // exact element-wise comparison of the first gold.size() outputs, at most 20 logged mismatches (:113-147)
  Correctness check_result(std::vector<SemiRingType> &gold) {
    if (gold.size() == 0) {
      std::cout << "Got gold of size " << gold.size() << " \n";
      return NOT_CHECKED;
    }
    const std::size_t output_length = _mem_manager._output_host_buffer.size() / sizeof(SemiRingType);
    const SemiRingType *res_ptr = reinterpret_cast<const SemiRingType *>(_mem_manager._output_host_buffer.data());
    if (output_length < gold.size()) return BAD_LENGTH;
    int error_count = 0;
    const int max_errors = 20;
    for (std::size_t i = 0; i < gold.size(); i++) {
      if (gold[i] != res_ptr[i]) {
        LOG_ERROR("Expected gold value ", gold[i], " at index ", i, " found ", res_ptr[i], " instead");
        if (++error_count == max_errors) break;
      }
    }
    return error_count > 0 ? BAD_VALUES : CORRECT;
  }

  // the runfile row as launch hints on top of the schedule the kernel JSON selected
  shb_launch_cfg launchCfgFor(const Run &run) const {
    shb_launch_cfg cfg = _args.launch;
    if (run.local1 == 64 || run.local1 == 128 || run.local1 == 256) cfg.block_threads = (int32_t)run.local1;
    const std::size_t groups = run.local1 ? run.global1 / run.local1 : run.global1;
    if (groups > 0 && groups < 2147483647u) cfg.grid_blocks = (int32_t)groups;
    return cfg;
  }

  // ONE launch of the hot path; returns the device time of the kernel only (:149-195)
  std::chrono::nanoseconds executeKernel(Run run) {
    start_timer(executeKernel, harness);
    const shb_launch_cfg cfg = launchCfgFor(run);
    std::cout << "Running kernel with queue: " << (void *)_ctx << " kernel : " << (void *)_mem_manager._matrix << "\n";
    uint64_t ns = 0;
    checkShbError(shb_spmv(_ctx, _mem_manager._matrix, _args.semiring, &cfg, _bound[2], _bound[3], &_args.alpha,
                           &_args.beta, _bound[_mem_manager._output_idx], &ns));
    LOG_DEBUG_INFO("Event CL_COMPLETE");
    report_timing(clEnqueueNDRangeKernel, harness, ns);
    return std::chrono::nanoseconds(ns);
  }

  // uploads the operator and the vectors, binds the positional arguments (:197-250)
  void allocateBuffers() {
    start_timer(allocateBuffers, Harness);
    LOG_DEBUG_INFO("setting matrix arguments");
    createOperator();
    LOG_DEBUG_INFO("setting vector arguments");
    _mem_manager._x_vect = createAndUploadGlobalArg(_args.x_vect, true);
    setGlobalArg(2, &_mem_manager._x_vect);
    _mem_manager._y_vect = createAndUploadGlobalArg(_args.y_vect, true);
    setGlobalArg(3, &_mem_manager._y_vect);
    LOG_DEBUG_INFO("setting constant arguments");
    setValueArg<SemiRingType>(4, &(_args.alpha));
    setValueArg<SemiRingType>(5, &(_args.beta));
    LOG_DEBUG_INFO("setting the output argument");
    _mem_manager._output_idx = 6;
    _mem_manager._output = createGlobalArg(_args.output);
    setGlobalArg(6, &_mem_manager._output);
    LOG_DEBUG_INFO("setting ", _args.temp_globals.size(), " temp global arguments (none allocated on CUDA)");
    LOG_DEBUG_INFO("setting ", _args.temp_locals.size(), " temp local arguments (shared memory is static on CUDA)");
    LOG_DEBUG_INFO("setting size arguments");
  }

  // the Lift kernels needed their global temporaries re-zeroed before every launch (:254-264);
  // the CUDA schedules keep partial results in registers / shared memory, so there is nothing to reset
  void resetTempBuffers() { start_timer(resetTempBuffers, Harness); }

  shb_buf *createAndUploadGlobalArg(std::vector<char> &arg, bool output = false) {
    start_timer(createAndUploadGlobalArg, harness);
    (void)output;
    LOG_DEBUG_INFO("Creating arg of size ", arg.size(), " from pointer ", static_cast<void *>(arg.data()));
    shb_buf *buffer = nullptr;
    checkShbError(shb_buf_create(_ctx, arg.size(), &buffer));
    uint64_t ns = 0;
    checkShbError(shb_buf_upload(_ctx, buffer, 0, arg.data(), arg.size(), &ns));
    report_timing(clEnqueueWriteBuffer, createAndUploadGlobalArg, ns);
    return buffer;
  }

  void writeToGlobalArg(std::vector<char> &arg, shb_buf *buffer) {
    start_timer(writeToGlobalArg, harness);
    LOG_DEBUG_INFO("uploading arg of size ", arg.size(), " from pointer ", static_cast<void *>(arg.data()));
    uint64_t ns = 0;
    checkShbError(shb_buf_upload(_ctx, buffer, 0, arg.data(), arg.size(), &ns));
    report_timing(clEnqueueWriteBuffer, writeToGlobalArg, ns);
  }

  void fillGlobalArg(size_t buffer_size, shb_buf *buffer) {
    start_timer(fillGlobalArg, harness);
    LOG_DEBUG_INFO("filling buffer with ", buffer_size, " bytes of zeros");
    uint64_t ns = 0;
    checkShbError(shb_buf_fill_zero(_ctx, buffer, 0, buffer_size, &ns));
    report_timing(clEnqueueFillBuffer, fillGlobalArg, ns);
  }

  void readFromGlobalArg(std::vector<char> &arg, shb_buf *buffer) {
    start_timer(readFromGlobalArg, harness);
    LOG_DEBUG_INFO("downloading arg of size: ", arg.size(), " into pointer ", static_cast<void *>(arg.data()));
    uint64_t ns = 0;
    checkShbError(shb_buf_download(_ctx, buffer, 0, arg.data(), arg.size(), &ns));
    report_timing(clEnqueueReadBuffer, readFromGlobalArg, ns);
  }

  shb_buf *createGlobalArg(uint64_t size) {
    start_timer(createGlobalArg, harness);
    LOG_DEBUG_INFO("creating global arg of size ", size);
    shb_buf *buffer = nullptr;
    checkShbError(shb_buf_create(_ctx, size, &buffer));
    checkShbError(shb_buf_fill_zero(_ctx, buffer, 0, size, nullptr));
    return buffer;
  }

  void setGlobalArg(int arg, shb_buf **mem) {
    start_timer(setGlobalArg, harness);
    LOG_DEBUG_INFO("setting global arg ", arg, " from memory ", static_cast<void *>(mem));
    if (arg < 0 || arg >= (int)(sizeof(_bound) / sizeof(_bound[0]))) checkShbError(SHB_ERR_INVALID);
    _bound[arg] = *mem;
  }

  template <typename ValueType> void setValueArg(unsigned int arg, ValueType *val) {
    start_timer(setValueArg, harness);
    LOG_DEBUG_INFO("setting value arg ", arg, " with value ", *val);
  }

  void setLocalArg(unsigned int arg, size_t size) {
    start_timer(setLocalArg, harness);
    LOG_DEBUG_INFO("setting local arg ", arg, " of size ", size);
  }

  shb_ctx *_ctx = nullptr;
  shb_buf *_bound[8];  // positional kernel arguments that are buffers: 2 x, 3 y, 6 out

  std::string _kernel_source;
  unsigned int _device;
  ArgContainer<SemiRingType> _args;
  DeviceMemoryManager<SemiRingType> _mem_manager;

  unsigned int _trials;
  std::chrono::milliseconds _timeout;
  double _delta;

private:
  // arguments 0 and 1: CSR from the host arrays, or one of BASELINE's synthetic operators built in HBM
  void createOperator() {
    start_timer(createAndUploadGlobalArg, harness);
    if (_args.synthetic.empty() && _args.coo) {
      // --gpu_encode: tuples up, stable sort by row + gather on the device (shb_coo.cu)
      const SparseMatrix<SemiRingType> &S = *_args.coo;
      uint64_t ns = 0;
      checkShbError(shb_csr_from_coo(_ctx, _args.matrix.n_rows, _args.matrix.n_cols, (int64_t)S.storedEntries(),
                                     S.tupleFirst().data(), S.tupleSecond().data(),
                                     _args.pattern ? nullptr : (const void *)S.tupleValues().data(),
                                     std::is_floating_point<SemiRingType>::value ? 0 : 1, _args.coo_truncate ? 1 : 0,
                                     &_mem_manager._matrix, &ns));
      report_timing(clEnqueueWriteBuffer, createAndUploadGlobalArg, ns);
      return;
    }
    if (_args.synthetic.empty()) {
      const CsrArrays<SemiRingType> &M = _args.matrix;
      uint64_t ns = 0;
      checkShbError(shb_csr_create(_ctx, M.n_rows, M.n_cols, M.nnz(), M.row_ptr.data(), M.col_idx.data(),
                                   _args.pattern ? nullptr : (const void *)M.vals.data(), (int)sizeof(SemiRingType),
                                   &_mem_manager._matrix, &ns));
      report_timing(clEnqueueWriteBuffer, createAndUploadGlobalArg, ns);
      return;
    }
    std::vector<std::string> f = CSV::tokenise_line(colonsToCommas(_args.synthetic));
    auto num = [&f](std::size_t i, long long dflt) { return i < f.size() ? std::atoll(f[i].c_str()) : dflt; };
    const std::string kind = f.empty() ? std::string() : f[0];
    if (kind == "uniform" || kind == "uniform-int") {
      const long long rows = num(1, 1 << 20), k = num(2, 16);
      checkShbError(shb_synth_uniform(_ctx, rows, rows, (int)k, 0x5EED0002ull, kind == "uniform" ? 0 : 1, 0, rows,
                                      &_mem_manager._matrix));
    } else if (kind == "rmat" || kind == "rmat-pr") {
      const long long scale = num(1, 16), ef = num(2, 16);
      const bool pr = kind == "rmat-pr";
      checkShbError(shb_synth_rmat(_ctx, (int)scale, (1ll << scale) * ef, pr ? 0x5EED0026ull : 0x5EED0024ull, pr ? 0 : 1,
                                   pr ? 1 : 0, 0.85f, &_mem_manager._matrix));
    } else if (kind == "grid") {
      checkShbError(shb_synth_grid2d(_ctx, num(1, 1024), num(2, 1024), 0x5EED0004ull, &_mem_manager._matrix));
    } else {
      LOG_ERROR("unknown synthetic operator '", _args.synthetic, "'");
      std::exit(1);
    }
  }
  static std::string colonsToCommas(std::string s) {
    for (auto &c : s)
      if (c == ':') c = ',';
    return s;
  }
};

template <typename TimingType, typename SemiRingType>
class IterativeHarness : public Harness<TimingType, SemiRingType> {
public:
  IterativeHarness(std::string &kernel_source, unsigned int platform, unsigned int device,
                   ArgContainer<SemiRingType> args, unsigned int trials, std::chrono::milliseconds timeout,
                   double delta)
      : Harness<TimingType, SemiRingType>(kernel_source, platform, device, std::move(args), trials, timeout, delta) {}

  // the reference's loops have no iteration cap and spin forever on some inputs
  void setIterationCap(unsigned int cap) { _max_iterations = cap ? cap : 1; }
  void setHostLoop(bool on) { _host_loop = on; }
  unsigned int lastIterationCount() const { return _last_iterations; }
  bool lastRunConverged() const { return _last_converged; }

protected:
  virtual bool should_terminate_iteration(std::vector<char> &input, std::vector<char> &output) = 0;

  // x_{k+1} = KERNEL(x_k, y = x_k) until the app's termination predicate holds, entirely on the device.
  // term_mode: SHB_TERM_ABSDIFF (sssp, pr: all |in - out| < delta) or SHB_TERM_EQUAL (bfs, scc).
  // Emits one RAW_RESULT row per iteration with that launch's device time, like app/sssp.cpp:124-126.
  std::vector<SqlStat> iterateOnDevice(Run run, unsigned int trial, int term_mode) {
    start_timer(iterateOnDevice, IterativeHarness);
    auto &mm = this->_mem_manager;
    const shb_launch_cfg cfg = this->launchCfgFor(run);
    std::vector<uint64_t> per_iter(_max_iterations, 0);
    uint32_t n_iters = 0;
    int in_work = 0;
    uint64_t total_ns = 0;
    shb_buf *x = this->_bound[2], *y = this->_bound[3], *work = this->_bound[mm._output_idx];
    const int status =
        shb_iterate(this->_ctx, mm._matrix, this->_args.semiring, &cfg, x, y, &this->_args.alpha, &this->_args.beta, work,
                    this->_delta, term_mode, /*y_rebind=*/1, _max_iterations, /*check_every=*/16, per_iter.data(), &n_iters,
                    &in_work, &total_ns);
    _last_converged = status == SHB_OK;
    if (status == SHB_ERR_NOT_CONVERGED)
      LOG_WARNING("iteration cap of ", _max_iterations, " reached before the termination test passed");
    else
      checkShbError(status);
    _last_iterations = n_iters;
    report_timing(iterateOnDevice, IterativeHarness, total_ns);
    std::vector<SqlStat> runtimes;
    for (uint32_t it = 0; it < n_iters; ++it)
      runtimes.push_back(SqlStat(std::chrono::nanoseconds(per_iter[it]), NOT_CHECKED, run.global1, run.local1, RAW_RESULT,
                                 trial, it));
    // leave the handles as the reference's pointer swaps would: "input" = the last output
    shb_buf *last = in_work ? work : x, *prev = in_work ? x : work;
    this->readFromGlobalArg(mm._output_host_buffer, last);
    if (mm._input_host_buffer.size() == mm._output_host_buffer.size())
      this->readFromGlobalArg(mm._input_host_buffer, prev);
    assertBuffersNotEqual(mm._output_host_buffer, mm._temp_out_buffer);
    this->setGlobalArg(mm._input_idx, &last);
    this->setGlobalArg(mm._output_idx, &prev);
    this->setGlobalArg(3, &last);
    return runtimes;
  }

  // the reference's own loop shape (app/sssp.cpp:97-155): launch, blocking read-back of the whole vector,
  // host-side termination scan, three argument re-binds -- every iteration
  std::vector<SqlStat> iterateOnHost(Run run, unsigned int trial) {
    start_timer(iterateOnHost, IterativeHarness);
    auto &mm = this->_mem_manager;
    std::vector<SqlStat> runtimes;
    shb_buf *input_mem = this->_bound[2], *output_mem = this->_bound[mm._output_idx];
    std::vector<char> *input_host = &mm._input_host_buffer, *output_host = &mm._output_host_buffer;
    bool should_terminate = false;
    unsigned int iteration = 0;
    do {
      LOG_DEBUG_INFO("Iteration: ", iteration);
      mm._temp_out_buffer = *output_host;
      this->resetTempBuffers();
      auto time = this->executeKernel(run);
      runtimes.push_back(SqlStat(time, NOT_CHECKED, run.global1, run.local1, RAW_RESULT, trial, iteration));
      this->readFromGlobalArg(*output_host, output_mem);
      assertBuffersNotEqual(*output_host, mm._temp_out_buffer);
      should_terminate = should_terminate_iteration(*input_host, *output_host);
      LOG_DEBUG_INFO("Should terminate iteration: ", should_terminate ? "true" : "false");
      std::swap(input_mem, output_mem);
      std::swap(input_host, output_host);
      this->setGlobalArg(mm._input_idx, &input_mem);
      this->setGlobalArg(mm._output_idx, &output_mem);
      this->setGlobalArg(3, &input_mem);
      iteration++;
    } while (!should_terminate && iteration < _max_iterations);
    _last_converged = should_terminate;
    _last_iterations = iteration;
    if (!should_terminate)
      LOG_WARNING("iteration cap of ", _max_iterations, " reached before the termination test passed");
    // the last output sits in *input_host after the final swap; keep it where outputVector() looks
    if (input_host != &mm._output_host_buffer) std::swap(mm._input_host_buffer, mm._output_host_buffer);
    return runtimes;
  }

  std::vector<SqlStat> iterate(Run run, unsigned int trial, int term_mode) {
    return _host_loop ? iterateOnHost(run, trial) : iterateOnDevice(run, trial, term_mode);
  }

  // restore x, y, out and the argument bindings for the next trial (:455-501).  The operator is
  // immutable on the device, so unlike the reference it is not uploaded again; the host mirror of x is
  // reset too (the reference forgets to).
  void resetInputs() {
    start_timer(allocateBuffers, Harness);
    auto &mm = this->_mem_manager;
    LOG_DEBUG_INFO("setting vector arguments");
    this->setGlobalArg(2, &mm._x_vect);
    this->writeToGlobalArg(this->_args.x_vect, mm._x_vect);
    this->setGlobalArg(3, &mm._y_vect);
    this->writeToGlobalArg(this->_args.y_vect, mm._y_vect);
    LOG_DEBUG_INFO("setting the output argument");
    this->setGlobalArg(6, &mm._output);
    this->fillGlobalArg(this->_args.output, mm._output);
    mm._input_host_buffer.assign(this->_args.x_vect.begin(), this->_args.x_vect.end());
    this->resetTempBuffers();
  }

  unsigned int _max_iterations = 100000;
  bool _host_loop = false;
  unsigned int _last_iterations = 0;
  bool _last_converged = false;
};
