// scc_harness -- component labelling under the max-min semiring (reference: app/scc.cpp).
//   matrix.scc_normalise(): diagonal -> INT_MIN, off-diagonal -> the entry's row index   (app/scc.cpp:225, src/sparse_matrix.cpp:433-456)
//   x[i] = i, y = INT_MIN, alpha = INT_MAX, beta = INT_MIN, zero = INT_MIN                  (app/scc.cpp:179-222)
//   x_{k+1} = KERNEL(x_k, y = x_k) until x_{k+1} == x_k                                     (:96-176)
// There is no separate canonical mode: the reference recurrence is the algorithm.
#include <limits>

#include "iterative_app.h"

typedef int SemiRingType;
typedef IterativeAppHarness<SemiRingType, SHB_TERM_EQUAL> HarnessSCC;

int main(int argc, char *argv[]) {
  start_timer(main, global);
  HarnessSetup<SemiRingType> s = harness_preamble<SemiRingType>(argc, argv);
  SparseMatrix<SemiRingType> &matrix = *s.matrix;
  KernelConfig<SemiRingType> &kernel = *s.kernel;

  IotaVectorGenerator<SemiRingType, XVectorGenerator<SemiRingType>> x;
  ConstYVectorGenerator<SemiRingType> y(std::numeric_limits<SemiRingType>::min());
  const SemiRingType alpha = std::numeric_limits<SemiRingType>::max();
  const SemiRingType beta = std::numeric_limits<SemiRingType>::min();
  const SemiRingType zero = std::numeric_limits<SemiRingType>::min();

  unsigned long max_alloc = deviceGetMaxAllocSize(s.platform, s.device);
  std::cout << "Got max alloc: " << max_alloc << "\n";

  ArgContainer<SemiRingType> args;
  try {
    if (s.synthetic.empty()) {
      matrix.scc_normalise();
      args = executorEncodeMatrix(max_alloc, kernel, matrix, zero, x, y, alpha, beta, true, s.gpu_encode);
    } else {
      args = executorSyntheticArgs(kernel, s.synthetic, matrix.height(), x, y, alpha, beta);
    }
  } catch (unsigned long attempted_alloc_size) {
    LOG_ERROR("Attempted to allocate: ", attempted_alloc_size, " bytes, but this platform's max is ", max_alloc);
  }
  expect_semiring(args, SHB_MAX_MIN_I32, "scc_harness");
  apply_variant_override(s, args);

  HarnessSCC harness(kernel.getSource(), s.platform, s.device, std::move(args), s.trials,
                     std::chrono::milliseconds(s.timeout_ms), s.delta);
  harness.setIterationCap(s.max_iterations);
  harness.setHostLoop(s.host_loop);
  run_iterative_app(harness, s, kernel.getName());
  return 0;
}
