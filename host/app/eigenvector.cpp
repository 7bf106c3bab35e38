// eigenvector_harness -- power iteration under plus-times (reference: app/eigenvector.cpp, which is entirely
// commented out there and not built, CMakeLists.txt:50; its intent is kept).
//   x = 1, y = 0, alpha = 1, beta = 0                                                    (app/eigenvector.cpp:219-222)
//   faithful mode: the recurrence the dead code describes -- x_{k+1} = A x_k, stop when all |x_k - x_{k+1}| < delta,
//     at most 10 launches (`while (!should_terminate && itcnt < 10)`, :110), no normalisation
//   canonical mode: out = A x_k; s_k = max |out_i|; x_{k+1} = out / s_k until max |dx| < delta (shb_power_iteration):
//     x converges to the dominant eigenvector, s_k to the dominant eigenvalue's magnitude
#include "iterative_app.h"

typedef float SemiRingType;
typedef IterativeAppHarness<SemiRingType, SHB_TERM_ABSDIFF> HarnessEigenvector;

int main(int argc, char *argv[]) {
  start_timer(main, global);
  HarnessSetup<SemiRingType> s = harness_preamble<SemiRingType>(argc, argv);
  SparseMatrix<SemiRingType> &matrix = *s.matrix;
  KernelConfig<SemiRingType> &kernel = *s.kernel;

  ConstXVectorGenerator<SemiRingType> x(1.0f);
  ConstYVectorGenerator<SemiRingType> y(0);
  const SemiRingType alpha = 1.0f, beta = 0.0f;

  unsigned long max_alloc = deviceGetMaxAllocSize(s.platform, s.device);
  std::cout << "Got max alloc: " << max_alloc << "\n";

  ArgContainer<SemiRingType> args;
  try {
    args = s.synthetic.empty()
               ? executorEncodeMatrix(max_alloc, kernel, matrix, 0.0f, x, y, alpha, beta, /*truncate=*/!s.canonical(), s.gpu_encode)
               : executorSyntheticArgs(kernel, s.synthetic, matrix.height(), x, y, alpha, beta);
  } catch (unsigned long attempted_alloc_size) {
    LOG_ERROR("Attempted to allocate: ", attempted_alloc_size, " bytes, but this platform's max is ", max_alloc);
  }
  expect_semiring(args, SHB_PLUS_TIMES_F32, "eigenvector_harness");
  apply_variant_override(s, args);
  const shb_launch_cfg launch = args.launch;

  HarnessEigenvector harness(kernel.getSource(), s.platform, s.device, std::move(args), s.trials,
                             std::chrono::milliseconds(s.timeout_ms), s.delta);
  const unsigned int cap = std::min(10u, std::max(1u, s.max_iterations));  // the reference's `itcnt < 10`
  harness.setIterationCap(s.canonical() ? s.max_iterations : cap);
  harness.setHostLoop(s.host_loop);

  if (!s.canonical()) {
    run_iterative_app(harness, s, kernel.getName());
    return 0;
  }
  const int64_t n = matrix.height();
  shb_buf *xb = nullptr, *work = nullptr;
  checkShbError(shb_buf_create(harness.context(), (uint64_t)n * 4 + 4096, &xb));
  checkShbError(shb_buf_create(harness.context(), (uint64_t)n * 4 + 4096, &work));
  int in_work = 0;
  float eigenvalue = 0.0f;
  run_canonical_app(s, kernel.getName(), harness.getDeviceName(), "Eigenvector", [&](const Run &) {
    checkShbError(shb_buf_fill_u32(harness.context(), xb, 0, (uint64_t)n, 0x3f800000u));  // x_0 = 1
    shb_graph_stats st = {};
    const int status = shb_power_iteration(harness.context(), harness.matrixHandle(), &launch, xb, work, s.delta,
                                           s.max_iterations, &eigenvalue, &in_work, &st);
    if (status != SHB_ERR_NOT_CONVERGED) checkShbError(status);
    CanonicalResult r;
    r.ns = st.total_ns; r.iterations = st.iterations; r.edges = st.edges_traversed; r.reached = st.reached;
    r.converged = st.converged != 0;
    return r;
  });
  std::cout << "Dominant eigenvalue estimate: " << eigenvalue << ENDL;
  std::vector<SemiRingType> out((std::size_t)n);
  checkShbError(shb_buf_download(harness.context(), in_work ? work : xb, 0, out.data(), (uint64_t)n * 4, nullptr));
  dump_vector(s.dump, out);
  shb_buf_destroy(xb);
  shb_buf_destroy(work);
  return 0;
}
