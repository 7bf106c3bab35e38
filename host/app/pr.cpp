// pr_harness -- PageRank under plus-times (reference: app/pr.cpp).
//   matrix.pagerank_normalise(0.85, 0): v <- (|v| / colsum) * d               (app/pr.cpp:199, src/sparse_matrix.cpp:409-431)
//   x = 1/N, y = 1, alpha = 1, beta = (1 - 0.85)/N, zero = 0                   (app/pr.cpp:182-189)
//   faithful mode: the reference's recurrence, quirks included -- the normalised values (< 1) all truncate
//     to 0 through `int` (Q2) and y is re-bound to x after the first launch (Q9), so
//     x_1 = beta, x_{k+1} = beta * x_k; stops when all |x_k - x_{k+1}| < delta
//   canonical mode: float values kept, y == 1 constant: x_{k+1} = d * A^ x_k + (1-d)/N until max |dx| < delta
#include "iterative_app.h"
#include "group_app.h"

typedef float SemiRingType;
typedef IterativeAppHarness<SemiRingType, SHB_TERM_ABSDIFF> HarnessPR;

int main(int argc, char *argv[]) {
  start_timer(main, global);
  HarnessSetup<SemiRingType> s = harness_preamble<SemiRingType>(argc, argv);
  SparseMatrix<SemiRingType> &matrix = *s.matrix;
  KernelConfig<SemiRingType> &kernel = *s.kernel;

  const SemiRingType dampingFactor = 0.85f;
  ConstXVectorGenerator<SemiRingType> x(1.0f / (SemiRingType)matrix.height());
  ConstYVectorGenerator<SemiRingType> y(1.0f);
  const SemiRingType alpha = 1.0f;
  const SemiRingType beta = (1.0f - dampingFactor) / (SemiRingType)matrix.height();

  unsigned long max_alloc = deviceGetMaxAllocSize(s.platform, s.device);
  std::cout << "Got max alloc: " << max_alloc << "\n";

  ArgContainer<SemiRingType> args;
  try {
    if (s.synthetic.empty()) {
      matrix.pagerank_normalise(dampingFactor, 0.0f);
      args = executorEncodeMatrix(max_alloc, kernel, matrix, 0.0f, x, y, alpha, beta, /*truncate=*/!s.canonical(), s.gpu_encode);
    } else {
      args = executorSyntheticArgs(kernel, s.synthetic, matrix.height(), x, y, alpha, beta);
    }
  } catch (unsigned long attempted_alloc_size) {
    LOG_ERROR("Attempted to allocate: ", attempted_alloc_size, " bytes, but this platform's max is ", max_alloc);
  }
  expect_semiring(args, SHB_PLUS_TIMES_F32, "pr_harness");
  apply_variant_override(s, args);
  const shb_launch_cfg launch = args.launch;

  HarnessPR harness(kernel.getSource(), s.platform, s.device, std::move(args), s.trials,
                    std::chrono::milliseconds(s.timeout_ms), s.delta);
  harness.setIterationCap(s.max_iterations);
  harness.setHostLoop(s.host_loop);

  if (!s.canonical()) {
    run_iterative_app(harness, s, kernel.getName());
    return 0;
  }
  const int64_t n = matrix.height();
  if (s.gpus > 1) {  // row-partitioned over s.gpus devices, x exchanged by the SpMV epilogue's peer stores
    GroupOperator go;
    make_group_operator(s, harness.context(), harness.matrixHandle(), /*no out-edge copy: PageRank only pulls*/ true, go);
    std::vector<SemiRingType> out((std::size_t)n);
    run_canonical_app(s, kernel.getName(), go.device_names, "PageRank", [&](const Run &) {
      shb_graph_stats st = {};
      const int status = shb_group_pagerank(go.grp, go.part, &launch, 1.0f / (SemiRingType)n, (1.0f - dampingFactor) / (SemiRingType)n,
                                            s.delta, s.max_iterations, 8, out.data(), nullptr, &st);
      if (status != SHB_ERR_NOT_CONVERGED) checkShbError(status);
      return canonical_from(st);
    });
    dump_vector(s.dump, out);
    return 0;
  }
  shb_buf *xb = nullptr, *work = nullptr;
  checkShbError(shb_buf_create(harness.context(), (uint64_t)n * 4 + 4096, &xb));
  checkShbError(shb_buf_create(harness.context(), (uint64_t)n * 4 + 4096, &work));
  int in_work = 0;
  run_canonical_app(s, kernel.getName(), harness.getDeviceName(), "PageRank", [&](const Run &) {
    shb_graph_stats st = {};
    const int status = shb_pagerank(harness.context(), harness.matrixHandle(), &launch, xb, work,
                                    1.0f / (SemiRingType)n, (1.0f - dampingFactor) / (SemiRingType)n, s.delta,
                                    s.max_iterations, 8, &in_work, &st);
    if (status != SHB_ERR_NOT_CONVERGED) checkShbError(status);
    CanonicalResult r;
    r.ns = st.total_ns; r.iterations = st.iterations; r.edges = st.edges_traversed; r.reached = st.reached;
    r.converged = st.converged != 0;
    return r;
  });
  std::vector<SemiRingType> out((std::size_t)n);
  checkShbError(shb_buf_download(harness.context(), in_work ? work : xb, 0, out.data(), (uint64_t)n * 4, nullptr));
  dump_vector(s.dump, out);
  shb_buf_destroy(xb);
  shb_buf_destroy(work);
  return 0;
}
