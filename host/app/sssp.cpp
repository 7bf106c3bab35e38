// sssp_harness -- single-source shortest paths under the min-plus semiring (reference: app/sssp.cpp).
//   x = y = (0 at index 0, FLT_MAX elsewhere), alpha = beta = 0, zero = FLT_MAX   (app/sssp.cpp:179-232)
//   faithful mode: x_{k+1} = KERNEL(x_k, y = x_k) until all |x_k - x_{k+1}| < delta  (:97-176)
//   canonical mode: libshb's worklist SSSP to the exact fixed point, float edge weights kept
#include <limits>

#include "iterative_app.h"
#include "group_app.h"

typedef float SemiRingType;
typedef IterativeAppHarness<SemiRingType, SHB_TERM_ABSDIFF> HarnessSSSP;

int main(int argc, char *argv[]) {
  start_timer(main, global);
  HarnessSetup<SemiRingType> s = harness_preamble<SemiRingType>(argc, argv);
  SparseMatrix<SemiRingType> &matrix = *s.matrix;
  KernelConfig<SemiRingType> &kernel = *s.kernel;

  const SemiRingType inf = std::numeric_limits<SemiRingType>::max();
  SourceVectorGenerator<SemiRingType, XVectorGenerator<SemiRingType>> x(0.0f, inf);
  SourceVectorGenerator<SemiRingType, YVectorGenerator<SemiRingType>> y(0.0f, inf);
  const SemiRingType alpha = 0.0f, beta = 0.0f;

  unsigned long max_alloc = deviceGetMaxAllocSize(s.platform, s.device);
  std::cout << "Got max alloc: " << max_alloc << "\n";

  ArgContainer<SemiRingType> args;
  try {
    args = s.synthetic.empty()
               ? executorEncodeMatrix(max_alloc, kernel, matrix, inf, x, y, alpha, beta, /*truncate=*/!s.canonical(), s.gpu_encode)
               : executorSyntheticArgs(kernel, s.synthetic, matrix.height(), x, y, alpha, beta);
  } catch (unsigned long attempted_alloc_size) {
    LOG_ERROR("Attempted to allocate: ", attempted_alloc_size, " bytes, but this platform's max is ", max_alloc);
  }
  expect_semiring(args, SHB_MIN_PLUS_F32, "sssp_harness");
  apply_variant_override(s, args);

  HarnessSSSP harness(kernel.getSource(), s.platform, s.device, std::move(args), s.trials,
                      std::chrono::milliseconds(s.timeout_ms), s.delta);
  harness.setIterationCap(s.max_iterations);
  harness.setHostLoop(s.host_loop);

  if (!s.canonical()) {
    run_iterative_app(harness, s, kernel.getName());
    return 0;
  }
  const int64_t n = matrix.height();
  if (s.gpus > 1) {  // block-asynchronous near-far in distance windows, boundary candidates as peer red.min
    GroupOperator go;
    const bool sym = s.symmetric || s.synthetic.compare(0, 5, "rmat:") == 0 || s.synthetic.compare(0, 5, "grid:") == 0;
    make_group_operator(s, harness.context(), harness.matrixHandle(), sym, go);
    std::vector<SemiRingType> out((std::size_t)n);
    run_canonical_app(s, kernel.getName(), go.device_names, "SSSP", [&](const Run &) {
      shb_graph_stats st = {};
      const int status = shb_group_sssp(go.grp, go.part, 0, out.data(), -1.0f, 0, 8, &st);
      if (status != SHB_ERR_NOT_CONVERGED) checkShbError(status);
      return canonical_from(st);
    });
    dump_vector(s.dump, out);
    return 0;
  }
  shb_buf *dist = nullptr;
  checkShbError(shb_buf_create(harness.context(), (uint64_t)n * 4 + 4096, &dist));
  shb_csr *AT = nullptr;
  checkShbError(shb_csr_transpose(harness.context(), harness.matrixHandle(), &AT));
  run_canonical_app(s, kernel.getName(), harness.getDeviceName(), "SSSP", [&](const Run &) {
    shb_graph_stats st = {};
    const int status = shb_sssp(harness.context(), harness.matrixHandle(), AT, 0, dist, 0, &st);
    if (status != SHB_ERR_NOT_CONVERGED) checkShbError(status);
    CanonicalResult r;
    r.ns = st.total_ns; r.iterations = st.iterations; r.edges = st.edges_traversed; r.reached = st.reached;
    r.converged = st.converged != 0;
    return r;
  });
  std::vector<SemiRingType> out((std::size_t)n);
  checkShbError(shb_buf_download(harness.context(), dist, 0, out.data(), (uint64_t)n * 4, nullptr));
  dump_vector(s.dump, out);
  shb_csr_destroy(AT);
  shb_buf_destroy(dist);
  return 0;
}
