// bfs_harness -- breadth-first search under the or-and semiring (reference: app/bfs.cpp).
//   x = y = (1 at index 0, 0 elsewhere), alpha = 1, beta = 0, zero = 0        (app/bfs.cpp:177-227)
//   faithful mode: x_{k+1} = KERNEL(x_k, y = x_k) until x_{k+1} == x_k (:94-174).  With beta = 0 this is the
//     set of vertices reachable by walks of length exactly k+1; it never settles on a bipartite component
//     (the reference spins forever there, TODO.md:7-9) -- the iteration cap ends it.
//   canonical mode: integer BFS levels, level[v] = first k with v in the frontier, -1 when unreached
//     (direction-optimising, device resident); stored zero values are no edges, as in bool_and.
#include "iterative_app.h"
#include "group_app.h"

typedef int SemiRingType;
typedef IterativeAppHarness<SemiRingType, SHB_TERM_EQUAL> HarnessBFS;

int main(int argc, char *argv[]) {
  start_timer(main, global);
  HarnessSetup<SemiRingType> s = harness_preamble<SemiRingType>(argc, argv);
  SparseMatrix<SemiRingType> &matrix = *s.matrix;
  KernelConfig<SemiRingType> &kernel = *s.kernel;

  SourceVectorGenerator<SemiRingType, XVectorGenerator<SemiRingType>> x(1, 0);
  SourceVectorGenerator<SemiRingType, YVectorGenerator<SemiRingType>> y(1, 0);
  const SemiRingType alpha = 1, beta = 0;

  unsigned long max_alloc = deviceGetMaxAllocSize(s.platform, s.device);
  std::cout << "Got max alloc: " << max_alloc << "\n";

  ArgContainer<SemiRingType> args;
  try {
    args = s.synthetic.empty() ? executorEncodeMatrix(max_alloc, kernel, matrix, 0, x, y, alpha, beta, true, s.gpu_encode)
                               : executorSyntheticArgs(kernel, s.synthetic, matrix.height(), x, y, alpha, beta);
  } catch (unsigned long attempted_alloc_size) {
    LOG_ERROR("Attempted to allocate: ", attempted_alloc_size, " bytes, but this platform's max is ", max_alloc);
  }
  expect_semiring(args, SHB_OR_AND_I32, "bfs_harness");
  apply_variant_override(s, args);

  HarnessBFS harness(kernel.getSource(), s.platform, s.device, std::move(args), s.trials,
                     std::chrono::milliseconds(s.timeout_ms), s.delta);
  harness.setIterationCap(s.max_iterations);
  harness.setHostLoop(s.host_loop);

  if (!s.canonical()) {
    run_iterative_app(harness, s, kernel.getName());
    return 0;
  }
  const int64_t n = matrix.height();
  if (s.gpus > 1) {  // row-partitioned, direction-optimising; frontier bits and counts travel as the kernels' peer stores
    GroupOperator go;
    const bool sym = s.symmetric || s.synthetic.compare(0, 5, "rmat:") == 0 || s.synthetic.compare(0, 5, "grid:") == 0;
    make_group_operator(s, harness.context(), harness.matrixHandle(), sym, go);
    std::vector<SemiRingType> out((std::size_t)n);
    run_canonical_app(s, kernel.getName(), go.device_names, "BFS", [&](const Run &) {
      shb_graph_stats st = {};
      const int status = shb_group_bfs(go.grp, go.part, 0, out.data(), 0, 8, &st);
      if (status != SHB_ERR_NOT_CONVERGED) checkShbError(status);
      return canonical_from(st);
    });
    dump_vector(s.dump, out);
    return 0;
  }
  shb_buf *levels = nullptr;
  checkShbError(shb_buf_create(harness.context(), (uint64_t)n * 4 + 4096, &levels));
  shb_csr *AT = nullptr;
  checkShbError(shb_csr_transpose(harness.context(), harness.matrixHandle(), &AT));
  run_canonical_app(s, kernel.getName(), harness.getDeviceName(), "BFS", [&](const Run &) {
    shb_graph_stats st = {};
    const int status = shb_bfs(harness.context(), harness.matrixHandle(), AT, 0, levels, 0, &st);
    if (status != SHB_ERR_NOT_CONVERGED) checkShbError(status);
    CanonicalResult r;
    r.ns = st.total_ns; r.iterations = st.iterations; r.edges = st.edges_traversed; r.reached = st.reached;
    r.converged = st.converged != 0;
    return r;
  });
  std::vector<SemiRingType> out((std::size_t)n);
  checkShbError(shb_buf_download(harness.context(), levels, 0, out.data(), (uint64_t)n * 4, nullptr));
  dump_vector(s.dump, out);
  shb_csr_destroy(AT);
  shb_buf_destroy(levels);
  return 0;
}
