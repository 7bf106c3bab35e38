// group_app.h -- `--gpus N` for the canonical modes of pr_harness / bfs_harness / sssp_harness: the operator is
// row-partitioned over the first N devices of the box and the round / level loops run through shb_group_*
// (include/shb.h), one host thread driving all devices.  The reference has one queue on one device
// (inc/harness.h:78-81); its host loops (app/{pr,bfs,sssp}.cpp executeRun) are what the group drivers generalise.
// The rows printed stay the reference's (inc/sql_stat.h:28-79): one RAW_RESULT per trial + the MEDIAN_RESULT.
#pragma once

#include <string>
#include <vector>

#include "app_common.h"
#include "device_utils.h"

struct GroupOperator {
  shb_group *grp = nullptr;
  shb_gpart *part = nullptr;
  int n_dev = 0;
  std::string device_names;
  ~GroupOperator() {
    if (part) shb_gpart_destroy(part);
    if (grp) shb_group_destroy(grp);
  }
};

// Builds the partition.  Synthetic operators are generated on every device (bit-identical generator) and each
// device keeps its nnz-balanced, 32-aligned row block; a Matrix Market operator is taken from the harness's
// device copy (A0 on ctx0) and split on the host.  `symmetric`: the operator equals its transpose, or the driver
// never looks at out-edges (PageRank): no out-edge copy is made.
template <typename T>
void make_group_operator(const HarnessSetup<T> &s, shb_ctx *ctx0, const shb_csr *A0, bool symmetric, GroupOperator &out) {
  start_timer(make_group_operator, group_app);
  const int N = (int)s.gpus;
  out.n_dev = N;
  checkShbError(shb_group_create(N, nullptr, &out.grp));
  for (int i = 0; i < N; ++i) {
    shb_ctx *c = nullptr;
    checkShbError(shb_group_ctx(out.grp, i, &c));
    char name[256] = {0};
    checkShbError(shb_ctx_device_name(c, name, sizeof name));
    out.device_names += (i ? " + " : "") + std::string(name);
  }
  shb_csr_info info;
  checkShbError(shb_csr_get_info(A0, &info));
  if (s.synthetic.empty()) {
    std::vector<int64_t> rp((std::size_t)info.n_rows + 1);
    std::vector<int32_t> col((std::size_t)std::max<int64_t>(info.nnz, 1));
    std::vector<uint32_t> val(info.has_values ? (std::size_t)std::max<int64_t>(info.nnz, 1) : 0);
    checkShbError(shb_csr_download(ctx0, A0, 0, info.n_rows, rp.data(), col.data(), info.has_values ? val.data() : nullptr));
    checkShbError(shb_group_partition_host(out.grp, info.n_rows, info.n_cols, rp.data(), col.data(),
                                           info.has_values ? (const void *)val.data() : nullptr, symmetric ? 1 : 0, &out.part));
    return;
  }
  std::string spec = s.synthetic;
  for (auto &ch : spec)
    if (ch == ':') ch = ',';
  const CSV::csv_line f = CSV::tokenise_line(spec);
  auto num = [&f](std::size_t i, long long dflt) { return i < f.size() ? std::atoll(f[i].c_str()) : dflt; };
  const std::string kind = f.empty() ? std::string() : f[0];
  std::vector<int64_t> bounds((std::size_t)N + 1);
  checkShbError(shb_csr_balance_rows(ctx0, A0, N, bounds.data()));
  for (int k = 1; k < N; ++k) {
    bounds[(std::size_t)k] = std::min<int64_t>(info.n_rows, (bounds[(std::size_t)k] + 31) / 32 * 32);
    bounds[(std::size_t)k] = std::max(bounds[(std::size_t)k], bounds[(std::size_t)k - 1]);
  }
  std::vector<shb_csr *> blocks((std::size_t)N, nullptr), at_blocks((std::size_t)N, nullptr);
  for (int i = 0; i < N; ++i) {
    shb_ctx *c = nullptr;
    checkShbError(shb_group_ctx(out.grp, i, &c));
    shb_csr *full = nullptr;
    if (kind == "rmat" || kind == "rmat-pr") {
      const long long scale = num(1, 16), ef = num(2, 16);
      const bool pr = kind == "rmat-pr";
      checkShbError(shb_synth_rmat(c, (int)scale, (1ll << scale) * ef, pr ? 0x5EED0026ull : 0x5EED0024ull, pr ? 0 : 1, pr ? 1 : 0,
                                   0.85f, &full));
    } else if (kind == "grid") {
      checkShbError(shb_synth_grid2d(c, num(1, 1024), num(2, 1024), 0x5EED0004ull, &full));
    } else if (kind == "uniform" || kind == "uniform-int") {
      const long long rows = num(1, 1 << 20), k = num(2, 16);
      checkShbError(shb_synth_uniform(c, rows, rows, (int)k, 0x5EED0002ull, kind == "uniform" ? 0 : 1, 0, rows, &full));
    } else {
      LOG_ERROR("unknown synthetic operator '", s.synthetic, "'");
      std::exit(1);
    }
    const int64_t lo = bounds[(std::size_t)i], hi = bounds[(std::size_t)i + 1];
    checkShbError(shb_csr_slice_rows(c, full, lo, hi - lo, &blocks[(std::size_t)i]));
    if (!symmetric) {
      shb_csr *t = nullptr;
      checkShbError(shb_csr_transpose(c, full, &t));
      checkShbError(shb_csr_slice_rows(c, t, lo, hi - lo, &at_blocks[(std::size_t)i]));
      shb_csr_destroy(t);
    }
    shb_csr_destroy(full);
  }
  checkShbError(shb_group_adopt(out.grp, blocks.data(), symmetric ? nullptr : at_blocks.data(), bounds.data(), info.n_cols, &out.part));
}

inline CanonicalResult canonical_from(const shb_graph_stats &st) {
  CanonicalResult r;
  r.ns = st.total_ns;
  r.iterations = st.iterations;
  r.edges = st.edges_traversed;
  r.reached = st.reached;
  r.converged = st.converged != 0;
  return r;
}
