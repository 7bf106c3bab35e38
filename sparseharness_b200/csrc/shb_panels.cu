// shb_panels.cu -- column panels for operators whose x does not fit the L2.
//
// PageRank on BASELINE.json config 5 (2^26 columns: x = 268 MB against a 126 MB L2) spends most of its DRAM traffic
// on 32-byte sectors fetched for 4-byte gathers that miss the L2 (profiles/r01_summary.md: ~34 GB of sectors for
// 9.4 GB of algorithmic bytes per sweep).  A column panel is the sub-operator of the columns
// [p * panel_cols, (p + 1) * panel_cols): its gathers fall into an x slice small enough to stay L2 resident.
// The SpMV then runs panel after panel, each a CSR of its own (global column ids, all the rows):
//     acc_0 = epi(A_0 (x) x, alpha, y, beta)             the caller's epilogue, once
//     acc_p = epi(A_p (x) x, alpha, acc_{p-1}, id)       id = the beta that makes h(v, id) = v
// which is the caller's out = epi((+)_p A_p (x) x, alpha, y, beta) because in all four semirings the epilogue is
// f(dp, alpha) (+) h(y, beta) with f distributing over (+) (example/<alg>/kernel5.json:3 user functions):
//   plus-times id = 1, min-plus id = 0 (all operands are |.|), or-and id = 1, max-min id = INT_MAX.
// Exact for min-plus / or-and / max-min and for integer-valued floats; general floats differ by the summation
// order (1e-5 relative, like every other schedule).  The last panel writes the caller's destinations and runs the
// convergence test; the ones before write a scratch accumulator owned by the operator.
#include <algorithm>
#include <cub/device/device_scan.cuh>
#include <limits.h>

#include "shb_internal.cuh"

// per row: how many entries fall into each panel
__global__ void panel_count_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col, int64_t n_rows,
                                   uint32_t n_cols, uint32_t panel_cols, int n_panels, int32_t *__restrict__ counts) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    int cnt[SHB_MAX_PANELS];
#pragma unroll
    for (int p = 0; p < SHB_MAX_PANELS; ++p) cnt[p] = 0;
    for (int k = row_ptr[r]; k < row_ptr[r + 1]; ++k) {
      const uint32_t c = (uint32_t)col[k];
      if (c < n_cols) {
        const int p = (int)(c / panel_cols);
#pragma unroll
        for (int q = 0; q < SHB_MAX_PANELS; ++q) cnt[q] += (q == p);
      }
    }
#pragma unroll
    for (int p = 0; p < SHB_MAX_PANELS; ++p)
      if (p < n_panels) counts[(int64_t)p * (n_rows + 1) + r] = cnt[p];
  }
}

struct PanelPtrs {
  int32_t *row_ptr[SHB_MAX_PANELS];
  int32_t *col[SHB_MAX_PANELS];
  uint32_t *val[SHB_MAX_PANELS];
};

// per row: copy the entries into their panels, keeping the row's order
__global__ void panel_fill_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                                  const uint32_t *__restrict__ val, int64_t n_rows, uint32_t n_cols, uint32_t panel_cols,
                                  const __grid_constant__ PanelPtrs P) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    int pos[SHB_MAX_PANELS];
#pragma unroll
    for (int p = 0; p < SHB_MAX_PANELS; ++p) pos[p] = P.row_ptr[p] ? P.row_ptr[p][r] : 0;
    for (int k = row_ptr[r]; k < row_ptr[r + 1]; ++k) {
      const uint32_t c = (uint32_t)col[k];
      if (c >= n_cols) continue;  // out-of-range column: gathers ZERO, contributes the additive identity
      const int p = (int)(c / panel_cols);
      int dst = 0;
#pragma unroll
      for (int q = 0; q < SHB_MAX_PANELS; ++q)
        if (q == p) dst = pos[q]++;
      SHB_DEV_CHECK(dst >= 0 && dst < P.row_ptr[p][r + 1]);
      P.col[p][dst] = (int32_t)c;
      if (val) P.val[p][dst] = val[k];
    }
  }
}

void shb_panels_free(shb_csr *A) {
  for (int p = 0; p < A->n_panels; ++p) shb_csr_destroy(A->panels[p]);
  A->n_panels = 0;
  cudaFree(A->panel_acc);
  A->panel_acc = nullptr;
}

SHB_API int shb_csr_build_panels(shb_ctx *c, shb_csr *A, int n_panels, int *n_built) {
  SHB_REQUIRE(c && A, "bad arguments");
  SHB_CUDA(cudaSetDevice(c->device));
  if (A->n_panels) {
    for (int p = 0; p < A->n_panels; ++p) A->device_bytes -= A->panels[p]->device_bytes;
    A->device_bytes -= (uint64_t)A->n_rows * 4;
    shb_panels_free(A);
  }
  if (n_panels <= 0) {  // automatic: x slices of at most 64 MB (half the L2; measured best on config 5: 16 / 32 / 64 / 128 MB
                        // slices give 12.5 / 8.4 / 7.0 / 7.4 ms per sweep against 9.3 without panels)
    int64_t slice = 64ll << 20;
    if (const char *e = getenv("SHB_PANEL_BYTES")) slice = std::max<int64_t>(1 << 16, atoll(e));
    n_panels = (int)std::min<int64_t>(SHB_MAX_PANELS, (A->n_cols * 4 + slice - 1) / slice);
  }
  n_panels = std::min(n_panels, SHB_MAX_PANELS);
  if (n_built) *n_built = n_panels <= 1 ? 0 : n_panels;
  if (n_panels <= 1 || A->nnz == 0 || A->n_rows == 0) return SHB_OK;
  const uint32_t panel_cols = (uint32_t)((A->n_cols + n_panels - 1) / n_panels);
  const int64_t n = A->n_rows;
  int32_t *counts = nullptr;
  void *tmp = nullptr;
  size_t tmp_bytes = 0;
  int status = SHB_OK;
  std::vector<int32_t> totals((size_t)n_panels, 0);
  PanelPtrs P;
  memset(&P, 0, sizeof P);
#define PN_CUDA(call)                                                                                     \
  do {                                                                                                    \
    cudaError_t _e = (call);                                                                              \
    if (_e != cudaSuccess) {                                                                              \
      shb_set_error("%s failed: %s (code: %d) in %s at line %d", #call, cudaGetErrorString(_e), (int)_e, \
                    __FILE__, __LINE__);                                                                  \
      status = _e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;                              \
      goto done;                                                                                          \
    }                                                                                                     \
  } while (0)
  {
    PN_CUDA(cudaMalloc(&counts, (size_t)n_panels * (size_t)(n + 1) * 4));
    PN_CUDA(cudaMemsetAsync(counts, 0, (size_t)n_panels * (size_t)(n + 1) * 4, c->stream));
    const unsigned grid = grid_for(n, 256, c->prop.multiProcessorCount * 16);
    panel_count_kernel<<<grid, 256, 0, c->stream>>>(A->row_ptr, A->col, n, (uint32_t)A->n_cols, panel_cols, n_panels, counts);
    PN_CUDA(cudaGetLastError());
    PN_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts, counts, (int)(n + 1), c->stream));
    PN_CUDA(cudaMalloc(&tmp, tmp_bytes));
    for (int p = 0; p < n_panels; ++p) {  // in place: counts[p][0..n] becomes the panel's row_ptr
      int32_t *cp = counts + (int64_t)p * (n + 1);
      PN_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, cp, cp, (int)(n + 1), c->stream));
      PN_CUDA(cudaMemcpyAsync(&totals[(size_t)p], cp + n, 4, cudaMemcpyDeviceToHost, c->stream));
    }
    PN_CUDA(cudaStreamSynchronize(c->stream));
    for (int p = 0; p < n_panels; ++p) {
      shb_csr *B = nullptr;
      status = shb_csr_alloc(c, n, A->n_cols, totals[(size_t)p], A->val != nullptr, &B);
      if (status != SHB_OK) goto done;
      A->panels[A->n_panels++] = B;
      PN_CUDA(cudaMemcpyAsync(B->row_ptr, counts + (int64_t)p * (n + 1), (size_t)(n + 1) * 4, cudaMemcpyDeviceToDevice, c->stream));
      P.row_ptr[p] = B->row_ptr;
      P.col[p] = B->col;
      P.val[p] = (uint32_t *)B->val;
    }
    panel_fill_kernel<<<grid, 256, 0, c->stream>>>(A->row_ptr, A->col, (const uint32_t *)A->val, n, (uint32_t)A->n_cols, panel_cols, P);
    PN_CUDA(cudaGetLastError());
    for (int p = 0; p < A->n_panels; ++p) {
      status = shb_csr_analyse(c, A->panels[p]);
      if (status != SHB_OK) goto done;
      A->device_bytes += A->panels[p]->device_bytes;
    }
    PN_CUDA(cudaMalloc(&A->panel_acc, (size_t)n * 4));
    A->device_bytes += (uint64_t)n * 4;
    A->panel_cols = panel_cols;
  }
done:
#undef PN_CUDA
  cudaFree(counts);
  cudaFree(tmp);
  if (status != SHB_OK) shb_panels_free(A);
  return status;
}

// the panel loop (called from shb_launch_spmv when the operator has panels and the caller asked for AUTO)
int shb_launch_spmv_panels(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const void *x, const void *y,
                           const void *alpha, const void *beta, void *const *outs, int n_out, int64_t row_offset,
                           const IterCtl &ctl) {
  union {
    float f;
    int32_t i;
  } id;
  switch (semiring) {
    case SHB_PLUS_TIMES_F32: id.f = 1.0f; break;
    case SHB_MIN_PLUS_F32: id.f = 0.0f; break;
    case SHB_OR_AND_I32: id.i = 1; break;
    case SHB_MAX_MIN_I32: id.i = INT_MAX; break;
    default: shb_set_error("unknown semiring %d", semiring); return SHB_ERR_UNSUPPORTED;
  }
  // y and the destinations are indexed by row_offset + r: shift the scratch so that acc[row_offset + r] is its word r
  char *acc = (char *)A->panel_acc - 4 * row_offset;
  for (int p = 0; p < A->n_panels; ++p) {
    const bool first = p == 0, last = p == A->n_panels - 1;
    shb_launch_cfg lc = {};
    if (cfg) lc = *cfg;
    lc.variant = SHB_VARIANT_AUTO;
    // tuning aid: SHB_PANEL_LANES=L runs every panel under the vector schedule with L lanes per row
    static const int panel_lanes = [] {
      const char *e = getenv("SHB_PANEL_LANES");
      return e ? atoi(e) : 0;
    }();
    if (panel_lanes > 0) {
      lc.variant = SHB_VARIANT_VECTOR;
      lc.lanes = panel_lanes;
    }
    if (!first) {
      lc.flags &= ~(uint32_t)(SHB_F_Y_IS_ONE | SHB_F_INTERNAL_Y_ONE);
      lc.flags |= SHB_F_NO_Y_SKIP;
    }
    IterCtl cp = ctl;
    if (!last) cp.changed = nullptr;
    void *acc_out[1] = {acc};
    SHB_TRY(shb_launch_spmv(c, A->panels[p], semiring, &lc, x, first ? y : (const void *)acc, alpha, first ? beta : (const void *)&id,
                            last ? outs : acc_out, last ? n_out : 1, row_offset, cp));
  }
  return SHB_OK;
}
