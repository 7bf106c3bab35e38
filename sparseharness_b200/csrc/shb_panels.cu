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
// order (1e-5 relative, like every other schedule).
//
// A panel keeps only the rows it HAS entries in (a compressed row list: row r of the panel is row row_ids[r] of the
// vectors): on a power-law operator most (row, panel) pairs are empty, and a panel launch that visits every row pays
// for them -- two row_ptr loads, a read and a write of the accumulator each (measured: the row visits, not the entries,
// bound a config-5 panel launch; profiles/r02_summary.md section 2).  So the loop is
//     acc   = epi(ZERO, alpha, y, beta)                 every row, one elementwise pass (the h(y, beta) term)
//     acc   = epi(A_p (x) x, alpha, acc, id)            the rows of panel p only, p = 0 .. P-1, in place
//     out_i = acc, convergence test                     every row, one elementwise pass (skipped when the caller has one
//                                                       destination and no test: the panels then work on it directly)
#include <algorithm>
#include <cub/device/device_scan.cuh>
#include <limits.h>

#include "spmv.cuh"

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

// after the in-place scan counts[] is the panel's row_ptr over ALL rows; flag the rows that have entries
__global__ void panel_flag_kernel(const int32_t *__restrict__ rp_full, int64_t n_rows, int32_t *__restrict__ flag) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r <= n_rows; r += (int64_t)gridDim.x * blockDim.x)
    flag[r] = (r < n_rows && rp_full[r + 1] > rp_full[r]) ? 1 : 0;
}

// pos = exclusive scan of the flags: the compressed row list and its row_ptr
__global__ void panel_compact_kernel(const int32_t *__restrict__ rp_full, const int32_t *__restrict__ pos, int64_t n_rows,
                                     int32_t *__restrict__ row_ids, int32_t *__restrict__ rp_compact) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r <= n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    if (r == n_rows) {
      rp_compact[pos[r]] = rp_full[r];  // the end of the last row
    } else if (rp_full[r + 1] > rp_full[r]) {
      SHB_DEV_CHECK(pos[r] >= 0 && pos[r] < pos[n_rows]);
      row_ids[pos[r]] = (int32_t)r;
      rp_compact[pos[r]] = rp_full[r];
    }
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
  if (n_panels <= 0) {  // automatic: x slices of at most 40 MB.  Measured on config 5 with compressed-row panels (ms per
                        // sweep by slice size): 16 MB 6.32, 24 MB 6.10, 32 MB 5.87, 38 MB 5.81, 45 MB 5.82, 54 MB 6.06,
                        // 64 MB 6.72; 9.33 without panels
    int64_t slice = 40ll << 20;
    if (const char *e = getenv("SHB_PANEL_BYTES")) slice = std::max<int64_t>(1 << 16, atoll(e));
    n_panels = (int)std::min<int64_t>(SHB_MAX_PANELS, (A->n_cols * 4 + slice - 1) / slice);
  }
  n_panels = std::min(n_panels, SHB_MAX_PANELS);
  if (n_built) *n_built = n_panels <= 1 ? 0 : n_panels;
  if (n_panels <= 1 || A->nnz == 0 || A->n_rows == 0) return SHB_OK;
  const uint32_t panel_cols = (uint32_t)((A->n_cols + n_panels - 1) / n_panels);
  const int64_t n = A->n_rows;
  int32_t *counts = nullptr, *pos = nullptr;
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
    PN_CUDA(cudaMalloc(&pos, (size_t)(n + 1) * 4));
    for (int p = 0; p < n_panels; ++p) {
      // the rows this panel has entries in: flags -> positions -> compressed row list + its row_ptr
      const int32_t *rp_full = counts + (int64_t)p * (n + 1);
      int32_t m = 0;
      panel_flag_kernel<<<grid, 256, 0, c->stream>>>(rp_full, n, pos);
      PN_CUDA(cudaGetLastError());
      PN_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, pos, pos, (int)(n + 1), c->stream));
      PN_CUDA(cudaMemcpyAsync(&m, pos + n, 4, cudaMemcpyDeviceToHost, c->stream));
      PN_CUDA(cudaStreamSynchronize(c->stream));
      shb_csr *B = nullptr;
      status = shb_csr_alloc(c, m, A->n_cols, totals[(size_t)p], A->val != nullptr, &B);
      if (status != SHB_OK) goto done;
      A->panels[A->n_panels++] = B;
      PN_CUDA(cudaMalloc(&B->row_ids, (size_t)std::max(m, 1) * 4));
      B->device_bytes += (uint64_t)std::max(m, 1) * 4;
      panel_compact_kernel<<<grid, 256, 0, c->stream>>>(rp_full, pos, n, B->row_ids, B->row_ptr);
      PN_CUDA(cudaGetLastError());
      P.row_ptr[p] = const_cast<int32_t *>(rp_full);  // the fill below places entries by the row_ptr over ALL rows
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
  cudaFree(pos);
  cudaFree(tmp);
  if (status != SHB_OK) shb_panels_free(A);
  return status;
}

// acc[r] = epi(ZERO, alpha, y[r], beta) for every row: the h(y, beta) term the panels then add their products to
template <class S>
__global__ void panel_init_kernel(const __grid_constant__ SpmvArgs<typename S::T> a, typename S::T *__restrict__ acc) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < a.n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t gr = a.row_offset + r;
    const T yv = a.use_y == 1 ? a.y[gr] : (a.use_y == 2 ? S::one() : S::zero());
    acc[gr] = S::epi(S::zero(), a.alpha, yv, a.beta);
  }
}

// every destination receives acc; the convergence test of the iterative drivers.  VEC: four rows per thread with 128-bit
// accesses (the caller checked that acc, x and every destination are 16-byte aligned at row 0 of the block); the scalar
// form ran at 2.2 TB/s (ncu: 360 us for 0.78 GB on config 5), a fifth of a sweep's non-panel time.
template <class S, bool VEC>
__global__ void __launch_bounds__(256) panel_finish_kernel(const __grid_constant__ SpmvArgs<typename S::T> a,
                                                           const typename S::T *__restrict__ acc) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  bool differs = false;
  if (VEC) {
    const int64_t n4 = a.n_rows >> 2;
    const int64_t n4_padded = (n4 + 31) / 32 * 32;  // whole warps reach the vote
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n4_padded; q += (int64_t)gridDim.x * blockDim.x) {
      if (q < n4) {
        const int64_t gr = a.row_offset + 4 * q;
        const uint4 o = *reinterpret_cast<const uint4 *>(acc + gr);
#pragma unroll 1
        for (int i = 0; i < a.n_out; ++i) *reinterpret_cast<uint4 *>(a.out[i] + gr) = o;
        if (a.ctl.changed) {
          const uint4 xo = *reinterpret_cast<const uint4 *>(a.x + gr);
          differs |= shb_row_differs(shb_from_bits<T>(xo.x), shb_from_bits<T>(o.x), a.ctl.delta, a.ctl.term_mode);
          differs |= shb_row_differs(shb_from_bits<T>(xo.y), shb_from_bits<T>(o.y), a.ctl.delta, a.ctl.term_mode);
          differs |= shb_row_differs(shb_from_bits<T>(xo.z), shb_from_bits<T>(o.z), a.ctl.delta, a.ctl.term_mode);
          differs |= shb_row_differs(shb_from_bits<T>(xo.w), shb_from_bits<T>(o.w), a.ctl.delta, a.ctl.term_mode);
        }
      }
    }
    // the last n_rows % 4 rows: lanes 0..2 of the first warp of block 0
    if (blockIdx.x == 0 && threadIdx.x < 32) {
      const int64_t r = 4 * n4 + threadIdx.x;
      if (r < a.n_rows) {
        const int64_t gr = a.row_offset + r;
        const T o = acc[gr];
#pragma unroll 1
        for (int i = 0; i < a.n_out; ++i) a.out[i][gr] = o;
        if (a.ctl.changed) differs |= shb_row_differs(a.x[gr], o, a.ctl.delta, a.ctl.term_mode);
      }
    }
  } else {
    const int64_t n_padded = (a.n_rows + 31) / 32 * 32;  // whole warps reach the vote
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_padded; r += (int64_t)gridDim.x * blockDim.x) {
      if (r < a.n_rows) {
        const int64_t gr = a.row_offset + r;
        const T o = acc[gr];
#pragma unroll 1
        for (int i = 0; i < a.n_out; ++i) a.out[i][gr] = o;
        if (a.ctl.changed) differs |= shb_row_differs(a.x[gr], o, a.ctl.delta, a.ctl.term_mode);
      }
    }
  }
  shb_publish_changed(a.ctl, differs);
}

template <class S>
static int panels_run(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const void *x, const void *y,
                      const void *alpha, const void *beta, void *const *outs, int n_out, int64_t row_offset, const IterCtl &ctl,
                      const void *identity_beta) {
  typedef typename S::T T;
  const uint32_t flags = cfg ? cfg->flags : 0u;
  SpmvArgs<T> a;
  memset(&a, 0, sizeof a);
  a.x = (const T *)x;
  a.y = (const T *)y;
  a.alpha = *(const T *)alpha;
  a.beta = *(const T *)beta;
  for (int i = 0; i < n_out; ++i) a.out[i] = (T *)outs[i];
  a.n_out = n_out;
  a.row_offset = row_offset;
  a.n_rows = A->n_rows;
  a.vlength = (uint32_t)A->n_cols;
  a.ctl = ctl;
  a.use_y = (flags & (SHB_F_INTERNAL_Y_ONE | SHB_F_Y_IS_ONE)) ? 2 : (flags & SHB_F_NO_Y_SKIP) ? 1 : (S::needs_y(a.beta) ? 1 : 0);
  // one destination and no convergence test: the panels accumulate in the destination itself.  y may alias out (the
  // init pass reads y[r] and writes out[r] in the same thread); x never does.
  const bool direct = n_out == 1 && ctl.changed == nullptr;
  // y and the destinations are indexed by row_offset + r: shift the scratch so that acc[row_offset + r] is its word r
  T *acc = direct ? (T *)outs[0] : reinterpret_cast<T *>((char *)A->panel_acc - 4 * row_offset);
  const unsigned grid = grid_for(A->n_rows, 256, c->prop.multiProcessorCount * 16);
  panel_init_kernel<S><<<grid, 256, 0, c->stream>>>(a, acc);
  SHB_CUDA(cudaGetLastError());
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
  lc.flags &= ~(uint32_t)(SHB_F_Y_IS_ONE | SHB_F_INTERNAL_Y_ONE);
  lc.flags |= SHB_F_NO_Y_SKIP;  // the accumulator is always read
  IterCtl cp = ctl;
  cp.changed = nullptr;  // the test runs in the finish pass
  void *acc_out[1] = {acc};
  for (int p = 0; p < A->n_panels; ++p)
    SHB_TRY(shb_launch_spmv(c, A->panels[p], semiring, &lc, x, acc, alpha, identity_beta, acc_out, 1, row_offset, cp));
  if (!direct) {
    bool vec = (reinterpret_cast<uintptr_t>(acc + row_offset) & 15u) == 0 && (reinterpret_cast<uintptr_t>(a.x + row_offset) & 15u) == 0;
    for (int i = 0; i < n_out; ++i) vec = vec && (reinterpret_cast<uintptr_t>(a.out[i] + row_offset) & 15u) == 0;
    if (vec) panel_finish_kernel<S, true><<<grid_for(A->n_rows / 4 + 32, 256, c->prop.multiProcessorCount * 16), 256, 0, c->stream>>>(a, acc);
    else panel_finish_kernel<S, false><<<grid, 256, 0, c->stream>>>(a, acc);
    SHB_CUDA(cudaGetLastError());
  }
  return SHB_OK;
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
    case SHB_PLUS_TIMES_F32:
      id.f = 1.0f;
      return panels_run<PlusTimesF32>(c, A, semiring, cfg, x, y, alpha, beta, outs, n_out, row_offset, ctl, &id);
    case SHB_MIN_PLUS_F32:
      id.f = 0.0f;
      return panels_run<MinPlusF32>(c, A, semiring, cfg, x, y, alpha, beta, outs, n_out, row_offset, ctl, &id);
    case SHB_OR_AND_I32:
      id.i = 1;
      return panels_run<OrAndI32>(c, A, semiring, cfg, x, y, alpha, beta, outs, n_out, row_offset, ctl, &id);
    case SHB_MAX_MIN_I32:
      id.i = INT_MAX;
      return panels_run<MaxMinI32>(c, A, semiring, cfg, x, y, alpha, beta, outs, n_out, row_offset, ctl, &id);
    default: shb_set_error("unknown semiring %d", semiring); return SHB_ERR_UNSUPPORTED;
  }
}
