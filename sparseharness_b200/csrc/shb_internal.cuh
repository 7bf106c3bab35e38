// shb_internal.cuh -- shared definitions of libshb.so (not part of the public ABI, see include/shb.h)
#pragma once
// Checked build (make -C sparseharness_b200/csrc CHECKED=1 -> lib/libshb_checked.so, selected with SHB_LIB=checked):
// every index a kernel derives from operator data before it touches shared memory, a ring slot or a list is tested
// on the device; a failure prints the condition and traps the launch.  This pool keeps compute-sanitizer closed
// (profiles/r02_sanitize_closed.txt), so the checked build run over the GPU test suite is the memory-safety evidence.
#ifdef SHB_CHECKED
#include <cstdio>
#define SHB_DEV_CHECK(cond)                                                                                      \
  do {                                                                                                           \
    if (!(cond)) {                                                                                               \
      printf("SHB_DEV_CHECK failed: %s at %s:%d (block %d thread %d)\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, \
             (int)threadIdx.x);                                                                                  \
      __trap();                                                                                                  \
    }                                                                                                            \
  } while (0)
#else
#define SHB_DEV_CHECK(cond) \
  do {                      \
  } while (0)
#endif
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "../../include/shb.h"

#define SHB_API extern "C" __attribute__((visibility("default")))

void shb_set_error(const char *fmt, ...);

#define SHB_CUDA(call)                                                                              \
  do {                                                                                              \
    cudaError_t _e = (call);                                                                        \
    if (_e != cudaSuccess) {                                                                        \
      shb_set_error("%s failed: %s (code: %d) in %s at line %d", #call, cudaGetErrorString(_e),     \
                    (int)_e, __FILE__, __LINE__);                                                   \
      return _e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;                          \
    }                                                                                               \
  } while (0)

#define SHB_REQUIRE(cond, msg)                                                                      \
  do {                                                                                              \
    if (!(cond)) {                                                                                  \
      shb_set_error("invalid argument: %s (%s) in %s at line %d", msg, #cond, __FILE__, __LINE__);  \
      return SHB_ERR_INVALID;                                                                       \
    }                                                                                               \
  } while (0)

#define SHB_TRY(expr)                                                                               \
  do {                                                                                              \
    int _s = (expr);                                                                                \
    if (_s != SHB_OK) return _s;                                                                    \
  } while (0)

struct shb_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool owns_stream = true;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaDeviceProp prop;
  // small device scratch: convergence flags etc.
  uint32_t *d_flags = nullptr;  // SHB_FLAG_SLOTS words
  uint32_t *h_flags = nullptr;  // pinned mirror
  std::vector<cudaEvent_t> ev_ring;
  // grow-only device staging for the host-vector entry point (x, y, out)
  void *stage[3] = {nullptr, nullptr, nullptr};
  size_t stage_bytes[3] = {0, 0, 0};
  cudaEvent_t hv[4] = {nullptr, nullptr, nullptr, nullptr};
  // second stream + events of the host-vector entry point: result slabs go back while the next slab computes
  cudaStream_t copy_stream = nullptr, h2d_stream = nullptr;
  cudaEvent_t chunk_ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};
static const int SHB_HOST_CHUNKS = 8;
static const int SHB_FLAG_SLOTS = 1024;

struct shb_buf {
  shb_ctx *ctx = nullptr;  // creator; may be destroyed before the buffer is (destroy paths use `device` only)
  int device = 0;
  void *ptr = nullptr;
  uint64_t bytes = 0;
  bool owned = true;
  bool ipc = false;
};

// Stream-kernel tile: nnz per CTA tile
static const int SHB_TILE_NNZ = 4096;

struct shb_pb;  // blocked copy of the operator (pb.cuh)
void shb_pb_free(shb_pb *pb);
#define SHB_MAX_PANELS 16

struct shb_csr {
  shb_ctx *ctx = nullptr;  // creator; may be destroyed before the operator is (destroy uses `device` only)
  int device = 0;
  int64_t n_rows = 0, n_cols = 0, nnz = 0;
  int32_t *row_ptr = nullptr;  // n_rows + 1 (+ padding)
  int32_t *col = nullptr;      // nnz rounded up to a tile multiple, pad = -1
  void *val = nullptr;         // same capacity, pad = 0 bits; nullptr = pattern
  int elem_bytes = 4;
  // schedule metadata (stream variant)
  int32_t *tile_row = nullptr;  // n_tiles + 1: first row touched by each tile
  // compressed row list (column panels): row r of THIS operator is row row_ids[r] of the vectors; nullptr = identity
  int32_t *row_ids = nullptr;
  int64_t n_tiles = 0;
  // statistics
  int64_t max_row_len = 0, min_row_len = 0, empty_rows = 0, straddling_rows = 0;
  int auto_variant = SHB_VARIANT_VECTOR, auto_lanes = 4;
  double gather_lines = 0.0;        // distinct 128-byte lines of x per 32 consecutive entries (sampled)
  bool auto_blocked_tried = false;  // the blocked copy is built at most once by the analysis
  uint64_t device_bytes = 0;
  // carry buffers for rows straddling stream tiles (element-typed, n_tiles each)
  void *carry_tail = nullptr;
  void *carry_head = nullptr;
  shb_pb *pb = nullptr;  // built on request (shb_csr_build_blocked)
  // column panels (shb_panels.cu): sub-operators whose gathers stay inside an L2-sized slice of x; built on request
  // (shb_csr_build_panels) or by the PageRank drivers when x exceeds the L2
  shb_csr *panels[SHB_MAX_PANELS] = {};
  int n_panels = 0;
  void *panel_acc = nullptr;  // n_rows words: the row sums carried from panel to panel
  uint32_t panel_cols = 0;
  // int32 length of every row, built on first use by shb_bfs (the push degrees of Beamer's direction rule)
  mutable int32_t *row_len = nullptr;
  // bit r set <=> row r is empty (vertices a BFS pull can never find), built on first use by shb_bfs_block_init
  mutable uint32_t *empty_row_bits = nullptr;
};

struct ScopedTimer {
  shb_ctx *c;
  uint64_t *ns;
  ScopedTimer(shb_ctx *ctx, uint64_t *out) : c(ctx), ns(out) {
    if (ns) cudaEventRecord(c->ev0, c->stream);
  }
  int finish() {
    if (!ns) return SHB_OK;
    SHB_CUDA(cudaEventRecord(c->ev1, c->stream));
    SHB_CUDA(cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    SHB_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    *ns = (uint64_t)((double)ms * 1.0e6 + 0.5);
    return SHB_OK;
  }
};

static inline int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }
// grid size of a grid-stride kernel: enough blocks for n items, capped at a few waves of the 148 SMs
static inline unsigned grid_for(int64_t n, int per_block = 256, int cap = 148 * 16) {
  const int64_t b = (n + per_block - 1) / per_block;
  return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}

// implemented in shb_core.cu, used across translation units
int shb_csr_alloc(shb_ctx *ctx, int64_t n_rows, int64_t n_cols, int64_t nnz, bool has_values, shb_csr **out);
int shb_csr_analyse(shb_ctx *ctx, shb_csr *A);

// internal launch flag: the y vector is the semiring one everywhere (canonical PageRank, y == 1)
#define SHB_F_INTERNAL_Y_ONE 0x10000u

// iteration control passed by value to the SpMV kernels (all fields optional)
struct IterCtl {
  const uint32_t *prev_changed;  // if non-null and *prev_changed == 0: the fixed point was reached -> exit
  uint32_t *changed;             // set to 1 when some row fails the convergence test
  double delta;
  int term_mode;                 // SHB_TERM_*
};

// type-erased SpMV launch shared by every entry point (shb_core.cu)
int shb_launch_spmv(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const void *x, const void *y,
                    const void *alpha, const void *beta, void *const *outs, int n_out, int64_t row_offset,
                    const IterCtl &ctl);
// column panels (shb_panels.cu)
void shb_panels_free(shb_csr *A);
// x vectors above this size make the PageRank drivers build column panels: 3/4 of the device's L2 (SHB_PANEL_MIN_BYTES overrides)
static inline uint64_t shb_panel_threshold_bytes(const shb_ctx *c) {
  if (const char *e = getenv("SHB_PANEL_MIN_BYTES")) return (uint64_t)atoll(e);
  return (uint64_t)c->prop.l2CacheSize / 4 * 3;
}
int shb_launch_spmv_panels(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const void *x, const void *y,
                           const void *alpha, const void *beta, void *const *outs, int n_out, int64_t row_offset,
                           const IterCtl &ctl);
// the fixed-point loop behind shb_iterate / shb_pagerank; y_mode 0: y0 every iteration, 1: re-bind y to
// the current input after iteration 0 (reference behaviour), 2: y is the semiring one
int shb_iterate_impl(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, void *x, const void *y0,
                     const void *alpha, const void *beta, void *work, double delta, int term_mode, int y_mode,
                     uint32_t max_iters, uint32_t check_every, uint64_t *per_iter_ns, uint32_t *n_iters,
                     int *result_in_work, uint64_t *total_ns, uint64_t *launches);
