// shb_group.cu -- multi-GPU inside the boundary: ONE process drives the N devices of a box (a parked host thread per
// device enqueues that device's share of every batch of steps; the caller's thread takes device 0).
//
// The reference is single-device (one queue on one device, inc/harness.h:78-81); the row-partitioned
// iteration is the north_star's generalisation of its host loops (app/{pr,bfs,sssp}.cpp executeRun).  This file is
// the C++ driver behind shb_group_* (include/shb.h): contexts and streams for every device, peer access instead
// of CUDA IPC, the round / level loops, and the per-iteration agreement of the devices.  It is written on top of
// the per-block primitives of the same ABI (shb_spmv_scatter, shb_bfs_block_*, shb_sssp_block_*) -- the kernels
// that do the work are theirs; what is added here is the loop and `grp_publish_kernel`:
//
//   every device launches it after its step; a thread per peer applies this device's words (add / min / max) to
//   the peer's copy with system-scope atomics over NVLink, fences, bumps the peer's arrival word; thread 0 then
//   spins until all N arrivals of this sequence number have landed locally.  One launch = all-reduce of a few
//   words + inter-device barrier, stream-ordered, no host involvement, no collective library.
//
// The host looks at device 0's copy of the reduced words once per `check_every` iterations.
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "shb_internal.cuh"

#define GRP_MAX_DEV 8
#define GRP_OP_ADD 0
#define GRP_OP_MIN 1
#define GRP_OP_MAX 2
#define GRP_WORDS 2

struct GrpPubArgs {
  int n, self, n_words;
  int op[GRP_WORDS];
  int src_is_u32[GRP_WORDS];
  const void *src[GRP_WORDS];             // local operand of each word
  long long *dst[GRP_WORDS][GRP_MAX_DEV];  // every device's copy of each word (dst[w][self] is local)
  unsigned long long *arrive[GRP_MAX_DEV]; // every device's arrival word
  unsigned long long target;               // arrivals expected locally once everyone has published
  uint32_t *status;                        // set to 1 on a barrier time-out
};

__global__ void grp_publish_kernel(const __grid_constant__ GrpPubArgs a) {
  const int t = threadIdx.x;
  if (t < a.n) {
    for (int w = 0; w < a.n_words; ++w) {
      const long long v = a.src_is_u32[w] ? (long long)*(const volatile uint32_t *)a.src[w]
                                          : *(const volatile long long *)a.src[w];
      if (a.op[w] == GRP_OP_ADD) {
        if (v != 0) atomicAdd_system((unsigned long long *)a.dst[w][t], (unsigned long long)v);
      } else if (a.op[w] == GRP_OP_MIN) {
        atomicMin_system(a.dst[w][t], v);
      } else {
        atomicMax_system(a.dst[w][t], v);
      }
    }
    __threadfence_system();
    atomicAdd_system(a.arrive[t], 1ull);
  }
  if (t == 0 && a.target) {
    const long long t0 = clock64();
    volatile unsigned long long *mine = a.arrive[a.self];
    while (*mine < a.target) {
      __nanosleep(200);
      if (clock64() - t0 > 8000000000ll) {  // ~4 s at 2 GHz: a peer died; report instead of hanging
        *a.status = 1u;
        break;
      }
    }
    __threadfence_system();
  }
}

// One parked host thread per device beyond the first.  With a single enqueueing thread eight devices x five launches
// per BFS level made the traversal launch-bound (1.25 ms at 8 GPUs against 0.49 ms with a process per GPU); the loops
// below hand every device's share of a batch of steps to its own thread -- the devices agree among themselves through
// the publish kernels, so the host threads never need to meet inside a batch.
struct GrpPool {
  std::vector<std::thread> th;
  std::mutex m;
  std::condition_variable go, done_cv;
  std::function<int(int)> job;
  uint64_t gen = 0;
  int pending = 0;
  bool stop = false;
  int st[8];
  std::string err[8];
};

struct shb_group {
  GrpPool *pool = nullptr;
  int n = 0;
  int dev[GRP_MAX_DEV];
  shb_ctx *ctx[GRP_MAX_DEV];
  cudaEvent_t ev0[GRP_MAX_DEV], ev1[GRP_MAX_DEV];
  unsigned long long *arrive[GRP_MAX_DEV];  // one arrival word per device (monotonic)
  uint32_t *status[GRP_MAX_DEV];
  unsigned long long seq = 0;
};

struct shb_gpart {
  shb_group *grp = nullptr;
  int64_t n_rows = 0, n_cols = 0, nnz = 0;
  int64_t bounds[GRP_MAX_DEV + 1];
  shb_csr *A[GRP_MAX_DEV];
  shb_csr *AT[GRP_MAX_DEV];  // == A[i] for symmetric operators
  bool symmetric = false;
};

static void grp_pool_start(shb_group *g) {
  if (g->n <= 1) return;
  GrpPool *P = new GrpPool();
  g->pool = P;
  for (int d = 1; d < g->n; ++d)
    P->th.emplace_back([P, d] {
      uint64_t seen = 0;
      for (;;) {
        std::function<int(int)> job;
        {
          std::unique_lock<std::mutex> lk(P->m);
          P->go.wait(lk, [&] { return P->stop || P->gen != seen; });
          if (P->stop) return;
          seen = P->gen;
          job = P->job;
        }
        const int st = job(d);
        std::string err = st != SHB_OK ? std::string(shb_last_error()) : std::string();
        {
          std::lock_guard<std::mutex> lk(P->m);
          P->st[d] = st;
          P->err[d] = err;
          if (--P->pending == 0) P->done_cv.notify_one();
        }
      }
    });
}

static void grp_pool_stop(shb_group *g) {
  GrpPool *P = g->pool;
  if (!P) return;
  {
    std::lock_guard<std::mutex> lk(P->m);
    P->stop = true;
  }
  P->go.notify_all();
  for (auto &t : P->th) t.join();
  delete P;
  g->pool = nullptr;
}

// f(d) for every device d, each on its own thread (device 0 on the caller's); the first failure wins
static int grp_parallel(shb_group *g, const std::function<int(int)> &f) {
  GrpPool *P = g->pool;
  if (!P) {
    for (int d = 0; d < g->n; ++d) SHB_TRY(f(d));
    return SHB_OK;
  }
  {
    std::lock_guard<std::mutex> lk(P->m);
    P->job = f;
    P->pending = g->n - 1;
    P->gen += 1;
  }
  P->go.notify_all();
  const int st0 = f(0);
  {
    std::unique_lock<std::mutex> lk(P->m);
    P->done_cv.wait(lk, [&] { return P->pending == 0; });
  }
  if (st0 != SHB_OK) return st0;
  for (int d = 1; d < g->n; ++d)
    if (P->st[d] != SHB_OK) {
      shb_set_error("%s", P->err[d].c_str());
      return P->st[d];
    }
  return SHB_OK;
}

#define GRP_CUDA SHB_CUDA

SHB_API int shb_group_create(int n_dev, const int *devices, shb_group **out) {
  SHB_REQUIRE(out && n_dev >= 1 && n_dev <= GRP_MAX_DEV, "1..8 devices");
  int count = 0;
  SHB_CUDA(cudaGetDeviceCount(&count));
  SHB_REQUIRE(count >= n_dev, "fewer CUDA devices than requested");
  shb_group *g = new shb_group();
  g->n = n_dev;
  for (int i = 0; i < n_dev; ++i) {
    g->dev[i] = devices ? devices[i] : i;
    g->ctx[i] = nullptr;
    g->arrive[i] = nullptr;
    g->status[i] = nullptr;
  }
  for (int i = 0; i < n_dev; ++i) {
    int s = shb_ctx_create(g->dev[i], &g->ctx[i]);
    if (s != SHB_OK) return s;
    GRP_CUDA(cudaSetDevice(g->dev[i]));
    for (int j = 0; j < n_dev; ++j) {
      if (j == i) continue;
      int can = 0;
      GRP_CUDA(cudaDeviceCanAccessPeer(&can, g->dev[i], g->dev[j]));
      if (!can) {
        shb_set_error("device %d cannot access device %d as a peer", g->dev[i], g->dev[j]);
        return SHB_ERR_UNSUPPORTED;
      }
      cudaError_t e = cudaDeviceEnablePeerAccess(g->dev[j], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else GRP_CUDA(e);
    }
    GRP_CUDA(cudaEventCreate(&g->ev0[i]));
    GRP_CUDA(cudaEventCreate(&g->ev1[i]));
    GRP_CUDA(cudaMalloc(&g->arrive[i], 8));
    GRP_CUDA(cudaMemset(g->arrive[i], 0, 8));
    GRP_CUDA(cudaMalloc(&g->status[i], 4));
    GRP_CUDA(cudaMemset(g->status[i], 0, 4));
  }
  if (!getenv("SHB_GROUP_ONE_THREAD")) grp_pool_start(g);  // (the variable keeps the single enqueueing thread, for A/B)
  *out = g;
  return SHB_OK;
}

SHB_API int shb_group_destroy(shb_group *g) {
  if (!g) return SHB_OK;
  grp_pool_stop(g);
  for (int i = 0; i < g->n; ++i) {
    if (!g->ctx[i]) continue;
    cudaSetDevice(g->dev[i]);
    cudaDeviceSynchronize();
    cudaEventDestroy(g->ev0[i]);
    cudaEventDestroy(g->ev1[i]);
    cudaFree(g->arrive[i]);
    cudaFree(g->status[i]);
    shb_ctx_destroy(g->ctx[i]);
  }
  delete g;
  return SHB_OK;
}

SHB_API int shb_group_size(const shb_group *g, int *n_dev) {
  SHB_REQUIRE(g && n_dev, "bad arguments");
  *n_dev = g->n;
  return SHB_OK;
}

SHB_API int shb_group_ctx(shb_group *g, int i, shb_ctx **ctx) {
  SHB_REQUIRE(g && ctx && i >= 0 && i < g->n, "bad arguments");
  *ctx = g->ctx[i];
  return SHB_OK;
}

static int grp_sync_all(shb_group *g) {
  for (int i = 0; i < g->n; ++i) SHB_TRY(shb_ctx_synchronize(g->ctx[i]));
  return SHB_OK;
}

// one publish per device: words[w] of every device receives op over all devices' src, then the barrier
// device d's publish with sequence number `seq` (1-based count of publishes on this group)
static int grp_publish_dev(shb_group *g, int d, unsigned long long seq, int n_words, const int *op, const int *src_is_u32,
                           const void *const src[][GRP_MAX_DEV], long long *const dst[][GRP_MAX_DEV]) {
  {
    GrpPubArgs a;
    memset(&a, 0, sizeof a);
    a.n = g->n;
    a.self = d;
    a.n_words = n_words;
    for (int w = 0; w < n_words; ++w) {
      a.op[w] = op[w];
      a.src_is_u32[w] = src_is_u32[w];
      a.src[w] = src[w][d];
      for (int j = 0; j < g->n; ++j) a.dst[w][j] = dst[w][j];
    }
    for (int j = 0; j < g->n; ++j) a.arrive[j] = g->arrive[j];
    a.target = (unsigned long long)g->n * seq;
    a.status = g->status[d];
    SHB_CUDA(cudaSetDevice(g->dev[d]));
    grp_publish_kernel<<<1, 32, 0, g->ctx[d]->stream>>>(a);
    SHB_CUDA(cudaGetLastError());
  }
  return SHB_OK;
}

static int grp_publish(shb_group *g, int n_words, const int *op, const int *src_is_u32, const void *const src[][GRP_MAX_DEV],
                       long long *const dst[][GRP_MAX_DEV]) {
  g->seq += 1;
  for (int d = 0; d < g->n; ++d) SHB_TRY(grp_publish_dev(g, d, g->seq, n_words, op, src_is_u32, src, dst));
  return SHB_OK;
}

static int grp_barrier(shb_group *g) { return grp_publish(g, 0, nullptr, nullptr, nullptr, nullptr); }

static int grp_check_status(shb_group *g) {
  for (int d = 0; d < g->n; ++d) {
    uint32_t st = 0;
    SHB_CUDA(cudaSetDevice(g->dev[d]));
    SHB_CUDA(cudaMemcpy(&st, g->status[d], 4, cudaMemcpyDeviceToHost));
    if (st) {
      shb_set_error("device %d timed out waiting for its peers at a group barrier", g->dev[d]);
      return SHB_ERR_CUDA;
    }
  }
  return SHB_OK;
}

// timing: events on every device between two host-synchronised points; the job's time is the slowest device's
static int grp_time_begin(shb_group *g) {
  SHB_TRY(grp_sync_all(g));
  for (int d = 0; d < g->n; ++d) {
    SHB_CUDA(cudaSetDevice(g->dev[d]));
    SHB_CUDA(cudaEventRecord(g->ev0[d], g->ctx[d]->stream));
  }
  return SHB_OK;
}
static int grp_time_end(shb_group *g, uint64_t *ns) {
  for (int d = 0; d < g->n; ++d) {
    SHB_CUDA(cudaSetDevice(g->dev[d]));
    SHB_CUDA(cudaEventRecord(g->ev1[d], g->ctx[d]->stream));
  }
  double worst = 0;
  for (int d = 0; d < g->n; ++d) {
    SHB_CUDA(cudaSetDevice(g->dev[d]));
    SHB_CUDA(cudaEventSynchronize(g->ev1[d]));
    float ms = 0.f;
    SHB_CUDA(cudaEventElapsedTime(&ms, g->ev0[d], g->ev1[d]));
    worst = std::max(worst, (double)ms);
  }
  if (ns) *ns = (uint64_t)(worst * 1.0e6 + 0.5);
  return SHB_OK;
}

// ---------------------------------------------------------------------------------------------------------
// partitions
// ---------------------------------------------------------------------------------------------------------
static void grp_balanced_bounds(int64_t n_rows, const int64_t *row_ptr, int parts, int64_t align, int64_t *bounds) {
  const int64_t nnz = row_ptr[n_rows];
  bounds[0] = 0;
  for (int k = 1; k < parts; ++k) {
    const int64_t want = (int64_t)((__int128)nnz * k / parts);
    int64_t r = std::lower_bound(row_ptr, row_ptr + n_rows + 1, want) - row_ptr;
    r = std::min(n_rows, (r + align - 1) / align * align);
    bounds[k] = std::max(r, bounds[k - 1]);
  }
  bounds[parts] = n_rows;
}

SHB_API int shb_gpart_destroy(shb_gpart *p) {
  if (!p) return SHB_OK;
  for (int i = 0; i < p->grp->n; ++i) {
    if (p->AT[i] && p->AT[i] != p->A[i]) shb_csr_destroy(p->AT[i]);
    if (p->A[i]) shb_csr_destroy(p->A[i]);
  }
  delete p;
  return SHB_OK;
}

SHB_API int shb_group_partition_host(shb_group *g, int64_t n_rows, int64_t n_cols, const int64_t *row_ptr,
                                     const int32_t *col_idx, const void *vals, int symmetric, shb_gpart **out) {
  SHB_REQUIRE(g && row_ptr && out && n_rows >= 1 && n_cols >= 1, "bad arguments");
  SHB_REQUIRE(row_ptr[n_rows] == 0 || col_idx, "col_idx is null");
  shb_gpart *p = new shb_gpart();
  p->grp = g;
  p->n_rows = n_rows;
  p->n_cols = n_cols;
  p->nnz = row_ptr[n_rows];
  p->symmetric = symmetric != 0;
  for (int i = 0; i < GRP_MAX_DEV; ++i) p->A[i] = p->AT[i] = nullptr;
  grp_balanced_bounds(n_rows, row_ptr, g->n, 32, p->bounds);
  // the transpose (out-edges of every vertex) for operators that are not symmetric: host counting sort, sliced
  // by the same vertex ownership
  std::vector<int64_t> t_ptr;
  std::vector<int32_t> t_col;
  std::vector<uint32_t> t_val;
  if (!symmetric && n_rows == n_cols) {
    const int64_t nnz = p->nnz;
    t_ptr.assign((size_t)n_cols + 1, 0);
    for (int64_t k = 0; k < nnz; ++k)
      if ((uint32_t)col_idx[k] < (uint64_t)n_cols) t_ptr[(size_t)col_idx[k] + 1]++;
    for (int64_t i = 0; i < n_cols; ++i) t_ptr[(size_t)i + 1] += t_ptr[(size_t)i];
    t_col.resize((size_t)std::max<int64_t>(t_ptr[(size_t)n_cols], 1));
    if (vals) t_val.resize(t_col.size());
    std::vector<int64_t> cur(t_ptr.begin(), t_ptr.end() - 1);
    for (int64_t r = 0; r < n_rows; ++r)
      for (int64_t k = row_ptr[r]; k < row_ptr[r + 1]; ++k) {
        if ((uint32_t)col_idx[k] >= (uint64_t)n_cols) continue;
        const int64_t q = cur[(size_t)col_idx[k]]++;
        t_col[(size_t)q] = (int32_t)r;
        if (vals) t_val[(size_t)q] = ((const uint32_t *)vals)[k];
      }
  }
  int status = SHB_OK;
  for (int i = 0; i < g->n && status == SHB_OK; ++i) {
    const int64_t lo = p->bounds[i], hi = p->bounds[i + 1];
    std::vector<int64_t> rp((size_t)(hi - lo) + 1);
    for (int64_t r = lo; r <= hi; ++r) rp[(size_t)(r - lo)] = row_ptr[r] - row_ptr[lo];
    status = shb_csr_create(g->ctx[i], hi - lo, n_cols, rp.back(), rp.data(), col_idx ? col_idx + row_ptr[lo] : nullptr,
                            vals ? (const char *)vals + 4 * row_ptr[lo] : nullptr, 4, &p->A[i], nullptr);
    if (status != SHB_OK) break;
    if (symmetric || n_rows != n_cols) {
      p->AT[i] = symmetric ? p->A[i] : nullptr;
    } else {
      for (int64_t r = lo; r <= hi; ++r) rp[(size_t)(r - lo)] = t_ptr[(size_t)r] - t_ptr[(size_t)lo];
      status = shb_csr_create(g->ctx[i], hi - lo, n_rows, rp.back(), rp.data(), t_col.data() + t_ptr[(size_t)lo],
                              vals ? (const void *)(t_val.data() + t_ptr[(size_t)lo]) : nullptr, 4, &p->AT[i], nullptr);
    }
  }
  if (status != SHB_OK) {
    shb_gpart_destroy(p);
    return status;
  }
  *out = p;
  return SHB_OK;
}

SHB_API int shb_group_adopt(shb_group *g, shb_csr *const *blocks, shb_csr *const *at_blocks, const int64_t *bounds,
                            int64_t n_cols, shb_gpart **out) {
  SHB_REQUIRE(g && blocks && bounds && out, "bad arguments");
  shb_gpart *p = new shb_gpart();
  p->grp = g;
  for (int i = 0; i < GRP_MAX_DEV; ++i) p->A[i] = p->AT[i] = nullptr;
  p->n_rows = bounds[g->n];
  p->n_cols = n_cols;
  p->symmetric = at_blocks == nullptr;
  for (int i = 0; i <= g->n; ++i) p->bounds[i] = bounds[i];
  for (int i = 0; i < g->n; ++i) {
    if (!blocks[i] || blocks[i]->device != g->dev[i] || blocks[i]->n_rows != bounds[i + 1] - bounds[i]) {
      delete p;
      shb_set_error("block %d is missing, on another device, or does not match the bounds", i);
      return SHB_ERR_INVALID;
    }
    p->A[i] = blocks[i];
    p->AT[i] = at_blocks ? at_blocks[i] : blocks[i];
    p->nnz += blocks[i]->nnz;
  }
  *out = p;
  return SHB_OK;
}

SHB_API int shb_gpart_info(const shb_gpart *p, int64_t *n_rows, int64_t *n_cols, int64_t *nnz, int64_t *bounds) {
  SHB_REQUIRE(p, "bad arguments");
  if (n_rows) *n_rows = p->n_rows;
  if (n_cols) *n_cols = p->n_cols;
  if (nnz) *nnz = p->nnz;
  if (bounds)
    for (int i = 0; i <= p->grp->n; ++i) bounds[i] = p->bounds[i];
  return SHB_OK;
}

SHB_API int shb_gpart_block(const shb_gpart *p, int i, shb_csr **A, shb_csr **AT) {
  SHB_REQUIRE(p && i >= 0 && i < p->grp->n, "bad arguments");
  if (A) *A = p->A[i];
  if (AT) *AT = p->AT[i];
  return SHB_OK;
}

// small RAII helper: buffers of one device freed on scope exit
struct GrpBufs {
  std::vector<shb_buf *> all;
  ~GrpBufs() {
    for (shb_buf *b : all) shb_buf_destroy(b);
  }
  int make(shb_ctx *c, uint64_t bytes, shb_buf **out) {
    SHB_TRY(shb_buf_create(c, std::max<uint64_t>(bytes, 8), out));
    all.push_back(*out);
    return SHB_OK;
  }
  int wrap(shb_ctx *c, void *ptr, uint64_t bytes, shb_buf **out) {
    SHB_TRY(shb_buf_wrap(c, ptr, bytes, out));
    all.push_back(*out);
    return SHB_OK;
  }
};

// ---------------------------------------------------------------------------------------------------------
// fixed point x_{k+1} = KERNEL(x_k) on the row partition (PageRank, faithful recurrences): every device keeps two
// full-length x buffers; the SpMV epilogue stores each finished out[r] into EVERY device's next x (shb_spmv_scatter:
// compute + exchange in one kernel), then one publish = MAX of the "some row changed" flags + barrier.
// y_mode 0: y0 every iteration, 1: y0 for the first launch then y = x (the reference's re-binding), 2: y == one.
// ---------------------------------------------------------------------------------------------------------
SHB_API int shb_group_iterate(shb_group *g, const shb_gpart *p, int semiring, const shb_launch_cfg *cfg, const void *x0_host,
                              uint32_t x0_fill_bits, const void *y0_host, const void *alpha, const void *beta, double delta,
                              int term_mode, int y_mode, uint32_t max_iters, uint32_t check_every, void *result_host,
                              void *prev_host, uint32_t *n_iters, uint64_t *total_ns) {
  SHB_REQUIRE(g && p && p->grp == g && alpha && beta && max_iters >= 1, "bad arguments");
  SHB_REQUIRE(p->n_rows == p->n_cols, "Matrix is not square. Failing computation.");
  check_every = std::max(1u, std::min(check_every, max_iters));
  const int N = g->n;
  const int64_t n = p->n_rows;
  GrpBufs bufs;
  shb_buf *x[GRP_MAX_DEV][2], *y0[GRP_MAX_DEV], *lf[GRP_MAX_DEV], *gw[GRP_MAX_DEV];
  shb_buf *peer_x[GRP_MAX_DEV][2][GRP_MAX_DEV];  // peer_x[d][k][j]: device j's x[k] as seen from context d
  for (int d = 0; d < N; ++d) {
    for (int k = 0; k < 2; ++k) SHB_TRY(bufs.make(g->ctx[d], (uint64_t)n * 4, &x[d][k]));
    SHB_TRY(bufs.make(g->ctx[d], (uint64_t)(max_iters + 1) * 4, &lf[d]));
    SHB_TRY(bufs.make(g->ctx[d], (uint64_t)(max_iters + 1) * 8, &gw[d]));
    SHB_TRY(shb_buf_fill_zero(g->ctx[d], lf[d], 0, (uint64_t)(max_iters + 1) * 4, nullptr));
    SHB_TRY(shb_buf_fill_zero(g->ctx[d], gw[d], 0, (uint64_t)(max_iters + 1) * 8, nullptr));
    y0[d] = nullptr;
    if (y0_host && y_mode != 2) {
      SHB_TRY(bufs.make(g->ctx[d], (uint64_t)n * 4, &y0[d]));
      SHB_TRY(shb_buf_upload(g->ctx[d], y0[d], 0, y0_host, (uint64_t)n * 4, nullptr));
    }
    if (x0_host) SHB_TRY(shb_buf_upload(g->ctx[d], x[d][0], 0, x0_host, (uint64_t)n * 4, nullptr));
    else SHB_TRY(shb_buf_fill_u32(g->ctx[d], x[d][0], 0, (uint64_t)n, x0_fill_bits));
  }
  for (int d = 0; d < N; ++d)
    for (int k = 0; k < 2; ++k)
      for (int j = 0; j < N; ++j) {
        if (j == d) peer_x[d][k][j] = x[d][k];
        else SHB_TRY(bufs.wrap(g->ctx[d], x[j][k]->ptr, (uint64_t)n * 4, &peer_x[d][k][j]));
      }
  shb_launch_cfg lc = {};
  if (cfg) lc = *cfg;
  if (y_mode == 2) lc.flags |= SHB_F_Y_IS_ONE;
  SHB_TRY(grp_time_begin(g));
  uint32_t done = 0;
  int cur = 0;
  bool converged = false;
  std::vector<long long> flags(check_every);
  while (done < max_iters && !converged) {
    const uint32_t batch = std::min(check_every, max_iters - done);
    const unsigned long long seq0 = g->seq;
    const int cur0 = cur;
    // every device's thread enqueues its `batch` steps: SpMV with the peer scatter in its epilogue, then the publish
    SHB_TRY(grp_parallel(g, [&](int d) -> int {
      for (uint32_t b = 0; b < batch; ++b) {
        const uint32_t it = done + b;
        const int c = cur0 ^ (int)(b & 1u);
        shb_buf changed, prev;
        changed = *lf[d];
        changed.ptr = (char *)lf[d]->ptr + 4 * (size_t)it;
        changed.bytes = 4;
        changed.owned = false;
        prev = *gw[d];
        prev.ptr = (char *)gw[d]->ptr + 8 * (size_t)(it ? it - 1 : 0);  // low word of the reduced flag (little endian)
        prev.bytes = 4;
        prev.owned = false;
        const shb_buf *yb = (y_mode == 0 || (y_mode == 1 && it == 0)) ? y0[d] : nullptr;
        SHB_TRY(shb_spmv_scatter(g->ctx[d], p->A[d], semiring, &lc, x[d][c], yb, alpha, beta, peer_x[d][c ^ 1], N,
                                 p->bounds[d], delta, term_mode, &changed, it ? &prev : nullptr, nullptr));
        int op[1] = {GRP_OP_MAX}, isu32[1] = {1};
        const void *src[1][GRP_MAX_DEV];
        long long *dst[1][GRP_MAX_DEV];
        for (int j = 0; j < N; ++j) {
          src[0][j] = (const char *)lf[j]->ptr + 4 * (size_t)it;
          dst[0][j] = (long long *)gw[j]->ptr + it;
        }
        SHB_TRY(grp_publish_dev(g, d, seq0 + b + 1, 1, op, isu32, src, dst));
      }
      return SHB_OK;
    }));
    g->seq = seq0 + batch;
    if (batch & 1u) cur ^= 1;
    SHB_TRY(shb_buf_download(g->ctx[0], gw[0], (uint64_t)done * 8, flags.data(), (uint64_t)batch * 8, nullptr));
    uint32_t ran = batch;
    for (uint32_t b = 0; b < batch; ++b)
      if (flags[b] == 0) {  // iteration done+b changed nothing anywhere: fixed point; later launches returned at once
        ran = b + 1;
        converged = true;
        break;
      }
    if ((batch - ran) & 1u) cur ^= 1;
    done += ran;
  }
  SHB_TRY(grp_time_end(g, total_ns));
  SHB_TRY(grp_check_status(g));
  if (result_host) SHB_TRY(shb_buf_download(g->ctx[0], x[0][cur], 0, result_host, (uint64_t)n * 4, nullptr));
  if (prev_host) SHB_TRY(shb_buf_download(g->ctx[0], x[0][cur ^ 1], 0, prev_host, (uint64_t)n * 4, nullptr));
  if (n_iters) *n_iters = done;
  return converged ? SHB_OK : SHB_ERR_NOT_CONVERGED;
}

SHB_API int shb_group_pagerank(shb_group *g, const shb_gpart *p, const shb_launch_cfg *cfg, float x_init, float beta,
                               double delta, uint32_t max_iters, uint32_t check_every, void *result_host, void *prev_host,
                               shb_graph_stats *stats) {
  SHB_REQUIRE(g && p, "bad arguments");
  uint32_t bits;
  memcpy(&bits, &x_init, 4);
  const float alpha = 1.0f;
  uint32_t iters = 0;
  uint64_t total = 0;
  // every device gathers from the full-length x: above the L2 size the sweeps run over column panels.  With
  // compressed-row panels the default slices (40 MB) are the best measured setting at every device count (config 5, ms
  // per sweep, no panels / 40 MB / 64 MB / 128 MB slices): 1 GPU 9.33 / 5.81 / 6.72 / 7.40, 2 GPUs 5.40 / 3.09 / 3.53 /
  // 4.42, 4 GPUs 2.50 / 1.85 / 2.01 / 2.42, 8 GPUs 1.44 / 1.25 / 1.28 (89 MB) / -.
  // SHB_GROUP_PANEL_BYTES=b overrides the slice size (0 = no panels): the tuning switch
  int64_t slice = 0;  // shb_csr_build_panels' default
  if (const char *e = getenv("SHB_GROUP_PANEL_BYTES")) slice = atoll(e) > 0 ? atoll(e) : -1;
  for (int d = 0; d < g->n; ++d)
    if (p->A[d]->n_panels == 0 && (uint64_t)p->n_cols * 4 > shb_panel_threshold_bytes(g->ctx[d]) &&
        (!cfg || cfg->variant == SHB_VARIANT_AUTO) && slice >= 0) {
      int n_panels = 0;  // automatic
      if (slice > 0) n_panels = (int)std::max<int64_t>(2, (p->n_cols * 4 + slice - 1) / slice);
      SHB_TRY(shb_csr_build_panels(g->ctx[d], p->A[d], n_panels, nullptr));
    }
  const int s = shb_group_iterate(g, p, SHB_PLUS_TIMES_F32, cfg, nullptr, bits, nullptr, &alpha, &beta, delta, SHB_TERM_ABSDIFF,
                                  2, max_iters, check_every, result_host, prev_host, &iters, &total);
  if (stats) {
    stats->iterations = iters;
    stats->converged = s == SHB_OK;
    stats->edges_traversed = p->nnz * (int64_t)iters;
    stats->reached = p->n_rows;
    stats->total_ns = total;
    stats->launches = (uint64_t)iters * 2 * g->n;
  }
  return s;
}

// ---------------------------------------------------------------------------------------------------------
// BFS levels: the shb_bfs_block_* kernels and their own peer publish (counts + barrier), one stream per device
// ---------------------------------------------------------------------------------------------------------
SHB_API int shb_group_bfs(shb_group *g, const shb_gpart *p, int64_t root, int32_t *levels_host, uint32_t max_levels,
                          uint32_t check_every, shb_graph_stats *stats) {
  SHB_REQUIRE(g && p && p->grp == g && root >= 0 && root < p->n_rows, "bad arguments");
  SHB_REQUIRE(p->n_rows == p->n_cols, "Matrix is not square. Failing computation.");
  const int N = g->n;
  const int64_t n = p->n_rows;
  for (int i = 1; i < N; ++i) SHB_REQUIRE(p->bounds[i] % 32 == 0, "vertex ownership must be word aligned (multiples of 32)");
  for (int i = 0; i < N; ++i) SHB_REQUIRE(p->AT[i] != nullptr, "BFS needs the out-edge blocks (AT)");
  if (max_levels == 0) max_levels = 1u << 30;
  const uint64_t n_words = (uint64_t)std::max<int64_t>((n + 31) / 32, 1);
  uint64_t cbytes = 0, sbytes = 0;
  uint32_t ring = 0;
  SHB_TRY(shb_bfs_block_counter_bytes(&cbytes, &ring));
  SHB_TRY(shb_bfs_block_state_bytes(&sbytes));
  check_every = std::max(1u, std::min(check_every, std::min(128u, ring / 2)));
  GrpBufs bufs;
  shb_buf *bm[GRP_MAX_DEV][3], *visited[GRP_MAX_DEV], *levels[GRP_MAX_DEV], *counters[GRP_MAX_DEV], *state[GRP_MAX_DEV], *deg[GRP_MAX_DEV];
  shb_buf *peer_bm[GRP_MAX_DEV][3][GRP_MAX_DEV], *peer_cnt[GRP_MAX_DEV][GRP_MAX_DEV];
  for (int d = 0; d < N; ++d) {
    for (int k = 0; k < 3; ++k) SHB_TRY(bufs.make(g->ctx[d], n_words * 4, &bm[d][k]));
    SHB_TRY(bufs.make(g->ctx[d], n_words * 4, &visited[d]));
    SHB_TRY(bufs.make(g->ctx[d], (uint64_t)std::max<int64_t>(p->A[d]->n_rows, 1) * 4, &levels[d]));
    SHB_TRY(bufs.make(g->ctx[d], cbytes, &counters[d]));
    SHB_TRY(shb_buf_fill_zero(g->ctx[d], counters[d], 0, cbytes, nullptr));
    SHB_TRY(bufs.make(g->ctx[d], sbytes, &state[d]));
    SHB_TRY(bufs.make(g->ctx[d], (uint64_t)n * 4, &deg[d]));
  }
  // push degree of every vertex (length of its AT row), full length on every device: each owner computes its slice,
  // the slices are copied to the peers
  for (int d = 0; d < N; ++d) {
    shb_buf slice = *deg[d];
    slice.ptr = (char *)deg[d]->ptr + 4 * p->bounds[d];
    slice.bytes = (uint64_t)std::max<int64_t>(p->AT[d]->n_rows, 1) * 4;
    slice.owned = false;
    if (p->AT[d]->n_rows > 0) SHB_TRY(shb_csr_row_lengths(g->ctx[d], p->AT[d], &slice));
  }
  SHB_TRY(grp_sync_all(g));
  for (int d = 0; d < N; ++d)
    for (int j = 0; j < N; ++j) {
      if (j == d || p->AT[j]->n_rows == 0) continue;
      SHB_CUDA(cudaMemcpyPeer((char *)deg[d]->ptr + 4 * p->bounds[j], g->dev[d], (char *)deg[j]->ptr + 4 * p->bounds[j], g->dev[j],
                              (size_t)p->AT[j]->n_rows * 4));
    }
  for (int d = 0; d < N; ++d)
    for (int j = 0; j < N; ++j) {
      for (int k = 0; k < 3; ++k) {
        if (j == d) peer_bm[d][k][j] = bm[d][k];
        else SHB_TRY(bufs.wrap(g->ctx[d], bm[j][k]->ptr, n_words * 4, &peer_bm[d][k][j]));
      }
      if (j == d) peer_cnt[d][j] = counters[d];
      else SHB_TRY(bufs.wrap(g->ctx[d], counters[j]->ptr, cbytes, &peer_cnt[d][j]));
    }
  uint64_t bseq = 0;  // publishes so far on these counters (their arrival word is monotonic)
  const bool peer = N > 1;
  for (int d = 0; d < N; ++d)
    SHB_TRY(shb_bfs_block_init(g->ctx[d], p->A[d], p->AT[d], p->bounds[d], n, p->nnz, root, deg[d], levels[d], bm[d][0], bm[d][1],
                               bm[d][2], visited[d], counters[d], state[d], peer ? 1 : 0));
  SHB_TRY(grp_sync_all(g));  // every device has zeroed its slots before anyone adds to them
  if (peer) {
    bseq += 1;
    for (int d = 0; d < N; ++d)
      SHB_TRY(shb_bfs_block_publish(g->ctx[d], 0, counters[d], state[d], peer_cnt[d], N, d, (uint64_t)N * bseq));
  }
  SHB_TRY(grp_time_begin(g));
  const int SLOT0 = 6;  // int64 word of counter slot 0 (include/shb.h: slot F = words 6 + 2 (F mod ring), + 1)
  uint32_t level = 0;
  bool done = false;
  int64_t reached_beyond_root = 0;
  std::vector<long long> cw(cbytes / 8);
  uint64_t launches = 0;
  while (!done && level < max_levels) {
    const uint32_t batch = std::min(check_every, max_levels - level);
    const uint64_t bseq0 = bseq;
    SHB_TRY(grp_parallel(g, [&](int d) -> int {
      for (uint32_t b = 0; b < batch; ++b) {
        const uint32_t L = level + b;
        SHB_TRY(shb_bfs_block_level(g->ctx[d], p->A[d], p->AT[d], p->bounds[d], n, L, deg[d], bm[d][L % 3], peer_bm[d][(L + 1) % 3], N, d,
                                    bm[d][(L + 2) % 3], visited[d], levels[d], counters[d], state[d], peer ? peer_cnt[d] : nullptr,
                                    peer ? (uint64_t)N * (bseq0 + b + 1) : 0));
      }
      return SHB_OK;
    }));
    if (peer) bseq += batch;
    launches += (uint64_t)batch * N;
    SHB_TRY(shb_buf_download(g->ctx[0], counters[0], 0, cw.data(), cbytes, nullptr));
    bool broke = false;
    for (uint32_t b = 0; b < batch; ++b) {
      const long long found = cw[(size_t)(SLOT0 + 2 * ((level + b + 1) % ring))];
      if (found == 0) {
        level += b + 1;
        done = true;
        broke = true;
        break;
      }
      reached_beyond_root += found;
    }
    if (!broke) level += batch;
  }
  if (!done)  // the level cap cut the traversal: the last frontier built still has to be labelled (as shb_bfs does)
    for (int d = 0; d < N; ++d)
      SHB_TRY(shb_bfs_block_label_frontier(g->ctx[d], p->A[d], p->bounds[d], n, level, bm[d][level % 3], bm[d][(level + 2) % 3],
                                           visited[d], levels[d], counters[d], state[d]));
  uint64_t total = 0;
  SHB_TRY(grp_time_end(g, &total));
  int64_t reached = 0, edges = 0;
  for (int d = 0; d < N; ++d) {
    SHB_TRY(shb_bfs_block_finish(g->ctx[d], p->A[d], levels[d], counters[d], state[d]));
    long long w[2] = {0, 0};
    SHB_TRY(shb_buf_download(g->ctx[d], counters[d], 0, w, 16, nullptr));
    reached += w[0];
    edges += w[1];
    if (levels_host && p->A[d]->n_rows > 0)
      SHB_TRY(shb_buf_download(g->ctx[d], levels[d], 0, levels_host + p->bounds[d], (uint64_t)p->A[d]->n_rows * 4, nullptr));
  }
  if (stats) {
    stats->iterations = done ? level : level + 1;
    stats->converged = done;
    stats->edges_traversed = edges;
    stats->reached = reached;
    stats->total_ns = total;
    stats->launches = launches;
  }
  (void)reached_beyond_root;
  return done ? SHB_OK : SHB_ERR_NOT_CONVERGED;
}

// ---------------------------------------------------------------------------------------------------------
// SSSP: block-asynchronous near-far in distance windows (shb_sssp_block_*); per round one publish carries the seed
// count (ADD) and the smallest pending distance (MIN), a second one is the barrier after the solves
// ---------------------------------------------------------------------------------------------------------
SHB_API int shb_group_sssp(shb_group *g, const shb_gpart *p, int64_t source, float *dist_host, float window_buckets,
                           uint32_t max_rounds, uint32_t check_every, shb_graph_stats *stats) {
  SHB_REQUIRE(g && p && p->grp == g && source >= 0 && source < p->n_rows, "bad arguments");
  SHB_REQUIRE(p->n_rows == p->n_cols, "Matrix is not square. Failing computation.");
  const int N = g->n;
  const int64_t n = p->n_rows;
  for (int i = 0; i < N; ++i) SHB_REQUIRE(p->AT[i] != nullptr, "SSSP needs the out-edge blocks (AT)");
  if (max_rounds == 0) max_rounds = 4096;
  if (window_buckets < 0.f) window_buckets = N == 1 ? 0.f : 8.f;
  check_every = std::max(1u, std::min(check_every, max_rounds));
  GrpBufs bufs;
  shb_buf *dist[GRP_MAX_DEV], *prev[GRP_MAX_DEV], *work[GRP_MAX_DEV], *lseed[GRP_MAX_DEV], *lmin[GRP_MAX_DEV], *gseed[GRP_MAX_DEV],
      *gmin[GRP_MAX_DEV], *st[GRP_MAX_DEV];
  shb_buf *peer_dist[GRP_MAX_DEV][GRP_MAX_DEV];
  const uint64_t wbytes = (uint64_t)(max_rounds + 2) * 8;
  for (int d = 0; d < N; ++d) {
    const int64_t nl = p->AT[d]->n_rows;
    uint64_t work_bytes = 0;
    SHB_TRY(shb_sssp_block_work_bytes(nl, &work_bytes));
    SHB_TRY(bufs.make(g->ctx[d], (uint64_t)n * 4, &dist[d]));
    SHB_TRY(bufs.make(g->ctx[d], (uint64_t)std::max<int64_t>(nl, 1) * 4, &prev[d]));
    SHB_TRY(bufs.make(g->ctx[d], work_bytes, &work[d]));
    SHB_TRY(bufs.make(g->ctx[d], wbytes, &lseed[d]));
    SHB_TRY(bufs.make(g->ctx[d], wbytes, &lmin[d]));
    SHB_TRY(bufs.make(g->ctx[d], wbytes, &gseed[d]));
    SHB_TRY(bufs.make(g->ctx[d], wbytes, &gmin[d]));
    SHB_TRY(bufs.make(g->ctx[d], 16, &st[d]));
    SHB_TRY(shb_buf_fill_zero(g->ctx[d], lseed[d], 0, wbytes, nullptr));
    SHB_TRY(shb_buf_fill_zero(g->ctx[d], gseed[d], 0, wbytes, nullptr));
    SHB_TRY(shb_buf_fill_u32(g->ctx[d], lmin[d], 0, wbytes / 4, 0x7F7FFFFFu));  // as int64 words: larger than any distance
    SHB_TRY(shb_buf_fill_u32(g->ctx[d], gmin[d], 0, wbytes / 4, 0x7F7FFFFFu));
    SHB_TRY(shb_sssp_block_init(g->ctx[d], p->AT[d], p->bounds[d], n, source, dist[d], prev[d], work[d]));
  }
  for (int d = 0; d < N; ++d)
    for (int j = 0; j < N; ++j) {
      if (j == d) peer_dist[d][j] = dist[d];
      else SHB_TRY(bufs.wrap(g->ctx[d], dist[j]->ptr, (uint64_t)n * 4, &peer_dist[d][j]));
    }
  SHB_TRY(grp_time_begin(g));
  uint32_t done = 0, rounds = 0;
  bool converged = false;
  std::vector<long long> seeds(check_every);
  while (done < max_rounds && !converged) {
    const uint32_t batch = std::min(check_every, max_rounds - done);
    const unsigned long long seq0 = g->seq;
    SHB_TRY(grp_parallel(g, [&](int d) -> int {
      for (uint32_t b = 0; b < batch; ++b) {
        const uint32_t r = done + b;
        SHB_TRY(shb_sssp_block_seed(g->ctx[d], p->AT[d], p->bounds[d], n, dist[d], prev[d], work[d], N > 1 ? lseed[d] : gseed[d],
                                    N > 1 ? lmin[d] : gmin[d], r));
        if (N > 1) {
          int op[2] = {GRP_OP_ADD, GRP_OP_MIN}, isu32[2] = {0, 0};
          const void *src[2][GRP_MAX_DEV];
          long long *dst[2][GRP_MAX_DEV];
          for (int j = 0; j < N; ++j) {
            src[0][j] = (const long long *)lseed[j]->ptr + r;
            src[1][j] = (const long long *)lmin[j]->ptr + r;
            dst[0][j] = (long long *)gseed[j]->ptr + r;
            dst[1][j] = (long long *)gmin[j]->ptr + r;
          }
          SHB_TRY(grp_publish_dev(g, d, seq0 + 2ull * b + 1, 2, op, isu32, src, dst));
        }
        SHB_TRY(shb_sssp_block_solve(g->ctx[d], p->AT[d], p->bounds[d], n, dist[d], peer_dist[d], N, d, p->bounds, prev[d], work[d],
                                     gmin[d], r, window_buckets, 0));
        // all peer candidates of this round have landed before anyone's next seed pass
        if (N > 1) SHB_TRY(grp_publish_dev(g, d, seq0 + 2ull * b + 2, 0, nullptr, nullptr, nullptr, nullptr));
      }
      return SHB_OK;
    }));
    if (N > 1) g->seq = seq0 + 2ull * batch;
    SHB_TRY(shb_buf_download(g->ctx[0], gseed[0], (uint64_t)done * 8, seeds.data(), (uint64_t)batch * 8, nullptr));
    uint32_t ran = batch;
    for (uint32_t b = 0; b < batch; ++b)
      if (seeds[b] == 0) {
        ran = b;
        converged = true;
        break;
      }
    rounds = done + ran;
    done += batch;
  }
  uint64_t total = 0;
  SHB_TRY(grp_time_end(g, &total));
  SHB_TRY(grp_check_status(g));
  int64_t reached = 0, edges = 0;
  for (int d = 0; d < N; ++d) {
    SHB_TRY(shb_sssp_block_finish(g->ctx[d], p->A[d], p->bounds[d], dist[d], work[d], st[d]));
    long long w[2] = {0, 0};
    SHB_TRY(shb_buf_download(g->ctx[d], st[d], 0, w, 16, nullptr));
    reached += w[0];
    edges += w[1];
    if (dist_host && p->AT[d]->n_rows > 0)
      SHB_TRY(shb_buf_download(g->ctx[d], dist[d], (uint64_t)p->bounds[d] * 4, dist_host + p->bounds[d], (uint64_t)p->AT[d]->n_rows * 4,
                               nullptr));
  }
  if (stats) {
    stats->iterations = rounds;
    stats->converged = converged;
    stats->edges_traversed = edges;
    stats->reached = reached;
    stats->total_ns = total;
    stats->launches = (uint64_t)rounds * N * 4;
  }
  return converged ? SHB_OK : SHB_ERR_NOT_CONVERGED;
}
