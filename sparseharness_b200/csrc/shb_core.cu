// shb_core.cu -- context, buffers, the CSR operator and the SpMV / iterate entry points of libshb.so.
// Public contract and the reference interfaces each entry point replaces: include/shb.h.
#include <stdarg.h>
#include <stdlib.h>

#include <algorithm>

#include "pb.cuh"
#include "spmv.cuh"

// ------------------------------------------------------------------------------------------------
// errors (<- inc/opencl_utils.h:15-171)
// ------------------------------------------------------------------------------------------------
static thread_local char g_err[1024] = "";

void shb_set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

SHB_API int shb_abi_version(void) { return SHB_ABI_VERSION; }
SHB_API const char *shb_last_error(void) { return g_err; }
SHB_API const char *shb_status_string(int s) {
  switch (s) {
    case SHB_OK: return "SHB_OK";
    case SHB_ERR_INVALID: return "SHB_ERR_INVALID";
    case SHB_ERR_CUDA: return "SHB_ERR_CUDA";
    case SHB_ERR_NO_DEVICE: return "SHB_ERR_NO_DEVICE";
    case SHB_ERR_OOM: return "SHB_ERR_OOM";
    case SHB_ERR_UNSUPPORTED: return "SHB_ERR_UNSUPPORTED";
    case SHB_ERR_TOO_LARGE: return "SHB_ERR_TOO_LARGE";
    case SHB_ERR_NOT_CONVERGED: return "SHB_ERR_NOT_CONVERGED";
    default: return "unknown error";
  }
}

// ------------------------------------------------------------------------------------------------
// device / context
// ------------------------------------------------------------------------------------------------
SHB_API int shb_device_count(int *count) {
  SHB_REQUIRE(count, "count is null");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  *count = n;
  return SHB_OK;
}

SHB_API int shb_device_max_alloc_bytes(int device, uint64_t *bytes) {
  SHB_REQUIRE(bytes, "bytes is null");
  int n = 0;
  shb_device_count(&n);
  if (device < 0 || device >= n) {
    shb_set_error("No CUDA devices found! (asked for device %d of %d)", device, n);
    return SHB_ERR_NO_DEVICE;
  }
  cudaDeviceProp p;
  SHB_CUDA(cudaGetDeviceProperties(&p, device));
  *bytes = (uint64_t)p.totalGlobalMem;
  return SHB_OK;
}

SHB_API int shb_device_props(int device, char *name, size_t name_cap, int *sm_count, int *cc_major, int *cc_minor,
                             uint64_t *total_mem, uint64_t *l2_bytes) {
  int n = 0;
  shb_device_count(&n);
  if (device < 0 || device >= n) {
    shb_set_error("No CUDA devices found! (asked for device %d of %d)", device, n);
    return SHB_ERR_NO_DEVICE;
  }
  cudaDeviceProp p;
  SHB_CUDA(cudaGetDeviceProperties(&p, device));
  if (name && name_cap) snprintf(name, name_cap, "%s", p.name);
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (total_mem) *total_mem = (uint64_t)p.totalGlobalMem;
  if (l2_bytes) *l2_bytes = (uint64_t)p.l2CacheSize;
  return SHB_OK;
}

SHB_API int shb_ctx_create(int device, shb_ctx **out) {
  SHB_REQUIRE(out, "ctx out is null");
  *out = nullptr;
  int n = 0;
  shb_device_count(&n);
  if (n == 0) {
    shb_set_error("No CUDA devices found!");
    return SHB_ERR_NO_DEVICE;
  }
  SHB_REQUIRE(device >= 0 && device < n, "device index out of range");
  SHB_CUDA(cudaSetDevice(device));
  shb_ctx *c = new shb_ctx();
  c->device = device;
  SHB_CUDA(cudaGetDeviceProperties(&c->prop, device));
  SHB_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  SHB_CUDA(cudaEventCreate(&c->ev0));
  SHB_CUDA(cudaEventCreate(&c->ev1));
  SHB_CUDA(cudaMalloc(&c->d_flags, SHB_FLAG_SLOTS * sizeof(uint32_t)));
  SHB_CUDA(cudaMemset(c->d_flags, 0, SHB_FLAG_SLOTS * sizeof(uint32_t)));
  SHB_CUDA(cudaMallocHost(&c->h_flags, SHB_FLAG_SLOTS * sizeof(uint32_t)));
  *out = c;
  return SHB_OK;
}

SHB_API int shb_ctx_destroy(shb_ctx *c) {
  if (!c) return SHB_OK;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  for (cudaEvent_t e : c->ev_ring) cudaEventDestroy(e);
  cudaEventDestroy(c->ev0);
  cudaEventDestroy(c->ev1);
  cudaFree(c->d_flags);
  cudaFreeHost(c->h_flags);
  for (int i = 0; i < 3; ++i) cudaFree(c->stage[i]);
  for (int i = 0; i < 4; ++i)
    if (c->hv[i]) cudaEventDestroy(c->hv[i]);
  for (int i = 0; i < 8; ++i)
    if (c->chunk_ev[i]) cudaEventDestroy(c->chunk_ev[i]);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->h2d_stream) cudaStreamDestroy(c->h2d_stream);
  if (c->owns_stream) cudaStreamDestroy(c->stream);
  delete c;
  return SHB_OK;
}

SHB_API int shb_ctx_set_stream(shb_ctx *c, void *stream) {
  SHB_REQUIRE(c, "ctx is null");
  if (c->owns_stream && c->stream) {
    cudaStreamSynchronize(c->stream);
    cudaStreamDestroy(c->stream);
  }
  c->stream = (cudaStream_t)stream;
  c->owns_stream = false;
  return SHB_OK;
}

SHB_API int shb_ctx_synchronize(shb_ctx *c) {
  SHB_REQUIRE(c, "ctx is null");
  SHB_CUDA(cudaSetDevice(c->device));
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  return SHB_OK;
}

SHB_API int shb_ctx_device_name(shb_ctx *c, char *name, size_t cap) {
  SHB_REQUIRE(c && name && cap, "bad arguments");
  snprintf(name, cap, "%s", c->prop.name);
  return SHB_OK;
}

SHB_API int shb_ctx_device(shb_ctx *c, int *device) {
  SHB_REQUIRE(c && device, "bad arguments");
  *device = c->device;
  return SHB_OK;
}

// ------------------------------------------------------------------------------------------------
// buffers (<- inc/harness.h:266-391)
// ------------------------------------------------------------------------------------------------
SHB_API int shb_buf_create(shb_ctx *c, uint64_t bytes, shb_buf **out) {
  SHB_REQUIRE(c && out, "bad arguments");
  *out = nullptr;
  SHB_CUDA(cudaSetDevice(c->device));
  shb_buf *b = new shb_buf();
  b->ctx = c;
  b->device = c->device;
  b->bytes = bytes;
  cudaError_t e = cudaMalloc(&b->ptr, bytes ? (size_t)round_up((int64_t)bytes, 256) : 256);
  if (e != cudaSuccess) {
    delete b;
    cudaGetLastError();
    shb_set_error("cudaMalloc of %llu bytes failed: %s (code: %d)", (unsigned long long)bytes, cudaGetErrorString(e),
                  (int)e);
    return e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;
  }
  *out = b;
  return SHB_OK;
}

SHB_API int shb_buf_wrap(shb_ctx *c, void *ptr, uint64_t bytes, shb_buf **out) {
  SHB_REQUIRE(c && out && ptr, "bad arguments");
  shb_buf *b = new shb_buf();
  b->ctx = c;
  b->device = c->device;
  b->ptr = ptr;
  b->bytes = bytes;
  b->owned = false;
  *out = b;
  return SHB_OK;
}

SHB_API int shb_buf_destroy(shb_buf *b) {
  if (!b) return SHB_OK;
  // the creating context may already be gone (handles are freed in any order by garbage-collected callers)
  if (b->ipc) {
    cudaSetDevice(b->device);
    cudaIpcCloseMemHandle(b->ptr);
  } else if (b->owned && b->ptr) {
    cudaSetDevice(b->device);
    cudaFree(b->ptr);
  }
  cudaGetLastError();  // a failed release must not surface as the next call's error
  delete b;
  return SHB_OK;
}

SHB_API int shb_buf_upload(shb_ctx *c, shb_buf *b, uint64_t offset, const void *host, uint64_t bytes, uint64_t *ns) {
  SHB_REQUIRE(c && b && (host || !bytes), "bad arguments");
  SHB_REQUIRE(offset + bytes <= b->bytes, "upload exceeds buffer");
  SHB_CUDA(cudaSetDevice(c->device));
  ScopedTimer t(c, ns);
  SHB_CUDA(cudaMemcpyAsync((char *)b->ptr + offset, host, bytes, cudaMemcpyHostToDevice, c->stream));
  if (!ns) SHB_CUDA(cudaStreamSynchronize(c->stream));  // blocking like CL_TRUE (inc/harness.h:311)
  return t.finish();
}

SHB_API int shb_buf_download(shb_ctx *c, const shb_buf *b, uint64_t offset, void *host, uint64_t bytes, uint64_t *ns) {
  SHB_REQUIRE(c && b && (host || !bytes), "bad arguments");
  SHB_REQUIRE(offset + bytes <= b->bytes, "download exceeds buffer");
  SHB_CUDA(cudaSetDevice(c->device));
  ScopedTimer t(c, ns);
  SHB_CUDA(cudaMemcpyAsync(host, (const char *)b->ptr + offset, bytes, cudaMemcpyDeviceToHost, c->stream));
  if (!ns) SHB_CUDA(cudaStreamSynchronize(c->stream));
  return t.finish();
}

SHB_API int shb_buf_fill_zero(shb_ctx *c, shb_buf *b, uint64_t offset, uint64_t bytes, uint64_t *ns) {
  SHB_REQUIRE(c && b, "bad arguments");
  SHB_REQUIRE(offset + bytes <= b->bytes, "fill exceeds buffer");
  SHB_CUDA(cudaSetDevice(c->device));
  ScopedTimer t(c, ns);
  SHB_CUDA(cudaMemsetAsync((char *)b->ptr + offset, 0, bytes, c->stream));
  return t.finish();
}

__global__ void fill_u32_kernel(uint32_t *p, uint64_t n, uint32_t v) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) p[i] = v;
}

SHB_API int shb_buf_fill_u32(shb_ctx *c, shb_buf *b, uint64_t offset_elems, uint64_t n, uint32_t pattern) {
  SHB_REQUIRE(c && b, "bad arguments");
  SHB_REQUIRE((offset_elems + n) * 4 <= b->bytes, "fill exceeds buffer");
  SHB_CUDA(cudaSetDevice(c->device));
  if (n == 0) return SHB_OK;
  const int blocks = (int)std::min<uint64_t>((n + 255) / 256, c->prop.multiProcessorCount * 16);
  fill_u32_kernel<<<blocks, 256, 0, c->stream>>>((uint32_t *)b->ptr + offset_elems, n, pattern);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

SHB_API int shb_buf_copy(shb_ctx *c, shb_buf *dst, const shb_buf *src, uint64_t bytes) {
  SHB_REQUIRE(c && dst && src, "bad arguments");
  SHB_REQUIRE(bytes <= dst->bytes && bytes <= src->bytes, "copy exceeds buffer");
  SHB_CUDA(cudaSetDevice(c->device));
  SHB_CUDA(cudaMemcpyAsync(dst->ptr, src->ptr, bytes, cudaMemcpyDeviceToDevice, c->stream));
  return SHB_OK;
}

SHB_API int shb_buf_size(const shb_buf *b, uint64_t *bytes) {
  SHB_REQUIRE(b && bytes, "bad arguments");
  *bytes = b->bytes;
  return SHB_OK;
}

SHB_API int shb_buf_device_ptr(const shb_buf *b, void **ptr) {
  SHB_REQUIRE(b && ptr, "bad arguments");
  *ptr = b->ptr;
  return SHB_OK;
}

SHB_API int shb_host_alloc_pinned(uint64_t bytes, void **ptr) {
  SHB_REQUIRE(ptr, "ptr is null");
  SHB_CUDA(cudaMallocHost(ptr, bytes ? bytes : 1));
  return SHB_OK;
}

SHB_API int shb_host_free_pinned(void *ptr) {
  if (ptr) SHB_CUDA(cudaFreeHost(ptr));
  return SHB_OK;
}

// ------------------------------------------------------------------------------------------------
// the CSR operator
// ------------------------------------------------------------------------------------------------
int shb_csr_alloc(shb_ctx *c, int64_t n_rows, int64_t n_cols, int64_t nnz, bool has_values, shb_csr **out) {
  SHB_REQUIRE(c && out, "bad arguments");
  *out = nullptr;
  SHB_REQUIRE(n_rows >= 0 && n_cols >= 0 && nnz >= 0, "negative size");
  if (nnz >= (int64_t)2147483647 - SHB_TILE_NNZ || n_rows >= (int64_t)2147483647 || n_cols >= (int64_t)2147483647) {
    shb_set_error("operator too large for one partition: rows=%lld cols=%lld nnz=%lld (limit 2^31-1); "
                  "row-partition it across devices",
                  (long long)n_rows, (long long)n_cols, (long long)nnz);
    return SHB_ERR_TOO_LARGE;
  }
  SHB_CUDA(cudaSetDevice(c->device));
  shb_csr *A = new shb_csr();
  A->ctx = c;
  A->device = c->device;
  A->n_rows = n_rows;
  A->n_cols = n_cols;
  A->nnz = nnz;
  const int64_t cap = std::max<int64_t>(round_up(nnz, SHB_TILE_NNZ), SHB_TILE_NNZ);
  A->n_tiles = (nnz + SHB_TILE_NNZ - 1) / SHB_TILE_NNZ;
  const size_t rp_bytes = (size_t)round_up((n_rows + 1) * 4, 256);
  size_t total = rp_bytes + (size_t)cap * 4 * (has_values ? 2 : 1);
  cudaError_t e = cudaMalloc(&A->row_ptr, rp_bytes);
  if (e == cudaSuccess) e = cudaMalloc(&A->col, (size_t)cap * 4);
  if (e == cudaSuccess && has_values) e = cudaMalloc(&A->val, (size_t)cap * 4);
  if (e == cudaSuccess) e = cudaMalloc(&A->tile_row, (size_t)(A->n_tiles + 1) * 4);
  if (e == cudaSuccess) e = cudaMalloc(&A->carry_tail, (size_t)(A->n_tiles + 1) * 4);
  if (e == cudaSuccess) e = cudaMalloc(&A->carry_head, (size_t)(A->n_tiles + 1) * 4);
  total += 3 * (size_t)(A->n_tiles + 1) * 4;
  if (e != cudaSuccess) {
    cudaGetLastError();
    shb_set_error("Attempted to allocate: %llu bytes for the operator: %s (code: %d)", (unsigned long long)total,
                  cudaGetErrorString(e), (int)e);
    shb_csr_destroy(A);
    return e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;
  }
  A->device_bytes = total;
  // padding: col = -1 (gathers ZERO), val bits = 0
  SHB_CUDA(cudaMemsetAsync(A->col + nnz, 0xff, (size_t)(cap - nnz) * 4, c->stream));
  if (has_values) SHB_CUDA(cudaMemsetAsync((int32_t *)A->val + nnz, 0, (size_t)(cap - nnz) * 4, c->stream));
  *out = A;
  return SHB_OK;
}

SHB_API int shb_csr_destroy(shb_csr *A) {
  if (!A) return SHB_OK;
  cudaSetDevice(A->device);
  cudaFree(A->row_ptr);
  cudaFree(A->col);
  cudaFree(A->val);
  cudaFree(A->tile_row);
  cudaFree(A->row_ids);
  cudaFree(A->carry_tail);
  cudaFree(A->carry_head);
  cudaFree(A->row_len);
  cudaFree(A->empty_row_bits);
  shb_pb_free(A->pb);
  shb_panels_free(A);
  cudaGetLastError();
  delete A;
  return SHB_OK;
}

// row-length statistics + tile -> first-row map, computed once per operator on the device
struct CsrStats {
  unsigned long long max_len, min_len, empty, straddle;
};

__global__ void csr_stats_kernel(const int32_t *row_ptr, int64_t n_rows, CsrStats *st) {
  unsigned long long mx = 0, mn = ~0ull, em = 0;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long len = (unsigned long long)(row_ptr[r + 1] - row_ptr[r]);
    mx = max(mx, len);
    mn = min(mn, len);
    em += (len == 0);
  }
  for (int off = 16; off > 0; off >>= 1) {
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, off));
    em += __shfl_xor_sync(0xffffffffu, em, off);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(&st->max_len, mx);
    atomicMin(&st->min_len, mn);
    atomicAdd(&st->empty, em);
  }
}

__global__ void csr_tile_rows_kernel(const int32_t *row_ptr, int64_t n_rows, int32_t nnz, int32_t *tile_row,
                                     int64_t n_tiles, CsrStats *st) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t > n_tiles) return;
  if (t == n_tiles) {
    tile_row[t] = (int32_t)(n_rows > 0 ? n_rows - 1 : 0);
    return;
  }
  const int32_t idx = (int32_t)(t * SHB_TILE_NNZ);
  // last r in [0, n_rows) with row_ptr[r] <= idx  (upper_bound - 1): the non-empty row holding idx
  int64_t lo = 0, hi = n_rows;  // invariant: row_ptr[lo] <= idx, answer in [lo, hi)
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (row_ptr[mid] <= idx) lo = mid; else hi = mid;
  }
  tile_row[t] = t == 0 ? 0 : (int32_t)lo;
  if (t >= 1) {
    const int32_t s = row_ptr[lo], e = row_ptr[lo + 1];
    if (s < idx && e > idx) atomicAdd(&st->straddle, 1ull);
  }
  (void)nnz;
}

// gather locality: distinct 128-byte lines of x touched by 32 consecutive stored entries (what one warp-wide gather
// of a row-major schedule costs in L1TEX -> XBAR requests), averaged over evenly spaced sample windows.
// ~32 = every gather its own line (uniform-random columns), ~8 = the 2-D grid.
__global__ void csr_gather_lines_kernel(const int32_t *__restrict__ col, int64_t nnz, int n_windows, unsigned long long *sum) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (warp >= n_windows) return;
  const int64_t base = (nnz - 32) / max(n_windows, 1) * warp;
  const int32_t c = col[base + lane];
  const unsigned peers = __match_any_sync(0xffffffffu, c >> 5);
  const unsigned leaders = __ballot_sync(0xffffffffu, (__ffs(peers) - 1) == lane);
  if (lane == 0) atomicAdd(sum, (unsigned long long)__popc(leaders));
}

// large, balanced, gather-bound operators: the propagation-blocked schedule (pb.cuh) is the one that gets off the
// L1TEX request ceiling; everything else keeps its row-major schedule
static bool blocked_candidate(const shb_csr *A) {
  static const bool enabled = getenv("SHB_AUTO_BLOCKED") && atoi(getenv("SHB_AUTO_BLOCKED")) != 0;  // off until measured
  if (!enabled || A->row_ids) return false;  // (a compressed-row panel is never blocked)
  const double mean = A->n_rows > 0 ? (double)A->nnz / (double)A->n_rows : 0.0;
  return A->nnz >= (1ll << 24) && A->n_cols >= (1ll << 18) && mean >= 2.0 && A->gather_lines >= 24.0 &&
         (double)A->max_row_len <= 32.0 * std::max(mean, 1.0) + 64.0;
}

static void choose_auto(shb_csr *A) {
  const double mean = A->n_rows > 0 ? (double)A->nnz / (double)A->n_rows : 0.0;
  // power-law / very skewed rows: nnz-split tiles keep every CTA equally loaded
  if (A->nnz > 0 && A->max_row_len > 2048 && (double)A->max_row_len > 64.0 * std::max(mean, 1.0)) {
    A->auto_variant = SHB_VARIANT_STREAM;
    A->auto_lanes = 0;
    return;
  }
  // short rows on a large operator: a sub-warp per row leaves most lanes idle and fetches row bounds per row; the
  // nnz-split tiles stream col / val fully coalesced whatever the row length (8192^2 grid, 4 nnz/row, min-plus sweep:
  // 5.35 TB/s against 3.5 TB/s for the 2-lane vector schedule on B200)
  if (mean < 8.0 && A->nnz >= (1 << 20)) {
    A->auto_variant = SHB_VARIANT_STREAM;
    A->auto_lanes = 0;
    return;
  }
  A->auto_variant = SHB_VARIANT_VECTOR;
  if (mean <= 3.0) A->auto_lanes = 1;
  else if (mean <= 6.0) A->auto_lanes = 2;
  else if (mean <= 24.0) A->auto_lanes = 4;
  else if (mean <= 48.0) A->auto_lanes = 8;
  else if (mean <= 96.0) A->auto_lanes = 16;
  else A->auto_lanes = 32;
}

int shb_csr_analyse(shb_ctx *c, shb_csr *A) {
  SHB_CUDA(cudaSetDevice(c->device));
  CsrStats *d_st = nullptr;
  SHB_CUDA(cudaMalloc(&d_st, sizeof(CsrStats)));
  CsrStats h = {0ull, ~0ull, 0ull, 0ull};
  SHB_CUDA(cudaMemcpyAsync(d_st, &h, sizeof h, cudaMemcpyHostToDevice, c->stream));
  if (A->n_rows > 0) {
    const int blocks = (int)std::min<int64_t>((A->n_rows + 255) / 256, c->prop.multiProcessorCount * 8);
    csr_stats_kernel<<<blocks, 256, 0, c->stream>>>(A->row_ptr, A->n_rows, d_st);
    const int64_t nt = A->n_tiles + 1;
    csr_tile_rows_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, c->stream>>>(A->row_ptr, A->n_rows, (int32_t)A->nnz,
                                                                             A->tile_row, A->n_tiles, d_st);
  }
  SHB_CUDA(cudaGetLastError());
  const int n_windows = A->nnz >= 4096 ? 2048 : 0;
  unsigned long long *d_lines = nullptr, h_lines = 0;
  if (n_windows) {
    SHB_CUDA(cudaMalloc(&d_lines, 8));
    SHB_CUDA(cudaMemsetAsync(d_lines, 0, 8, c->stream));
    csr_gather_lines_kernel<<<(n_windows * 32 + 255) / 256, 256, 0, c->stream>>>(A->col, A->nnz, n_windows, d_lines);
    SHB_CUDA(cudaGetLastError());
    SHB_CUDA(cudaMemcpyAsync(&h_lines, d_lines, 8, cudaMemcpyDeviceToHost, c->stream));
  }
  SHB_CUDA(cudaMemcpyAsync(&h, d_st, sizeof h, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  SHB_CUDA(cudaFree(d_st));
  cudaFree(d_lines);
  A->max_row_len = A->n_rows > 0 ? (int64_t)h.max_len : 0;
  A->min_row_len = A->n_rows > 0 ? (int64_t)h.min_len : 0;
  A->empty_rows = (int64_t)h.empty;
  A->straddling_rows = (int64_t)h.straddle;
  A->gather_lines = n_windows ? (double)h_lines / (double)n_windows : 0.0;
  choose_auto(A);
  // the encode step of the blocked schedule (the analogue of the reference re-encoding a matrix for its kernel,
  // inc/kernel_utils.h:35-168) happens here, once, when the statistics ask for it; an operator whose rows or columns
  // are too skewed for the layout keeps its row-major schedule
  if (!A->pb && !A->auto_blocked_tried && blocked_candidate(A)) {
    A->auto_blocked_tried = true;
    if (shb_csr_build_blocked(c, A, 0, 1) != SHB_OK) shb_set_error("%s", "");
  }
  return SHB_OK;
}

SHB_API int shb_csr_create(shb_ctx *c, int64_t n_rows, int64_t n_cols, int64_t nnz, const int64_t *row_ptr,
                           const int32_t *col_idx, const void *vals, int elem_bytes, shb_csr **out, uint64_t *upload_ns) {
  SHB_REQUIRE(c && out && row_ptr && (col_idx || nnz == 0), "bad arguments");
  SHB_REQUIRE(elem_bytes == 4, "only 4-byte elements (float / int32) are supported");
  SHB_REQUIRE(row_ptr[0] == 0 && row_ptr[n_rows] == nnz, "row_ptr does not span [0, nnz]");
  shb_csr *A = nullptr;
  SHB_TRY(shb_csr_alloc(c, n_rows, n_cols, nnz, vals != nullptr, &A));
  std::vector<int32_t> rp32((size_t)n_rows + 1);
  for (int64_t i = 0; i <= n_rows; ++i) {
    if (i > 0 && row_ptr[i] < row_ptr[i - 1]) {
      shb_csr_destroy(A);
      shb_set_error("invalid argument: row_ptr is not non-decreasing at row %lld", (long long)i);
      return SHB_ERR_INVALID;
    }
    rp32[(size_t)i] = (int32_t)row_ptr[i];
  }
  ScopedTimer t(c, upload_ns);
  SHB_CUDA(cudaMemcpyAsync(A->row_ptr, rp32.data(), (size_t)(n_rows + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  if (nnz) SHB_CUDA(cudaMemcpyAsync(A->col, col_idx, (size_t)nnz * 4, cudaMemcpyHostToDevice, c->stream));
  if (nnz && vals) SHB_CUDA(cudaMemcpyAsync(A->val, vals, (size_t)nnz * 4, cudaMemcpyHostToDevice, c->stream));
  SHB_CUDA(cudaStreamSynchronize(c->stream));  // rp32 goes out of scope
  SHB_TRY(t.finish());
  int s = shb_csr_analyse(c, A);
  if (s != SHB_OK) {
    shb_csr_destroy(A);
    return s;
  }
  *out = A;
  return SHB_OK;
}

SHB_API int shb_csr_get_info(const shb_csr *A, shb_csr_info *info) {
  SHB_REQUIRE(A && info, "bad arguments");
  memset(info, 0, sizeof *info);
  info->n_rows = A->n_rows;
  info->n_cols = A->n_cols;
  info->nnz = A->nnz;
  info->max_row_len = A->max_row_len;
  info->min_row_len = A->min_row_len;
  info->empty_rows = A->empty_rows;
  info->mean_row_len = A->n_rows ? (double)A->nnz / (double)A->n_rows : 0.0;
  info->has_values = A->val != nullptr;
  info->elem_bytes = A->elem_bytes;
  info->auto_variant = A->auto_variant;
  info->auto_lanes = A->auto_lanes;
  info->stream_tiles = A->n_tiles;
  info->straddling_rows = A->straddling_rows;
  info->device_bytes = A->device_bytes;
  info->algorithmic_bytes_spmv = (uint64_t)A->nnz * (4 + (A->val ? 4 : 0)) + (uint64_t)(A->n_rows + 1) * 4 +
                                 (uint64_t)A->n_cols * 4 + (uint64_t)A->n_rows * 4;
  return SHB_OK;
}

__global__ void rebase_rows_kernel(const int32_t *row_ptr, int64_t row_begin, int64_t count, int64_t *out) {
  const int32_t base = row_ptr[row_begin];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= count; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = (int64_t)(row_ptr[row_begin + i] - base);
}

SHB_API int shb_csr_download(shb_ctx *c, const shb_csr *A, int64_t row_begin, int64_t row_count, int64_t *row_ptr,
                             int32_t *col_idx, void *vals) {
  SHB_REQUIRE(c && A && row_ptr, "bad arguments");
  SHB_REQUIRE(row_begin >= 0 && row_count >= 0 && row_begin + row_count <= A->n_rows, "row range out of bounds");
  SHB_CUDA(cudaSetDevice(c->device));
  int64_t *d_rp = nullptr;
  SHB_CUDA(cudaMalloc(&d_rp, (size_t)(row_count + 1) * 8));
  rebase_rows_kernel<<<(unsigned)std::min<int64_t>((row_count + 256) / 256, c->prop.multiProcessorCount * 8), 256, 0, c->stream>>>(
      A->row_ptr, row_begin, row_count, d_rp);
  SHB_CUDA(cudaGetLastError());
  SHB_CUDA(cudaMemcpyAsync(row_ptr, d_rp, (size_t)(row_count + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  int32_t be[2] = {0, 0};
  SHB_CUDA(cudaMemcpyAsync(&be[0], A->row_ptr + row_begin, 4, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaMemcpyAsync(&be[1], A->row_ptr + row_begin + row_count, 4, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  SHB_CUDA(cudaFree(d_rp));
  const size_t n = (size_t)(be[1] - be[0]);
  if (col_idx && n) SHB_CUDA(cudaMemcpyAsync(col_idx, A->col + be[0], n * 4, cudaMemcpyDeviceToHost, c->stream));
  if (vals && n && A->val)
    SHB_CUDA(cudaMemcpyAsync(vals, (const int32_t *)A->val + be[0], n * 4, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  return SHB_OK;
}

// ---- 1-D row partition helpers (multi-GPU) ---------------------------------------------
__global__ void balance_rows_kernel(const int32_t *row_ptr, int64_t n_rows, int64_t nnz, int parts, int64_t *bounds) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > parts) return;
  if (k == parts) {
    bounds[k] = n_rows;
    return;
  }
  const int64_t target = (nnz * (int64_t)k) / parts;
  int64_t lo = 0, hi = n_rows;  // first r with row_ptr[r] >= target
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if ((int64_t)row_ptr[mid] < target) lo = mid + 1; else hi = mid;
  }
  bounds[k] = k == 0 ? 0 : lo;
}

SHB_API int shb_csr_balance_rows(shb_ctx *c, const shb_csr *A, int parts, int64_t *row_bounds) {
  SHB_REQUIRE(c && A && row_bounds, "bad arguments");
  SHB_REQUIRE(parts >= 1 && parts <= 1024, "parts must be in 1..1024");
  SHB_CUDA(cudaSetDevice(c->device));
  int64_t *d = nullptr;
  SHB_CUDA(cudaMalloc(&d, (size_t)(parts + 1) * 8));
  balance_rows_kernel<<<(parts + 1 + 127) / 128, 128, 0, c->stream>>>(A->row_ptr, A->n_rows, A->nnz, parts, d);
  SHB_CUDA(cudaGetLastError());
  SHB_CUDA(cudaMemcpyAsync(row_bounds, d, (size_t)(parts + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  SHB_CUDA(cudaFree(d));
  return SHB_OK;
}

__global__ void rebase_rows32_kernel(const int32_t *row_ptr, int64_t row_begin, int64_t count, int32_t *out) {
  const int32_t base = row_ptr[row_begin];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= count; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = row_ptr[row_begin + i] - base;
}

SHB_API int shb_csr_slice_rows(shb_ctx *c, const shb_csr *A, int64_t row_begin, int64_t row_count, shb_csr **out) {
  SHB_REQUIRE(c && A && out, "bad arguments");
  SHB_REQUIRE(row_begin >= 0 && row_count >= 0 && row_begin + row_count <= A->n_rows, "row range out of bounds");
  SHB_CUDA(cudaSetDevice(c->device));
  int32_t be[2] = {0, 0};
  SHB_CUDA(cudaMemcpyAsync(&be[0], A->row_ptr + row_begin, 4, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaMemcpyAsync(&be[1], A->row_ptr + row_begin + row_count, 4, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  const int64_t nnz = (int64_t)be[1] - (int64_t)be[0];
  shb_csr *S = nullptr;
  SHB_TRY(shb_csr_alloc(c, row_count, A->n_cols, nnz, A->val != nullptr, &S));
  rebase_rows32_kernel<<<(unsigned)std::min<int64_t>((row_count + 256) / 256, c->prop.multiProcessorCount * 8), 256, 0, c->stream>>>(
      A->row_ptr, row_begin, row_count, S->row_ptr);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && nnz)
    e = cudaMemcpyAsync(S->col, A->col + be[0], (size_t)nnz * 4, cudaMemcpyDeviceToDevice, c->stream);
  if (e == cudaSuccess && nnz && A->val)
    e = cudaMemcpyAsync(S->val, (const int32_t *)A->val + be[0], (size_t)nnz * 4, cudaMemcpyDeviceToDevice, c->stream);
  if (e != cudaSuccess) {
    shb_csr_destroy(S);
    shb_set_error("slicing rows failed: %s (code: %d)", cudaGetErrorString(e), (int)e);
    return SHB_ERR_CUDA;
  }
  int st = shb_csr_analyse(c, S);
  if (st != SHB_OK) {
    shb_csr_destroy(S);
    return st;
  }
  *out = S;
  return SHB_OK;
}

// ------------------------------------------------------------------------------------------------
// kernel selection + launch  (<- Harness::executeKernel, inc/harness.h:149-195)
// ------------------------------------------------------------------------------------------------
template <class S, int L, int R, bool HAS_VAL, bool HINTS>
static int launch_vector(shb_ctx *c, const SpmvArgs<typename S::T> &a, int block) {
  constexpr int rows_per_warp = (32 / L) * R;
  const int64_t warps = (a.n_rows + rows_per_warp - 1) / rows_per_warp;
  const int wpb = block / 32;
  const int64_t blocks = (warps + wpb - 1) / wpb;
  if (blocks == 0) return SHB_OK;
  SHB_REQUIRE(blocks < 2147483647LL, "grid too large");
  spmv_vector_kernel<S, L, R, HAS_VAL, HINTS><<<(unsigned)blocks, block, 0, c->stream>>>(a);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

template <class S, bool HAS_VAL>
static int dispatch_vector(shb_ctx *c, const SpmvArgs<typename S::T> &a, int lanes, int rpg, int block, bool hints) {
#define SHB_V(LL, RR)                                                            \
  if (lanes == LL && rpg == RR)                                                  \
    return hints ? launch_vector<S, LL, RR, HAS_VAL, true>(c, a, block)          \
                 : launch_vector<S, LL, RR, HAS_VAL, false>(c, a, block);
  SHB_V(1, 1) SHB_V(1, 2) SHB_V(2, 2) SHB_V(4, 1) SHB_V(4, 2) SHB_V(4, 4) SHB_V(8, 2) SHB_V(8, 4) SHB_V(16, 2)
  SHB_V(32, 1) SHB_V(32, 2)
#undef SHB_V
  shb_set_error("vector variant lanes=%d rows_per_group=%d is not instantiated", lanes, rpg);
  return SHB_ERR_UNSUPPORTED;
}

static int default_rpg(int lanes) {
  switch (lanes) {
    case 1: return 2;
    case 2: return 2;
    case 4: return 2;  // measured on B200, config 2: R=2 1.070 ms vs R=4 1.085 ms
    case 8: return 2;
    case 16: return 2;
    default: return 1;
  }
}

#ifndef SHB_STREAM_SCAN_DEFAULT
#define SHB_STREAM_SCAN_DEFAULT false  // measured: no gain
#endif
#ifndef SHB_STREAM_TMA_DEFAULT
#define SHB_STREAM_TMA_DEFAULT false  // until measured
#endif

// BLOCKED: one persistent cooperative kernel, a CTA per SM, producer and consumer warps (pb.cuh).
// SHB_PB_SPLIT=1 runs the two roles as two launches over a full-size ring (debug / profiling aid).
template <class S, bool HAS_VAL, int LQ>
static int launch_blocked(shb_ctx *c, shb_pb *pb, const SpmvArgs<typename S::T> &a) {
  typedef typename S::T T;
  const int n_groups = pb->g.n_pass * pb->g.n_panel;
  PbArgs<T> pa;
  pa.g = pb->g;
  pa.lcol = pb->lcol;
  pa.val = (const T *)pb->val;
  pa.chunk_desc = pb->chunk_desc;
  pa.delta = pb->delta;
  pa.uchunk = pb->uchunk;
  pa.slot = pb->slot;
  pa.cons_start = pb->cons_start;
  pa.cdesc = pb->cdesc;
  pa.long_bits = pb->long_bits;
  pa.partial = (T *)pb->partial;
  pa.ready = pb->sync;
  pa.consumed = pb->sync + n_groups;
  pa.exit_count = pb->sync + 2 * n_groups;
  static const char *split_env = getenv("SHB_PB_SPLIT");  // "1": both roles as two launches, "p" / "c": one role only
  const bool split = split_env != nullptr;
  if (split) {
    SHB_TRY(shb_pb_reserve_ring(c, pb, n_groups));
    pa.g.depth = n_groups;
  }
  pa.ring = (T *)pb->ring;
  const size_t smem = shb_pb_smem_bytes(pb);
  // the plain call (one destination, no y, no convergence test) has its own instantiation of the consumer
  const bool plain = a.n_out == 1 && a.use_y == 0 && a.ctl.changed == nullptr;
  const void *kernel = plain ? (const void *)pb_fused_kernel<S, HAS_VAL, LQ, true> : (const void *)pb_fused_kernel<S, HAS_VAL, LQ, false>;
  SHB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SpmvArgs<T> aa = a;
  void *args[] = {&pa, &aa};
  const dim3 grid((unsigned)pb->g.n_prod), block(PB_THREADS);
  if (!split) {
    pa.roles = PB_ROLE_PRODUCE | PB_ROLE_CONSUME;
    pa.reset_sync = 1;
    SHB_CUDA(cudaLaunchCooperativeKernel(kernel, grid, block, args, smem, c->stream));
    if (pb->n_long > 0) {  // the rows the blocked copy leaves out, whole, from the operator's own CSR arrays
      const int blocks = (int)std::min<int64_t>((pb->n_long + 7) / 8, (int64_t)c->prop.multiProcessorCount * 16);
      pb_long_rows_kernel<S, HAS_VAL><<<blocks, 256, 0, c->stream>>>(aa, pb->long_rows, (int)pb->n_long);
      SHB_CUDA(cudaGetLastError());
    }
  } else {
    if (split_env[0] != 'c') {
      pa.roles = PB_ROLE_PRODUCE;
      pa.reset_sync = split_env[0] == 'p';
      SHB_CUDA(cudaLaunchCooperativeKernel(kernel, grid, block, args, smem, c->stream));
    }
    if (split_env[0] != 'p') {
      pa.roles = PB_ROLE_CONSUME;
      pa.reset_sync = 1;
      SHB_CUDA(cudaLaunchCooperativeKernel(kernel, grid, block, args, smem, c->stream));
      if (pb->n_long > 0) {
        const int blocks = (int)std::min<int64_t>((pb->n_long + 7) / 8, (int64_t)c->prop.multiProcessorCount * 16);
        pb_long_rows_kernel<S, HAS_VAL><<<blocks, 256, 0, c->stream>>>(aa, pb->long_rows, (int)pb->n_long);
        SHB_CUDA(cudaGetLastError());
      }
    }
  }
  return SHB_OK;
}

template <class S>
static int run_blocked(shb_ctx *c, const shb_csr *A, const SpmvArgs<typename S::T> &a) {
  SHB_REQUIRE(A->pb != nullptr, "SHB_VARIANT_BLOCKED needs shb_csr_build_blocked on this operator first");
  SHB_REQUIRE(A->pb->g.n_prod == c->prop.multiProcessorCount, "blocked layout was built for another device");
  if (A->pb->g.lmax == 8) return A->pb->val ? launch_blocked<S, true, 2>(c, A->pb, a) : launch_blocked<S, false, 2>(c, A->pb, a);
  return A->pb->val ? launch_blocked<S, true, 4>(c, A->pb, a) : launch_blocked<S, false, 4>(c, A->pb, a);
}

template <class S>
static int run_spmv(shb_ctx *c, const shb_csr *A, const shb_launch_cfg *cfg, SpmvArgs<typename S::T> &a) {
  typedef typename S::T T;
  shb_launch_cfg d = {};
  if (cfg) d = *cfg;
  int variant = d.variant == SHB_VARIANT_AUTO ? A->auto_variant : d.variant;
  int lanes = d.lanes > 0 ? d.lanes : (A->auto_lanes > 0 ? A->auto_lanes : 4);
  if (A->nnz == 0) variant = SHB_VARIANT_SCALAR;
  const bool hints = !(d.flags & SHB_F_NO_CACHE_HINTS);
  int block = d.block_threads;
  if (block != 64 && block != 128 && block != 256) block = 256;  // Run.local1 outside this set: default
  a.use_y = (d.flags & (SHB_F_INTERNAL_Y_ONE | SHB_F_Y_IS_ONE)) ? 2 : (d.flags & SHB_F_NO_Y_SKIP) ? 1 : (S::needs_y(a.beta) ? 1 : 0);
  a.reversed = (d.flags & SHB_F_REVERSED_ORDER) ? 1 : 0;
  const bool has_val = A->val != nullptr;
  if (A->n_rows == 0) return SHB_OK;
  switch (variant) {
    case SHB_VARIANT_SCALAR: {
      int64_t blocks = (A->n_rows + block - 1) / block;
      if (d.grid_blocks > 0) blocks = std::min<int64_t>(blocks, d.grid_blocks);
      blocks = std::min<int64_t>(blocks, 2147483647LL);
      if (has_val) spmv_scalar_kernel<S, true><<<(unsigned)blocks, block, 0, c->stream>>>(a);
      else spmv_scalar_kernel<S, false><<<(unsigned)blocks, block, 0, c->stream>>>(a);
      SHB_CUDA(cudaGetLastError());
      return SHB_OK;
    }
    case SHB_VARIANT_VECTOR: {
      SHB_REQUIRE(lanes == 1 || lanes == 2 || lanes == 4 || lanes == 8 || lanes == 16 || lanes == 32,
                  "lanes must be 1,2,4,8,16 or 32");
      const int rpg = d.rows_per_group > 0 ? d.rows_per_group : default_rpg(lanes);
      return has_val ? dispatch_vector<S, true>(c, a, lanes, rpg, block, hints)
                     : dispatch_vector<S, false>(c, a, lanes, rpg, block, hints);
    }
    case SHB_VARIANT_STREAM: {
      const int nt = (int)A->n_tiles;
      T *ct = (T *)A->carry_tail, *ch = (T *)A->carry_head;
      // pieces longer than this leave the thread-per-row loop for a warp each (SHB_STREAM_LONG: tuning aid)
      static const int long_piece = [] {
        const char *e = getenv("SHB_STREAM_LONG");
        return e ? std::max(SHB_LONG_PIECE_MIN, std::min(4096, atoi(e))) : SHB_LONG_PIECE;
      }();
      // the merge-path row phase (spmv.cuh), meant for short skewed rows (column panels of power-law operators)
      static const int merge_env = [] {
        const char *e = getenv("SHB_STREAM_MERGE");
        return e ? atoi(e) : -1;
      }();
      // (off unless asked for: measured 2.6x SLOWER than the thread-per-row loop on the config-5 panels and on RMAT-22 --
      //  shared-memory latency chains of the walk and three more CTA barriers per tile, profiles/r02_summary.md section 2)
      bool merge = false;
      if (merge_env >= 0) merge = merge_env != 0;
      if (d.flags & SHB_F_STREAM_MERGE) merge = true;
      if (d.flags & SHB_F_STREAM_NO_MERGE) merge = false;
      // the segmented-scan row phase (spmv.cuh, MODE 2): SHB_STREAM_SCAN=0|1 overrides the choice, as do the flags
      static const int scan_env = [] {
        const char *e = getenv("SHB_STREAM_SCAN");
        return e ? atoi(e) : -1;
      }();
      bool scan = SHB_STREAM_SCAN_DEFAULT;
      if (scan_env >= 0) scan = scan_env != 0;
      if (d.flags & SHB_F_STREAM_SCAN) scan = true;
      if (d.flags & (SHB_F_STREAM_NO_MERGE | SHB_F_STREAM_MERGE)) scan = false;
      // phase 1 through the TMA unit (spmv.cuh, MODE 3): SHB_STREAM_TMA=0|1, SHB_F_STREAM_TMA
      static const int tma_env = [] {
        const char *e = getenv("SHB_STREAM_TMA");
        return e ? atoi(e) : -1;
      }();
      bool tma = SHB_STREAM_TMA_DEFAULT;
      if (tma_env >= 0) tma = tma_env != 0;
      if (d.flags & SHB_F_STREAM_TMA) tma = true;
      if (d.flags & (SHB_F_STREAM_NO_MERGE | SHB_F_STREAM_MERGE | SHB_F_STREAM_SCAN)) tma = false;
      const int mode = scan ? 2 : (merge ? 1 : (tma ? 3 : 0));
#define SHB_LAUNCH_STREAM(HV, HN)                                                                                              \
  do {                                                                                                                         \
    if (mode == 2) spmv_stream_kernel<S, HV, HN, 2><<<nt, SHB_STREAM_THREADS, 0, c->stream>>>(a, A->tile_row, nt, (int32_t)A->nnz, ct, ch, long_piece); \
    else if (mode == 3) spmv_stream_kernel<S, HV, HN, 3><<<nt, SHB_STREAM_THREADS, 0, c->stream>>>(a, A->tile_row, nt, (int32_t)A->nnz, ct, ch, long_piece); \
    else if (mode == 1) spmv_stream_kernel<S, HV, HN, 1><<<nt, SHB_STREAM_THREADS, 0, c->stream>>>(a, A->tile_row, nt, (int32_t)A->nnz, ct, ch, long_piece); \
    else spmv_stream_kernel<S, HV, HN, 0><<<nt, SHB_STREAM_THREADS, 0, c->stream>>>(a, A->tile_row, nt, (int32_t)A->nnz, ct, ch, long_piece);       \
  } while (0)
      if (has_val) {
        if (hints) SHB_LAUNCH_STREAM(true, true);
        else SHB_LAUNCH_STREAM(true, false);
      } else {
        if (hints) SHB_LAUNCH_STREAM(false, true);
        else SHB_LAUNCH_STREAM(false, false);
      }
#undef SHB_LAUNCH_STREAM
      SHB_CUDA(cudaGetLastError());
      if (A->straddling_rows > 0) {
        const int blocks = std::min((nt + 255) / 256, c->prop.multiProcessorCount * 4);
        spmv_fixup_kernel<S><<<blocks, 256, 0, c->stream>>>(a, A->tile_row, nt, (int32_t)A->nnz, ct, ch);
        SHB_CUDA(cudaGetLastError());
      }
      return SHB_OK;
    }
    case SHB_VARIANT_BLOCKED:
      return run_blocked<S>(c, A, a);
    default:
      shb_set_error("unknown kernel variant %d", variant);
      return SHB_ERR_UNSUPPORTED;
  }
}

template <class S>
static int fill_args(const shb_csr *A, const void *x, const void *y, const void *alpha, const void *beta,
                     void *const *outs, int n_out, int64_t row_offset, const IterCtl &ctl,
                     SpmvArgs<typename S::T> &a) {
  typedef typename S::T T;
  memset(&a, 0, sizeof a);
  a.row_ptr = A->row_ptr;
  a.row_ids = A->row_ids;
  a.col = A->col;
  a.val = (const T *)A->val;
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
  return SHB_OK;
}

// type-erased launch used by every entry point (spmv, iterate, pagerank, scatter)
int shb_launch_spmv(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const void *x, const void *y,
                    const void *alpha, const void *beta, void *const *outs, int n_out, int64_t row_offset,
                    const IterCtl &ctl) {
  SHB_REQUIRE(n_out >= 1 && n_out <= SHB_MAX_DST, "1..8 destinations");
  // an operator with column panels runs panel after panel unless the caller names a schedule
  if (A->n_panels > 1 && (!cfg || cfg->variant == SHB_VARIANT_AUTO) && !(cfg && (cfg->flags & SHB_F_NO_PANELS)))
    return shb_launch_spmv_panels(c, A, semiring, cfg, x, y, alpha, beta, outs, n_out, row_offset, ctl);
  switch (semiring) {
    case SHB_PLUS_TIMES_F32: {
      SpmvArgs<float> a;
      fill_args<PlusTimesF32>(A, x, y, alpha, beta, outs, n_out, row_offset, ctl, a);
      return run_spmv<PlusTimesF32>(c, A, cfg, a);
    }
    case SHB_MIN_PLUS_F32: {
      SpmvArgs<float> a;
      fill_args<MinPlusF32>(A, x, y, alpha, beta, outs, n_out, row_offset, ctl, a);
      return run_spmv<MinPlusF32>(c, A, cfg, a);
    }
    case SHB_OR_AND_I32: {
      SpmvArgs<int32_t> a;
      fill_args<OrAndI32>(A, x, y, alpha, beta, outs, n_out, row_offset, ctl, a);
      return run_spmv<OrAndI32>(c, A, cfg, a);
    }
    case SHB_MAX_MIN_I32: {
      SpmvArgs<int32_t> a;
      fill_args<MaxMinI32>(A, x, y, alpha, beta, outs, n_out, row_offset, ctl, a);
      return run_spmv<MaxMinI32>(c, A, cfg, a);
    }
    default:
      shb_set_error("unknown semiring %d", semiring);
      return SHB_ERR_UNSUPPORTED;
  }
}

static int check_vectors(const shb_csr *A, const shb_buf *x, const shb_buf *y, const shb_buf *out) {
  SHB_REQUIRE(x && out, "x / out buffer is null");
  SHB_REQUIRE(x->bytes >= (uint64_t)A->n_cols * 4, "x shorter than n_cols");
  SHB_REQUIRE(out->bytes >= (uint64_t)A->n_rows * 4, "out shorter than n_rows");
  SHB_REQUIRE(!y || y->bytes >= (uint64_t)A->n_rows * 4, "y shorter than n_rows");
  SHB_REQUIRE(out->ptr != x->ptr, "out must not alias x");
  return SHB_OK;
}

SHB_API int shb_spmv(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const shb_buf *x,
                     const shb_buf *y, const void *alpha, const void *beta, shb_buf *out, uint64_t *kernel_ns) {
  SHB_REQUIRE(c && A && alpha && beta, "bad arguments");
  SHB_TRY(check_vectors(A, x, y, out));
  SHB_CUDA(cudaSetDevice(c->device));
  const void *yp = y ? y->ptr : x->ptr;  // never dereferenced unless the epilogue needs y
  if (!y) {
    // y omitted: only legal when the semiring's epilogue ignores it for this beta
    bool needs = true;
    switch (semiring) {
      case SHB_PLUS_TIMES_F32: needs = PlusTimesF32::needs_y(*(const float *)beta); break;
      case SHB_MIN_PLUS_F32: needs = true; break;
      case SHB_OR_AND_I32: needs = OrAndI32::needs_y(*(const int32_t *)beta); break;
      case SHB_MAX_MIN_I32: needs = MaxMinI32::needs_y(*(const int32_t *)beta); break;
    }
    SHB_REQUIRE(!needs, "y is required for this semiring / beta");
  }
  IterCtl ctl = {nullptr, nullptr, 0.0, 0};
  void *outs[1] = {out->ptr};
  ScopedTimer t(c, kernel_ns);
  SHB_TRY(shb_launch_spmv(c, A, semiring, cfg, x->ptr, yp, alpha, beta, outs, 1, 0, ctl));
  return t.finish();
}

static int stage_reserve(shb_ctx *c, int slot, size_t bytes) {
  if (c->stage_bytes[slot] >= bytes) return SHB_OK;
  if (c->stage[slot]) SHB_CUDA(cudaFree(c->stage[slot]));
  c->stage[slot] = nullptr;
  c->stage_bytes[slot] = 0;
  SHB_CUDA(cudaMalloc(&c->stage[slot], bytes));
  c->stage_bytes[slot] = bytes;
  return SHB_OK;
}

SHB_API int shb_spmv_host(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const void *x_host,
                          const void *y_host, const void *alpha, const void *beta, void *out_host, uint64_t *total_ns,
                          uint64_t *kernel_ns) {
  SHB_REQUIRE(c && A && x_host && out_host && alpha && beta, "bad arguments");
  SHB_CUDA(cudaSetDevice(c->device));
  const size_t xb = (size_t)std::max<int64_t>(A->n_cols, 1) * 4, rb = (size_t)std::max<int64_t>(A->n_rows, 1) * 4;
  SHB_TRY(stage_reserve(c, 0, xb));
  SHB_TRY(stage_reserve(c, 2, rb));
  if (y_host) SHB_TRY(stage_reserve(c, 1, rb));
  for (int i = 0; i < 4; ++i)
    if (!c->hv[i]) SHB_CUDA(cudaEventCreate(&c->hv[i]));
  void *x = c->stage[0], *y = y_host ? c->stage[1] : c->stage[0], *out = c->stage[2];
  SHB_CUDA(cudaEventRecord(c->hv[0], c->stream));
  SHB_CUDA(cudaMemcpyAsync(x, x_host, (size_t)A->n_cols * 4, cudaMemcpyHostToDevice, c->stream));
  if (y_host) SHB_CUDA(cudaMemcpyAsync(y, y_host, (size_t)A->n_rows * 4, cudaMemcpyHostToDevice, c->stream));
  SHB_CUDA(cudaEventRecord(c->hv[1], c->stream));
  IterCtl ctl = {nullptr, nullptr, 0.0, 0};
  void *outs[1] = {out};
  if (!y_host) {
    bool needs = true;
    switch (semiring) {
      case SHB_PLUS_TIMES_F32: needs = PlusTimesF32::needs_y(*(const float *)beta); break;
      case SHB_OR_AND_I32: needs = OrAndI32::needs_y(*(const int32_t *)beta); break;
      case SHB_MAX_MIN_I32: needs = MaxMinI32::needs_y(*(const int32_t *)beta); break;
      default: needs = true;
    }
    SHB_REQUIRE(!needs, "y is required for this semiring / beta");
  }
  // Row-major schedules on a large operator: the rows are launched in SHB_HOST_CHUNKS slabs and each slab's result
  // goes back to the host on a second stream while the next slab computes (the D2H copy hides behind the kernel).
  shb_launch_cfg d = {};
  if (cfg) d = *cfg;
  const int variant = d.variant == SHB_VARIANT_AUTO ? A->auto_variant : d.variant;
  const bool sliceable = (variant == SHB_VARIANT_VECTOR || variant == SHB_VARIANT_SCALAR) && A->nnz > 0;
  const int n_chunks = (sliceable && A->n_rows >= (1 << 20) && !getenv("SHB_HOST_NO_OVERLAP")) ? SHB_HOST_CHUNKS : 1;
  if (n_chunks == 1) {
    SHB_TRY(shb_launch_spmv(c, A, semiring, cfg, x, y, alpha, beta, outs, 1, 0, ctl));
    SHB_CUDA(cudaEventRecord(c->hv[2], c->stream));
    SHB_CUDA(cudaMemcpyAsync(out_host, out, (size_t)A->n_rows * 4, cudaMemcpyDeviceToHost, c->stream));
    SHB_CUDA(cudaEventRecord(c->hv[3], c->stream));
  } else {
    if (!c->copy_stream) SHB_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < SHB_HOST_CHUNKS; ++i)
      if (!c->chunk_ev[i]) SHB_CUDA(cudaEventCreateWithFlags(&c->chunk_ev[i], cudaEventDisableTiming));
    const int64_t per = (A->n_rows + n_chunks - 1) / n_chunks / 32 * 32;
    for (int k = 0; k < n_chunks; ++k) {
      const int64_t r0 = per * k, r1 = k == n_chunks - 1 ? A->n_rows : per * (k + 1);
      shb_csr S = *A;  // shallow view of rows [r0, r1): row_ptr values stay absolute, outputs land at r0 + r
      S.row_ptr = A->row_ptr + r0;
      S.n_rows = r1 - r0;
      shb_launch_cfg dk = d;
      dk.variant = variant;
      if (dk.lanes <= 0) dk.lanes = A->auto_lanes;
      SHB_TRY(shb_launch_spmv(c, &S, semiring, &dk, x, y, alpha, beta, outs, 1, r0, ctl));
      SHB_CUDA(cudaEventRecord(c->chunk_ev[k], c->stream));
      SHB_CUDA(cudaStreamWaitEvent(c->copy_stream, c->chunk_ev[k], 0));
      SHB_CUDA(cudaMemcpyAsync((char *)out_host + r0 * 4, (char *)out + r0 * 4, (size_t)(r1 - r0) * 4, cudaMemcpyDeviceToHost,
                               c->copy_stream));
    }
    SHB_CUDA(cudaEventRecord(c->hv[2], c->stream));  // end of the last slab's kernel
    SHB_CUDA(cudaEventRecord(c->chunk_ev[0], c->copy_stream));
    SHB_CUDA(cudaStreamWaitEvent(c->stream, c->chunk_ev[0], 0));  // the call's stream owns the completion again
    SHB_CUDA(cudaEventRecord(c->hv[3], c->stream));
  }
  SHB_CUDA(cudaEventSynchronize(c->hv[3]));
  float ms = 0.f, kms = 0.f;
  SHB_CUDA(cudaEventElapsedTime(&ms, c->hv[0], c->hv[3]));
  SHB_CUDA(cudaEventElapsedTime(&kms, c->hv[1], c->hv[2]));
  if (total_ns) *total_ns = (uint64_t)((double)ms * 1e6 + 0.5);
  if (kernel_ns) *kernel_ns = (uint64_t)((double)kms * 1e6 + 0.5);
  return SHB_OK;
}

// A batch of host vectors through the same operator (the harness's trial loop with a fresh x per trial, or several
// right-hand sides): x_{i+1} is uploaded while SpMV i runs and out_{i-1} goes back -- three streams, double-buffered
// device staging, PCIe busy in both directions under the kernel.  Every step still moves its own x to the device
// and its own result to the host; *total_ns covers the first upload to the last download.
SHB_API int shb_spmv_host_batch(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, int n_vec,
                                const void *const *x_hosts, const void *const *y_hosts, const void *alpha, const void *beta,
                                void *const *out_hosts, uint64_t *total_ns) {
  SHB_REQUIRE(c && A && x_hosts && out_hosts && alpha && beta && n_vec >= 1, "bad arguments");
  SHB_CUDA(cudaSetDevice(c->device));
  if (!y_hosts) {
    bool needs = true;
    switch (semiring) {
      case SHB_PLUS_TIMES_F32: needs = PlusTimesF32::needs_y(*(const float *)beta); break;
      case SHB_OR_AND_I32: needs = OrAndI32::needs_y(*(const int32_t *)beta); break;
      case SHB_MAX_MIN_I32: needs = MaxMinI32::needs_y(*(const int32_t *)beta); break;
      default: needs = true;
    }
    SHB_REQUIRE(!needs, "y is required for this semiring / beta");
  }
  const size_t xb = ((size_t)std::max<int64_t>(A->n_cols, 1) * 4 + 255) / 256 * 256;
  const size_t rb = ((size_t)std::max<int64_t>(A->n_rows, 1) * 4 + 255) / 256 * 256;
  SHB_TRY(stage_reserve(c, 0, 2 * xb));
  SHB_TRY(stage_reserve(c, 2, 2 * rb));
  if (y_hosts) SHB_TRY(stage_reserve(c, 1, 2 * rb));
  if (!c->copy_stream) SHB_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
  if (!c->h2d_stream) SHB_CUDA(cudaStreamCreateWithFlags(&c->h2d_stream, cudaStreamNonBlocking));
  for (int i = 0; i < 4; ++i)
    if (!c->hv[i]) SHB_CUDA(cudaEventCreate(&c->hv[i]));
  for (int i = 0; i < 8; ++i)
    if (!c->chunk_ev[i]) SHB_CUDA(cudaEventCreateWithFlags(&c->chunk_ev[i], cudaEventDisableTiming));
  cudaEvent_t *x_ready = &c->chunk_ev[0], *k_done = &c->chunk_ev[2], *d_done = &c->chunk_ev[4];  // [2] each
  IterCtl ctl = {nullptr, nullptr, 0.0, 0};
  SHB_CUDA(cudaEventRecord(c->hv[0], c->stream));
  SHB_CUDA(cudaStreamWaitEvent(c->h2d_stream, c->hv[0], 0));  // earlier work of the context precedes the batch
  for (int i = 0; i < n_vec; ++i) {
    SHB_REQUIRE(x_hosts[i] && out_hosts[i] && (!y_hosts || y_hosts[i]), "null host vector in the batch");
    const int b = i & 1;
    char *x = (char *)c->stage[0] + b * xb, *out = (char *)c->stage[2] + b * rb;
    char *y = y_hosts ? (char *)c->stage[1] + b * rb : x;
    if (i >= 2) SHB_CUDA(cudaStreamWaitEvent(c->h2d_stream, k_done[b], 0));  // SpMV i-2 has read this x / y slot
    SHB_CUDA(cudaMemcpyAsync(x, x_hosts[i], (size_t)A->n_cols * 4, cudaMemcpyHostToDevice, c->h2d_stream));
    if (y_hosts) SHB_CUDA(cudaMemcpyAsync(y, y_hosts[i], (size_t)A->n_rows * 4, cudaMemcpyHostToDevice, c->h2d_stream));
    SHB_CUDA(cudaEventRecord(x_ready[b], c->h2d_stream));
    SHB_CUDA(cudaStreamWaitEvent(c->stream, x_ready[b], 0));
    if (i >= 2) SHB_CUDA(cudaStreamWaitEvent(c->stream, d_done[b], 0));  // result i-2 has left this out slot
    void *outs[1] = {out};
    SHB_TRY(shb_launch_spmv(c, A, semiring, cfg, x, y, alpha, beta, outs, 1, 0, ctl));
    SHB_CUDA(cudaEventRecord(k_done[b], c->stream));
    SHB_CUDA(cudaStreamWaitEvent(c->copy_stream, k_done[b], 0));
    SHB_CUDA(cudaMemcpyAsync(out_hosts[i], out, (size_t)A->n_rows * 4, cudaMemcpyDeviceToHost, c->copy_stream));
    SHB_CUDA(cudaEventRecord(d_done[b], c->copy_stream));
  }
  SHB_CUDA(cudaStreamWaitEvent(c->stream, d_done[(n_vec - 1) & 1], 0));
  if (n_vec >= 2) SHB_CUDA(cudaStreamWaitEvent(c->stream, d_done[(n_vec - 2) & 1], 0));
  SHB_CUDA(cudaEventRecord(c->hv[3], c->stream));
  SHB_CUDA(cudaEventSynchronize(c->hv[3]));
  float ms = 0.f;
  SHB_CUDA(cudaEventElapsedTime(&ms, c->hv[0], c->hv[3]));
  if (total_ns) *total_ns = (uint64_t)((double)ms * 1e6 + 0.5);
  return SHB_OK;
}

// ------------------------------------------------------------------------------------------------
// device-resident fixed-point iteration  (<- app/sssp.cpp:97-155 and siblings)
// ------------------------------------------------------------------------------------------------
SHB_API int shb_iterate(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, shb_buf *x,
                        const shb_buf *y0, const void *alpha, const void *beta, shb_buf *work, double delta,
                        int term_mode, int y_rebind, uint32_t max_iters, uint32_t check_every, uint64_t *per_iter_ns,
                        uint32_t *n_iters, int *result_in_work, uint64_t *total_ns) {
  SHB_REQUIRE(c && A && x && work && alpha && beta && n_iters && result_in_work, "bad arguments");
  SHB_REQUIRE(A->n_rows == A->n_cols, "Matrix is not square. Failing computation.");
  SHB_TRY(check_vectors(A, x, y0, work));
  return shb_iterate_impl(c, A, semiring, cfg, x->ptr, y0 ? y0->ptr : nullptr, alpha, beta, work->ptr, delta, term_mode,
                          y_rebind ? 1 : 0, max_iters, check_every, per_iter_ns, n_iters, result_in_work, total_ns,
                          nullptr);
}

int shb_iterate_impl(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, void *x_ptr,
                     const void *y0_ptr, const void *alpha, const void *beta, void *work_ptr, double delta,
                     int term_mode, int y_mode, uint32_t max_iters, uint32_t check_every, uint64_t *per_iter_ns,
                     uint32_t *n_iters, int *result_in_work, uint64_t *total_ns, uint64_t *launches) {
  SHB_REQUIRE(max_iters >= 1, "max_iters must be >= 1");
  SHB_CUDA(cudaSetDevice(c->device));
  shb_launch_cfg lcfg = {};
  if (cfg) lcfg = *cfg;
  if (y_mode == 2) lcfg.flags |= SHB_F_INTERNAL_Y_ONE;
  cfg = &lcfg;
  if (check_every == 0) check_every = 16;
  check_every = std::min<uint32_t>(check_every, SHB_FLAG_SLOTS);
  const bool timing = per_iter_ns != nullptr;
  if (timing) {
    while (c->ev_ring.size() < (size_t)check_every + 1) {
      cudaEvent_t e;
      SHB_CUDA(cudaEventCreate(&e));
      c->ev_ring.push_back(e);
    }
  }
  void *bufs[2] = {x_ptr, work_ptr};
  int in = 0;  // index of the current input
  uint32_t done = 0;
  bool converged = false;
  uint64_t n_launch = 0;
  SHB_CUDA(cudaEventRecord(c->ev0, c->stream));
  while (done < max_iters && !converged) {
    const uint32_t batch = std::min<uint32_t>(check_every, max_iters - done);
    SHB_CUDA(cudaMemsetAsync(c->d_flags, 0, batch * sizeof(uint32_t), c->stream));
    if (timing) SHB_CUDA(cudaEventRecord(c->ev_ring[0], c->stream));
    for (uint32_t b = 0; b < batch; ++b) {
      IterCtl ctl;
      ctl.prev_changed = b == 0 ? nullptr : c->d_flags + (b - 1);
      ctl.changed = c->d_flags + b;
      ctl.delta = delta;
      ctl.term_mode = term_mode;
      const void *yp = (done + b == 0 || y_mode != 1) ? (y0_ptr ? y0_ptr : bufs[in]) : bufs[in];
      void *outs[1] = {bufs[in ^ 1]};
      SHB_TRY(shb_launch_spmv(c, A, semiring, cfg, bufs[in], yp, alpha, beta, outs, 1, 0, ctl));
      ++n_launch;
      if (timing) SHB_CUDA(cudaEventRecord(c->ev_ring[b + 1], c->stream));
      in ^= 1;
    }
    SHB_CUDA(cudaMemcpyAsync(c->h_flags, c->d_flags, batch * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    SHB_CUDA(cudaStreamSynchronize(c->stream));
    // the first launch whose flag stayed 0 reached the fixed point; launches after it returned at once
    uint32_t ran = batch;
    for (uint32_t b = 0; b < batch; ++b)
      if (c->h_flags[b] == 0) {
        ran = b + 1;
        converged = true;
        break;
      }
    if (timing)
      for (uint32_t b = 0; b < ran; ++b) {
        float ms = 0.f;
        SHB_CUDA(cudaEventElapsedTime(&ms, c->ev_ring[b], c->ev_ring[b + 1]));
        per_iter_ns[done + b] = (uint64_t)((double)ms * 1e6 + 0.5);
      }
    // buffers were flipped `batch` times but only `ran` launches wrote: undo the extra flips
    if ((batch - ran) & 1u) in ^= 1;
    done += ran;
  }
  SHB_CUDA(cudaEventRecord(c->ev1, c->stream));
  SHB_CUDA(cudaEventSynchronize(c->ev1));
  if (total_ns) {
    float ms = 0.f;
    SHB_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    *total_ns = (uint64_t)((double)ms * 1e6 + 0.5);
  }
  *n_iters = done;
  if (launches) *launches = n_launch;
  *result_in_work = in;  // the last output became the "input" after the final flip
  if (!converged) {
    shb_set_error("iteration cap of %u reached without convergence", max_iters);
    return SHB_ERR_NOT_CONVERGED;
  }
  return SHB_OK;
}
