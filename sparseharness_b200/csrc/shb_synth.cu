// shb_synth.cu -- synthetic workloads generated on the device and CSR transpose.
//
// The reference only ingests Matrix Market text (src/sparse_matrix.cpp:11-70), which cannot hold
// BASELINE.json's configs 2-5 (up to 2^30 non-zeros).  These generators build the same operators
// directly in HBM.  The integer hash arithmetic is identical to oracle/shb_oracle.c
// (orc_gen_uniform / orc_rmat_edge / orc_gen_grid2d) so tests can compare bit for bit.
// cub (CUDA toolkit) is used for the one-off sorts/scans of workload construction only; it is not
// on the SpMV hot path.
#include <cub/cub.cuh>

#include <algorithm>

#include "shb_internal.cuh"

__host__ __device__ static inline uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
__device__ static inline float u24_to_unit(uint32_t u) { return __fsub_rn(__fmul_rn((float)u, 1.0f / 8388608.0f), 1.0f); }

// ---- C2 family: exactly k entries per row, uniform random columns --------------------------------
__global__ void gen_uniform_kernel(int64_t n_cols, int k, uint64_t seed, int val_mode, int64_t row_begin,
                                   int64_t row_count, int32_t *row_ptr, int32_t *col, float *val) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > row_count) return;
  row_ptr[i] = (int32_t)(i * k);
  if (i == row_count) return;
  const int64_t r = row_begin + i;
  uint64_t key[32];
  uint32_t uu[32];
  for (int j = 0; j < k; ++j) {
    const uint64_t h = splitmix64(seed ^ (uint64_t)(r * k + j));
    const uint64_t c = ((h >> 32) * (uint64_t)n_cols) >> 32;
    key[j] = (c << 5) | (uint64_t)j;
    uu[j] = (uint32_t)(h & 0xFFFFFFu);
  }
  for (int a = 1; a < k; ++a) {
    const uint64_t kk = key[a];
    int b = a - 1;
    while (b >= 0 && key[b] > kk) {
      key[b + 1] = key[b];
      --b;
    }
    key[b + 1] = kk;
  }
  for (int j = 0; j < k; ++j) {
    const int src = (int)(key[j] & 31);
    col[i * k + j] = (int32_t)(key[j] >> 5);
    const uint32_t u = uu[src];
    val[i * k + j] = val_mode == 0 ? u24_to_unit(u) : val_mode == 1 ? (float)((int)(u % 17u) - 8) : 1.0f;
  }
}

SHB_API int shb_synth_uniform(shb_ctx *c, int64_t n_rows, int64_t n_cols, int k, uint64_t seed, int val_mode,
                              int64_t row_begin, int64_t row_count, shb_csr **out) {
  SHB_REQUIRE(c && out, "bad arguments");
  SHB_REQUIRE(k >= 1 && k <= 32, "k must be in 1..32");
  SHB_REQUIRE(row_begin >= 0 && row_count >= 0 && row_begin + row_count <= n_rows, "row range out of bounds");
  shb_csr *A = nullptr;
  SHB_TRY(shb_csr_alloc(c, row_count, n_cols, row_count * k, true, &A));
  const int64_t n = row_count + 1;
  gen_uniform_kernel<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(n_cols, k, seed, val_mode, row_begin, row_count,
                                                                        A->row_ptr, A->col, (float *)A->val);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    shb_csr_destroy(A);
    shb_set_error("gen_uniform_kernel launch failed: %s", cudaGetErrorString(e));
    return SHB_ERR_CUDA;
  }
  int s = shb_csr_analyse(c, A);
  if (s != SHB_OK) {
    shb_csr_destroy(A);
    return s;
  }
  *out = A;
  return SHB_OK;
}

__global__ void gen_vector_kernel(float *x, int64_t n, uint64_t seed, int mode) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    x[i] = mode == 0 ? u24_to_unit((uint32_t)(splitmix64(seed ^ (uint64_t)i) & 0xFFFFFFu)) : 1.0f;
}

SHB_API int shb_synth_vector(shb_ctx *c, shb_buf *x, int64_t n, uint64_t seed, int mode) {
  SHB_REQUIRE(c && x, "bad arguments");
  SHB_REQUIRE((uint64_t)n * 4 <= x->bytes, "vector exceeds buffer");
  SHB_CUDA(cudaSetDevice(c->device));
  if (n == 0) return SHB_OK;
  gen_vector_kernel<<<(unsigned)std::min<int64_t>((n + 255) / 256, c->prop.multiProcessorCount * 16), 256, 0, c->stream>>>((float *)x->ptr, n, seed,
                                                                                                mode);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

// ---- RMAT (Graph500-style a,b,c,d = .57,.19,.19,.05 as 24-bit thresholds) + label scramble -----------
struct Scramble {
  uint32_t m1, m2, add, mask;
  int shift;
};
static Scramble make_scramble(int scale, uint64_t seed) {
  Scramble s;
  s.mask = scale >= 32 ? 0xFFFFFFFFu : ((1u << scale) - 1u);
  s.m1 = (uint32_t)(splitmix64(seed ^ 0xA5A5A5A5ULL) | 1ULL);
  s.m2 = (uint32_t)(splitmix64(seed ^ 0x5A5A5A5AULL) | 1ULL);
  s.add = (uint32_t)splitmix64(seed ^ 0x1234567ULL);
  s.shift = (scale + 1) / 2;
  return s;
}
__device__ static inline uint32_t scramble(uint32_t v, const Scramble &s) {
  v = (v * s.m1 + s.add) & s.mask;
  v ^= v >> s.shift;
  v = (v * s.m2) & s.mask;
  v ^= v >> s.shift;
  return v & s.mask;
}

__global__ void gen_rmat_keys_kernel(int scale, int64_t n_edges, uint64_t seed, int symmetrise, Scramble sc, uint64_t *keys) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (int64_t)gridDim.x * blockDim.x) {
    uint32_t s = 0, d = 0;
    const uint64_t base = splitmix64(seed ^ (uint64_t)e);
    for (int l = 0; l < scale; ++l) {
      const uint32_t r = (uint32_t)(splitmix64(base + (uint64_t)l) >> 40);
      uint32_t bi, bj;
      if (r < 9563013u) { bi = 0; bj = 0; }
      else if (r < 12750684u) { bi = 0; bj = 1; }
      else if (r < 15938355u) { bi = 1; bj = 0; }
      else { bi = 1; bj = 1; }
      s = (s << 1) | bi;
      d = (d << 1) | bj;
    }
    s = scramble(s, sc);
    d = scramble(d, sc);
    if (symmetrise) {
      const bool loop = s == d;
      keys[2 * e] = loop ? ~0ull : (((uint64_t)d << 32) | s);
      keys[2 * e + 1] = loop ? ~0ull : (((uint64_t)s << 32) | d);
    } else {
      keys[e] = ((uint64_t)d << 32) | s;  // pull CSR: row = destination, column = source
    }
  }
}

// row_ptr[r] = first index whose key's row >= r (keys sorted, m valid entries)
__global__ void keys_to_row_ptr_kernel(const uint64_t *keys, int64_t m, int64_t n_rows, int32_t *row_ptr) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r <= n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    int64_t lo = 0, hi = m;  // first i with (keys[i] >> 32) >= r
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if ((int64_t)(keys[mid] >> 32) < r) lo = mid + 1; else hi = mid;
    }
    row_ptr[r] = (int32_t)lo;
  }
}

__global__ void keys_to_cols_kernel(const uint64_t *keys, int64_t m, int32_t *col) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x)
    col[i] = (int32_t)(keys[i] & 0xFFFFFFFFull);
}

__global__ void count_valid_kernel(const uint64_t *keys, int64_t m, unsigned long long *count) {
  unsigned long long local = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x)
    local += keys[i] != ~0ull;
  for (int off = 16; off > 0; off >>= 1) local += __shfl_xor_sync(0xffffffffu, local, off);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(count, local);
}

__global__ void col_histogram_kernel(const int32_t *col, int64_t nnz, uint32_t *hist) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(&hist[col[i]], 1u);
}

// pagerank_normalise on an all-ones pattern (src/sparse_matrix.cpp:409-431): val = (|1| / colsum) * d,
// colsum = float count of stored entries in that column; quotient and product in double, rounded once
__global__ void pagerank_values_kernel(const int32_t *col, int64_t nnz, const uint32_t *hist, float damping, float *val) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
    const float cs = (float)hist[col[i]];
    val[i] = (float)((1.0 / (double)cs) * (double)damping);
  }
}

#define synth_grid(n) grid_for((n), 256, c->prop.multiProcessorCount * 32)  // (a context `c` is in scope at every use)

SHB_API int shb_synth_rmat(shb_ctx *c, int scale, int64_t n_edges, uint64_t seed, int symmetrise, int value_mode,
                           float damping, shb_csr **out) {
  SHB_REQUIRE(c && out, "bad arguments");
  SHB_REQUIRE(scale >= 1 && scale <= 30, "scale must be in 1..30");
  SHB_REQUIRE(n_edges >= 0, "negative edge count");
  SHB_CUDA(cudaSetDevice(c->device));
  const int64_t n = (int64_t)1 << scale;
  const int64_t cap = symmetrise ? 2 * n_edges : n_edges;
  if (cap >= 2147483647LL - SHB_TILE_NNZ) {
    shb_set_error("RMAT with %lld stored entries exceeds one partition (2^31-1)", (long long)cap);
    return SHB_ERR_TOO_LARGE;
  }
  uint64_t *keys = nullptr, *keys2 = nullptr;
  void *tmp = nullptr;
  unsigned long long *d_count = nullptr;
  shb_csr *A = nullptr;
  int status = SHB_OK;
#define RM_CUDA(call)                                                                                     \
  do {                                                                                                    \
    cudaError_t _e = (call);                                                                              \
    if (_e != cudaSuccess) {                                                                              \
      shb_set_error("%s failed: %s (code: %d)", #call, cudaGetErrorString(_e), (int)_e);                  \
      status = _e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;                              \
      goto done;                                                                                          \
    }                                                                                                     \
  } while (0)
  {
    const size_t kb = (size_t)std::max<int64_t>(cap, 1) * 8;
    RM_CUDA(cudaMalloc(&keys, kb));
    RM_CUDA(cudaMalloc(&keys2, kb));
    RM_CUDA(cudaMalloc(&d_count, 16));
    RM_CUDA(cudaMemsetAsync(d_count, 0, 16, c->stream));
    const Scramble sc = make_scramble(scale, seed);
    if (n_edges > 0) gen_rmat_keys_kernel<<<synth_grid(n_edges), 256, 0, c->stream>>>(scale, n_edges, seed, symmetrise, sc, keys);
    RM_CUDA(cudaGetLastError());
    // sort by (row, col); self loops (all-ones keys) go to the end
    size_t tb = 0;
    cub::DoubleBuffer<uint64_t> db(keys, keys2);
    RM_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tb, db, (int64_t)cap, 0, 64, c->stream));
    RM_CUDA(cudaMalloc(&tmp, std::max<size_t>(tb, 256)));
    RM_CUDA(cub::DeviceRadixSort::SortKeys(tmp, tb, db, (int64_t)cap, 0, 64, c->stream));
    uint64_t *sorted = db.Current();
    uint64_t *other = db.Alternate();
    int64_t m = cap;
    if (symmetrise) {
      // de-duplicate, then drop the self-loop sentinel
      size_t tb2 = 0;
      RM_CUDA(cub::DeviceSelect::Unique(nullptr, tb2, sorted, other, (int64_t *)d_count, (int64_t)cap, c->stream));
      if (tb2 > tb) {
        RM_CUDA(cudaFree(tmp));
        tmp = nullptr;
        RM_CUDA(cudaMalloc(&tmp, tb2));
      }
      RM_CUDA(cub::DeviceSelect::Unique(tmp, tb2, sorted, other, (int64_t *)d_count, (int64_t)cap, c->stream));
      int64_t uniq = 0;
      RM_CUDA(cudaMemcpyAsync(&uniq, d_count, 8, cudaMemcpyDeviceToHost, c->stream));
      RM_CUDA(cudaStreamSynchronize(c->stream));
      std::swap(sorted, other);
      RM_CUDA(cudaMemsetAsync(d_count, 0, 16, c->stream));
      if (uniq > 0) count_valid_kernel<<<synth_grid(uniq), 256, 0, c->stream>>>(sorted, uniq, d_count);
      RM_CUDA(cudaGetLastError());
      unsigned long long valid = 0;
      RM_CUDA(cudaMemcpyAsync(&valid, d_count, 8, cudaMemcpyDeviceToHost, c->stream));
      RM_CUDA(cudaStreamSynchronize(c->stream));
      m = (int64_t)valid;  // the sentinel, if present, is the single last unique key
    }
    status = shb_csr_alloc(c, n, n, m, value_mode == 1, &A);
    if (status != SHB_OK) goto done;
    keys_to_row_ptr_kernel<<<synth_grid(n + 1), 256, 0, c->stream>>>(sorted, m, n, A->row_ptr);
    if (m > 0) keys_to_cols_kernel<<<synth_grid(m), 256, 0, c->stream>>>(sorted, m, A->col);
    RM_CUDA(cudaGetLastError());
    if (value_mode == 1 && m > 0) {
      uint32_t *hist = (uint32_t *)other;  // reuse: n * 4 bytes <= cap * 8
      if ((size_t)n * 4 > kb) {
        shb_set_error("internal: histogram scratch too small");
        status = SHB_ERR_INVALID;
        goto done;
      }
      RM_CUDA(cudaMemsetAsync(hist, 0, (size_t)n * 4, c->stream));
      col_histogram_kernel<<<synth_grid(m), 256, 0, c->stream>>>(A->col, m, hist);
      pagerank_values_kernel<<<synth_grid(m), 256, 0, c->stream>>>(A->col, m, hist, damping, (float *)A->val);
      RM_CUDA(cudaGetLastError());
    }
    RM_CUDA(cudaStreamSynchronize(c->stream));
    status = shb_csr_analyse(c, A);
  }
done:
  cudaFree(keys);
  cudaFree(keys2);
  cudaFree(tmp);
  cudaFree(d_count);
  if (status != SHB_OK) {
    shb_csr_destroy(A);
    return status;
  }
  *out = A;
  return SHB_OK;
#undef RM_CUDA
}

// ---- 2-D grid (C4) -------------------------------------------------------------------------------------
__device__ static inline float grid_weight(uint64_t lo, int dir, uint64_t seed) {
  return (float)(1u + (uint32_t)(splitmix64(seed ^ (lo * 2u + (uint64_t)dir)) % 255u));
}

__global__ void gen_grid_kernel(int64_t W, int64_t H, uint64_t seed, int32_t *row_ptr, int32_t *col, float *val) {
  const int64_t n = W * H;
  for (int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u <= n; u += (int64_t)gridDim.x * blockDim.x) {
    const int64_t y = u / W, x = u % W;
    // entries of the complete grid rows above (row y' holds 4W-2 entries, minus W on the top and on
    // the bottom row) + entries of the vertices to the left in this grid row; u == n gives nnz
    int64_t p = y * (4 * W - 2) - (y > 0 ? W : 0) - (y > H - 1 ? W : 0);
    if (u == n) {
      row_ptr[u] = (int32_t)p;
      continue;
    }
    const int top = y == 0, bottom = y == H - 1;
    p += x * (4 - top - bottom) - (x > 0 ? 1 : 0);
    row_ptr[u] = (int32_t)p;
    if (y > 0) { col[p] = (int32_t)(u - W); val[p] = grid_weight((uint64_t)(u - W), 1, seed); ++p; }
    if (x > 0) { col[p] = (int32_t)(u - 1); val[p] = grid_weight((uint64_t)(u - 1), 0, seed); ++p; }
    if (x + 1 < W) { col[p] = (int32_t)(u + 1); val[p] = grid_weight((uint64_t)u, 0, seed); ++p; }
    if (y + 1 < H) { col[p] = (int32_t)(u + W); val[p] = grid_weight((uint64_t)u, 1, seed); ++p; }
  }
}

SHB_API int shb_synth_grid2d(shb_ctx *c, int64_t W, int64_t H, uint64_t seed, shb_csr **out) {
  SHB_REQUIRE(c && out, "bad arguments");
  SHB_REQUIRE(W >= 2 && H >= 2, "grid must be at least 2x2");
  const int64_t n = W * H;
  const int64_t nnz = 4 * n - 2 * W - 2 * H;
  shb_csr *A = nullptr;
  SHB_TRY(shb_csr_alloc(c, n, n, nnz, true, &A));
  gen_grid_kernel<<<synth_grid(n + 1), 256, 0, c->stream>>>(W, H, seed, A->row_ptr, A->col, (float *)A->val);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    shb_csr_destroy(A);
    shb_set_error("gen_grid_kernel launch failed: %s", cudaGetErrorString(e));
    return SHB_ERR_CUDA;
  }
  int s = shb_csr_analyse(c, A);
  if (s != SHB_OK) {
    shb_csr_destroy(A);
    return s;
  }
  *out = A;
  return SHB_OK;
}

// ---- transpose: out-adjacency for push steps (deterministic: sort by (col,row)) ----------------------
__global__ void transpose_keys_kernel(const int32_t *row_ptr, const int32_t *col, int64_t n_rows, uint64_t *keys,
                                      uint32_t *src_idx) {
  // one warp per row keeps the writes coalesced for long rows
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n_rows; r += n_warps) {
    const int s = row_ptr[r], e = row_ptr[r + 1];
    for (int k = s + lane; k < e; k += 32) {
      keys[k] = ((uint64_t)(uint32_t)col[k] << 32) | (uint64_t)(uint32_t)r;
      src_idx[k] = (uint32_t)k;
    }
  }
}

__global__ void permute_vals_kernel(const uint32_t *src_idx, const uint32_t *val_in, int64_t nnz, uint32_t *val_out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x)
    val_out[i] = val_in[src_idx[i]];
}

SHB_API int shb_csr_transpose(shb_ctx *c, const shb_csr *A, shb_csr **out) {
  SHB_REQUIRE(c && A && out, "bad arguments");
  SHB_CUDA(cudaSetDevice(c->device));
  shb_csr *T = nullptr;
  SHB_TRY(shb_csr_alloc(c, A->n_cols, A->n_rows, A->nnz, A->val != nullptr, &T));
  const int64_t nnz = A->nnz;
  uint64_t *k1 = nullptr, *k2 = nullptr;
  uint32_t *i1 = nullptr, *i2 = nullptr;
  void *tmp = nullptr;
  int status = SHB_OK;
#define TR_CUDA(call)                                                                                     \
  do {                                                                                                    \
    cudaError_t _e = (call);                                                                              \
    if (_e != cudaSuccess) {                                                                              \
      shb_set_error("%s failed: %s (code: %d)", #call, cudaGetErrorString(_e), (int)_e);                  \
      status = _e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;                              \
      goto done;                                                                                          \
    }                                                                                                     \
  } while (0)
  {
    const size_t nn = (size_t)std::max<int64_t>(nnz, 1);
    TR_CUDA(cudaMalloc(&k1, nn * 8));
    TR_CUDA(cudaMalloc(&k2, nn * 8));
    TR_CUDA(cudaMalloc(&i1, nn * 4));
    TR_CUDA(cudaMalloc(&i2, nn * 4));
    if (nnz > 0 && A->n_rows > 0)
      transpose_keys_kernel<<<synth_grid(A->n_rows * 32), 256, 0, c->stream>>>(A->row_ptr, A->col, A->n_rows, k1, i1);
    TR_CUDA(cudaGetLastError());
    cub::DoubleBuffer<uint64_t> dk(k1, k2);
    cub::DoubleBuffer<uint32_t> dv(i1, i2);
    size_t tb = 0;
    TR_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (int64_t)nnz, 0, 64, c->stream));
    TR_CUDA(cudaMalloc(&tmp, std::max<size_t>(tb, 256)));
    TR_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, dk, dv, (int64_t)nnz, 0, 64, c->stream));
    keys_to_row_ptr_kernel<<<synth_grid(T->n_rows + 1), 256, 0, c->stream>>>(dk.Current(), nnz, T->n_rows, T->row_ptr);
    if (nnz > 0) {
      keys_to_cols_kernel<<<synth_grid(nnz), 256, 0, c->stream>>>(dk.Current(), nnz, T->col);
      if (A->val)
        permute_vals_kernel<<<synth_grid(nnz), 256, 0, c->stream>>>(dv.Current(), (const uint32_t *)A->val, nnz,
                                                                   (uint32_t *)T->val);
    }
    TR_CUDA(cudaGetLastError());
    TR_CUDA(cudaStreamSynchronize(c->stream));
    status = shb_csr_analyse(c, T);
  }
done:
  cudaFree(k1);
  cudaFree(k2);
  cudaFree(i1);
  cudaFree(i2);
  cudaFree(tmp);
  if (status != SHB_OK) {
    shb_csr_destroy(T);
    return status;
  }
  *out = T;
  return SHB_OK;
#undef TR_CUDA
}
