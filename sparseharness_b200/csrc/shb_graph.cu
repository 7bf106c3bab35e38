// shb_graph.cu -- device-resident graph drivers on the semirings + the multi-GPU scatter entry points.
//
// Reference counterparts (all host-driven, one blocking read-back of the whole vector per step):
//   app/bfs.cpp:94-152  HarnessBFS::executeRun    -> shb_bfs      (or-and, integer levels; Q8 decision)
//   app/sssp.cpp:97-155 HarnessSSSP::executeRun   -> shb_sssp     (min-plus to the exact fixed point)
//   app/pr.cpp:96-154   HarnessPR::executeRun     -> shb_pagerank (plus-times, canonical mode; Q9)
// Here the frontier, the vectors and the termination state never leave the device: the host
// enqueues a batch of steps and looks at one small state record per batch.
#include <algorithm>

#include "ldst.cuh"
#include "semiring.cuh"
#include "shb_internal.cuh"

static unsigned grid_for(int64_t n, int per_block = 256, int cap = 148 * 16) {
  return (unsigned)std::max<int64_t>(1, std::min<int64_t>((n + per_block - 1) / per_block, cap));
}

// =================================================================================================
// BFS: level-synchronous, direction-optimising (push over A^T for small frontiers, pull with early
// exit over A otherwise).  frontier_{k+1} = (A (x) frontier_k) AND NOT visited under or-and.
// =================================================================================================
struct BfsState {
  uint32_t level;        // level of the current frontier
  uint32_t mode;         // 0 = push (top-down), 1 = pull (bottom-up)
  uint32_t done;
  uint32_t have_at;
  uint32_t n_frontier;   // vertices in the current frontier
  uint32_t prev_frontier;
  uint32_t next_count;   // accumulated while building the next frontier
  uint32_t pad;
  unsigned long long next_edges;   // sum of push-degrees of the next frontier
  unsigned long long unexplored;   // edges not yet scanned (Beamer's m_u)
  unsigned long long reached, edges_traversed;
};

__global__ void bfs_init_kernel(BfsState *st, int32_t *levels, uint32_t *bm0, int64_t root, const int32_t *a_row_ptr,
                                const int32_t *at_row_ptr, unsigned long long nnz, int have_at) {
  levels[root] = 0;
  bm0[root >> 5] = 1u << (root & 31);
  st->level = 0;
  st->mode = have_at ? 0 : 1;
  st->done = 0;
  st->have_at = have_at;
  st->n_frontier = 1;
  st->prev_frontier = 0;
  st->next_count = 0;
  st->next_edges = 0;
  const int32_t *rp = have_at ? at_row_ptr : a_row_ptr;
  st->unexplored = nnz - (unsigned long long)(rp[root + 1] - rp[root]);
  st->reached = 1;
  st->edges_traversed = (unsigned long long)(a_row_ptr[root + 1] - a_row_ptr[root]);
}

__global__ void bfs_clear_kernel(const BfsState *st, uint32_t *bm0, uint32_t *bm1, int64_t n_words) {
  if (st->done) return;
  uint32_t *next = ((st->level + 1) & 1) ? bm1 : bm0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (int64_t)gridDim.x * blockDim.x)
    next[i] = 0u;
}

// push: one warp per bitmap word of the current frontier; lanes stride over each frontier vertex's
// out-adjacency (row of A^T)
template <bool HAS_VAL>
__global__ void bfs_push_kernel(BfsState *st, const int32_t *__restrict__ at_row_ptr, const int32_t *__restrict__ at_col,
                                const int32_t *__restrict__ at_val, const int32_t *__restrict__ a_row_ptr,
                                int32_t *levels, uint32_t *bm0, uint32_t *bm1, int64_t n_words) {
  if (st->done || st->mode != 0) return;
  const uint32_t level = st->level;
  const uint32_t *cur = (level & 1) ? bm1 : bm0;
  uint32_t *next = (level & 1) ? bm0 : bm1;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  uint32_t found = 0;
  unsigned long long push_deg = 0, pull_deg = 0;
  for (int64_t w = warp; w < n_words; w += n_warps) {
    uint32_t word = cur[w];
    while (word) {
      const int b = __ffs(word) - 1;
      word &= word - 1;
      const int64_t u = w * 32 + b;
      const int s = at_row_ptr[u], e = at_row_ptr[u + 1];
      for (int k = s + lane; k < e; k += 32) {
        if (HAS_VAL && at_val[k] == 0) continue;  // bool_and(x, a): a stored 0 is no edge
        const int32_t v = at_col[k];
        if (levels[v] == -1) {
          const int32_t old = atomicCAS(&levels[v], -1, (int32_t)level + 1);
          if (old == -1) {
            atomicOr(&next[v >> 5], 1u << (v & 31));
            ++found;
            push_deg += (unsigned long long)(at_row_ptr[v + 1] - at_row_ptr[v]);
            pull_deg += (unsigned long long)(a_row_ptr[v + 1] - a_row_ptr[v]);
          }
        }
      }
    }
  }
  for (int off = 16; off > 0; off >>= 1) {
    found += __shfl_xor_sync(0xffffffffu, found, off);
    push_deg += __shfl_xor_sync(0xffffffffu, push_deg, off);
    pull_deg += __shfl_xor_sync(0xffffffffu, pull_deg, off);
  }
  if (lane == 0 && found) {
    atomicAdd(&st->next_count, found);
    atomicAdd(&st->next_edges, push_deg);
    atomicAdd(&st->edges_traversed, pull_deg);
  }
}

// pull: thread per unvisited row, stop at the first in-neighbour that is in the frontier bitmap
template <bool HAS_VAL>
__global__ void bfs_pull_kernel(BfsState *st, const int32_t *__restrict__ a_row_ptr, const int32_t *__restrict__ a_col,
                                const int32_t *__restrict__ a_val, const int32_t *__restrict__ deg_row_ptr,
                                int32_t *levels, uint32_t *bm0, uint32_t *bm1, int64_t n) {
  if (st->done || st->mode != 1) return;
  const uint32_t level = st->level;
  const uint32_t *cur = (level & 1) ? bm1 : bm0;
  uint32_t *next = (level & 1) ? bm0 : bm1;
  const int lane = threadIdx.x & 31;
  uint32_t found_n = 0;
  unsigned long long push_deg = 0, pull_deg = 0;
  const int64_t n_padded = (n + 31) / 32 * 32;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_padded; r += (int64_t)gridDim.x * blockDim.x) {
    bool found = false;
    if (r < n && levels[r] == -1) {
      const int s = a_row_ptr[r], e = a_row_ptr[r + 1];
      for (int k = s; k < e; ++k) {
        if (HAS_VAL && a_val[k] == 0) continue;
        const int32_t c = a_col[k];
        if ((cur[c >> 5] >> (c & 31)) & 1u) {
          found = true;
          break;
        }
      }
      if (found) {
        levels[r] = (int32_t)level + 1;
        ++found_n;
        push_deg += (unsigned long long)(deg_row_ptr[r + 1] - deg_row_ptr[r]);
        pull_deg += (unsigned long long)(e - s);
      }
    }
    const uint32_t ballot = __ballot_sync(0xffffffffu, found);
    if (lane == 0 && ballot) next[r >> 5] = ballot;
  }
  for (int off = 16; off > 0; off >>= 1) {
    found_n += __shfl_xor_sync(0xffffffffu, found_n, off);
    push_deg += __shfl_xor_sync(0xffffffffu, push_deg, off);
    pull_deg += __shfl_xor_sync(0xffffffffu, pull_deg, off);
  }
  if (lane == 0 && found_n) {
    atomicAdd(&st->next_count, found_n);
    atomicAdd(&st->next_edges, push_deg);
    atomicAdd(&st->edges_traversed, pull_deg);
  }
}

__global__ void bfs_decide_kernel(BfsState *st, unsigned long long n) {
  if (st->done) return;
  const uint32_t nf = st->next_count;
  if (nf == 0) {
    st->done = 1;
    return;
  }
  st->reached += nf;
  const unsigned long long mf = st->next_edges;
  st->unexplored = st->unexplored > mf ? st->unexplored - mf : 0ull;
  if (st->have_at) {
    // Beamer's direction heuristic (alpha = 14, beta = 24)
    if (st->mode == 0) {
      if (mf > st->unexplored / 14ull) st->mode = 1;
    } else {
      if ((unsigned long long)nf < n / 24ull && nf < st->n_frontier) st->mode = 0;
    }
  }
  st->prev_frontier = st->n_frontier;
  st->n_frontier = nf;
  st->next_count = 0;
  st->next_edges = 0;
  st->level += 1;
}

SHB_API int shb_bfs(shb_ctx *c, const shb_csr *A, const shb_csr *AT, int64_t root, shb_buf *levels, uint32_t max_levels,
                    shb_graph_stats *stats) {
  SHB_REQUIRE(c && A && levels, "bad arguments");
  SHB_REQUIRE(A->n_rows == A->n_cols, "Matrix is not square. Failing computation.");
  SHB_REQUIRE(root >= 0 && root < A->n_rows, "root out of range");
  SHB_REQUIRE(levels->bytes >= (uint64_t)A->n_rows * 4, "levels shorter than n_rows");
  SHB_REQUIRE(!AT || (AT->n_rows == A->n_rows && AT->nnz == A->nnz), "transpose does not match");
  SHB_CUDA(cudaSetDevice(c->device));
  const int64_t n = A->n_rows;
  const int64_t n_words = (n + 31) / 32;
  if (max_levels == 0) max_levels = 1u << 30;
  BfsState *st = nullptr;
  uint32_t *bm = nullptr;
  SHB_CUDA(cudaMalloc(&st, sizeof(BfsState)));
  cudaError_t me = cudaMalloc(&bm, (size_t)n_words * 8);
  if (me != cudaSuccess) {
    cudaFree(st);
    shb_set_error("cudaMalloc of the frontier bitmaps failed: %s", cudaGetErrorString(me));
    return SHB_ERR_OOM;
  }
  uint32_t *bm0 = bm, *bm1 = bm + n_words;
  int status = SHB_OK;
  BfsState h;
  memset(&h, 0, sizeof h);
  uint64_t launches = 0;
  cudaEventRecord(c->ev0, c->stream);
  cudaMemsetAsync(levels->ptr, 0xff, (size_t)n * 4, c->stream);
  cudaMemsetAsync(bm, 0, (size_t)n_words * 8, c->stream);
  bfs_init_kernel<<<1, 1, 0, c->stream>>>(st, (int32_t *)levels->ptr, bm0, root, A->row_ptr, AT ? AT->row_ptr : nullptr,
                                          (unsigned long long)A->nnz, AT != nullptr);
  const int BATCH = 8;
  uint32_t issued = 0;
  const bool a_val = A->val != nullptr;
  while (true) {
    for (int b = 0; b < BATCH && issued < max_levels; ++b, ++issued) {
      bfs_clear_kernel<<<grid_for(n_words), 256, 0, c->stream>>>(st, bm0, bm1, n_words);
      if (AT) {
        if (a_val)
          bfs_push_kernel<true><<<grid_for(n_words * 32, 256, 148 * 8), 256, 0, c->stream>>>(
              st, AT->row_ptr, AT->col, (const int32_t *)AT->val, A->row_ptr, (int32_t *)levels->ptr, bm0, bm1, n_words);
        else
          bfs_push_kernel<false><<<grid_for(n_words * 32, 256, 148 * 8), 256, 0, c->stream>>>(
              st, AT->row_ptr, AT->col, nullptr, A->row_ptr, (int32_t *)levels->ptr, bm0, bm1, n_words);
      }
      if (a_val)
        bfs_pull_kernel<true><<<grid_for(n, 256, 148 * 32), 256, 0, c->stream>>>(
            st, A->row_ptr, A->col, (const int32_t *)A->val, AT ? AT->row_ptr : A->row_ptr, (int32_t *)levels->ptr, bm0, bm1, n);
      else
        bfs_pull_kernel<false><<<grid_for(n, 256, 148 * 32), 256, 0, c->stream>>>(
            st, A->row_ptr, A->col, nullptr, AT ? AT->row_ptr : A->row_ptr, (int32_t *)levels->ptr, bm0, bm1, n);
      bfs_decide_kernel<<<1, 1, 0, c->stream>>>(st, (unsigned long long)n);
      launches += AT ? 4 : 3;
    }
    cudaMemcpyAsync(&h, st, sizeof h, cudaMemcpyDeviceToHost, c->stream);
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) {
      shb_set_error("shb_bfs failed: %s (code: %d)", cudaGetErrorString(e), (int)e);
      status = SHB_ERR_CUDA;
      break;
    }
    if (h.done) break;
    if (issued >= max_levels) {
      shb_set_error("level cap of %u reached before the frontier emptied", max_levels);
      status = SHB_ERR_NOT_CONVERGED;
      break;
    }
  }
  cudaEventRecord(c->ev1, c->stream);
  cudaEventSynchronize(c->ev1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  if (stats) {
    stats->iterations = h.level + 1;  // levels 0..h.level are populated
    stats->converged = h.done;
    stats->edges_traversed = (int64_t)h.edges_traversed;
    stats->reached = (int64_t)h.reached;
    stats->total_ns = (uint64_t)((double)ms * 1e6 + 0.5);
    stats->launches = launches;
  }
  cudaFree(st);
  cudaFree(bm);
  return status;
}

// =================================================================================================
// SSSP: worklist Bellman-Ford under min-plus in ONE persistent cooperative kernel.
// d[v] <- min(d[v], |d[u]| + |w(u,v)|) is applied with atomicMin on the float bits (all values are
// non-negative, so integer order == float order).  fl(a + w) is monotone in a, hence any relaxation
// schedule reaches the same fixed point as the reference's dense sweeps, bit for bit.
// =================================================================================================
struct SsspState {
  uint32_t q_size[3];
  uint32_t bar_count;
  uint32_t bar_gen;
  uint32_t rounds;
  uint32_t done;
  uint32_t pad;
};

__device__ __forceinline__ void grid_barrier(SsspState *st, uint32_t n_blocks) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    volatile uint32_t *gen = &st->bar_gen;
    const uint32_t g = *gen;
    if (atomicAdd(&st->bar_count, 1u) == n_blocks - 1) {
      st->bar_count = 0;
      __threadfence();
      *gen = g + 1;
    } else {
      while (*gen == g) __nanosleep(32);
    }
    __threadfence();
  }
  __syncthreads();
}

#define SSSP_THREADS 256
#define SSSP_LOCAL_Q 2048

template <bool HAS_VAL>
__global__ void __launch_bounds__(SSSP_THREADS)
    sssp_persistent_kernel(SsspState *st, const int32_t *__restrict__ at_row_ptr, const int32_t *__restrict__ at_col,
                           const float *__restrict__ at_val, uint32_t *dist, int32_t *q0, int32_t *q1, uint32_t *stamp,
                           uint32_t max_rounds) {
  __shared__ int32_t local_q[SSSP_LOCAL_Q];
  __shared__ uint32_t local_n, global_base;
  const uint32_t n_blocks = gridDim.x;
  const int64_t n_threads = (int64_t)gridDim.x * blockDim.x;
  uint32_t round = 0;
  while (true) {
    const uint32_t qs = *(volatile uint32_t *)&st->q_size[round % 3];
    if (qs == 0 || round >= max_rounds) break;
    const int32_t *cur = (round & 1) ? q1 : q0;
    int32_t *next = (round & 1) ? q0 : q1;
    uint32_t *next_size = &st->q_size[(round + 1) % 3];
    if (blockIdx.x == 0 && threadIdx.x == 0) st->q_size[(round + 2) % 3] = 0;  // dead since the last barrier
    if (threadIdx.x == 0) local_n = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (int64_t)qs; i += n_threads) {
      const int32_t u = __ldcg(cur + i);
      const float du = __uint_as_float(__ldcg(dist + u));
      const int s = at_row_ptr[u], e = at_row_ptr[u + 1];
      for (int k = s; k < e; ++k) {
        const int32_t v = at_col[k];
        const float w = HAS_VAL ? at_val[k] : 1.0f;
        const uint32_t cb = __float_as_uint(MinPlusF32::mult(du, w));
        if (cb < __ldcg(dist + v)) {
          const uint32_t old = atomicMin(dist + v, cb);
          if (cb < old && atomicExch(stamp + v, round + 1) != round + 1) {
            const uint32_t slot = atomicAdd(&local_n, 1u);
            if (slot < SSSP_LOCAL_Q) local_q[slot] = v;
            else next[atomicAdd(next_size, 1u)] = v;  // block-local staging full: straight to HBM
          }
        }
      }
    }
    __syncthreads();
    const uint32_t ln = min(local_n, (uint32_t)SSSP_LOCAL_Q);
    if (threadIdx.x == 0 && ln) global_base = atomicAdd(next_size, ln);
    __syncthreads();
    for (uint32_t j = threadIdx.x; j < ln; j += blockDim.x) next[global_base + j] = local_q[j];
    grid_barrier(st, n_blocks);
    ++round;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->rounds = round;
    st->done = (*(volatile uint32_t *)&st->q_size[round % 3] == 0) ? 1u : 0u;
  }
}

__global__ void sssp_init_kernel(SsspState *st, uint32_t *dist, int32_t *q0, uint32_t *stamp, int64_t source) {
  dist[source] = 0u;  // bits of 0.0f
  q0[0] = (int32_t)source;
  stamp[source] = 0u;
  st->q_size[0] = 1;
  st->q_size[1] = 0;
  st->q_size[2] = 0;
  st->bar_count = 0;
  st->bar_gen = 0;
  st->rounds = 0;
  st->done = 0;
}

// reached vertices and the TEPS numerator (sum of in-degrees of reached vertices)
__global__ void reached_stats_kernel(const uint32_t *value, uint32_t unreached, const int32_t *a_row_ptr, int64_t n,
                                     unsigned long long *out) {
  unsigned long long cnt = 0, edges = 0;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x)
    if (value[r] != unreached) {
      ++cnt;
      edges += (unsigned long long)(a_row_ptr[r + 1] - a_row_ptr[r]);
    }
  for (int off = 16; off > 0; off >>= 1) {
    cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    edges += __shfl_xor_sync(0xffffffffu, edges, off);
  }
  if ((threadIdx.x & 31) == 0 && cnt) {
    atomicAdd(&out[0], cnt);
    atomicAdd(&out[1], edges);
  }
}

SHB_API int shb_sssp(shb_ctx *c, const shb_csr *A, const shb_csr *AT_in, int64_t source, shb_buf *dist, uint32_t max_rounds,
                     shb_graph_stats *stats) {
  SHB_REQUIRE(c && A && dist, "bad arguments");
  SHB_REQUIRE(A->n_rows == A->n_cols, "Matrix is not square. Failing computation.");
  SHB_REQUIRE(source >= 0 && source < A->n_rows, "source out of range");
  SHB_REQUIRE(dist->bytes >= (uint64_t)A->n_rows * 4, "dist shorter than n_rows");
  SHB_CUDA(cudaSetDevice(c->device));
  const int64_t n = A->n_rows;
  if (max_rounds == 0) max_rounds = 0xffffffffu;
  shb_csr *AT_own = nullptr;
  const shb_csr *AT = AT_in;
  if (!AT) {
    SHB_TRY(shb_csr_transpose(c, A, &AT_own));
    AT = AT_own;
  }
  SsspState *st = nullptr;
  int32_t *q = nullptr;
  uint32_t *stamp = nullptr;
  unsigned long long *d_out = nullptr;
  int status = SHB_OK;
  SsspState h;
  memset(&h, 0, sizeof h);
  unsigned long long h_out[2] = {0, 0};
  float ms = 0.f;
#define SS_CUDA(call)                                                                                     \
  do {                                                                                                    \
    cudaError_t _e = (call);                                                                              \
    if (_e != cudaSuccess) {                                                                              \
      shb_set_error("%s failed: %s (code: %d)", #call, cudaGetErrorString(_e), (int)_e);                  \
      status = _e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;                              \
      goto done;                                                                                          \
    }                                                                                                     \
  } while (0)
  {
    SS_CUDA(cudaMalloc(&st, sizeof(SsspState)));
    SS_CUDA(cudaMalloc(&q, (size_t)n * 8));
    SS_CUDA(cudaMalloc(&stamp, (size_t)n * 4));
    SS_CUDA(cudaMalloc(&d_out, 16));
    SS_CUDA(cudaMemsetAsync(d_out, 0, 16, c->stream));
    int32_t *q0 = q, *q1 = q + n;
    int per_sm = 0;
    const bool has_val = AT->val != nullptr;
    if (has_val) SS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sssp_persistent_kernel<true>, SSSP_THREADS, 0));
    else SS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sssp_persistent_kernel<false>, SSSP_THREADS, 0));
    per_sm = std::max(1, std::min(per_sm, 4));
    const int blocks = per_sm * c->prop.multiProcessorCount;
    SS_CUDA(cudaEventRecord(c->ev0, c->stream));
    // dist = ZERO of min-plus everywhere (3.4028235E38f = 0x7f7fffff), stamp = never
    {
      const uint32_t fmax_bits = 0x7f7fffffu;
      // fill with a kernel-free trick is not possible for this pattern; use the u32 fill kernel via API
      shb_buf tmp;
      tmp.ctx = c; tmp.ptr = dist->ptr; tmp.bytes = dist->bytes; tmp.owned = false;
      status = shb_buf_fill_u32(c, &tmp, 0, (uint64_t)n, fmax_bits);
      if (status != SHB_OK) goto done;
    }
    SS_CUDA(cudaMemsetAsync(stamp, 0xff, (size_t)n * 4, c->stream));
    sssp_init_kernel<<<1, 1, 0, c->stream>>>(st, (uint32_t *)dist->ptr, q0, stamp, source);
    SS_CUDA(cudaGetLastError());
    {
      const int32_t *a0 = AT->row_ptr, *a1 = AT->col;
      const float *a2 = (const float *)AT->val;
      uint32_t *a3 = (uint32_t *)dist->ptr;
      void *args[] = {&st, &a0, &a1, &a2, &a3, &q0, &q1, &stamp, &max_rounds};
      if (has_val)
        SS_CUDA(cudaLaunchCooperativeKernel((void *)sssp_persistent_kernel<true>, dim3(blocks), dim3(SSSP_THREADS), args, 0, c->stream));
      else
        SS_CUDA(cudaLaunchCooperativeKernel((void *)sssp_persistent_kernel<false>, dim3(blocks), dim3(SSSP_THREADS), args, 0, c->stream));
    }
    reached_stats_kernel<<<grid_for(n), 256, 0, c->stream>>>((const uint32_t *)dist->ptr, 0x7f7fffffu, A->row_ptr, n, d_out);
    SS_CUDA(cudaEventRecord(c->ev1, c->stream));
    SS_CUDA(cudaMemcpyAsync(&h, st, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(cudaMemcpyAsync(h_out, d_out, 16, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(cudaStreamSynchronize(c->stream));
    SS_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (!h.done) {
      shb_set_error("round cap of %u reached before the worklist emptied", max_rounds);
      status = SHB_ERR_NOT_CONVERGED;
    }
  }
done:
  if (stats) {
    stats->iterations = h.rounds;
    stats->converged = h.done;
    stats->reached = (int64_t)h_out[0];
    stats->edges_traversed = (int64_t)h_out[1];
    stats->total_ns = (uint64_t)((double)ms * 1e6 + 0.5);
    stats->launches = 4;
  }
  cudaFree(st);
  cudaFree(q);
  cudaFree(stamp);
  cudaFree(d_out);
  if (AT_own) shb_csr_destroy(AT_own);
  return status;
#undef SS_CUDA
}

// =================================================================================================
// PageRank, canonical mode (Q9): x_{k+1} = A x_k * 1 + 1 * beta with y == 1 never re-bound.
// =================================================================================================
SHB_API int shb_pagerank(shb_ctx *c, const shb_csr *A, const shb_launch_cfg *cfg, shb_buf *x, shb_buf *work, float x_init,
                         float beta, double delta, uint32_t max_iters, uint32_t check_every, int *result_in_work,
                         shb_graph_stats *stats) {
  SHB_REQUIRE(c && A && x && work && result_in_work, "bad arguments");
  SHB_REQUIRE(A->n_rows == A->n_cols, "Matrix is not square. Failing computation.");
  SHB_REQUIRE(x->bytes >= (uint64_t)A->n_rows * 4 && work->bytes >= (uint64_t)A->n_rows * 4, "vector shorter than n_rows");
  SHB_REQUIRE(x->ptr != work->ptr, "x and work must differ");
  SHB_CUDA(cudaSetDevice(c->device));
  uint32_t bits;
  memcpy(&bits, &x_init, 4);
  SHB_TRY(shb_buf_fill_u32(c, x, 0, (uint64_t)A->n_rows, bits));
  const float alpha = 1.0f;
  uint32_t iters = 0;
  uint64_t total = 0, launches = 0;
  int s = shb_iterate_impl(c, A, SHB_PLUS_TIMES_F32, cfg, x->ptr, nullptr, &alpha, &beta, work->ptr, delta,
                           SHB_TERM_ABSDIFF, 2, max_iters, check_every, nullptr, &iters, result_in_work, &total, &launches);
  if (stats) {
    stats->iterations = iters;
    stats->converged = s == SHB_OK;
    stats->edges_traversed = (int64_t)A->nnz * (int64_t)iters;
    stats->reached = A->n_rows;
    stats->total_ns = total;
    stats->launches = launches;
  }
  return s;
}

// =================================================================================================
// multi-GPU: CUDA IPC export/open + the fused SpMV -> scatter-to-all-peers launch
// =================================================================================================
SHB_API int shb_ipc_export(shb_buf *b, void *handle) {
  SHB_REQUIRE(b && handle, "bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == SHB_IPC_HANDLE_BYTES, "ipc handle size");
  SHB_CUDA(cudaSetDevice(b->ctx->device));
  cudaIpcMemHandle_t h;
  SHB_CUDA(cudaIpcGetMemHandle(&h, b->ptr));
  memcpy(handle, &h, sizeof h);
  return SHB_OK;
}

SHB_API int shb_ipc_open(shb_ctx *c, const void *handle, uint64_t bytes, shb_buf **out) {
  SHB_REQUIRE(c && handle && out, "bad arguments");
  SHB_CUDA(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  void *ptr = nullptr;
  SHB_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
  shb_buf *b = new shb_buf();
  b->ctx = c;
  b->ptr = ptr;
  b->bytes = bytes;
  b->owned = false;
  b->ipc = true;
  *out = b;
  return SHB_OK;
}

SHB_API int shb_ipc_close(shb_buf *b) { return shb_buf_destroy(b); }

SHB_API int shb_spmv_scatter(shb_ctx *c, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const shb_buf *x,
                             const shb_buf *y, const void *alpha, const void *beta, shb_buf *const *dst, int n_dst,
                             int64_t row_offset, double delta, int term_mode, shb_buf *changed_flag,
                             const shb_buf *prev_changed, uint64_t *kernel_ns) {
  SHB_REQUIRE(c && A && x && alpha && beta && dst, "bad arguments");
  SHB_REQUIRE(n_dst >= 1 && n_dst <= 8, "1..8 destinations");
  SHB_REQUIRE(row_offset >= 0, "negative row offset");
  SHB_REQUIRE(x->bytes >= (uint64_t)A->n_cols * 4, "x shorter than n_cols");
  void *outs[8];
  for (int i = 0; i < n_dst; ++i) {
    SHB_REQUIRE(dst[i] && dst[i]->bytes >= (uint64_t)(row_offset + A->n_rows) * 4, "destination too short");
    SHB_REQUIRE(dst[i]->ptr != x->ptr, "a destination aliases x");
    outs[i] = dst[i]->ptr;
  }
  SHB_REQUIRE(!y || y->bytes >= (uint64_t)(row_offset + A->n_rows) * 4, "y too short");
  SHB_REQUIRE(!changed_flag || (uint64_t)(row_offset + A->n_rows) * 4 <= x->bytes, "convergence test needs x[row_offset + r]");
  SHB_CUDA(cudaSetDevice(c->device));
  IterCtl ctl = {prev_changed ? (const uint32_t *)prev_changed->ptr : nullptr,
                 changed_flag ? (uint32_t *)changed_flag->ptr : nullptr, delta, term_mode};
  shb_launch_cfg lcfg = {};
  if (cfg) lcfg = *cfg;
  if (!y) {
    // y omitted: y is x, the iterative apps' re-bound y (app/sssp.cpp:150), unless SHB_F_Y_IS_ONE
    lcfg.flags &= ~SHB_F_NO_Y_SKIP;
  }
  ScopedTimer t(c, kernel_ns);
  SHB_TRY(shb_launch_spmv(c, A, semiring, &lcfg, x->ptr, y ? y->ptr : x->ptr, alpha, beta, outs, n_dst, row_offset, ctl));
  return t.finish();
}
