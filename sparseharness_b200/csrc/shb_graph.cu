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
// SSSP: near-far (delta-stepping) label correcting under min-plus in ONE persistent cooperative kernel.
// d[v] <- min(d[v], |d[u]| + |w(u,v)|) is applied with atomicMin on the float bits (all values are
// non-negative, so integer order == float order).  fl(a + w) is monotone in a, hence any relaxation
// schedule reaches the same least fixed point as the reference's dense sweeps (app/sssp.cpp:112-153),
// bit for bit; the schedule only decides how much work that takes:
//   * tentative distances below the bucket bound T go to the NEAR list as (vertex, distance) pairs; an
//     entry is live only while dist[vertex] still equals its distance, so a vertex improved twice is
//     relaxed once (the stale entry is skipped on pop);
//   * everything else waits in the FAR pile until the near list runs dry; the pile is then split
//     against the next bound T' = min(far) + delta.  A vertex enters the pile exactly once, on its first
//     touch (old distance == ZERO) with a distance >= T: a reached vertex whose distance is still >= T
//     is in the pile by induction, so later improvements need no second entry and no flag array;
//   * inside a bucket every CTA keeps relaxing what it discovers out of two shared-memory lists, with
//     no grid barrier between those local rounds (the dependent chain of a wave front costs L2 round
//     trips, not barriers); lists that overflow spill to the global near list of the next phase.
// delta = C * mean|w| / mean degree (near-far heuristic of Davidson et al.; C = 128 measured best on B200).
// =================================================================================================
struct SsspState {
  uint32_t near_size[3];  // global near lists; slot = phase % 3
  uint32_t far_size[3];   // far piles; slot = generation % 3
  uint32_t far_min[3];    // smallest distance (bits) recorded for that pile
  uint32_t bar_count;
  uint32_t bar_gen;
  uint32_t phases;
  uint32_t buckets;
  uint32_t done;
  uint32_t overflow;
  float delta;
  double w_sum;
  unsigned long long w_cnt;
  uint32_t cap_bits;  // partitioned solve: upper end of the previous round's distance window (kept across rounds)
  uint32_t pad3;
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
      while (*gen == g) __nanosleep(20);
    }
    __threadfence();
  }
  __syncthreads();
}

#define SSSP_THREADS 512
#define SSSP_CAP 2048     // entries per shared-memory near list (two lists)
#define SSSP_FSTAGE 2048  // far pushes staged per CTA between flushes
#define SSSP_BIG_ROW 32   // rows at least this long are relaxed by the whole warp

struct SsspShared {
  uint2 lst[2][SSSP_CAP];
  int32_t fstage[SSSP_FSTAGE];
  uint32_t lcount[2];
  uint32_t fcount, fbase, fmin;
};

struct SsspArgs {
  SsspState *st;
  const int32_t *at_row_ptr;
  const int32_t *at_col;
  const float *at_val;
  uint32_t *dist;
  uint2 *near_q[2];
  int32_t *far_q[2];
  uint32_t n;
  uint32_t near_cap;
  uint32_t max_phases;
};

// one successful relaxation: dist[v] just dropped to cb
__device__ __forceinline__ void sssp_improved(const SsspArgs &a, SsspShared &sh, int32_t v, uint32_t cb, uint32_t old,
                                              uint32_t t_hi,
                                              int nxt, uint2 *near_next, uint32_t *near_next_size, int32_t *far_cur,
                                              uint32_t *far_cur_size, uint32_t &fmin) {
  if (cb < t_hi) {
    // v is popped a round from now: have its row bounds in L2 by then
    asm volatile("prefetch.global.L2 [%0];" ::"l"(a.at_row_ptr + v));
    const uint32_t slot = atomicAdd(&sh.lcount[nxt], 1u);
    if (slot < SSSP_CAP) {
      sh.lst[nxt][slot] = make_uint2((uint32_t)v, cb);
    } else {
      const uint32_t g = atomicAdd(near_next_size, 1u);
      if (g < a.near_cap) near_next[g] = make_uint2((uint32_t)v, cb);
      else a.st->overflow = 1u;
    }
  } else {
    fmin = min(fmin, cb);
    if (old == 0x7f7fffffu) {  // first touch: the vertex's one and only far entry
      const uint32_t slot = atomicAdd(&sh.fcount, 1u);
      if (slot < SSSP_FSTAGE) sh.fstage[slot] = v;
      else far_cur[atomicAdd(far_cur_size, 1u)] = v;  // staging full: straight to HBM (one entry per vertex: fits)
    }
  }
}

__device__ __forceinline__ void sssp_flush_far(SsspShared &sh, int32_t *far_cur, uint32_t *far_cur_size) {
  // called by all threads of the CTA between two __syncthreads-separated rounds
  const uint32_t cnt = min(sh.fcount, (uint32_t)SSSP_FSTAGE);
  if (cnt == 0) return;  // uniform: fcount is stable here
  if (threadIdx.x == 0) sh.fbase = atomicAdd(far_cur_size, cnt);
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < cnt; j += blockDim.x) far_cur[sh.fbase + j] = sh.fstage[j];
  __syncthreads();
  if (threadIdx.x == 0) sh.fcount = 0;
}

template <bool HAS_VAL>
__global__ void __launch_bounds__(SSSP_THREADS, 1) sssp_nearfar_kernel(const __grid_constant__ SsspArgs a) {
  __shared__ SsspShared sh;
  SsspState *st = a.st;
  const uint32_t n_blocks = gridDim.x;
  const int lane = threadIdx.x & 31;
  const uint32_t n_threads = gridDim.x * blockDim.x;
  const float delta = *(volatile float *)&st->delta;
  uint32_t phase = 0, fgen = 0, buckets = 1;
  uint32_t t_hi = __float_as_uint(delta);  // bucket bound (exclusive) of the near list
  bool finished = false;
  if (threadIdx.x == 0) {
    sh.lcount[0] = sh.lcount[1] = 0;
    sh.fcount = 0;
    sh.fmin = 0xffffffffu;
  }
  __syncthreads();
  while (phase < a.max_phases) {
    const uint32_t ns = *(volatile uint32_t *)&st->near_size[phase % 3];
    const uint2 *near_cur = a.near_q[phase & 1];
    uint2 *near_next = a.near_q[(phase + 1) & 1];
    uint32_t *near_next_size = &st->near_size[(phase + 1) % 3];
    if (blockIdx.x == 0 && threadIdx.x == 0) st->near_size[(phase + 2) % 3] = 0;  // dead since the last barrier
    uint32_t fmin = 0xffffffffu;
    if (ns > 0) {
      // ---- near phase: relax the global near list, then whatever this CTA discovers, to exhaustion
      int32_t *far_cur = a.far_q[fgen & 1];
      uint32_t *far_cur_size = &st->far_size[fgen % 3];
      const uint32_t ch = max(32u, min((uint32_t)SSSP_CAP, (ns + n_blocks - 1) / n_blocks));
      for (uint32_t base = blockIdx.x * ch; base < ns; base += n_blocks * ch) {
        const uint32_t cnt = min(ch, ns - base);
        for (uint32_t j = threadIdx.x; j < cnt; j += blockDim.x) sh.lst[0][j] = near_cur[base + j];
        if (threadIdx.x == 0) sh.lcount[0] = cnt;
        int cur = 0;
        while (true) {
          __syncthreads();
          const uint32_t cn = sh.lcount[cur];
          if (cn == 0) break;
          const int nxt = cur ^ 1;
          for (uint32_t i = threadIdx.x; i < ((cn + 31u) & ~31u); i += blockDim.x) {
            int s = 0, e = 0;
            uint32_t du_bits = 0;
            if (i < cn) {
              const uint2 ent = sh.lst[cur][i];
              const uint32_t d_now = __ldcg(a.dist + ent.x);  // the three loads go out together
              const int s0 = __ldg(a.at_row_ptr + ent.x), e0 = __ldg(a.at_row_ptr + ent.x + 1);
              if (d_now == ent.y) {  // still the vertex's distance: live entry
                s = s0;
                e = e0;
                du_bits = ent.y;
              }
            }
            const float du = __uint_as_float(du_bits);
            const bool big = (e - s) >= SSSP_BIG_ROW;
            if (!big) {
              for (int k = s; k < e; k += 4) {
                int32_t vv[4];
                uint32_t cb[4], old[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  vv[j] = (k + j < e) ? __ldg(a.at_col + k + j) : -1;
                  const float w = (HAS_VAL && k + j < e) ? __ldg(a.at_val + k + j) : 1.0f;
                  cb[j] = __float_as_uint(MinPlusF32::mult(du, w));
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) old[j] = ((uint32_t)vv[j] < a.n) ? atomicMin(a.dist + vv[j], cb[j]) : 0u;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (cb[j] < old[j])
                    sssp_improved(a, sh, vv[j], cb[j], old[j], t_hi, nxt, near_next, near_next_size, far_cur, far_cur_size, fmin);
              }
            }
            uint32_t bigs = __ballot_sync(0xffffffffu, big);
            while (bigs) {
              const int src = __ffs(bigs) - 1;
              bigs &= bigs - 1;
              const int bs = __shfl_sync(0xffffffffu, s, src), be = __shfl_sync(0xffffffffu, e, src);
              const float bd = __uint_as_float(__shfl_sync(0xffffffffu, du_bits, src));
              for (int k = bs + lane; k < be; k += 32) {
                const int32_t v = __ldg(a.at_col + k);
                const float w = HAS_VAL ? __ldg(a.at_val + k) : 1.0f;
                const uint32_t c1 = __float_as_uint(MinPlusF32::mult(bd, w));
                if ((uint32_t)v < a.n && c1 < __ldcg(a.dist + v)) {
                  const uint32_t o1 = atomicMin(a.dist + v, c1);
                  if (c1 < o1) sssp_improved(a, sh, v, c1, o1, t_hi, nxt, near_next, near_next_size, far_cur, far_cur_size, fmin);
                }
              }
            }
          }
          __syncthreads();
          if (threadIdx.x == 0) {
            sh.lcount[cur] = 0;
            sh.lcount[nxt] = min(sh.lcount[nxt], (uint32_t)SSSP_CAP);  // the excess went to the global list
          }
          if (sh.fcount >= SSSP_FSTAGE / 2) sssp_flush_far(sh, far_cur, far_cur_size);
          cur = nxt;
        }
      }
      __syncthreads();
      sssp_flush_far(sh, far_cur, far_cur_size);
      for (int off = 16; off > 0; off >>= 1) fmin = min(fmin, __shfl_xor_sync(0xffffffffu, fmin, off));
      if (lane == 0 && fmin != 0xffffffffu) atomicMin(&sh.fmin, fmin);
      __syncthreads();
      if (threadIdx.x == 0 && sh.fmin != 0xffffffffu) {
        atomicMin(&st->far_min[fgen % 3], sh.fmin);
        sh.fmin = 0xffffffffu;
      }
    } else {
      // ---- the near list ran dry: split the far pile against the next bucket bound
      const uint32_t fs = *(volatile uint32_t *)&st->far_size[fgen % 3];
      if (fs == 0) {
        finished = true;
        break;
      }
      const uint32_t t_old = t_hi;
      t_hi = __float_as_uint(__fadd_rn(__uint_as_float(*(volatile uint32_t *)&st->far_min[fgen % 3]), delta));
      ++buckets;
      const int32_t *far_cur = a.far_q[fgen & 1];
      int32_t *far_next = a.far_q[(fgen + 1) & 1];
      uint32_t *far_next_size = &st->far_size[(fgen + 1) % 3];
      if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->far_size[(fgen + 2) % 3] = 0;  // the pile split one bucket ago
        st->far_min[(fgen + 2) % 3] = 0xffffffffu;
      }
      for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ((fs + 31u) & ~31u); i += n_threads) {
        int kind = 0;
        int32_t v = 0;
        uint32_t d = 0;
        if (i < fs) {
          v = far_cur[i];
          d = __ldcg(a.dist + v);
          if (d < t_old) {
            // went through the near list in an earlier bucket: stale, dropped
          } else if (d < t_hi) {
            kind = 1;
          } else {
            kind = 2;
            fmin = min(fmin, d);
          }
        }
        const uint32_t m1 = __ballot_sync(0xffffffffu, kind == 1), m2 = __ballot_sync(0xffffffffu, kind == 2);
        uint32_t b1 = 0, b2 = 0;
        if (lane == 0) {
          if (m1) b1 = atomicAdd(near_next_size, __popc(m1));
          if (m2) b2 = atomicAdd(far_next_size, __popc(m2));
        }
        b1 = __shfl_sync(0xffffffffu, b1, 0);
        b2 = __shfl_sync(0xffffffffu, b2, 0);
        const uint32_t below = (1u << lane) - 1u;
        if (kind == 1) {
          const uint32_t g = b1 + __popc(m1 & below);
          if (g < a.near_cap) near_next[g] = make_uint2((uint32_t)v, d);
          else st->overflow = 1u;
        } else if (kind == 2) {
          far_next[b2 + __popc(m2 & below)] = v;
        }
      }
      for (int off = 16; off > 0; off >>= 1) fmin = min(fmin, __shfl_xor_sync(0xffffffffu, fmin, off));
      if (lane == 0 && fmin != 0xffffffffu) atomicMin(&st->far_min[(fgen + 1) % 3], fmin);
      ++fgen;
    }
    grid_barrier(st, n_blocks);
    ++phase;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->phases = phase;
    st->buckets = buckets;
    st->done = (finished && st->overflow == 0) ? 1u : 0u;
  }
}

// mean |weight| over a strided sample of the stored values (the bucket width heuristic needs a scale)
__global__ void sssp_weight_stats_kernel(SsspState *st, const float *val, int64_t nnz, int64_t stride) {
  double s = 0.0;
  unsigned long long c = 0;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * stride; i < nnz; i += (int64_t)gridDim.x * blockDim.x * stride) {
    s += (double)fabsf(val[i]);
    ++c;
  }
  for (int off = 16; off > 0; off >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, off);
    c += __shfl_xor_sync(0xffffffffu, c, off);
  }
  if ((threadIdx.x & 31) == 0 && c) {
    atomicAdd(&st->w_sum, s);
    atomicAdd(&st->w_cnt, c);
  }
}

__global__ void sssp_init_kernel(SsspState *st, uint32_t *dist, uint2 *near0, int64_t source, double mean_degree,
                                 float c_scale, float delta_override) {
  dist[source] = 0u;  // bits of 0.0f
  near0[0] = make_uint2((uint32_t)source, 0u);
  st->near_size[0] = 1;
  st->near_size[1] = st->near_size[2] = 0;
  for (int i = 0; i < 3; ++i) {
    st->far_size[i] = 0;
    st->far_min[i] = 0xffffffffu;
  }
  st->bar_count = st->bar_gen = 0;
  st->phases = st->buckets = st->done = st->overflow = 0;
  const double mean_w = st->w_cnt ? st->w_sum / (double)st->w_cnt : 1.0;
  float delta = (float)(c_scale * mean_w / (mean_degree > 1.0 ? mean_degree : 1.0));
  if (delta_override > 0.f) delta = delta_override;
  if (!(delta > 0.f) || delta > 1.0e37f) delta = 1.0f;
  st->delta = delta;
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
  uint2 *near_q = nullptr;
  int32_t *far_q = nullptr;
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
    // bucket width: C * mean|w| / mean degree; SHB_SSSP_C and SHB_SSSP_DELTA are tuning overrides
    float c_scale = 128.0f, delta_override = 0.0f;
    if (const char *e = getenv("SHB_SSSP_C")) c_scale = (float)atof(e);
    if (const char *e = getenv("SHB_SSSP_DELTA")) delta_override = (float)atof(e);
    int per_sm_cap = 1;
    if (const char *e = getenv("SHB_SSSP_CTAS_PER_SM")) per_sm_cap = std::max(1, atoi(e));
    const uint64_t near_cap = (uint64_t)n + 1024;
    SS_CUDA(cudaMalloc(&st, sizeof(SsspState)));
    SS_CUDA(cudaMalloc(&near_q, (size_t)near_cap * 2 * sizeof(uint2)));
    SS_CUDA(cudaMalloc(&far_q, (size_t)n * 2 * sizeof(int32_t)));
    SS_CUDA(cudaMalloc(&d_out, 16));
    SS_CUDA(cudaMemsetAsync(d_out, 0, 16, c->stream));
    int per_sm = 0;
    const bool has_val = AT->val != nullptr;
    if (has_val) SS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sssp_nearfar_kernel<true>, SSSP_THREADS, 0));
    else SS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sssp_nearfar_kernel<false>, SSSP_THREADS, 0));
    per_sm = std::max(1, std::min(per_sm, per_sm_cap));
    const int blocks = per_sm * c->prop.multiProcessorCount;
    SS_CUDA(cudaEventRecord(c->ev0, c->stream));
    // dist = ZERO of min-plus everywhere (3.4028235E38f = 0x7f7fffff)
    {
      shb_buf tmp;
      tmp.ctx = c; tmp.device = c->device; tmp.ptr = dist->ptr; tmp.bytes = dist->bytes; tmp.owned = false;
      status = shb_buf_fill_u32(c, &tmp, 0, (uint64_t)n, 0x7f7fffffu);
      if (status != SHB_OK) goto done;
    }
    SS_CUDA(cudaMemsetAsync(st, 0, sizeof(SsspState), c->stream));
    if (has_val && AT->nnz > 0) {
      const int64_t stride = std::max<int64_t>(1, AT->nnz / (1 << 22));
      sssp_weight_stats_kernel<<<148, 256, 0, c->stream>>>(st, (const float *)AT->val, AT->nnz, stride);
    }
    sssp_init_kernel<<<1, 1, 0, c->stream>>>(st, (uint32_t *)dist->ptr, near_q, source,
                                             n > 0 ? (double)AT->nnz / (double)n : 1.0, c_scale, delta_override);
    SS_CUDA(cudaGetLastError());
    {
      SsspArgs a;
      a.st = st;
      a.at_row_ptr = AT->row_ptr;
      a.at_col = AT->col;
      a.at_val = (const float *)AT->val;
      a.dist = (uint32_t *)dist->ptr;
      a.near_q[0] = near_q;
      a.near_q[1] = near_q + near_cap;
      a.far_q[0] = far_q;
      a.far_q[1] = far_q + n;
      a.n = (uint32_t)n;
      a.near_cap = (uint32_t)near_cap;
      a.max_phases = max_rounds;
      void *args[] = {&a};
      if (has_val)
        SS_CUDA(cudaLaunchCooperativeKernel((void *)sssp_nearfar_kernel<true>, dim3(blocks), dim3(SSSP_THREADS), args, 0, c->stream));
      else
        SS_CUDA(cudaLaunchCooperativeKernel((void *)sssp_nearfar_kernel<false>, dim3(blocks), dim3(SSSP_THREADS), args, 0, c->stream));
    }
    reached_stats_kernel<<<grid_for(n), 256, 0, c->stream>>>((const uint32_t *)dist->ptr, 0x7f7fffffu, A->row_ptr, n, d_out);
    SS_CUDA(cudaEventRecord(c->ev1, c->stream));
    SS_CUDA(cudaMemcpyAsync(&h, st, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(cudaMemcpyAsync(h_out, d_out, 16, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(cudaStreamSynchronize(c->stream));
    SS_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (h.overflow) {
      shb_set_error("near list overflow (capacity %llu entries)", (unsigned long long)near_cap);
      status = SHB_ERR_CUDA;
    } else if (!h.done) {
      shb_set_error("phase cap of %u reached before the worklists emptied", max_rounds);
      status = SHB_ERR_NOT_CONVERGED;
    }
    if (getenv("SHB_SSSP_VERBOSE"))
      fprintf(stderr, "shb_sssp: delta %g, %u buckets, %u phases, %d CTAs, %.3f ms\n", (double)h.delta, h.buckets, h.phases, blocks, (double)ms);
  }
done:
  if (stats) {
    stats->iterations = h.phases;
    stats->converged = h.done;
    stats->reached = (int64_t)h_out[0];
    stats->edges_traversed = (int64_t)h_out[1];
    stats->total_ns = (uint64_t)((double)ms * 1e6 + 0.5);
    stats->launches = 5;
  }
  cudaFree(st);
  cudaFree(near_q);
  cudaFree(far_q);
  cudaFree(d_out);
  if (AT_own) shb_csr_destroy(AT_own);
  return status;
#undef SS_CUDA
}

// does row [s, e) have an in-neighbour in the frontier bitmap?  BFS_PROBE_BATCH entries per step: the column loads and the
// bitmap probes of a step are issued together, so a row costs e/BATCH dependent round trips instead of e (the early exit
// makes every step wait for its probes).
#ifndef BFS_PROBE_BATCH
#define BFS_PROBE_BATCH 4
#endif
template <bool HAS_VAL>
__device__ __forceinline__ bool bfs_row_hits(const int32_t *__restrict__ col, const int32_t *__restrict__ val,
                                             const uint32_t *__restrict__ cur, int s, int e) {
  for (int k = s; k < e; k += BFS_PROBE_BATCH) {
    int32_t c[BFS_PROBE_BATCH];
    uint32_t w[BFS_PROBE_BATCH];
#pragma unroll
    for (int j = 0; j < BFS_PROBE_BATCH; ++j) {
      const bool in = (k + j < e) && !(HAS_VAL && val[k + j] == 0);  // bool_and(x, a): a stored 0 is no edge
      c[j] = in ? col[k + j] : -1;
    }
#pragma unroll
    for (int j = 0; j < BFS_PROBE_BATCH; ++j) w[j] = c[j] >= 0 ? cur[c[j] >> 5] : 0u;
    bool hit = false;
#pragma unroll
    for (int j = 0; j < BFS_PROBE_BATCH; ++j) hit |= (c[j] >= 0) && ((w[j] >> (c[j] & 31)) & 1u);
    if (hit) return true;
  }
  return false;
}


// =================================================================================================
// Row-partitioned BFS (multi-GPU), direction-optimising.  Rank k owns the vertices
// [row_offset, row_offset + n_local): their rows of A (in-edges, pull), optionally their rows of A^T
// (out-edges, push) -- both with GLOBAL column ids --, its slice of `levels`, and full-length copies of the
// two frontier bitmaps and of the visited bitmap.  One level on every rank =
//   decide  (1 thread)  reads the ALL-REDUCED (n_f, m_f) of the current frontier and applies Beamer's
//                       heuristic; every rank computes the same mode from the same numbers;
//   prep                visited |= frontier (full length); levels of the owned frontier vertices;
//   push | pull         push: owned frontier vertices mark unvisited out-neighbours with atomicOr in the
//                       local next bitmap and, when the bit was new here, in every peer's next bitmap;
//                       pull: every unvisited owned row looks for an in-neighbour in the frontier (early
//                       exit; hub rows by the whole warp) and the warp's ballot word is stored into every
//                       rank's next bitmap.  Peer bitmaps are CUDA-IPC mappings: the exchange rides on the
//                       kernel's own stores over NVLink, no separate collective moves the frontier;
//   clear               the consumed frontier bitmap is zeroed for reuse two levels on.
// counters (int64): slot F = words [2+2F, 3+2F] = (vertices, sum of push degrees) of frontier F as seen by
// this rank; the caller all-reduces (SUM) slot F+1 after level F -- the barrier that orders the peer
// stores before the next level's reads.  Push may count a vertex once per discovering rank: the numbers
// steer the heuristic and the termination test (> 0) only; reached / TEPS come from the levels at the end.
// =================================================================================================
#define BFS_BLOCK_BIG_ROW 128
#define BFS_PUSH_BIG_ROW 1024  // a frontier vertex with at least this many out-edges is expanded by the whole grid
#define BFS_PUSH_BIG_CAP 1024
// counters layout (int64 words): 0 reached, 1 TEPS numerator (after finish), 2 arrival word of the peer barrier,
// 4..5 staging of the level's (vertices, degrees), slot F at BFS_SLOT(F), BFS_SLOT(F) + 1.  The slots are a ring of
// BFS_SLOT_RING frontiers (a traversal may have any number of levels): the prep kernel of level F zeroes slot F + 2,
// which nobody adds to before this rank has joined the barrier of level F.
#define BFS_FLAG_WORD 2
#define BFS_STAGE_WORD 4
#define BFS_SLOT_RING 256
#define BFS_SLOT(F) (6 + 2 * ((F) & (BFS_SLOT_RING - 1)))
#define BFS_COUNTER_WORDS (6 + 2 * BFS_SLOT_RING)

struct BfsPeers {
  uint32_t *bm[8];
  int n;
  int self;
};

struct BfsBlockState {
  uint32_t mode;  // 0 push, 1 pull
  uint32_t done;
  uint32_t have_push;
  uint32_t pad;
  unsigned long long n_prev;      // vertices of the previous frontier
  unsigned long long unexplored;  // Beamer's m_u
  unsigned long long n_total;
  uint32_t big_count;             // hub vertices of the current push level, expanded by the whole grid
  uint32_t pad2;
  int32_t big_list[BFS_PUSH_BIG_CAP];
};

// Start of a level, one launch: (a) thread 0 of block 0 applies Beamer's rule to the all-reduced (n_f, m_f) of the
// current frontier -- the later kernels of the level read the mode it leaves in the state; every other thread derives
// "the traversal is over" from the same counter slot, so nobody depends on when that store lands; (b) visited |=
// frontier over the whole bitmap, and the owned words give their vertices the level; (c) the spare bitmap (the
// frontier of two levels ago, the next frontier of the NEXT level) is zeroed: peers only start storing into it after
// this level's barrier, which this rank joins after this kernel.
__global__ void bfs_block_prep_kernel(BfsBlockState *st, long long *counters_rw, const uint32_t *__restrict__ cur,
                                      uint32_t *visited, uint32_t *spare, int32_t *levels, uint32_t level, int64_t n_words,
                                      int64_t word_lo, int64_t word_hi, int64_t row_offset, int64_t n_local) {
  const long long *counters = counters_rw;
  const unsigned long long nf = (unsigned long long)counters[BFS_SLOT(level)];
  const bool done = (*(volatile uint32_t *)&st->done != 0u) || nf == 0;  // either order of thread 0's store: same value
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->big_count = 0;
    counters_rw[BFS_SLOT(level + 2)] = 0;
    counters_rw[BFS_SLOT(level + 2) + 1] = 0;
    if (nf == 0) {
      st->done = 1;
    } else if (!done) {
      const unsigned long long mf = (unsigned long long)counters[BFS_SLOT(level) + 1];
      st->unexplored = st->unexplored > mf ? st->unexplored - mf : 0ull;
      if (st->have_push) {
        if (st->mode == 0) {
          if (mf > st->unexplored / 14ull) st->mode = 1;
        } else {
          if (nf < st->n_total / 24ull && nf < st->n_prev) st->mode = 0;
        }
      } else {
        st->mode = 1;
      }
      st->n_prev = nf;
    }
  }
  if (done) return;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_words; w += (int64_t)gridDim.x * blockDim.x) {
    spare[w] = 0u;
    uint32_t bits = cur[w];
    if (!bits) continue;
    visited[w] |= bits;
    if (w >= word_lo && w < word_hi) {
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        const int64_t r = w * 32 + b - row_offset;
        if (r >= 0 && r < n_local && levels[r] == -1) levels[r] = (int32_t)level;
      }
    }
  }
}

// one out-edge of a frontier vertex: mark the target in the local next bitmap and, if new here, in every peer's
__device__ __forceinline__ void bfs_push_visit(int32_t v, const uint32_t *__restrict__ visited, const BfsPeers &next,
                                               uint32_t *mine, const int32_t *__restrict__ deg, unsigned long long &found,
                                               unsigned long long &push_deg) {
  const uint32_t bit = 1u << (v & 31);
  if (visited[v >> 5] & bit) return;
  const uint32_t old = atomicOr(&mine[v >> 5], bit);
  if (!(old & bit)) {  // new on this rank: tell every peer once
    for (int i = 0; i < next.n; ++i)
      if (i != next.self) atomicOr(&next.bm[i][v >> 5], bit);
    ++found;
    push_deg += (unsigned long long)(deg ? deg[v] : 1);
  }
}

__device__ __forceinline__ void bfs_push_account(long long *acc, unsigned long long found, unsigned long long push_deg) {
  for (int off = 16; off > 0; off >>= 1) {
    found += __shfl_xor_sync(0xffffffffu, found, off);
    push_deg += __shfl_xor_sync(0xffffffffu, push_deg, off);
  }
  if ((threadIdx.x & 31) == 0 && found) {
    atomicAdd(reinterpret_cast<unsigned long long *>(&acc[0]), found);
    atomicAdd(reinterpret_cast<unsigned long long *>(&acc[1]), push_deg);
  }
}

// push: one warp per owned frontier word; lanes stride over each frontier vertex's out-adjacency.  Hub vertices
// are parked in the state's list and expanded by the whole grid in bfs_block_push_big_kernel (one warp walking a
// 10^5-entry adjacency would serialise the level).
template <bool HAS_VAL>
__global__ void bfs_block_push_kernel(BfsBlockState *st, const int32_t *__restrict__ at_row_ptr,
                                      const int32_t *__restrict__ at_col, const int32_t *__restrict__ at_val,
                                      const int32_t *__restrict__ deg, const uint32_t *__restrict__ cur,
                                      const uint32_t *__restrict__ visited, const __grid_constant__ BfsPeers next,
                                      long long *acc, int64_t word_lo, int64_t word_hi, int64_t row_offset, int64_t n_local) {
  if (st->done || st->mode != 0) return;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  uint32_t *mine = next.bm[next.self];
  unsigned long long found = 0, push_deg = 0;
  for (int64_t w = word_lo + warp; w < word_hi; w += n_warps) {
    uint32_t word = cur[w];
    while (word) {
      const int b = __ffs(word) - 1;
      word &= word - 1;
      const int64_t u = w * 32 + b - row_offset;
      if (u < 0 || u >= n_local) continue;
      const int s = at_row_ptr[u], e = at_row_ptr[u + 1];
      if (e - s >= BFS_PUSH_BIG_ROW) {
        uint32_t slot = 0;
        if (lane == 0) slot = atomicAdd(&st->big_count, 1u);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot < BFS_PUSH_BIG_CAP) {
          if (lane == 0) st->big_list[slot] = (int32_t)u;
          continue;
        }
      }
      for (int k = s + lane; k < e; k += 32) {
        if (HAS_VAL && at_val[k] == 0) continue;  // bool_and(x, a): a stored 0 is no edge
        bfs_push_visit(at_col[k], visited, next, mine, deg, found, push_deg);
      }
    }
  }
  bfs_push_account(acc, found, push_deg);
}

template <bool HAS_VAL>
__global__ void bfs_block_push_big_kernel(const BfsBlockState *st, const int32_t *__restrict__ at_row_ptr,
                                          const int32_t *__restrict__ at_col, const int32_t *__restrict__ at_val,
                                          const int32_t *__restrict__ deg, const uint32_t *__restrict__ visited,
                                          const __grid_constant__ BfsPeers next, long long *acc) {
  if (st->done || st->mode != 0) return;
  const uint32_t n_big = min(st->big_count, (uint32_t)BFS_PUSH_BIG_CAP);
  if (n_big == 0) return;
  uint32_t *mine = next.bm[next.self];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, n_threads = (int64_t)gridDim.x * blockDim.x;
  unsigned long long found = 0, push_deg = 0;
  for (uint32_t i = 0; i < n_big; ++i) {
    const int32_t u = st->big_list[i];
    const int s = at_row_ptr[u], e = at_row_ptr[u + 1];
    for (int64_t k = s + tid; k < e; k += n_threads) {
      if (HAS_VAL && at_val[k] == 0) continue;
      bfs_push_visit(at_col[k], visited, next, mine, deg, found, push_deg);
    }
  }
  bfs_push_account(acc, found, push_deg);
}

template <bool HAS_VAL>
__global__ void bfs_block_pull_kernel(const BfsBlockState *st, const int32_t *__restrict__ row_ptr,
                                      const int32_t *__restrict__ col, const int32_t *__restrict__ val,
                                      const int32_t *__restrict__ deg, const uint32_t *__restrict__ cur,
                                      const uint32_t *__restrict__ visited, const __grid_constant__ BfsPeers next,
                                      int32_t *levels, long long *acc, uint32_t level, int64_t n_local, int64_t row_offset) {
  if (st->done || st->mode != 1) return;
  const int lane = threadIdx.x & 31;
  const int64_t word_offset = row_offset >> 5;
  unsigned long long found_n = 0, push_deg = 0;
  const int64_t n_padded = (n_local + 31) / 32 * 32;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_padded; r += (int64_t)gridDim.x * blockDim.x) {
    bool found = false, big = false;
    int s = 0, e = 0;
    // one visited word per warp step instead of 32 level loads; vertices without in-edges were marked at init
    const uint32_t seen = visited[word_offset + (r >> 5)];
    if (r < n_local && !((seen >> lane) & 1u)) {
      s = row_ptr[r];
      e = row_ptr[r + 1];
      big = (e - s) >= BFS_BLOCK_BIG_ROW;
      if (!big) found = bfs_row_hits<HAS_VAL>(col, val, cur, s, e);
    }
    // hub rows: the whole warp scans, 32 entries per step, until one lane sees a frontier neighbour
    uint32_t bigs = __ballot_sync(0xffffffffu, big);
    while (bigs) {
      const int src = __ffs(bigs) - 1;
      bigs &= bigs - 1;
      const int bs = __shfl_sync(0xffffffffu, s, src), be = __shfl_sync(0xffffffffu, e, src);
      bool hit = false;
      for (int k0 = bs; k0 < be && !hit; k0 += 32) {
        const int k = k0 + lane;
        bool h = false;
        if (k < be && !(HAS_VAL && val[k] == 0)) {
          const int32_t c = col[k];
          h = ((cur[c >> 5] >> (c & 31)) & 1u) != 0u;
        }
        hit = __any_sync(0xffffffffu, h);
      }
      if (lane == src) found = hit;
    }
    if (found) {
      levels[r] = (int32_t)level + 1;
      ++found_n;
      push_deg += (unsigned long long)(deg ? deg[row_offset + r] : 1);
    }
    const uint32_t ballot = __ballot_sync(0xffffffffu, found);
    if (lane == 0) {
      const int64_t w = word_offset + (r >> 5);
      for (int i = 0; i < next.n; ++i) next.bm[i][w] = ballot;
    }
  }
  for (int off = 16; off > 0; off >>= 1) {
    found_n += __shfl_xor_sync(0xffffffffu, found_n, off);
    push_deg += __shfl_xor_sync(0xffffffffu, push_deg, off);
  }
  if (lane == 0 && found_n) {
    atomicAdd(reinterpret_cast<unsigned long long *>(&acc[0]), found_n);
    atomicAdd(reinterpret_cast<unsigned long long *>(&acc[1]), push_deg);
  }
}

// The exchange of a counter slot without a collective library: lane i adds this rank's staged (vertices,
// degrees) to rank i's slot with system-scope atomics over NVLink, fences, and bumps rank i's arrival word;
// lane 0 then waits until all `world` ranks have arrived here `target / world` times.  Every peer store of the
// level's push / pull kernels was issued by earlier kernels of this stream, so it is ordered before the arrival
// (fence + stream order) and visible to whoever observes the arrival with an acquire load.  target == 0: no
// wait (all blocks on one device, stream ordered).  A wait longer than ~4 s marks the traversal failed
// instead of hanging the device.
struct BfsComm {
  long long *counters[8];
  int n;
};

__global__ void bfs_block_publish_kernel(BfsBlockState *st, const __grid_constant__ BfsComm comm, long long *stage,
                                         int slot_word, int flag_word, int self, unsigned long long target) {
  const int lane = threadIdx.x;
  const long long nf = stage[0], mf = stage[1];
  if (lane < comm.n) {
    long long *c = comm.counters[lane];
    if (nf | mf) {
      atomicAdd_system(reinterpret_cast<unsigned long long *>(&c[slot_word]), (unsigned long long)nf);
      atomicAdd_system(reinterpret_cast<unsigned long long *>(&c[slot_word + 1]), (unsigned long long)mf);
    }
    __threadfence_system();
    atomicAdd_system(reinterpret_cast<unsigned long long *>(&c[flag_word]), 1ull);
  }
  __syncwarp();
  if (lane == 0) {
    stage[0] = 0;
    stage[1] = 0;
    if (target) {
      const unsigned long long *flag = reinterpret_cast<const unsigned long long *>(&comm.counters[self][flag_word]);
      unsigned long long t0 = 0, now = 0;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
      while (true) {
        unsigned long long v;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
        if (v >= target) break;
        __nanosleep(200);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (now - t0 > 4000000000ull) {
          st->done = 1;
          st->pad = 1;  // barrier timeout: reported by shb_bfs_block_finish
          break;
        }
      }
    }
    __threadfence_system();
  }
}

// owned vertices without in-edges can never be found by a pull: they start out "visited" (their level stays -1) so
// that the pull sweeps skip them with the visited word instead of fetching their row bounds every level
__global__ void bfs_block_mark_isolated_kernel(const int32_t *__restrict__ row_ptr, uint32_t *visited, int64_t n_local,
                                               int64_t word_offset) {
  const int lane = threadIdx.x & 31;
  const int64_t n_padded = (n_local + 31) / 32 * 32;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_padded; r += (int64_t)gridDim.x * blockDim.x) {
    const bool iso = r < n_local && row_ptr[r + 1] == row_ptr[r];
    const uint32_t m = __ballot_sync(0xffffffffu, iso);
    if (lane == 0 && m) visited[word_offset + (r >> 5)] = m;
  }
}

__global__ void bfs_block_init_kernel(BfsBlockState *st, int32_t *levels, uint32_t *bm0, long long *acc, int64_t root,
                                      int64_t row_offset, int64_t n_local, const int32_t *deg, int have_push,
                                      unsigned long long nnz_total, unsigned long long n_total) {
  bm0[root >> 5] = 1u << (root & 31);
  st->mode = have_push ? 0u : 1u;
  st->done = 0;
  st->have_push = (uint32_t)have_push;
  st->pad = 0;
  st->n_prev = 0;
  st->unexplored = nnz_total;
  st->n_total = n_total;
  st->big_count = 0;
  st->pad2 = 0;
  if (root >= row_offset && root < row_offset + n_local) {  // the owner reports frontier 0 (slot 0 is then exchanged)
    levels[root - row_offset] = 0;
    acc[0] = 1;
    acc[1] = deg ? (long long)deg[root] : 1;
  }
}

// reached vertices and the TEPS numerator (sum of in-degrees of reached vertices) of the owned slice
__global__ void bfs_block_stats_kernel(const int32_t *levels, const int32_t *row_ptr, int64_t n_local, long long *counters) {
  unsigned long long cnt = 0, edges = 0;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_local; r += (int64_t)gridDim.x * blockDim.x)
    if (levels[r] >= 0) {
      ++cnt;
      edges += (unsigned long long)(row_ptr[r + 1] - row_ptr[r]);
    }
  for (int off = 16; off > 0; off >>= 1) {
    cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    edges += __shfl_xor_sync(0xffffffffu, edges, off);
  }
  if ((threadIdx.x & 31) == 0 && cnt) {
    atomicAdd(reinterpret_cast<unsigned long long *>(&counters[0]), cnt);
    atomicAdd(reinterpret_cast<unsigned long long *>(&counters[1]), edges);
  }
}

__global__ void csr_row_lengths_kernel(const int32_t *row_ptr, int32_t *out, int64_t n) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x)
    out[r] = row_ptr[r + 1] - row_ptr[r];
}

SHB_API int shb_csr_row_lengths(shb_ctx *c, const shb_csr *A, shb_buf *out) {
  SHB_REQUIRE(c && A && out, "bad arguments");
  SHB_REQUIRE(out->bytes >= (uint64_t)A->n_rows * 4, "out shorter than n_rows");
  SHB_CUDA(cudaSetDevice(c->device));
  if (A->n_rows > 0) csr_row_lengths_kernel<<<grid_for(A->n_rows), 256, 0, c->stream>>>(A->row_ptr, (int32_t *)out->ptr, A->n_rows);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

static int bfs_block_check(const shb_csr *A, const shb_csr *AT, int64_t row_offset, int64_t n_total) {
  SHB_REQUIRE(A->n_cols == n_total, "the block's column space must be the whole vertex set");
  SHB_REQUIRE(row_offset >= 0 && row_offset + A->n_rows <= n_total, "row block outside the vertex set");
  SHB_REQUIRE(A->n_rows == 0 || (row_offset & 31) == 0, "row_offset must be a multiple of 32 (a bitmap word has one owner)");
  SHB_REQUIRE((A->n_rows & 31) == 0 || row_offset + A->n_rows == n_total, "only the last block may end off a word boundary");
  SHB_REQUIRE(!AT || (AT->n_rows == A->n_rows && AT->n_cols == n_total), "the push-side block does not match");
  return SHB_OK;
}

SHB_API int shb_bfs_block_state_bytes(uint64_t *bytes) {
  SHB_REQUIRE(bytes, "bad arguments");
  *bytes = sizeof(BfsBlockState);
  return SHB_OK;
}

SHB_API int shb_bfs_block_counter_bytes(uint64_t *bytes, uint32_t *slot_ring) {
  SHB_REQUIRE(bytes, "bad arguments");
  *bytes = (uint64_t)BFS_COUNTER_WORDS * 8;
  if (slot_ring) *slot_ring = BFS_SLOT_RING;
  return SHB_OK;
}

static int bfs_block_comm(shb_buf *counters, shb_buf *const *peer_counters, int n, int self, uint32_t max_slot, BfsComm *comm) {
  (void)max_slot;
  const uint64_t need = (uint64_t)BFS_COUNTER_WORDS * 8;
  SHB_REQUIRE(counters->bytes >= need, "counters shorter than shb_bfs_block_counter_bytes");
  comm->n = 0;
  for (int i = 0; i < 8; ++i) comm->counters[i] = nullptr;
  if (!peer_counters) return SHB_OK;
  SHB_REQUIRE(n >= 1 && n <= 8 && self >= 0 && self < n, "1..8 ranks, one of them local");
  for (int i = 0; i < n; ++i) {
    SHB_REQUIRE(peer_counters[i] && peer_counters[i]->bytes >= need, "a peer's counters are too short");
    comm->counters[i] = (long long *)peer_counters[i]->ptr;
  }
  SHB_REQUIRE(peer_counters[self]->ptr == counters->ptr, "peer_counters[self_index] must be the local counters");
  comm->n = n;
  return SHB_OK;
}

SHB_API int shb_bfs_block_init(shb_ctx *c, const shb_csr *A, const shb_csr *AT, int64_t row_offset, int64_t n_total,
                               int64_t nnz_total, int64_t root, const shb_buf *deg, shb_buf *levels, shb_buf *frontier0,
                               shb_buf *frontier1, shb_buf *frontier2, shb_buf *visited, shb_buf *counters, shb_buf *state,
                               int staged) {
  SHB_REQUIRE(c && A && levels && frontier0 && frontier1 && frontier2 && visited && counters && state, "bad arguments");
  SHB_TRY(bfs_block_check(A, AT, row_offset, n_total));
  SHB_REQUIRE(root >= 0 && root < n_total, "root out of range");
  const uint64_t bm_bytes = (uint64_t)((n_total + 31) / 32) * 4;
  SHB_REQUIRE(levels->bytes >= (uint64_t)A->n_rows * 4, "levels shorter than the row block");
  SHB_REQUIRE(frontier0->bytes >= bm_bytes && frontier1->bytes >= bm_bytes && frontier2->bytes >= bm_bytes &&
                  visited->bytes >= bm_bytes, "bitmap shorter than n_total bits");
  SHB_REQUIRE(!deg || deg->bytes >= (uint64_t)n_total * 4, "deg shorter than n_total");
  SHB_REQUIRE(state->bytes >= sizeof(BfsBlockState), "state too short");
  SHB_REQUIRE(counters->bytes >= (uint64_t)BFS_COUNTER_WORDS * 8, "counters shorter than shb_bfs_block_counter_bytes");
  SHB_CUDA(cudaSetDevice(c->device));
  long long *cnt = (long long *)counters->ptr;
  if (A->n_rows > 0) SHB_CUDA(cudaMemsetAsync(levels->ptr, 0xff, (size_t)A->n_rows * 4, c->stream));
  SHB_CUDA(cudaMemsetAsync(frontier0->ptr, 0, (size_t)bm_bytes, c->stream));
  SHB_CUDA(cudaMemsetAsync(frontier1->ptr, 0, (size_t)bm_bytes, c->stream));
  SHB_CUDA(cudaMemsetAsync(frontier2->ptr, 0, (size_t)bm_bytes, c->stream));
  SHB_CUDA(cudaMemsetAsync(visited->ptr, 0, (size_t)bm_bytes, c->stream));
  // everything but the arrival word, which counts monotonically over the life of the buffer
  SHB_CUDA(cudaMemsetAsync(cnt, 0, BFS_FLAG_WORD * 8, c->stream));
  SHB_CUDA(cudaMemsetAsync(cnt + BFS_STAGE_WORD, 0, (size_t)(BFS_COUNTER_WORDS - BFS_STAGE_WORD) * 8, c->stream));
  long long *acc = staged ? cnt + BFS_STAGE_WORD : cnt + BFS_SLOT(0);
  if (A->n_rows > 0)
    bfs_block_mark_isolated_kernel<<<grid_for(A->n_rows), 256, 0, c->stream>>>(A->row_ptr, (uint32_t *)visited->ptr, A->n_rows,
                                                                               row_offset >> 5);
  bfs_block_init_kernel<<<1, 1, 0, c->stream>>>((BfsBlockState *)state->ptr, (int32_t *)levels->ptr, (uint32_t *)frontier0->ptr,
                                                acc, root, row_offset, A->n_rows, deg ? (const int32_t *)deg->ptr : nullptr,
                                                AT != nullptr, (unsigned long long)nnz_total, (unsigned long long)n_total);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

// second half of init: the exchange of slot 0.  Separate so that every rank has zeroed its counters (the caller's
// barrier between the two calls) before any peer adds to them.
SHB_API int shb_bfs_block_publish(shb_ctx *c, uint32_t slot, shb_buf *counters, shb_buf *state, shb_buf *const *peer_counters,
                                  int n_ranks, int self_index, uint64_t barrier_target) {
  SHB_REQUIRE(c && counters && state && peer_counters, "bad arguments");
  BfsComm comm;
  SHB_TRY(bfs_block_comm(counters, peer_counters, n_ranks, self_index, slot, &comm));
  SHB_CUDA(cudaSetDevice(c->device));
  long long *cnt = (long long *)counters->ptr;
  bfs_block_publish_kernel<<<1, 32, 0, c->stream>>>((BfsBlockState *)state->ptr, comm, cnt + BFS_STAGE_WORD, BFS_SLOT(slot),
                                                    BFS_FLAG_WORD, self_index, (unsigned long long)barrier_target);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

SHB_API int shb_bfs_block_level(shb_ctx *c, const shb_csr *A, const shb_csr *AT, int64_t row_offset, int64_t n_total,
                                uint32_t level, const shb_buf *deg, const shb_buf *frontier, shb_buf *const *next, int n_next,
                                int self_index, shb_buf *spare, shb_buf *visited, shb_buf *levels, shb_buf *counters,
                                shb_buf *state, shb_buf *const *peer_counters, uint64_t barrier_target) {
  SHB_REQUIRE(c && A && frontier && next && spare && visited && levels && counters && state, "bad arguments");
  SHB_REQUIRE(n_next >= 1 && n_next <= 8 && self_index >= 0 && self_index < n_next, "1..8 destinations, one of them local");
  SHB_TRY(bfs_block_check(A, AT, row_offset, n_total));
  const int64_t n_words = (n_total + 31) / 32;
  const uint64_t bm_bytes = (uint64_t)n_words * 4;
  SHB_REQUIRE(frontier->bytes >= bm_bytes && visited->bytes >= bm_bytes && spare->bytes >= bm_bytes, "bitmap shorter than n_total bits");
  SHB_REQUIRE(spare->ptr != frontier->ptr, "the spare bitmap aliases the current frontier");
  SHB_REQUIRE(levels->bytes >= (uint64_t)A->n_rows * 4, "levels shorter than the row block");
  SHB_REQUIRE(state->bytes >= sizeof(BfsBlockState), "state too short");
  BfsComm comm;
  SHB_TRY(bfs_block_comm(counters, peer_counters, n_next, self_index, level + 1, &comm));
  BfsPeers peers;
  peers.n = n_next;
  peers.self = self_index;
  for (int i = 0; i < 8; ++i) peers.bm[i] = nullptr;
  for (int i = 0; i < n_next; ++i) {
    SHB_REQUIRE(next[i] && next[i]->bytes >= bm_bytes, "next-frontier bitmap too short");
    SHB_REQUIRE(next[i]->ptr != frontier->ptr && next[i]->ptr != spare->ptr, "a next-frontier bitmap aliases the current / spare one");
    peers.bm[i] = (uint32_t *)next[i]->ptr;
  }
  SHB_CUDA(cudaSetDevice(c->device));
  BfsBlockState *st = (BfsBlockState *)state->ptr;
  long long *cnt = (long long *)counters->ptr;
  long long *acc = comm.n ? cnt + BFS_STAGE_WORD : cnt + BFS_SLOT(level + 1);
  const int32_t *dg = deg ? (const int32_t *)deg->ptr : nullptr;
  const uint32_t *cur = (const uint32_t *)frontier->ptr;
  const int64_t n_local = A->n_rows, word_lo = row_offset >> 5, word_hi = (row_offset + n_local + 31) >> 5;
  bfs_block_prep_kernel<<<grid_for(n_words), 256, 0, c->stream>>>(st, cnt, cur, (uint32_t *)visited->ptr, (uint32_t *)spare->ptr,
                                                                  (int32_t *)levels->ptr, level, n_words, word_lo, word_hi,
                                                                  row_offset, n_local);
  if (n_local > 0) {
    if (AT) {
      const unsigned grid = grid_for((word_hi - word_lo) * 32, 256, 148 * 8);
      if (AT->val)
        bfs_block_push_kernel<true><<<grid, 256, 0, c->stream>>>(st, AT->row_ptr, AT->col, (const int32_t *)AT->val, dg, cur,
                                                                 (const uint32_t *)visited->ptr, peers, acc, word_lo, word_hi,
                                                                 row_offset, n_local);
      else
        bfs_block_push_kernel<false><<<grid, 256, 0, c->stream>>>(st, AT->row_ptr, AT->col, nullptr, dg, cur,
                                                                  (const uint32_t *)visited->ptr, peers, acc, word_lo, word_hi,
                                                                  row_offset, n_local);
      if (AT->val)
        bfs_block_push_big_kernel<true><<<148 * 4, 256, 0, c->stream>>>(st, AT->row_ptr, AT->col, (const int32_t *)AT->val, dg,
                                                                        (const uint32_t *)visited->ptr, peers, acc);
      else
        bfs_block_push_big_kernel<false><<<148 * 4, 256, 0, c->stream>>>(st, AT->row_ptr, AT->col, nullptr, dg,
                                                                         (const uint32_t *)visited->ptr, peers, acc);
    }
    const unsigned grid = grid_for(n_local, 256, 148 * 32);
    if (A->val)
      bfs_block_pull_kernel<true><<<grid, 256, 0, c->stream>>>(st, A->row_ptr, A->col, (const int32_t *)A->val, dg, cur,
                                                               (const uint32_t *)visited->ptr, peers, (int32_t *)levels->ptr,
                                                               acc, level, n_local, row_offset);
    else
      bfs_block_pull_kernel<false><<<grid, 256, 0, c->stream>>>(st, A->row_ptr, A->col, nullptr, dg, cur,
                                                                (const uint32_t *)visited->ptr, peers, (int32_t *)levels->ptr,
                                                                acc, level, n_local, row_offset);
  }
  if (comm.n)
    bfs_block_publish_kernel<<<1, 32, 0, c->stream>>>(st, comm, cnt + BFS_STAGE_WORD, BFS_SLOT(level + 1), BFS_FLAG_WORD,
                                                      self_index, (unsigned long long)barrier_target);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

// per-rank reached / edge counts from the levels; SHB_ERR_CUDA if a peer barrier timed out during the traversal
SHB_API int shb_bfs_block_finish(shb_ctx *c, const shb_csr *A, const shb_buf *levels, shb_buf *counters, const shb_buf *state) {
  SHB_REQUIRE(c && A && levels && counters && state, "bad arguments");
  SHB_REQUIRE(levels->bytes >= (uint64_t)A->n_rows * 4 && counters->bytes >= 16, "buffers too short");
  SHB_REQUIRE(state->bytes >= sizeof(BfsBlockState), "state too short");
  SHB_CUDA(cudaSetDevice(c->device));
  SHB_CUDA(cudaMemsetAsync(counters->ptr, 0, 16, c->stream));
  if (A->n_rows > 0)
    bfs_block_stats_kernel<<<grid_for(A->n_rows), 256, 0, c->stream>>>((const int32_t *)levels->ptr, A->row_ptr, A->n_rows,
                                                                       (long long *)counters->ptr);
  SHB_CUDA(cudaGetLastError());
  BfsBlockState h;
  SHB_CUDA(cudaMemcpyAsync(&h, state->ptr, sizeof h, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  if (h.pad) {
    shb_set_error("a peer barrier of the partitioned BFS timed out (a rank stopped taking part)");
    return SHB_ERR_CUDA;
  }
  return SHB_OK;
}

// =================================================================================================
// shb_bfs: the single-device driver (<- app/bfs.cpp:94-152 HarnessBFS::executeRun, integer levels, Q8 decision) is
// the row-partitioned code above run on ONE block that owns every vertex: same kernels, no peers, the level's counts
// accumulate straight into the slot.  The host enqueues batches of levels and polls the count ring once per batch.
// =================================================================================================
SHB_API int shb_bfs(shb_ctx *c, const shb_csr *A, const shb_csr *AT, int64_t root, shb_buf *levels, uint32_t max_levels,
                    shb_graph_stats *stats) {
  SHB_REQUIRE(c && A && levels, "bad arguments");
  SHB_REQUIRE(A->n_rows == A->n_cols, "Matrix is not square. Failing computation.");
  SHB_REQUIRE(root >= 0 && root < A->n_rows, "root out of range");
  SHB_REQUIRE(levels->bytes >= (uint64_t)A->n_rows * 4, "levels shorter than n_rows");
  SHB_REQUIRE(!AT || (AT->n_rows == A->n_rows && AT->nnz == A->nnz), "transpose does not match");
  SHB_CUDA(cudaSetDevice(c->device));
  const int64_t n = A->n_rows;
  const size_t bm_bytes = ((size_t)((n + 31) / 32) * 4 + 255) / 256 * 256;
  const size_t cnt_bytes = (size_t)BFS_COUNTER_WORDS * 8, st_bytes = (sizeof(BfsBlockState) + 255) / 256 * 256;
  if (max_levels == 0) max_levels = 1u << 30;
  // push degrees of every vertex, cached on the operator (one pass the first time it is traversed)
  if (AT && !AT->row_len && n > 0) {
    int32_t *rl = nullptr;
    SHB_CUDA(cudaMalloc(&rl, (size_t)n * 4));
    csr_row_lengths_kernel<<<grid_for(n), 256, 0, c->stream>>>(AT->row_ptr, rl, n);
    AT->row_len = rl;
  }
  char *scratch = nullptr;
  cudaError_t me = cudaMalloc(&scratch, 4 * bm_bytes + cnt_bytes + st_bytes);
  if (me != cudaSuccess) {
    shb_set_error("cudaMalloc of the frontier bitmaps failed: %s", cudaGetErrorString(me));
    return SHB_ERR_OOM;
  }
  auto view = [&](size_t off, size_t bytes) {
    shb_buf b;
    b.ctx = c; b.device = c->device; b.ptr = scratch + off; b.bytes = bytes; b.owned = false;
    return b;
  };
  shb_buf bm[3] = {view(0, bm_bytes), view(bm_bytes, bm_bytes), view(2 * bm_bytes, bm_bytes)};
  shb_buf visited = view(3 * bm_bytes, bm_bytes), counters = view(4 * bm_bytes, cnt_bytes);
  shb_buf state = view(4 * bm_bytes + cnt_bytes, st_bytes), deg = view(0, 0);
  if (AT) {
    deg.ptr = AT->row_len;
    deg.bytes = (uint64_t)n * 4;
  }
  int status = SHB_OK;
  uint64_t launches = 0;
  uint32_t issued = 0, populated = 0;
  bool done = false;
  long long ring[BFS_COUNTER_WORDS];
  cudaMemsetAsync(scratch + 4 * bm_bytes, 0, cnt_bytes, c->stream);  // the arrival word too (unused without peers)
  cudaEventRecord(c->ev0, c->stream);
  status = shb_bfs_block_init(c, A, AT, 0, n, A->nnz, root, AT ? &deg : nullptr, levels, &bm[0], &bm[1], &bm[2], &visited,
                              &counters, &state, 0);
  const uint32_t BATCH = 8;
  while (status == SHB_OK && !done) {
    const uint32_t first = issued;
    for (uint32_t b = 0; b < BATCH && issued < max_levels && status == SHB_OK; ++b, ++issued) {
      shb_buf *next[1] = {&bm[(issued + 1) % 3]};
      status = shb_bfs_block_level(c, A, AT, 0, n, issued, AT ? &deg : nullptr, &bm[issued % 3], next, 1, 0, &bm[(issued + 2) % 3],
                                   &visited, levels, &counters, &state, nullptr, 0);
      launches += AT ? 4 : 2;
    }
    if (status != SHB_OK) break;
    cudaMemcpyAsync(ring, counters.ptr, cnt_bytes, cudaMemcpyDeviceToHost, c->stream);
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) {
      shb_set_error("shb_bfs failed: %s (code: %d)", cudaGetErrorString(e), (int)e);
      status = SHB_ERR_CUDA;
      break;
    }
    for (uint32_t f = first; f < issued; ++f) {
      if (ring[BFS_SLOT(f + 1)] == 0) {  // launch f found nothing: frontiers 0..f are the populated levels
        populated = f + 1;
        done = true;
        break;
      }
    }
    if (!done && issued >= max_levels) {
      populated = issued + 1;
      // the vertices of the last frontier get their level (a push level leaves that to the next level's prep)
      const int64_t n_words = (n + 31) / 32;
      bfs_block_prep_kernel<<<grid_for(n_words), 256, 0, c->stream>>>(
          (BfsBlockState *)state.ptr, (long long *)counters.ptr, (const uint32_t *)bm[issued % 3].ptr, (uint32_t *)visited.ptr,
          (uint32_t *)bm[(issued + 2) % 3].ptr, (int32_t *)levels->ptr, issued, n_words, 0, n_words, 0, n);
      shb_set_error("level cap of %u reached before the frontier emptied", max_levels);
      status = SHB_ERR_NOT_CONVERGED;
      break;
    }
  }
  int fin = shb_bfs_block_finish(c, A, levels, &counters, &state);
  cudaEventRecord(c->ev1, c->stream);
  cudaEventSynchronize(c->ev1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, c->ev0, c->ev1);
  long long tot[2] = {0, 0};
  cudaMemcpy(tot, counters.ptr, 16, cudaMemcpyDeviceToHost);
  if (status == SHB_OK && fin != SHB_OK) status = fin;
  if (stats) {
    stats->iterations = populated;
    stats->converged = done ? 1u : 0u;
    stats->reached = tot[0];
    stats->edges_traversed = tot[1];
    stats->total_ns = (uint64_t)((double)ms * 1e6 + 0.5);
    stats->launches = launches;
  }
  cudaFree(scratch);
  cudaGetLastError();
  return status;
}

// =================================================================================================
// Row-partitioned SSSP (multi-GPU): block-asynchronous near-far.  Rank k owns the vertices
// [lo, lo + n_local): their out-edges (rows lo.. of A^T, GLOBAL target ids), the authoritative slice of `dist`
// (a full-length array on every rank; the entries of other ranks' vertices are only a filter cache) and
// `prev[v - lo]` = the distance v was last expanded with.  One round on every rank:
//   seed    every owned v with dist[v] < prev[v] -- lowered by a peer since its last expansion, or the source --
//           enters the far pile;
//   solve   the near-far kernel of shb_sssp relaxes to exhaustion INSIDE the block: an improved owned target is
//           queued as usual; an improved target of another rank is sent to its owner as one fire-and-forget
//           red.min on the owner's dist over the CUDA-IPC peer mapping (NVLink) and not expanded here.
// The caller puts a barrier between solve and the next seed (all peer reds have landed) and stops when a seed
// pass finds nothing on any rank.  fl(a + w) is monotone, so this schedule reaches the same least fixed point
// as shb_sssp and the reference's dense sweeps, bit for bit.  Only vertices on a block boundary ever cross
// NVLink; no collective moves distances.
// =================================================================================================
struct SsspBlockArgs {
  SsspState *st;
  const int32_t *at_row_ptr;  // n_local + 1, row r = out-edges of vertex lo + r
  const int32_t *at_col;
  const float *at_val;
  uint32_t *dist;             // full length, local
  uint32_t *peer_dist[8];     // every rank's dist (peer_dist[self] == dist)
  long long bounds[9];        // vertex ownership: rank i owns [bounds[i], bounds[i+1])
  int n_ranks, self;
  uint32_t *prev;             // n_local
  uint32_t *pile_bits;        // n_local bits: entered the far pile in this round
  uint2 *near_q[2];
  int32_t *far_q[2];
  uint32_t lo, n_local, n_total;
  uint32_t near_cap;
  uint32_t max_phases;
  const long long *gmin;  // all-reduced smallest pending distance (bits) of this round; nullptr = no window
  float window_buckets;   // the round expands distances below gmin + window_buckets * delta only
};

__device__ __forceinline__ int sssp_owner(const SsspBlockArgs &a, uint32_t v) {
  int o = 0;
  while (o + 1 < a.n_ranks && (long long)v >= a.bounds[o + 1]) ++o;
  return o;
}

// dist[v] (local copy) just dropped from `old` to cb
__device__ __forceinline__ void sssp_block_improved(const SsspBlockArgs &a, SsspShared &sh, uint32_t v, uint32_t cb, uint32_t old,
                                                    uint32_t t_hi,
                                                    int nxt, uint2 *near_next, uint32_t *near_next_size, int32_t *far_cur,
                                                    uint32_t *far_cur_size, uint32_t &fmin) {
  const uint32_t r = v - a.lo;
  if (r >= a.n_local) {  // another rank's vertex: its owner learns the candidate, nobody expands it here
    uint32_t *p = a.peer_dist[sssp_owner(a, v)] + v;
    asm volatile("red.relaxed.sys.global.min.u32 [%0], %1;" ::"l"(p), "r"(cb) : "memory");
    return;
  }
  if (cb < t_hi) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(a.at_row_ptr + r));
    const uint32_t slot = atomicAdd(&sh.lcount[nxt], 1u);
    if (slot < SSSP_CAP) {
      sh.lst[nxt][slot] = make_uint2(v, cb);
    } else {
      const uint32_t g = atomicAdd(near_next_size, 1u);
      if (g < a.near_cap) near_next[g] = make_uint2(v, cb);
      else a.st->overflow = 1u;
    }
  } else {
    fmin = min(fmin, cb);
    const uint32_t bit = 1u << (r & 31);
    uint32_t *word = &a.pile_bits[r >> 5];
    bool enter;
    if (old == 0x7f7fffffu) {  // never reached before: cannot be in the pile; mark it without waiting for the answer
      asm volatile("red.relaxed.gpu.global.or.b32 [%0], %1;" ::"l"(word), "r"(bit) : "memory");
      enter = true;
    } else if (__ldcg(word) & bit) {  // the common case of a vertex improved again while it waits in the pile
      enter = false;
    } else {
      enter = !(atomicOr(word, bit) & bit);  // first time in this round's pile
    }
    if (enter) {
      const uint32_t slot = atomicAdd(&sh.fcount, 1u);
      if (slot < SSSP_FSTAGE) sh.fstage[slot] = (int32_t)v;
      else far_cur[atomicAdd(far_cur_size, 1u)] = (int32_t)v;
    }
  }
}

// round set-up: reset the work lists (delta is kept), then every owned vertex whose distance dropped below the
// value it was last expanded with goes to far pile 0
__global__ void sssp_block_reset_kernel(SsspState *st) {
  for (int i = 0; i < 3; ++i) {
    st->near_size[i] = 0;
    st->far_size[i] = 0;
    st->far_min[i] = 0xffffffffu;
  }
  st->bar_count = st->bar_gen = 0;
  st->phases = st->buckets = st->done = st->overflow = 0;
}

__global__ void sssp_block_seed_kernel(SsspState *st, const uint32_t *dist, const uint32_t *prev, uint32_t *pile_bits,
                                       int32_t *far0, uint32_t lo, uint32_t n_local, long long *count_out, long long *min_out) {
  const int lane = threadIdx.x & 31;
  uint32_t fmin = 0xffffffffu, fmin_front = 0xffffffffu, cnt = 0;
  // the window advances with the FRONT: its lower end is the smallest pending distance at or beyond the previous
  // window's upper end; stragglers below it (corrections sent by a peer) are expanded along the way and no longer
  // hold the whole sweep back
  const uint32_t last_cap = st->cap_bits;
  const uint32_t n_padded = (n_local + 31u) & ~31u;
  for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_padded; r += gridDim.x * blockDim.x) {
    bool seed = false;
    uint32_t d = 0;
    if (r < n_local) {
      d = dist[lo + r];
      seed = d < prev[r];
    }
    const uint32_t m = __ballot_sync(0xffffffffu, seed);
    if (lane == 0) pile_bits[r >> 5] = m;  // the seeds are in the pile; everything else starts outside
    if (m) {
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(&st->far_size[0], __popc(m));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (seed) {
        far0[base + __popc(m & ((1u << lane) - 1u))] = (int32_t)(lo + r);
        fmin = min(fmin, d);
        if (d >= last_cap) fmin_front = min(fmin_front, d);
        ++cnt;
      }
    }
  }
  for (int off = 16; off > 0; off >>= 1) {
    fmin = min(fmin, __shfl_xor_sync(0xffffffffu, fmin, off));
    fmin_front = min(fmin_front, __shfl_xor_sync(0xffffffffu, fmin_front, off));
    cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
  }
  if (lane == 0 && cnt) {
    atomicMin(&st->far_min[0], fmin);
    atomicAdd(reinterpret_cast<unsigned long long *>(count_out), (unsigned long long)cnt);
    if (min_out && fmin_front != 0xffffffffu) atomicMin(min_out, (long long)fmin_front);
  }
}

template <bool HAS_VAL>
__global__ void __launch_bounds__(SSSP_THREADS, 1) sssp_block_kernel(const __grid_constant__ SsspBlockArgs a) {
  __shared__ SsspShared sh;
  SsspState *st = a.st;
  const uint32_t n_blocks = gridDim.x;
  const int lane = threadIdx.x & 31;
  const uint32_t n_threads = gridDim.x * blockDim.x;
  const float delta = *(volatile float *)&st->delta;
  uint32_t phase = 0, fgen = 0, buckets = 0;
  uint32_t t_hi = 0;  // nothing has been through the near list yet: the first phase splits the seeded pile
  // distance window of this round: vertices at or beyond cap stay unexpanded (prev > dist) and are seeded again
  uint32_t cap = 0xffffffffu;
  if (a.gmin) {
    const long long g = *(volatile const long long *)a.gmin;
    const uint32_t gb = g < 0x7f7fffffll ? (uint32_t)g : 0x7f7fffffu;
    cap = __float_as_uint(__fadd_rn(__uint_as_float(gb), a.window_buckets * delta));
    if (blockIdx.x == 0 && threadIdx.x == 0) st->cap_bits = cap;  // read by the next round's seed pass
  }
  bool finished = false;
  if (threadIdx.x == 0) {
    sh.lcount[0] = sh.lcount[1] = 0;
    sh.fcount = 0;
    sh.fmin = 0xffffffffu;
  }
  __syncthreads();
  while (phase < a.max_phases) {
    const uint32_t ns = *(volatile uint32_t *)&st->near_size[phase % 3];
    const uint2 *near_cur = a.near_q[phase & 1];
    uint2 *near_next = a.near_q[(phase + 1) & 1];
    uint32_t *near_next_size = &st->near_size[(phase + 1) % 3];
    if (blockIdx.x == 0 && threadIdx.x == 0) st->near_size[(phase + 2) % 3] = 0;
    uint32_t fmin = 0xffffffffu;
    if (ns > 0) {
      int32_t *far_cur = a.far_q[fgen & 1];
      uint32_t *far_cur_size = &st->far_size[fgen % 3];
      const uint32_t ch = max(32u, min((uint32_t)SSSP_CAP, (ns + n_blocks - 1) / n_blocks));
      for (uint32_t base = blockIdx.x * ch; base < ns; base += n_blocks * ch) {
        const uint32_t cnt = min(ch, ns - base);
        for (uint32_t j = threadIdx.x; j < cnt; j += blockDim.x) sh.lst[0][j] = near_cur[base + j];
        if (threadIdx.x == 0) sh.lcount[0] = cnt;
        int cur = 0;
        while (true) {
          __syncthreads();
          const uint32_t cn = sh.lcount[cur];
          if (cn == 0) break;
          const int nxt = cur ^ 1;
          for (uint32_t i = threadIdx.x; i < ((cn + 31u) & ~31u); i += blockDim.x) {
            int s = 0, e = 0;
            uint32_t du_bits = 0;
            if (i < cn) {
              const uint2 ent = sh.lst[cur][i];
              const uint32_t r = ent.x - a.lo;
              const uint32_t d_now = __ldcg(a.dist + ent.x);
              const int s0 = __ldg(a.at_row_ptr + r), e0 = __ldg(a.at_row_ptr + r + 1);
              if (d_now == ent.y) {  // still the vertex's distance: live entry, expanded with this value
                s = s0;
                e = e0;
                du_bits = ent.y;
                a.prev[r] = ent.y;
              }
            }
            const float du = __uint_as_float(du_bits);
            const bool big = (e - s) >= SSSP_BIG_ROW;
            if (!big) {
              for (int k = s; k < e; k += 4) {
                int32_t vv[4];
                uint32_t cb[4], old[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  vv[j] = (k + j < e) ? __ldg(a.at_col + k + j) : -1;
                  const float w = (HAS_VAL && k + j < e) ? __ldg(a.at_val + k + j) : 1.0f;
                  cb[j] = __float_as_uint(MinPlusF32::mult(du, w));
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) old[j] = ((uint32_t)vv[j] < a.n_total) ? atomicMin(a.dist + vv[j], cb[j]) : 0u;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (cb[j] < old[j])
                    sssp_block_improved(a, sh, (uint32_t)vv[j], cb[j], old[j], t_hi, nxt, near_next, near_next_size, far_cur, far_cur_size, fmin);
              }
            }
            uint32_t bigs = __ballot_sync(0xffffffffu, big);
            while (bigs) {
              const int src = __ffs(bigs) - 1;
              bigs &= bigs - 1;
              const int bs = __shfl_sync(0xffffffffu, s, src), be = __shfl_sync(0xffffffffu, e, src);
              const float bd = __uint_as_float(__shfl_sync(0xffffffffu, du_bits, src));
              for (int k = bs + lane; k < be; k += 32) {
                const int32_t v = __ldg(a.at_col + k);
                const float w = HAS_VAL ? __ldg(a.at_val + k) : 1.0f;
                const uint32_t c1 = __float_as_uint(MinPlusF32::mult(bd, w));
                if ((uint32_t)v < a.n_total && c1 < __ldcg(a.dist + v)) {
                  const uint32_t o1 = atomicMin(a.dist + v, c1);
                  if (c1 < o1) sssp_block_improved(a, sh, (uint32_t)v, c1, o1, t_hi, nxt, near_next, near_next_size, far_cur, far_cur_size, fmin);
                }
              }
            }
          }
          __syncthreads();
          if (threadIdx.x == 0) {
            sh.lcount[cur] = 0;
            sh.lcount[nxt] = min(sh.lcount[nxt], (uint32_t)SSSP_CAP);
          }
          if (sh.fcount >= SSSP_FSTAGE / 2) sssp_flush_far(sh, far_cur, far_cur_size);
          cur = nxt;
        }
      }
      __syncthreads();
      sssp_flush_far(sh, far_cur, far_cur_size);
      for (int off = 16; off > 0; off >>= 1) fmin = min(fmin, __shfl_xor_sync(0xffffffffu, fmin, off));
      if (lane == 0 && fmin != 0xffffffffu) atomicMin(&sh.fmin, fmin);
      __syncthreads();
      if (threadIdx.x == 0 && sh.fmin != 0xffffffffu) {
        atomicMin(&st->far_min[fgen % 3], sh.fmin);
        sh.fmin = 0xffffffffu;
      }
    } else {
      const uint32_t fs = *(volatile uint32_t *)&st->far_size[fgen % 3];
      const uint32_t fm = *(volatile uint32_t *)&st->far_min[fgen % 3];
      if (fs == 0 || fm >= cap) {  // nothing left, or nothing left inside this round's window
        finished = true;
        break;
      }
      const uint32_t t_old = t_hi;
      t_hi = min(__float_as_uint(__fadd_rn(__uint_as_float(fm), delta)), cap);
      ++buckets;
      const int32_t *far_cur = a.far_q[fgen & 1];
      int32_t *far_next = a.far_q[(fgen + 1) & 1];
      uint32_t *far_next_size = &st->far_size[(fgen + 1) % 3];
      if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->far_size[(fgen + 2) % 3] = 0;
        st->far_min[(fgen + 2) % 3] = 0xffffffffu;
      }
      for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ((fs + 31u) & ~31u); i += n_threads) {
        int kind = 0;
        int32_t v = 0;
        uint32_t d = 0;
        if (i < fs) {
          v = far_cur[i];
          d = __ldcg(a.dist + v);
          if (d < t_old) {
            // went through the near list in an earlier bucket -- or a peer lowered it just now, in which case
            // prev[v] is still above d and the next round's seed pass picks it up
          } else if (d < t_hi) {
            kind = 1;
          } else {
            kind = 2;
            fmin = min(fmin, d);
          }
        }
        const uint32_t m1 = __ballot_sync(0xffffffffu, kind == 1), m2 = __ballot_sync(0xffffffffu, kind == 2);
        uint32_t b1 = 0, b2 = 0;
        if (lane == 0) {
          if (m1) b1 = atomicAdd(near_next_size, __popc(m1));
          if (m2) b2 = atomicAdd(far_next_size, __popc(m2));
        }
        b1 = __shfl_sync(0xffffffffu, b1, 0);
        b2 = __shfl_sync(0xffffffffu, b2, 0);
        const uint32_t below = (1u << lane) - 1u;
        if (kind == 1) {
          const uint32_t g = b1 + __popc(m1 & below);
          if (g < a.near_cap) near_next[g] = make_uint2((uint32_t)v, d);
          else st->overflow = 1u;
        } else if (kind == 2) {
          far_next[b2 + __popc(m2 & below)] = v;
        }
      }
      for (int off = 16; off > 0; off >>= 1) fmin = min(fmin, __shfl_xor_sync(0xffffffffu, fmin, off));
      if (lane == 0 && fmin != 0xffffffffu) atomicMin(&st->far_min[(fgen + 1) % 3], fmin);
      ++fgen;
    }
    grid_barrier(st, n_blocks);
    ++phase;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->phases = phase;
    st->buckets = buckets;
    st->done = (finished && st->overflow == 0) ? 1u : 0u;
  }
}

struct SsspBlockLayout {
  size_t near_off, far_off, pile_off, total;
  uint64_t near_cap;
};
static SsspBlockLayout sssp_block_layout(int64_t n_local) {
  SsspBlockLayout L;
  L.near_cap = (uint64_t)n_local * 2 + 4096;
  L.near_off = (sizeof(SsspState) + 255) / 256 * 256;
  L.far_off = L.near_off + (size_t)L.near_cap * 2 * sizeof(uint2);
  L.pile_off = L.far_off + ((size_t)n_local * 2 * sizeof(int32_t) + 255) / 256 * 256;
  L.total = L.pile_off + ((size_t)(n_local + 31) / 32 * 4 + 255) / 256 * 256;
  return L;
}

SHB_API int shb_sssp_block_work_bytes(int64_t n_local, uint64_t *bytes) {
  SHB_REQUIRE(bytes && n_local >= 0, "bad arguments");
  *bytes = sssp_block_layout(n_local).total;
  return SHB_OK;
}

static int sssp_block_check(const shb_csr *AT, int64_t row_offset, int64_t n_total, const shb_buf *dist, const shb_buf *prev,
                            const shb_buf *work) {
  SHB_REQUIRE(AT->n_cols == n_total && n_total < (1ll << 31), "the block's column space must be the whole vertex set");
  SHB_REQUIRE(row_offset >= 0 && row_offset + AT->n_rows <= n_total, "row block outside the vertex set");
  SHB_REQUIRE(dist->bytes >= (uint64_t)n_total * 4, "dist shorter than n_total");
  SHB_REQUIRE(prev->bytes >= (uint64_t)AT->n_rows * 4, "prev shorter than the row block");
  SHB_REQUIRE(work->bytes >= sssp_block_layout(AT->n_rows).total, "work shorter than shb_sssp_block_work_bytes");
  return SHB_OK;
}

SHB_API int shb_sssp_block_init(shb_ctx *c, const shb_csr *AT, int64_t row_offset, int64_t n_total, int64_t source,
                                shb_buf *dist, shb_buf *prev, shb_buf *work) {
  SHB_REQUIRE(c && AT && dist && prev && work, "bad arguments");
  SHB_TRY(sssp_block_check(AT, row_offset, n_total, dist, prev, work));
  SHB_REQUIRE(source >= 0 && source < n_total, "source out of range");
  SHB_CUDA(cudaSetDevice(c->device));
  SHB_TRY(shb_buf_fill_u32(c, dist, 0, (uint64_t)n_total, 0x7f7fffffu));
  if (AT->n_rows > 0) SHB_TRY(shb_buf_fill_u32(c, prev, 0, (uint64_t)AT->n_rows, 0x7f7fffffu));
  SsspState *st = (SsspState *)work->ptr;
  SHB_CUDA(cudaMemsetAsync(st, 0, sizeof(SsspState), c->stream));
  if (AT->val && AT->nnz > 0) {
    const int64_t stride = std::max<int64_t>(1, AT->nnz / (1 << 22));
    sssp_weight_stats_kernel<<<148, 256, 0, c->stream>>>(st, (const float *)AT->val, AT->nnz, stride);
  }
  float c_scale = 128.0f, delta_override = 0.0f;
  if (const char *e = getenv("SHB_SSSP_C")) c_scale = (float)atof(e);
  if (const char *e = getenv("SHB_SSSP_DELTA")) delta_override = (float)atof(e);
  // sssp_init_kernel also plants the source: dist[source] = 0 on EVERY rank (only the owner will seed it) and
  // writes one near entry that the first round's reset discards
  const SsspBlockLayout L = sssp_block_layout(AT->n_rows);
  sssp_init_kernel<<<1, 1, 0, c->stream>>>(st, (uint32_t *)dist->ptr, (uint2 *)((char *)work->ptr + L.near_off), source,
                                           AT->n_rows > 0 ? (double)AT->nnz / (double)AT->n_rows : 1.0, c_scale, delta_override);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

// first half of a round: reset the work lists and collect the owned vertices whose distance dropped below the
// value they were last expanded with.  Their number is ADDED to int64 word `word` of seed_count, their smallest
// distance (bits) is MIN-ed into word `word` of seed_min (optional; pre-set to a large value by the caller).
SHB_API int shb_sssp_block_seed(shb_ctx *c, const shb_csr *AT, int64_t row_offset, int64_t n_total, const shb_buf *dist,
                                const shb_buf *prev, shb_buf *work, shb_buf *seed_count, shb_buf *seed_min, uint64_t word) {
  SHB_REQUIRE(c && AT && dist && prev && work && seed_count, "bad arguments");
  SHB_TRY(sssp_block_check(AT, row_offset, n_total, dist, prev, work));
  SHB_REQUIRE(seed_count->bytes >= (word + 1) * 8 && (!seed_min || seed_min->bytes >= (word + 1) * 8), "seed buffers too short");
  SHB_CUDA(cudaSetDevice(c->device));
  const SsspBlockLayout L = sssp_block_layout(AT->n_rows);
  char *w = (char *)work->ptr;
  SsspState *st = (SsspState *)w;
  sssp_block_reset_kernel<<<1, 1, 0, c->stream>>>(st);
  if (AT->n_rows > 0)
    sssp_block_seed_kernel<<<grid_for(AT->n_rows), 256, 0, c->stream>>>(
        st, (const uint32_t *)dist->ptr, (const uint32_t *)prev->ptr, (uint32_t *)(w + L.pile_off), (int32_t *)(w + L.far_off),
        (uint32_t)row_offset, (uint32_t)AT->n_rows, (long long *)seed_count->ptr + word,
        seed_min ? (long long *)seed_min->ptr + word : nullptr);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

// second half: the near-far solve of the block.  window_buckets > 0 (needs seed_min, all-reduced MIN by the caller):
// only distances below seed_min[word] + window_buckets * delta are expanded in this round; vertices beyond stay
// pending and are seeded again, so that all ranks advance through the distance range together.
SHB_API int shb_sssp_block_solve(shb_ctx *c, const shb_csr *AT, int64_t row_offset, int64_t n_total, shb_buf *dist,
                                 shb_buf *const *peer_dist, int n_ranks, int self_index, const int64_t *bounds, shb_buf *prev,
                                 shb_buf *work, const shb_buf *seed_min, uint64_t word, float window_buckets,
                                 uint32_t max_phases) {
  SHB_REQUIRE(c && AT && dist && peer_dist && bounds && prev && work, "bad arguments");
  SHB_REQUIRE(n_ranks >= 1 && n_ranks <= 8 && self_index >= 0 && self_index < n_ranks, "1..8 ranks, one of them local");
  SHB_TRY(sssp_block_check(AT, row_offset, n_total, dist, prev, work));
  SHB_REQUIRE(bounds[0] == 0 && bounds[n_ranks] == n_total && bounds[self_index] == row_offset &&
                  bounds[self_index + 1] == row_offset + AT->n_rows, "bounds do not describe this block");
  SHB_REQUIRE(!(window_buckets > 0.f) || (seed_min && seed_min->bytes >= (word + 1) * 8), "a window needs the all-reduced seed_min");
  SHB_REQUIRE(peer_dist[self_index] && peer_dist[self_index]->ptr == dist->ptr, "peer_dist[self_index] must be the local dist");
  SHB_CUDA(cudaSetDevice(c->device));
  const SsspBlockLayout L = sssp_block_layout(AT->n_rows);
  char *w = (char *)work->ptr;
  SsspBlockArgs a;
  memset(&a, 0, sizeof a);
  a.st = (SsspState *)w;
  a.at_row_ptr = AT->row_ptr;
  a.at_col = AT->col;
  a.at_val = (const float *)AT->val;
  a.dist = (uint32_t *)dist->ptr;
  for (int i = 0; i < n_ranks; ++i) {
    SHB_REQUIRE(peer_dist[i] && peer_dist[i]->bytes >= (uint64_t)n_total * 4, "a peer's dist is too short");
    SHB_REQUIRE(bounds[i] <= bounds[i + 1], "bounds not monotone");
    a.peer_dist[i] = (uint32_t *)peer_dist[i]->ptr;
    a.bounds[i] = bounds[i];
  }
  a.bounds[n_ranks] = bounds[n_ranks];
  a.n_ranks = n_ranks;
  a.self = self_index;
  a.prev = (uint32_t *)prev->ptr;
  a.pile_bits = (uint32_t *)(w + L.pile_off);
  a.near_q[0] = (uint2 *)(w + L.near_off);
  a.near_q[1] = a.near_q[0] + L.near_cap;
  a.far_q[0] = (int32_t *)(w + L.far_off);
  a.far_q[1] = a.far_q[0] + AT->n_rows;
  a.lo = (uint32_t)row_offset;
  a.n_local = (uint32_t)AT->n_rows;
  a.n_total = (uint32_t)n_total;
  a.near_cap = (uint32_t)std::min<uint64_t>(L.near_cap, 0xffffffffu);
  a.max_phases = max_phases ? max_phases : 0xffffffffu;
  a.gmin = window_buckets > 0.f ? (const long long *)seed_min->ptr + word : nullptr;
  a.window_buckets = window_buckets;
  const bool has_val = AT->val != nullptr;
  const int blocks = c->prop.multiProcessorCount;  // one 512-thread CTA per SM, co-resident by construction
  void *args[] = {&a};
  if (has_val)
    SHB_CUDA(cudaLaunchCooperativeKernel((void *)sssp_block_kernel<true>, dim3(blocks), dim3(SSSP_THREADS), args, 0, c->stream));
  else
    SHB_CUDA(cudaLaunchCooperativeKernel((void *)sssp_block_kernel<false>, dim3(blocks), dim3(SSSP_THREADS), args, 0, c->stream));
  return SHB_OK;
}

// after the last round: solver status of the last round + (reached, sum of row lengths of reached) of the owned
// slice into stats_out[0..1] (int64).  `A_rows` supplies the TEPS degrees (the block's rows of A; for a symmetric
// graph the same operator as AT).
SHB_API int shb_sssp_block_finish(shb_ctx *c, const shb_csr *A_rows, int64_t row_offset, const shb_buf *dist,
                                  const shb_buf *work, shb_buf *stats_out) {
  SHB_REQUIRE(c && A_rows && dist && work && stats_out, "bad arguments");
  SHB_REQUIRE(dist->bytes >= (uint64_t)(row_offset + A_rows->n_rows) * 4 && stats_out->bytes >= 16, "buffers too short");
  SHB_REQUIRE(work->bytes >= sizeof(SsspState), "work too short");
  SHB_CUDA(cudaSetDevice(c->device));
  SHB_CUDA(cudaMemsetAsync(stats_out->ptr, 0, 16, c->stream));
  if (A_rows->n_rows > 0)
    reached_stats_kernel<<<grid_for(A_rows->n_rows), 256, 0, c->stream>>>((const uint32_t *)dist->ptr + row_offset, 0x7f7fffffu,
                                                                         A_rows->row_ptr, A_rows->n_rows,
                                                                         (unsigned long long *)stats_out->ptr);
  SHB_CUDA(cudaGetLastError());
  SsspState h;
  SHB_CUDA(cudaMemcpyAsync(&h, work->ptr, sizeof h, cudaMemcpyDeviceToHost, c->stream));
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  if (h.overflow) {
    shb_set_error("near list overflow in a partitioned SSSP round");
    return SHB_ERR_CUDA;
  }
  if (!h.done) {
    shb_set_error("phase cap reached before the block's worklists emptied");
    return SHB_ERR_NOT_CONVERGED;
  }
  return SHB_OK;
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
  SHB_CUDA(cudaSetDevice(b->device));
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
  b->device = c->device;
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
