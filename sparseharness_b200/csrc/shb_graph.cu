// shb_graph.cu -- device-resident graph drivers on the semirings + the multi-GPU scatter entry points.
//
// Reference counterparts (all host-driven, one blocking read-back of the whole vector per step):
//   app/bfs.cpp:94-152  HarnessBFS::executeRun    -> shb_bfs      (or-and, integer levels; Q8 decision) in shb_bfs.cu
//   app/sssp.cpp:97-155 HarnessSSSP::executeRun   -> shb_sssp     (min-plus to the exact fixed point)
//   app/pr.cpp:96-154   HarnessPR::executeRun     -> shb_pagerank (plus-times, canonical mode; Q9)
// Here the frontier, the vectors and the termination state never leave the device: the host
// enqueues a batch of steps and looks at one small state record per batch.
#include <algorithm>

#include "ldst.cuh"
#include "semiring.cuh"
#include "shb_internal.cuh"


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
    // (the size counter keeps counting past the list's capacity when it overflows -- the entries beyond it were dropped
    //  and `overflow` is set: never read past the end; the host starts over with a longer list or reports the overflow)
    const uint32_t ns = min(*(volatile uint32_t *)&st->near_size[phase % 3], a.near_cap);
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
        // every thread must have read the zero count that ended the loop before thread 0 refills
        // lcount[0] / lst[0] for the next chunk (cur can be 0 here): without this a slow warp could see
        // the new count, skip the break and run one barrier out of phase with its CTA
        __syncthreads();
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
      {  // far_min + delta, but never == far_min (delta below half an ulp of far_min would re-split the
         // same pile for ever): non-negative float bits order like integers, so +1 is the next float
        const uint32_t fmb = *(volatile uint32_t *)&st->far_min[fgen % 3];
        t_hi = max(__float_as_uint(__fadd_rn(__uint_as_float(fmb), delta)), fmb < 0x7f7fffffu ? fmb + 1u : fmb);
      }
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

// one attempt with a global near list of `near_cap` entries; *overflowed says the list was too small (dist is then
// lowered but not fully expanded: the caller starts over with a larger list)
static int sssp_attempt(shb_ctx *c, const shb_csr *A, const shb_csr *AT_in, int64_t source, shb_buf *dist, uint32_t max_rounds,
                        shb_graph_stats *stats, uint64_t near_cap, bool *overflowed);

SHB_API int shb_sssp(shb_ctx *c, const shb_csr *A, const shb_csr *AT_in, int64_t source, shb_buf *dist, uint32_t max_rounds,
                     shb_graph_stats *stats) {
  SHB_REQUIRE(c && A && dist, "bad arguments");
  // The global near list takes one entry per successful improvement beyond a CTA's shared list; a vertex can be improved
  // many times in a phase, so on dense graphs with few vertices n + 1024 entries can be too few.  An overflow is not an
  // error of the input: start over with a list four times as long (at most n + nnz + 1024: one entry per edge relaxation
  // of a phase).
  uint64_t near_cap = (uint64_t)std::max<int64_t>(A->n_rows, 0) + 1024;
  const uint64_t cap_max = std::min<uint64_t>((uint64_t)A->n_rows + (uint64_t)A->nnz + 1024, 0xfffffff0ull);
  for (;;) {
    bool overflowed = false;
    const int status = sssp_attempt(c, A, AT_in, source, dist, max_rounds, stats, near_cap, &overflowed);
    if (!overflowed || near_cap >= cap_max) return status;
    near_cap = std::min<uint64_t>(near_cap * 4, cap_max);
  }
}

static int sssp_attempt(shb_ctx *c, const shb_csr *A, const shb_csr *AT_in, int64_t source, shb_buf *dist, uint32_t max_rounds,
                        shb_graph_stats *stats, uint64_t near_cap, bool *overflowed) {
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
      sssp_weight_stats_kernel<<<c->prop.multiProcessorCount, 256, 0, c->stream>>>(st, (const float *)AT->val, AT->nnz, stride);
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
      *overflowed = true;
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
    cap = max(__float_as_uint(__fadd_rn(__uint_as_float(gb), a.window_buckets * delta)), gb < 0x7f7fffffu ? gb + 1u : gb);
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
    // (the size counter keeps counting past the list's capacity when it overflows -- the entries beyond it were dropped
    //  and `overflow` is set: never read past the end; the host starts over with a longer list or reports the overflow)
    const uint32_t ns = min(*(volatile uint32_t *)&st->near_size[phase % 3], a.near_cap);
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
        // every thread must have read the zero count that ended the loop before thread 0 refills
        // lcount[0] / lst[0] for the next chunk (cur can be 0 here): without this a slow warp could see
        // the new count, skip the break and run one barrier out of phase with its CTA
        __syncthreads();
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
      t_hi = min(max(__float_as_uint(__fadd_rn(__uint_as_float(fm), delta)), fm < 0x7f7fffffu ? fm + 1u : fm), cap);
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
    sssp_weight_stats_kernel<<<c->prop.multiProcessorCount, 256, 0, c->stream>>>(st, (const float *)AT->val, AT->nnz, stride);
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
  // x larger than the L2: the sweeps run over column panels (shb_panels.cu), built once and kept on the operator
  if (A->n_panels == 0 && (uint64_t)A->n_cols * 4 > shb_panel_threshold_bytes(c) && (!cfg || cfg->variant == SHB_VARIANT_AUTO))
    SHB_TRY(shb_csr_build_panels(c, const_cast<shb_csr *>(A), 0, nullptr));
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

// =================================================================================================
// Power iteration (the eigenvector app, plus-times; reference: app/eigenvector.cpp:42-137 -- dead code there:
// x_{k+1} = A x_k with a cap of 10 launches and no normalisation, which shb_iterate with max_iters = 10 reproduces).
// Canonical driver: out = A x_k; s_k = max_i |out_i|; x_{k+1} = out / s_k; stop when all |x_{k+1} - x_k| < delta.
// Two small kernels follow each SpMV: the infinity norm (float bits of |v| order like integers: one atomicMax
// per warp), and ONE fused pass that scales, stores and runs the convergence test.  Everything stays on the device;
// the host reads two words per iteration batch.
// =================================================================================================
__global__ void absmax_kernel(const float *__restrict__ v, int64_t n, uint32_t *out_bits) {
  uint32_t m = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    m = max(m, __float_as_uint(fabsf(v[i])));
  for (int off = 16; off > 0; off >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, off));
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out_bits, m);
}

__global__ void scale_test_kernel(float *__restrict__ out, const float *__restrict__ x_old, int64_t n, const uint32_t *norm_bits,
                                  double delta, uint32_t *changed) {
  const float s = __uint_as_float(*norm_bits);
  bool differs = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float o = s != 0.0f ? __fdiv_rn(out[i], s) : out[i];
    out[i] = o;
    differs |= shb_row_differs(x_old[i], o, delta, SHB_TERM_ABSDIFF);
  }
  if (__any_sync(0xffffffffu, differs) && (threadIdx.x & 31) == 0 && shb_ld_volatile_u32(changed) == 0) *changed = 1u;
}

SHB_API int shb_power_iteration(shb_ctx *c, const shb_csr *A, const shb_launch_cfg *cfg, shb_buf *x, shb_buf *work, double delta,
                                uint32_t max_iters, float *eigenvalue, int *result_in_work, shb_graph_stats *stats) {
  SHB_REQUIRE(c && A && x && work && result_in_work && max_iters >= 1, "bad arguments");
  SHB_REQUIRE(A->n_rows == A->n_cols, "Matrix is not square. Failing computation.");
  SHB_REQUIRE(x->bytes >= (uint64_t)A->n_rows * 4 && work->bytes >= (uint64_t)A->n_rows * 4, "vector shorter than n_rows");
  SHB_REQUIRE(x->ptr != work->ptr, "x and work must differ");
  SHB_CUDA(cudaSetDevice(c->device));
  const int64_t n = A->n_rows;
  const float alpha = 1.0f, beta = 0.0f;
  void *bufs[2] = {x->ptr, work->ptr};
  int in = 0;
  uint32_t done = 0;
  bool converged = false;
  uint32_t norm_bits = 0;
  IterCtl none;
  memset(&none, 0, sizeof none);
  const unsigned grid = grid_for(n, 256, c->prop.multiProcessorCount * 8);
  SHB_CUDA(cudaEventRecord(c->ev0, c->stream));
  while (done < max_iters && !converged) {
    // words 0 / 1 of the context's flag scratch: the norm bits and the "some element moved" flag
    SHB_CUDA(cudaMemsetAsync(c->d_flags, 0, 2 * sizeof(uint32_t), c->stream));
    void *outs[1] = {bufs[in ^ 1]};
    SHB_TRY(shb_launch_spmv(c, A, SHB_PLUS_TIMES_F32, cfg, bufs[in], bufs[in], &alpha, &beta, outs, 1, 0, none));
    absmax_kernel<<<grid, 256, 0, c->stream>>>((const float *)bufs[in ^ 1], n, c->d_flags);
    scale_test_kernel<<<grid, 256, 0, c->stream>>>((float *)bufs[in ^ 1], (const float *)bufs[in], n, c->d_flags, delta, c->d_flags + 1);
    SHB_CUDA(cudaGetLastError());
    SHB_CUDA(cudaMemcpyAsync(c->h_flags, c->d_flags, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    SHB_CUDA(cudaStreamSynchronize(c->stream));
    norm_bits = c->h_flags[0];
    converged = c->h_flags[1] == 0;
    in ^= 1;
    ++done;
  }
  SHB_CUDA(cudaEventRecord(c->ev1, c->stream));
  SHB_CUDA(cudaEventSynchronize(c->ev1));
  float ms = 0.f;
  SHB_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  *result_in_work = in;
  if (eigenvalue) memcpy(eigenvalue, &norm_bits, 4);
  if (stats) {
    stats->iterations = done;
    stats->converged = converged;
    stats->edges_traversed = (int64_t)A->nnz * (int64_t)done;
    stats->reached = n;
    stats->total_ns = (uint64_t)((double)ms * 1.0e6 + 0.5);
    stats->launches = (uint64_t)done * 3;
  }
  return converged ? SHB_OK : SHB_ERR_NOT_CONVERGED;
}
