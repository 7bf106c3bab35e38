// pb.cuh -- the BLOCKED schedule: a propagation-blocked semiring SpMV for gather-bound operators.
//
// Why: on an operator whose columns are spread over a vector much larger than L1 (BASELINE.json config 2:
// 2^28 uniform-random gathers into a 64 MB x) every gather of the row-major schedules (spmv.cuh) is its
// own L1TEX->XBAR request and the kernel sits on the SM's one-request-per-cycle limit at a third of the
// HBM roofline (profiles/r01_summary.md).  Here no thread ever gathers from or scatters to global memory.
// ONE persistent kernel, one CTA per SM, two warp-specialised roles per CTA:
//
//   producer warps  CTA p stages x[slice(pass, p)] (W <= 40960 columns) in shared memory once per pass,
//                   streams the non-zeros whose column falls in that slice -- stored column-slice-major
//                   as (16-bit local column, value) -- with 128-bit loads, gathers x from SHARED memory
//                   and writes the products mult(x, a) as a coalesced stream into a ring of DEPTH row
//                   panels that stays resident in the 126 MB L2;
//   consumer warps  CTA s owns RS consecutive rows of every row panel: it reads, from the ring, the one
//                   contiguous piece every producer wrote for those rows, drops each product into its
//                   precomputed slot of a row-sorted shared-memory buffer (plain st.shared, no atomics),
//                   then reduces each row's slots in order and runs the usual fused epilogue /
//                   convergence test / multi-destination store on the last pass.
//
// Producers and consumers of different CTAs meet through per-panel counters in global memory
// (release/acquire); a producer may run DEPTH panels ahead of the slowest consumer.  Every global access
// is a coalesced stream and the products never reach HBM, so DRAM traffic stays at the operator's
// 8 bytes per non-zero.  Entry order of the operator copy: (pass, panel, producer p, consumer s, row,
// column), built once per operator on the device (shb_pb.cu).  Row sums are formed in a fixed order
// (pass, then slice, then column): deterministic for every semiring, bit-exact for min-plus / or-and /
// max-min and for integer-valued floats, within the stated 1e-5 relative tolerance for general floats.
#pragma once
#include "spmv.cuh"

struct PbGeom {
  int n_prod;    // column slices per pass == consumers per row panel == CTAs (the SM count)
  int n_pass;    // passes over the column space
  int n_panel;   // row panels
  int W;         // columns per slice (multiple of 4, <= 65536)
  int RS;        // rows per consumer
  int cbuf_cap;  // entries of the consumer's row-sorted buffer (>= the largest consumer unit)
  int slot_cap;  // entries per ring slot (>= the largest (pass, panel) group, padding included)
  int depth;     // ring slots
};

struct shb_pb {
  PbGeom g;
  int64_t cap = 0;      // entries including padding (every unit starts at a multiple of 8)
  int64_t n_valid = 0;  // stored entries with an in-range column
  int max_unit = 0;     // largest consumer unit (entries of one (pass, panel, s))
  uint16_t *lcol = nullptr;
  void *val = nullptr;
  uint16_t *slot = nullptr;     // [cap] position of the entry in its consumer's row-sorted buffer
  int32_t *unit_ptr = nullptr;  // [n_pass * n_panel * n_prod + 1] padded start of unit (pass, panel, p)
  int2 *cons_desc = nullptr;    // [(pass, panel, s, p)] {start, length} of piece (pass, panel, p, s)
  uint16_t *crp = nullptr;      // [(pass, panel, s)][RS + 1] local row offsets into the row-sorted buffer
  void *ring = nullptr;         // [depth * slot_cap] the product ring
  void *partial = nullptr;      // [n_rows] row sums carried from pass to pass
  uint32_t *sync = nullptr;     // ready[n_groups], consumed[n_groups], exit counter
  int ring_depth_alloc = 0;
  uint64_t device_bytes = 0;
};

#define PB_THREADS 1024
#define PB_PROD_THREADS 512
#define PB_CONS_THREADS (PB_THREADS - PB_PROD_THREADS)
#define PB_MAX_W 40960
#define PB_SMEM_LIMIT (227 * 1024)
#define PB_ROLE_PRODUCE 1
#define PB_ROLE_CONSUME 2

template <class T> struct PbArgs {
  PbGeom g;
  const uint16_t *lcol;
  const T *val;  // nullptr = pattern
  const uint16_t *slot;
  const int32_t *unit_ptr;
  const int2 *cons_desc;
  const uint16_t *crp;
  T *ring;
  T *partial;
  uint32_t *ready;     // [n_groups]
  uint32_t *consumed;  // [n_groups]
  uint32_t *exit_count;
  int roles;          // PB_ROLE_* mask (both = the fused kernel; one at a time = the two-launch debug mode)
  int reset_sync;     // zero the counters when the last CTA leaves
};

__device__ __forceinline__ void pb_bar(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ uint32_t pb_ld_acquire(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void pb_red_release(uint32_t *p) {
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}
__device__ __forceinline__ void pb_prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void pb_bulk_prefetch_l2(const void *p, uint32_t bytes) {  // 16-byte aligned, multiple of 16
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t pb_ld_cg_b32(const void *p) {
  uint32_t v;
  asm volatile("ld.global.cg.b32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

// 8 products of one 16-byte group of local columns
template <class S, bool HAS_VAL>
__device__ __forceinline__ void pb_mult8(const typename S::T *xs, const uint4 &lc, const uint4 &v0, const uint4 &v1,
                                         uint4 &o0, uint4 &o1) {
  typedef typename S::T T;
  const T x0 = xs[lc.x & 0xffffu], x1 = xs[lc.x >> 16], x2 = xs[lc.y & 0xffffu], x3 = xs[lc.y >> 16];
  const T x4 = xs[lc.z & 0xffffu], x5 = xs[lc.z >> 16], x6 = xs[lc.w & 0xffffu], x7 = xs[lc.w >> 16];
  o0.x = shb_bits<T>(S::mult(x0, HAS_VAL ? shb_from_bits<T>(v0.x) : S::one()));
  o0.y = shb_bits<T>(S::mult(x1, HAS_VAL ? shb_from_bits<T>(v0.y) : S::one()));
  o0.z = shb_bits<T>(S::mult(x2, HAS_VAL ? shb_from_bits<T>(v0.z) : S::one()));
  o0.w = shb_bits<T>(S::mult(x3, HAS_VAL ? shb_from_bits<T>(v0.w) : S::one()));
  o1.x = shb_bits<T>(S::mult(x4, HAS_VAL ? shb_from_bits<T>(v1.x) : S::one()));
  o1.y = shb_bits<T>(S::mult(x5, HAS_VAL ? shb_from_bits<T>(v1.y) : S::one()));
  o1.z = shb_bits<T>(S::mult(x6, HAS_VAL ? shb_from_bits<T>(v1.z) : S::one()));
  o1.w = shb_bits<T>(S::mult(x7, HAS_VAL ? shb_from_bits<T>(v1.w) : S::one()));
}

// ---------------------------------------------------------------------------------------------
// producer role: PB_PROD_THREADS threads, named barrier 1
// ---------------------------------------------------------------------------------------------
template <class S, bool HAS_VAL>
__device__ __forceinline__ void pb_produce(const PbArgs<typename S::T> &c, const SpmvArgs<typename S::T> &a,
                                           typename S::T *xs, int tid) {
  typedef typename S::T T;
  const int p = blockIdx.x;
  const uint64_t pol_stream = shb_policy_evict_first();
  const uint32_t zbits = shb_bits<T>(S::zero());
  const bool x_aligned = ((reinterpret_cast<uintptr_t>(a.x) & 15u) == 0);
  for (int pass = 0; pass < c.g.n_pass; ++pass) {
    // the slice of x; columns past the end of the vector read as the semiring zero
    const int64_t c0 = ((int64_t)pass * c.g.n_prod + p) * c.g.W;
    for (int i = tid * 4; i < c.g.W; i += PB_PROD_THREADS * 4) {
      const int64_t cc = c0 + i;
      uint4 v;
      if (x_aligned && cc + 3 < (int64_t)a.vlength) {
        v = shb_ld_stream_v4<true>(a.x + cc, true, zbits, pol_stream);
      } else {
        v.x = cc + 0 < (int64_t)a.vlength ? shb_bits<T>(__ldg(a.x + cc + 0)) : zbits;
        v.y = cc + 1 < (int64_t)a.vlength ? shb_bits<T>(__ldg(a.x + cc + 1)) : zbits;
        v.z = cc + 2 < (int64_t)a.vlength ? shb_bits<T>(__ldg(a.x + cc + 2)) : zbits;
        v.w = cc + 3 < (int64_t)a.vlength ? shb_bits<T>(__ldg(a.x + cc + 3)) : zbits;
      }
      *reinterpret_cast<uint4 *>(xs + i) = v;
    }
    pb_bar(1, PB_PROD_THREADS);
    for (int panel = 0; panel < c.g.n_panel; ++panel) {
      const int g = pass * c.g.n_panel + panel;
      const int u = g * c.g.n_prod + p;
      const int b = __ldg(c.unit_ptr + u), e = __ldg(c.unit_ptr + u + 1);
      const int gstart = __ldg(c.unit_ptr + g * c.g.n_prod);
      if (g >= c.g.depth) {  // the ring slot must have been drained by every consumer
        if (tid == 0)
          while (pb_ld_acquire(c.consumed + (g - c.g.depth)) < (uint32_t)c.g.n_prod) __nanosleep(64);
        pb_bar(1, PB_PROD_THREADS);
      }
      T *ring = c.ring + (int64_t)(g % c.g.depth) * c.g.slot_cap - gstart;  // ring[i] for the global position i
      for (int i = b + tid * 8; i < e; i += PB_PROD_THREADS * 16) {
        const int i2 = i + PB_PROD_THREADS * 8;
        const bool p2 = i2 < e;
        if ((tid & 31) == 0) {  // pull this warp's part of the stream, two steps ahead, into L2
          const int ip = i + PB_PROD_THREADS * 32;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int ib = ip + h * PB_PROD_THREADS * 8;
            const int n = min(256, e - ib);  // entries of this warp's 16-byte groups (multiple of 8)
            if (n > 0) {
              pb_bulk_prefetch_l2(c.lcol + ib, (uint32_t)n * 2u);
              if (HAS_VAL) pb_bulk_prefetch_l2(c.val + ib, (uint32_t)n * 4u);
            }
          }
        }
        const uint4 lcA = shb_ld_stream_v4<true>(c.lcol + i, true, 0u, pol_stream);
        const uint4 lcB = shb_ld_stream_v4<true>(c.lcol + i2, p2, 0u, pol_stream);
        uint4 vA0 = make_uint4(0, 0, 0, 0), vA1 = vA0, vB0 = vA0, vB1 = vA0;
        if (HAS_VAL) {
          vA0 = shb_ld_stream_v4<true>(c.val + i, true, 0u, pol_stream);
          vA1 = shb_ld_stream_v4<true>(c.val + i + 4, true, 0u, pol_stream);
          vB0 = shb_ld_stream_v4<true>(c.val + i2, p2, 0u, pol_stream);
          vB1 = shb_ld_stream_v4<true>(c.val + i2 + 4, p2, 0u, pol_stream);
        }
        uint4 oA0, oA1, oB0, oB1;
        pb_mult8<S, HAS_VAL>(xs, lcA, vA0, vA1, oA0, oA1);
        pb_mult8<S, HAS_VAL>(xs, lcB, vB0, vB1, oB0, oB1);
        *reinterpret_cast<uint4 *>(ring + i) = oA0;
        *reinterpret_cast<uint4 *>(ring + i + 4) = oA1;
        if (p2) {
          *reinterpret_cast<uint4 *>(ring + i2) = oB0;
          *reinterpret_cast<uint4 *>(ring + i2 + 4) = oB1;
        }
      }
      pb_bar(1, PB_PROD_THREADS);  // every producer thread's stores are ordered before the release below
      if (tid == 0) pb_red_release(c.ready + g);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// consumer role: PB_CONS_THREADS threads, named barrier 2.  K = 32-entry steps of a piece that are
// issued together (Q = 12 / K pieces in flight per warp); longer pieces finish in a scalar tail loop.
// Per unit: [poll ready] bar | scatter ring -> row-sorted buffer | bar | reduce rows, store | ...
// The next unit's piece descriptors and row offsets are fetched while this one is scattered.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void pb_cp_async16(void *smem_dst, const void *gmem_src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void pb_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void pb_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// 16-bit words of one unit's row metadata: [RS] row lengths then [ceil(RS / 32)] group bases, a 16-byte multiple
__host__ __device__ __forceinline__ int pb_crp_stride(int RS) { return (RS + (RS + 31) / 32 + 7) & ~7; }

__device__ __forceinline__ uint4 pb_ld_cg_v4(const void *p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// four products into their slots of the row-sorted buffer (0xffff = padding entry)
template <class T> __device__ __forceinline__ void pb_scatter4(T *cbuf, const uint4 &pr, const uint2 &sl) {
  const uint32_t s0 = sl.x & 0xffffu, s1 = sl.x >> 16, s2 = sl.y & 0xffffu, s3 = sl.y >> 16;
  if (s0 != 0xffffu) cbuf[s0] = shb_from_bits<T>(pr.x);
  if (s1 != 0xffffu) cbuf[s1] = shb_from_bits<T>(pr.y);
  if (s2 != 0xffffu) cbuf[s2] = shb_from_bits<T>(pr.z);
  if (s3 != 0xffffu) cbuf[s3] = shb_from_bits<T>(pr.w);
}

// K = 128-entry steps of a piece issued together, Q = 4 / K pieces in flight per warp
template <class S, int K>
__device__ __forceinline__ void pb_consume(const PbArgs<typename S::T> &c, const SpmvArgs<typename S::T> &a,
                                           typename S::T *cbuf, uint16_t *crp_s, int tid, bool &differs) {
  typedef typename S::T T;
  constexpr int Q = 4 / K;
  constexpr int NW = PB_CONS_THREADS / 32;
  const int s = blockIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int n_units = c.g.n_pass * c.g.n_panel;
  const int stride = pb_crp_stride(c.g.RS);
  const int ngr = (c.g.RS + 31) / 32;
  const int cnt = (c.g.n_prod - warp + NW - 1) / NW;  // pieces of this warp: p = warp + l * NW
  const int my_p = warp + lane * NW;

  // prologue: unit 0's descriptors and row metadata
  int2 d_next = make_int2(0, 0);
  if (my_p < c.g.n_prod) d_next = __ldg(c.cons_desc + (int64_t)s * c.g.n_prod + my_p);
  {
    const uint16_t *src = c.crp + (int64_t)s * stride;
    for (int i = tid * 8; i < stride; i += PB_CONS_THREADS * 8) pb_cp_async16(crp_s + i, src + i);
    pb_cp_async_commit();
  }
  for (int g = 0; g < n_units; ++g) {
    const int pass = g / c.g.n_panel, panel = g - pass * c.g.n_panel;
    const int2 d = d_next;
    if (tid == 0)
      while (pb_ld_acquire(c.ready + g) < (uint32_t)c.g.n_prod) __nanosleep(32);
    pb_bar(2, PB_CONS_THREADS);  // ring slot g is complete; everyone is done reducing unit g - 1
    if (g + 1 < n_units) {
      const int64_t gs1 = (int64_t)(g + 1) * c.g.n_prod + s;
      if (my_p < c.g.n_prod) d_next = __ldg(c.cons_desc + gs1 * c.g.n_prod + my_p);
      const uint16_t *src = c.crp + gs1 * stride;
      uint16_t *dst = crp_s + ((g + 1) & 1) * stride;
      for (int i = tid * 8; i < stride; i += PB_CONS_THREADS * 8) pb_cp_async16(dst + i, src + i);
      if (g + 1 >= c.g.n_panel) {  // the next unit adds to row sums carried in global memory
        const int np = (g + 1) % c.g.n_panel;
        const int64_t r1 = ((int64_t)np * c.g.n_prod + s) * c.g.RS;
        for (int i = tid * 32; i < c.g.RS; i += PB_CONS_THREADS * 32)
          if (r1 + i < a.n_rows) pb_prefetch_l2(c.partial + r1 + i);
      }
    }
    pb_cp_async_commit();
    const int gstart = __ldg(c.unit_ptr + g * c.g.n_prod);
    const T *ring = c.ring + (int64_t)(g % c.g.depth) * c.g.slot_cap - gstart;
    for (int j0 = 0; j0 < cnt; j0 += Q) {
      int st[Q], ln[Q];
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        st[q] = __shfl_sync(0xffffffffu, d.x, (j0 + q) & 31);
        ln[q] = __shfl_sync(0xffffffffu, d.y, (j0 + q) & 31);
        if (j0 + q >= cnt) ln[q] = 0;
      }
      uint4 pr[Q][K];
      uint2 sl[Q][K];
#pragma unroll
      for (int q = 0; q < Q; ++q)
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const int i = (k * 32 + lane) * 4;  // pieces start at multiples of 4 and are padded to multiples of 4
          sl[q][k] = make_uint2(0xffffffffu, 0xffffffffu);
          pr[q][k] = make_uint4(0u, 0u, 0u, 0u);
          if (i < ln[q]) {
            pr[q][k] = pb_ld_cg_v4(ring + st[q] + i);
            sl[q][k] = __ldg(reinterpret_cast<const uint2 *>(c.slot + st[q] + i));
          }
        }
#pragma unroll
      for (int q = 0; q < Q; ++q)
#pragma unroll
        for (int k = 0; k < K; ++k) pb_scatter4<T>(cbuf, pr[q][k], sl[q][k]);
#pragma unroll
      for (int q = 0; q < Q; ++q)
        for (int i = (K * 32 + lane) * 4; i < ln[q]; i += 128)
          pb_scatter4<T>(cbuf, pb_ld_cg_v4(ring + st[q] + i), __ldg(reinterpret_cast<const uint2 *>(c.slot + st[q] + i)));
    }
    // the next unit's slot stream: pull this lane's piece towards L2 (d_next has long arrived)
    if (g + 1 < n_units && my_p < c.g.n_prod)
      for (int i = 0; i < d_next.y; i += 64) pb_prefetch_l2(c.slot + d_next.x + i);
    pb_cp_async_wait<1>();  // this unit's row metadata has landed (the next unit's may still be in flight)
    pb_bar(2, PB_CONS_THREADS);
    // the ring slot is drained (its values sit in shared memory): hand it back to the producers
    if (tid == 0) pb_red_release(c.consumed + g);
    const uint16_t *lens = crp_s + (g & 1) * stride;
    const uint16_t *gbase = lens + c.g.RS;
    const int64_t r0 = ((int64_t)panel * c.g.n_prod + s) * c.g.RS;
    const bool last = pass == c.g.n_pass - 1;
    // one warp per group of 32 rows; entry k of the group's rows sits in consecutive words
    T carried = S::zero();
    if (pass > 0 && warp < ngr && warp * 32 + lane < c.g.RS && r0 + warp * 32 + lane < a.n_rows)
      carried = c.partial[r0 + warp * 32 + lane];
    for (int m = warp; m < ngr; m += NW) {
      const int i = m * 32 + lane;
      const bool ok = i < c.g.RS && r0 + i < a.n_rows;
      T acc = carried;
      const int in = i + NW * 32;  // the carried sum of this lane's next row: in flight during the reduction
      if (pass > 0 && m + NW < ngr && in < c.g.RS && r0 + in < a.n_rows) carried = c.partial[r0 + in];
      const int len = i < c.g.RS ? (int)lens[i] : 0;
      int off = gbase[m];
      const int maxlen = __reduce_max_sync(0xffffffffu, len);
      for (int k = 0; k < maxlen; ++k) {
        const uint32_t act = __ballot_sync(0xffffffffu, k < len);
        if (k < len) acc = S::add(acc, cbuf[off + __popc(act & lt_mask)]);
        off += __popc(act);
      }
      if (ok) {
        if (last) shb_finish_row<S>(a, r0 + i, acc, differs);
        else c.partial[r0 + i] = acc;
      }
    }
  }
  pb_cp_async_wait<0>();
}

template <class S, bool HAS_VAL, int K>
__global__ void __launch_bounds__(PB_THREADS, 1) pb_fused_kernel(const __grid_constant__ PbArgs<typename S::T> c,
                                                                  const __grid_constant__ SpmvArgs<typename S::T> a) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  extern __shared__ __align__(16) unsigned char pb_smem[];
  T *xs = reinterpret_cast<T *>(pb_smem);
  T *cbuf = xs + c.g.W;
  uint16_t *crp_s = reinterpret_cast<uint16_t *>(cbuf + c.g.cbuf_cap);
  bool differs = false;
  if (threadIdx.x < PB_PROD_THREADS) {
    if (c.roles & PB_ROLE_PRODUCE) pb_produce<S, HAS_VAL>(c, a, xs, threadIdx.x);
  } else {
    if (c.roles & PB_ROLE_CONSUME) pb_consume<S, K>(c, a, cbuf, crp_s, threadIdx.x - PB_PROD_THREADS, differs);
  }
  shb_publish_changed(a.ctl, differs);
  __syncthreads();
  if (c.reset_sync && threadIdx.x == 0) {
    // the last CTA to leave zeroes the panel counters for the next launch on this operator
    __threadfence();
    if (atomicAdd(c.exit_count, 1u) == gridDim.x - 1) {
      const int n_groups = c.g.n_pass * c.g.n_panel;
      for (int i = 0; i < n_groups; ++i) {
        c.ready[i] = 0;
        c.consumed[i] = 0;
      }
      *c.exit_count = 0;
      __threadfence();
    }
  }
}

// host side (shb_pb.cu)
size_t shb_pb_smem_bytes(const shb_pb *pb);
int shb_pb_reserve_ring(shb_ctx *c, shb_pb *pb, int depth);
