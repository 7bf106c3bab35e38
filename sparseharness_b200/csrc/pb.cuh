// pb.cuh -- the BLOCKED schedule: a propagation-blocked semiring SpMV for gather-bound operators.
//
// Why: on an operator whose columns are spread over a vector much larger than L1 (BASELINE.json config 2:
// 2^28 uniform-random gathers into a 64 MB x) every gather of the row-major schedules (spmv.cuh) is its
// own L1TEX->XBAR request and the kernel sits on the SM's one-request-per-cycle limit at a third of the
// HBM roofline (profiles/r01_summary.md).  Here no thread ever gathers from or scatters to global memory:
// every global access is a coalesced 128-bit stream, the random accesses are shared-memory accesses.
//
// ONE persistent cooperative kernel, one CTA per SM, two warp-specialised roles per CTA:
//
//   producer warps  CTA p keeps x[slice(pass, p)] (W columns) in shared memory for a whole pass and streams
//                   the non-zeros whose column falls in that slice -- stored column-slice-major as
//                   (16-bit local column, value), ONE contiguous stream per (pass, p) -- with 128-bit loads
//                   issued two chunks ahead, gathers x from SHARED memory, and stores the products
//                   mult(x, a) with 128-bit stores into a ring of DEPTH row-panel slots that stays resident
//                   in the 126 MB L2.  The ring is laid out in CONSUMER order: the four products of an aligned
//                   quad go to `stream position + delta(piece)`; a 64-bit flag word per 256-entry chunk says
//                   where pieces start, so the piece of a quad is one popc away.  Warps never meet at a CTA
//                   barrier inside a pass: they walk the stream round-robin and count their arrivals per panel
//                   in shared memory; the last one publishes the panel (release) for the consumers.
//   consumer warps  CTA s owns RS consecutive rows of every row panel.  What the 148 producers wrote for those
//                   rows is ONE contiguous, 16-byte aligned region of the ring slot, and the matching stream of
//                   16-bit slots is contiguous too: 128-bit loads, then plain st.shared of each product into
//                   its precomputed slot of a row-sorted shared-memory buffer (padding goes to a dump word),
//                   then a thread per row sums its slots in order and runs the usual fused epilogue /
//                   convergence test / multi-destination store on the last pass (row sums of earlier passes are
//                   carried in global memory).
//
// Producers and consumers of different CTAs meet through per-panel counters in global memory
// (release/acquire); a producer may run DEPTH panels ahead of the slowest consumer.  The products never reach
// HBM, so DRAM traffic stays near the operator's 8 bytes per non-zero.  The operator copy is built once per
// operator on the device (shb_pb.cu).  Row sums are formed in a fixed order (pass, then slice, then the
// CSR order inside a slice): deterministic for every semiring, bit-exact for min-plus / or-and / max-min and
// for integer-valued floats, within the stated 1e-5 relative tolerance for general floats.
#pragma once
#include "spmv.cuh"

struct PbGeom {
  int n_prod;    // column slices per pass == consumers per row panel == CTAs (the SM count)
  int n_pass;    // passes over the column space
  int n_panel;   // row panels
  int W;         // columns per slice (multiple of 4, <= 65536)
  int RS;        // rows per consumer
  int cbuf_cap;  // words of the consumer's row-sorted buffer (>= the largest consumer unit); word cbuf_cap = dump
  int slot_cap;  // entries per ring slot (>= the largest (pass, panel) group, padding included); + PB_CHUNK trash
  int depth;     // ring slots
};

#define PB_CHUNK 256  // entries a producer warp handles per step (8 per lane); units are padded to it

struct shb_pb {
  PbGeom g;
  int64_t cap = 0;      // producer-order entries including padding (pieces to 4, units to PB_CHUNK)
  int64_t ccap = 0;     // consumer-order entries including padding (pieces to 4)
  int64_t n_valid = 0;  // stored entries with an in-range column
  int64_t n_pieces = 0; // non-empty pieces + one padding pseudo-piece per padded unit
  int max_unit = 0;     // largest consumer unit (valid entries of one (pass, panel, s))
  uint16_t *lcol = nullptr;      // [cap] producer order
  void *val = nullptr;           // [cap] or null (pattern)
  uint4 *chunk_desc = nullptr;   // [cap / PB_CHUNK] {flags lo, flags hi, pieces started before the chunk, 0}
  int32_t *delta = nullptr;      // [n_pieces] ring offset (inside the slot) minus stream position
  int32_t *uchunk = nullptr;     // [n_pass * n_prod * n_panel + 1] first chunk of unit (pass, p, panel)
  uint16_t *slot = nullptr;      // [ccap] consumer order: position in the consumer's row-sorted buffer
  int32_t *cons_start = nullptr; // [n_groups * n_prod + 1] start of consumer region (group, s) in `slot`
  uint16_t *crp = nullptr;       // [(group, s)][pb_crp_stride(RS)] local row offsets into the row-sorted buffer
  void *ring = nullptr;          // [depth * (slot_cap + PB_CHUNK)] the product ring
  void *partial = nullptr;       // [n_rows] row sums carried from pass to pass
  uint32_t *sync = nullptr;      // ready[n_groups], consumed[n_groups], exit counter
  int ring_depth_alloc = 0;
  uint64_t device_bytes = 0;
};

#define PB_THREADS 768        // launched with 80 registers per thread (768 x 80 = 61440: the CTA's pool)
#define PB_PROD_THREADS 256   // two producer warpgroups; setmaxnreg moves 4096 registers from the consumers to them
#define PB_PROD_REGS 96
#define PB_CONS_REGS 72
#define PB_CONS_THREADS (PB_THREADS - PB_PROD_THREADS)
#define PB_PROD_WARPS (PB_PROD_THREADS / 32)
#define PB_CBUF_MAX 12288     // entries of the row-sorted buffer: one round of consumer registers (PB_CU x 512 x 4)
#define PB_MAX_W 30720
#define PB_MAX_DEPTH 8
#define PB_ARR 16  // arrival counters in shared memory (> PB_MAX_DEPTH + 1 panels can be live in one CTA)
#define PB_SMEM_LIMIT (227 * 1024)
#define PB_ROLE_PRODUCE 1
#define PB_ROLE_CONSUME 2

template <class T> struct PbArgs {
  PbGeom g;
  const uint16_t *lcol;
  const T *val;  // nullptr = pattern
  const uint4 *chunk_desc;
  const int32_t *delta;
  const int32_t *uchunk;
  const uint16_t *slot;
  const int32_t *cons_start;
  const uint16_t *crp;
  T *ring;
  T *partial;
  uint32_t *ready;     // [n_groups]
  uint32_t *consumed;  // [n_groups]
  uint32_t *exit_count;
  int roles;          // PB_ROLE_* mask (both = the fused kernel; one at a time = the two-launch debug mode)
  int reset_sync;     // zero the counters when the last CTA leaves
};

// 16-bit words of one unit's row offsets (RS + 1), padded to a 16-byte multiple for cp.async
__host__ __device__ __forceinline__ int pb_crp_stride(int RS) { return (RS + 1 + 7) & ~7; }

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
__device__ __forceinline__ uint4 pb_ld_cg_v4(const void *p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint2 pb_ld_stream_v2(const void *p, uint64_t pol) {
  uint2 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.b32 {%0, %1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ void pb_st_v4(void *p, const uint4 &v) {
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- TMA (bulk async copy) staging of the x slice: one thread issues, the mbarrier counts the bytes -----------
__device__ __forceinline__ uint32_t pb_smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void pb_mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pb_smem_addr(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void pb_mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pb_smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void pb_bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(pb_smem_addr(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(pb_smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ void pb_mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok = 0;
  const uint32_t addr = pb_smem_addr(bar);
  while (!ok)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok)
                 : "r"(addr), "r"(parity)
                 : "memory");
}

// ---- the producer's stream pipeline: cp.async into a per-warp ring of PB_STAGES chunk slots --------------------------
// Register-staged loads do not pipeline here: ptxas has six scoreboards per warp, the loads of several chunks in flight
// end up sharing one, and the wait for the oldest chunk then waits for the newest (measured: one memory round trip per
// chunk whatever the depth).  cp.async completes in per-thread commit groups instead: `wait_group<N>` is a FIFO.
// Slot layout (PB_STAGE_BYTES): [32 lanes x 16 B local columns][32 x 16 B values 0-3][32 x 16 B values 4-7][16 B flags];
// every lane copies and later reads ITS OWN 16-byte words (conflict-free LDS.128), lane 0 copies the chunk's flag word.
#define PB_STAGES 4
#define PB_STAGE_BYTES 1552

__device__ __forceinline__ void pb_cp_async16_hint(void *smem_dst, const void *gmem_src, uint64_t pol) {
  (void)pol;  // (cp.async with an L2 cache-hint operand faulted as an illegal instruction on sm_100a: plain form)
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(pb_smem_addr(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void pb_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void pb_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <bool HAS_VAL, class T>
__device__ __forceinline__ void pb_stage_issue(unsigned char *st, const PbArgs<T> &c, int cc, bool pred, int lane, uint64_t pol) {
  if (pred) {
    const int64_t i0 = (int64_t)cc * PB_CHUNK + lane * 8;
    pb_cp_async16_hint(st + lane * 16, c.lcol + i0, pol);
    if (HAS_VAL) {
      pb_cp_async16_hint(st + 512 + lane * 16, c.val + i0, pol);
      pb_cp_async16_hint(st + 1024 + lane * 16, c.val + i0 + 4, pol);
    }
    if (lane == 0) pb_cp_async16_hint(st + 1536, c.chunk_desc + cc, pol);
  }
  pb_cp_async_commit();  // one group per step, empty or not: the wait below counts groups
}

// the piece of each of the lane's two quads (pieces started up to and including the quad, minus one) and the
// pieces' deltas: issued one chunk ahead of the stores that need them
template <class T>
__device__ __forceinline__ void pb_stage_deltas(const unsigned char *st, const PbArgs<T> &c, int lane, bool pred, int &d0, int &d1) {
  const uint4 d = *reinterpret_cast<const uint4 *>(st + 1536);
  const uint64_t flags = ((uint64_t)d.y << 32) | (uint64_t)d.x;
  const int q0 = lane * 2;
  const int ord0 = max(0, (int)d.z + __popcll(flags & ((2ull << q0) - 1ull)) - 1);
  const int ord1 = ord0 + (int)((flags >> (q0 + 1)) & 1ull);
  d0 = pred ? __ldg(c.delta + ord0) : 0;
  d1 = pred ? __ldg(c.delta + ord1) : 0;
}

// gather, multiply and store one chunk into ring slot `ring` (consumer order)
template <class S, bool HAS_VAL>
__device__ __forceinline__ void pb_stage_process(const unsigned char *st, const typename S::T *xs, typename S::T *ring, int cc,
                                                 int lane, int d0, int d1) {
  typedef typename S::T T;
  const uint4 lc = *reinterpret_cast<const uint4 *>(st + lane * 16);
  uint4 v0 = make_uint4(0u, 0u, 0u, 0u), v1 = v0;
  if (HAS_VAL) {
    v0 = *reinterpret_cast<const uint4 *>(st + 512 + lane * 16);
    v1 = *reinterpret_cast<const uint4 *>(st + 1024 + lane * 16);
  }
  const T x0 = xs[lc.x & 0xffffu], x1 = xs[lc.x >> 16], x2 = xs[lc.y & 0xffffu], x3 = xs[lc.y >> 16];
  const T x4 = xs[lc.z & 0xffffu], x5 = xs[lc.z >> 16], x6 = xs[lc.w & 0xffffu], x7 = xs[lc.w >> 16];
  uint4 o0, o1;
  o0.x = shb_bits<T>(S::mult(x0, HAS_VAL ? shb_from_bits<T>(v0.x) : S::one()));
  o0.y = shb_bits<T>(S::mult(x1, HAS_VAL ? shb_from_bits<T>(v0.y) : S::one()));
  o0.z = shb_bits<T>(S::mult(x2, HAS_VAL ? shb_from_bits<T>(v0.z) : S::one()));
  o0.w = shb_bits<T>(S::mult(x3, HAS_VAL ? shb_from_bits<T>(v0.w) : S::one()));
  o1.x = shb_bits<T>(S::mult(x4, HAS_VAL ? shb_from_bits<T>(v1.x) : S::one()));
  o1.y = shb_bits<T>(S::mult(x5, HAS_VAL ? shb_from_bits<T>(v1.y) : S::one()));
  o1.z = shb_bits<T>(S::mult(x6, HAS_VAL ? shb_from_bits<T>(v1.z) : S::one()));
  o1.w = shb_bits<T>(S::mult(x7, HAS_VAL ? shb_from_bits<T>(v1.w) : S::one()));
  const int i0 = cc * PB_CHUNK + lane * 8;  // < 2^31 (checked at build time)
  pb_st_v4(ring + (i0 + d0), o0);
  pb_st_v4(ring + (i0 + 4 + d1), o1);
}

// ---------------------------------------------------------------------------------------------
// producer role: PB_PROD_THREADS threads (two warpgroups), named barrier 1 (twice per pass, around the x slice
// load).  Per warp: the streams of chunks k+2 and k+3 are in flight, chunk k+1 has landed and its deltas are being
// fetched, chunk k is gathered, multiplied and stored.
// ---------------------------------------------------------------------------------------------
template <class S, bool HAS_VAL>
__device__ __forceinline__ void pb_produce(const PbArgs<typename S::T> &c, const SpmvArgs<typename S::T> &a,
                                           typename S::T *xs, uint32_t *arr, uint64_t *xbar, unsigned char *pstage, int tid) {
  typedef typename S::T T;
  const int p = blockIdx.x;
  uint32_t xphase = 0;  // completed phases of the x-slice mbarrier
  const int lane = tid & 31, warp = tid >> 5;
  const uint64_t pol_stream = shb_policy_evict_first();
  const uint32_t zbits = shb_bits<T>(S::zero());
  const bool x_aligned = ((reinterpret_cast<uintptr_t>(a.x) & 15u) == 0);
  const bool fused = (c.roles & PB_ROLE_CONSUME) != 0;
  const int64_t slot_stride = (int64_t)c.g.slot_cap + PB_CHUNK;
  constexpr int NW = PB_PROD_WARPS;
  unsigned char *const wst = pstage + (size_t)warp * PB_STAGES * PB_STAGE_BYTES;  // this warp's ring of chunk slots

  // the warp's arrival at the end of its share of a panel; the CTA's last arrival publishes the panel
  auto arrive = [&](int g) {
    __syncwarp();
    if (lane == 0) {
      __threadfence();
      const uint32_t old = atomicAdd(&arr[g % PB_ARR], 1u);
      if ((old % NW) == NW - 1) {
        __threadfence();
        pb_red_release(c.ready + g);
      }
    }
  };
  // ring slot g % depth must have been drained by every consumer (panel g - depth).  The counter of the NEXT
  // panel is read early (precheck, every lane the same word) so that the usual transition costs no round trip.
  const int n_groups = c.g.n_pass * c.g.n_panel;
  uint32_t pre = 0;
  int pre_g = -1;
  auto precheck = [&](int g) {
    if (fused && g >= c.g.depth && g < n_groups) {
      pre = pb_ld_acquire(c.consumed + (g - c.g.depth));
      pre_g = g;
    }
  };
  auto wait_slot = [&](int g) {
    if (fused && g >= c.g.depth && !(pre_g == g && pre >= (uint32_t)c.g.n_prod)) {
      if (lane == 0)
        while (pb_ld_acquire(c.consumed + (g - c.g.depth)) < (uint32_t)c.g.n_prod) __nanosleep(40);
      __syncwarp();
    }
  };

  for (int pass = 0; pass < c.g.n_pass; ++pass) {
    pb_bar(1, PB_PROD_THREADS);  // every warp is done gathering from the previous slice
    // the slice of x, staged by TMA: thread 0 arms the mbarrier with the byte count and issues 1-D bulk copies
    // (global -> shared, 32 KB each); the 16-byte aligned body arrives through the async proxy while the other
    // threads fill what a bulk copy cannot carry (an unaligned x, the tail, columns past the end of the vector,
    // which read as the semiring zero)
    const int64_t c0 = ((int64_t)pass * c.g.n_prod + p) * c.g.W;
    const int valid = (int)max((int64_t)0, min((int64_t)c.g.W, (int64_t)a.vlength - c0));
    const int bulk = x_aligned ? (valid & ~3) : 0;
    if (bulk > 0 && tid == 0) {
      pb_mbar_expect_tx(xbar, (uint32_t)bulk * 4u);
      for (int off = 0; off < bulk; off += 8192) pb_bulk_g2s(xs + off, a.x + c0 + off, (uint32_t)min(8192, bulk - off) * 4u, xbar);
    }
    for (int i = bulk + tid; i < c.g.W; i += PB_PROD_THREADS) xs[i] = i < valid ? a.x[c0 + i] : shb_from_bits<T>(zbits);

    const int ub = (pass * c.g.n_prod + p) * c.g.n_panel;  // first unit of this (pass, p) stream
    const int g0 = pass * c.g.n_panel;
    const int c_begin = __ldg(c.uchunk + ub), c_end = __ldg(c.uchunk + ub + c.g.n_panel);
    int panel = 0;
    int ce = __ldg(c.uchunk + ub + 1);                                 // end chunk of the current unit
    int ce_next = c.g.n_panel > 1 ? __ldg(c.uchunk + ub + 2) : c_end;  // ... and of the one after
    int cc = c_begin + warp;
    // the streams of the first three chunks go out while the slice is still landing
    pb_stage_issue<HAS_VAL>(wst + 0 * PB_STAGE_BYTES, c, cc, cc < c_end, lane, pol_stream);
    pb_stage_issue<HAS_VAL>(wst + 1 * PB_STAGE_BYTES, c, cc + NW, cc + NW < c_end, lane, pol_stream);
    pb_stage_issue<HAS_VAL>(wst + 2 * PB_STAGE_BYTES, c, cc + 2 * NW, cc + 2 * NW < c_end, lane, pol_stream);
    if (bulk > 0) {
      pb_mbar_wait(xbar, xphase & 1u);
      ++xphase;
    }
    pb_bar(1, PB_PROD_THREADS);
    int d0 = 0, d1 = 0;  // deltas of the chunk about to be processed
    pb_cp_async_wait<2>();  // the first chunk has landed
    __syncwarp();
    pb_stage_deltas(wst, c, lane, cc < c_end, d0, d1);

    bool need_wait = true;
    T *ring = c.ring;
    // move to the unit that holds chunk cc, arriving at every panel left behind
    auto advance = [&]() {
      while (cc >= ce && panel < c.g.n_panel - 1) {
        // (a warp with no chunk in a panel still waits for its slot: that keeps every warp of the CTA within
        //  `depth` panels of the others, which is what the PB_ARR arrival counters rely on)
        if (need_wait) wait_slot(g0 + panel);
        arrive(g0 + panel);
        ++panel;
        ce = ce_next;
        ce_next = panel + 2 <= c.g.n_panel ? __ldg(c.uchunk + ub + panel + 2) : c_end;
        need_wait = true;
      }
      if (need_wait) {
        wait_slot(g0 + panel);
        precheck(g0 + panel + 1);
        ring = c.ring + (int64_t)((g0 + panel) % c.g.depth) * slot_stride;
        need_wait = false;
      }
    };
    int k = 0;  // ring position of the chunk being processed
    while (cc < c_end) {
      // stream of chunk k + 3 -> the slot chunk k - 1 has just left (every lane is past its reads: the syncwarp below)
      pb_stage_issue<HAS_VAL>(wst + ((k + 3) % PB_STAGES) * PB_STAGE_BYTES, c, cc + 3 * NW, cc + 3 * NW < c_end, lane, pol_stream);
      advance();
      pb_cp_async_wait<2>();  // at most chunks k + 2 and k + 3 are still in flight: chunk k + 1 has landed
      __syncwarp();           // ... for every lane (lane 0 copied the flag word)
      int n0, n1;
      pb_stage_deltas(wst + ((k + 1) % PB_STAGES) * PB_STAGE_BYTES, c, lane, cc + NW < c_end, n0, n1);
      pb_stage_process<S, HAS_VAL>(wst + (k % PB_STAGES) * PB_STAGE_BYTES, xs, ring, cc, lane, d0, d1);
      d0 = n0;
      d1 = n1;
      cc += NW;
      k = (k + 1) % PB_STAGES;
    }
    pb_cp_async_wait<0>();
    // the panel of the warp's last chunk, then the panels it has no chunks in
    for (; panel < c.g.n_panel; ++panel) {
      if (need_wait) wait_slot(g0 + panel);
      arrive(g0 + panel);
      need_wait = true;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// consumer role: PB_CONS_THREADS threads, named barrier 2.
// Per unit:  bar | registers -> row-sorted buffer (st.shared per product) | bar | hand the ring slot back |
//            issue the NEXT unit's ring + slot loads into registers | thread per row: sum, store | ...
// so the loads of a unit fly while the previous unit is reduced, and the row offsets of a unit arrive by cp.async
// while the unit before it is scattered.
// ---------------------------------------------------------------------------------------------
template <class T> __device__ __forceinline__ void pb_scatter4(T *cbuf, const uint4 &pr, const uint2 &sl) {
  cbuf[sl.x & 0xffffu] = shb_from_bits<T>(pr.x);
  cbuf[sl.x >> 16] = shb_from_bits<T>(pr.y);
  cbuf[sl.y & 0xffffu] = shb_from_bits<T>(pr.z);
  cbuf[sl.y >> 16] = shb_from_bits<T>(pr.w);
}
__device__ __forceinline__ void pb_cp_async16(void *smem_dst, const void *gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(pb_smem_addr(smem_dst)), "l"(gmem_src) : "memory");
}

#define PB_CU 6  // quads a consumer thread holds in registers (6 x 512 threads x 4 = 12288 entries per round; 8 spills at 80 registers)

template <class S>
__device__ __forceinline__ void pb_consume(const PbArgs<typename S::T> &c, const SpmvArgs<typename S::T> &a,
                                           typename S::T *cbuf, uint16_t *crp_s, uint32_t *flag_s, int tid, bool &differs) {
  typedef typename S::T T;
  constexpr int NT = PB_CONS_THREADS;
  constexpr int U = PB_CU;
  const int s = blockIdx.x;
  const int n_groups = c.g.n_pass * c.g.n_panel;
  const int64_t slot_stride = (int64_t)c.g.slot_cap + PB_CHUNK;
  const bool fused = (c.roles & PB_ROLE_PRODUCE) != 0;
  const uint64_t pol_stream = shb_policy_evict_first();
  const int stride = pb_crp_stride(c.g.RS);  // 16-bit words per unit, a 16-byte multiple

  uint4 pr[U];
  uint2 sl[U];
  // first-round loads of unit (g, cs, ce, gs): up to U quads per thread
  auto issue = [&](int g, int cs, int ce, int gs) {
    const T *ring_u = c.ring + (int64_t)(g % c.g.depth) * slot_stride + (cs - gs);
    const uint16_t *slot_u = c.slot + cs;
    const int nq = (ce - cs) >> 2;
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const int q = tid + k * NT;
      if (q < nq) {
        pr[k] = pb_ld_cg_v4(ring_u + 4 * q);
        sl[k] = pb_ld_stream_v2(slot_u + 4 * q, pol_stream);
      }
    }
  };
  auto fetch_crp = [&](int g) {  // row offsets of unit g into its half of crp_s
    const uint16_t *src = c.crp + ((int64_t)g * c.g.n_prod + s) * stride;
    uint16_t *dst = crp_s + (g & 1) * stride;
    for (int i = tid * 8; i < stride; i += NT * 8) pb_cp_async16(dst + i, src + i);
    pb_cp_async_commit();
  };

  int cs = __ldg(c.cons_start + s), ce = __ldg(c.cons_start + s + 1), gs = __ldg(c.cons_start);
  bool have = false;  // the registers hold the first round of the current unit
  fetch_crp(0);
  for (int g = 0; g < n_groups; ++g) {
    const int pass = g / c.g.n_panel, panel = g - pass * c.g.n_panel;
    const int64_t u = (int64_t)g * c.g.n_prod + s;
    const int64_t r0 = ((int64_t)panel * c.g.n_prod + s) * c.g.RS;
    const int nr = (int)max((int64_t)0, min((int64_t)c.g.RS, a.n_rows - r0));
    // next unit's extents (their latency hides behind this unit)
    int cs_n = 0, ce_n = 0, gs_n = 0;
    if (g + 1 < n_groups) {
      cs_n = __ldg(c.cons_start + u + c.g.n_prod);
      ce_n = __ldg(c.cons_start + u + c.g.n_prod + 1);
      gs_n = __ldg(c.cons_start + (int64_t)(g + 1) * c.g.n_prod);
    }
    // carried sums of this unit: towards L2 while the products are scattered
    if (pass > 0)
      for (int i = tid * 32; i < nr; i += NT * 32) pb_prefetch_l2(c.partial + r0 + i);
    if (!have) {
      if (fused && tid == 0)
        while (pb_ld_acquire(c.ready + g) < (uint32_t)c.g.n_prod) __nanosleep(20);
      pb_bar(2, NT);  // ring slot g is complete (and everyone is done reducing unit g - 1)
      issue(g, cs, ce, gs);
    } else {
      pb_bar(2, NT);  // everyone is done reducing unit g - 1: the row-sorted buffer is free
    }
    // the next unit's row offsets go into the half unit g - 1 has just stopped reading
    if (g + 1 < n_groups) fetch_crp(g + 1);
    else pb_cp_async_commit();  // keeps the group count uniform for the wait below
    {
      const int nq = (ce - cs) >> 2;
#pragma unroll
      for (int k = 0; k < U; ++k)
        if (tid + k * NT < nq) pb_scatter4<T>(cbuf, pr[k], sl[k]);
      if (nq > NT * U) {  // units larger than one round (rare geometries): the rest, round by round
        const T *ring_u = c.ring + (int64_t)(g % c.g.depth) * slot_stride + (cs - gs);
        const uint16_t *slot_u = c.slot + cs;
        for (int q = tid + NT * U; q < nq; q += NT) pb_scatter4<T>(cbuf, pb_ld_cg_v4(ring_u + 4 * q), pb_ld_stream_v2(slot_u + 4 * q, pol_stream));
      }
    }
    // is the next unit's ring slot complete already?  one thread looks, the barrier spreads the verdict
    if (tid == 0) *flag_s = (g + 1 < n_groups && (!fused || pb_ld_acquire(c.ready + g + 1) >= (uint32_t)c.g.n_prod)) ? 1u : 0u;
    pb_cp_async_wait<1>();  // this unit's row offsets have landed (the next unit's may still be in flight)
    pb_bar(2, NT);
    // the ring slot is drained (its values sit in shared memory): hand it back to the producers
    if (fused && tid == 0) pb_red_release(c.consumed + g);
    have = *flag_s != 0u;
    const bool last = pass == c.g.n_pass - 1;
    const uint16_t *crp_u = crp_s + (g & 1) * stride;
    constexpr int RPT = 5;  // rows per thread per sweep (5 x 512 covers the usual unit in one sweep)
    T part[RPT];
    // the carried sums of the first sweep are fetched BEFORE the next unit's loads go out: a wait on them must not
    // end up behind those loads on a shared scoreboard
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
      const int i = tid + k * NT;
      part[k] = (pass > 0 && i < nr) ? c.partial[r0 + i] : S::zero();
    }
    if (have) issue(g + 1, cs_n, ce_n, gs_n);  // in flight during the reduction below
    for (int i0 = 0; i0 < nr; i0 += NT * RPT) {
      if (i0 > 0) {
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
          const int i = i0 + tid + k * NT;
          part[k] = (pass > 0 && i < nr) ? c.partial[r0 + i] : S::zero();
        }
      }
#pragma unroll
      for (int k = 0; k < RPT; ++k) {
        const int i = i0 + tid + k * NT;
        if (i < nr) {
          const int b = crp_u[i], e = crp_u[i + 1];
          T acc = part[k];
          int j = b;
          for (; j + 3 < e; j += 4) {
            const T v0 = cbuf[j], v1 = cbuf[j + 1], v2 = cbuf[j + 2], v3 = cbuf[j + 3];
            acc = S::add(S::add(S::add(S::add(acc, v0), v1), v2), v3);
          }
          for (; j < e; ++j) acc = S::add(acc, cbuf[j]);
          if (last) shb_finish_row<S>(a, r0 + i, acc, differs);
          else c.partial[r0 + i] = acc;
        }
      }
    }
    cs = cs_n;
    ce = ce_n;
    gs = gs_n;
  }
  pb_cp_async_wait<0>();
}

template <class S, bool HAS_VAL>
__global__ void __launch_bounds__(PB_THREADS, 1) pb_fused_kernel(const __grid_constant__ PbArgs<typename S::T> c,
                                                                  const __grid_constant__ SpmvArgs<typename S::T> a) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  extern __shared__ __align__(16) unsigned char pb_smem[];
  T *xs = reinterpret_cast<T *>(pb_smem);
  T *cbuf = xs + c.g.W;
  uint16_t *crp_s = reinterpret_cast<uint16_t *>(cbuf + c.g.cbuf_cap + 32);
  uint32_t *arr = reinterpret_cast<uint32_t *>(crp_s + 2 * pb_crp_stride(c.g.RS));
  uint32_t *flag_s = arr + PB_ARR;
  uint64_t *xbar = reinterpret_cast<uint64_t *>(arr + PB_ARR + 2);
  unsigned char *pstage = reinterpret_cast<unsigned char *>(arr + PB_ARR + 4 + 4);  // 16-byte aligned: 96 bytes past arr
  if (threadIdx.x < PB_ARR) arr[threadIdx.x] = 0;
  if (threadIdx.x == 0) pb_mbar_init(xbar, 1);
  __syncthreads();
  bool differs = false;
  if (threadIdx.x < PB_PROD_THREADS) {
    // the CTA's register pool is fixed at launch (768 threads x 80): the four consumer warpgroups hand 8 registers per
    // thread back (512 x 8 = 4096) and the two producer warpgroups take them (256 x 16 = 4096); inc blocks until then
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(PB_PROD_REGS));
    if (c.roles & PB_ROLE_PRODUCE) pb_produce<S, HAS_VAL>(c, a, xs, arr, xbar, pstage, threadIdx.x);
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(PB_CONS_REGS));
    if (c.roles & PB_ROLE_CONSUME) pb_consume<S>(c, a, cbuf, crp_s, flag_s, threadIdx.x - PB_PROD_THREADS, differs);
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
