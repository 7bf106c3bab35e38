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
//                   its precomputed slot of a shared-memory buffer with lmax words per row (padding goes to a
//                   dump word); then a thread per row reads its lmax words with 128-bit loads, puts the
//                   semiring zero back, adds them up in order and runs the usual fused epilogue / convergence
//                   test / multi-destination store on the last pass (row sums of earlier passes are carried in
//                   global memory).  No row offsets, no per-row loop, no divergence.  Rows with more than
//                   lmax (8 or 16) entries in one pass (3 % of BASELINE config 2's rows at 8) are left out of the blocked
//                   copy and computed whole by `pb_long_rows_kernel` (a warp per row) right after.
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
  int cbuf_cap;  // words of the consumer's row buffer = RS * lmax; word cbuf_cap = dump
  int slot_cap;  // entries per ring slot (>= the largest (pass, panel) group, padding included); + PB_CHUNK trash
  int depth;     // ring slots
  int lmax;      // words per row and pass in the consumer's buffer (8 or 16); rows with more entries in a pass are "long"
};

#define PB_CHUNK 256  // entries a producer warp handles per step (8 per lane); units are padded to it
#define PB_LMAX_MAX 16  // words per row and pass in the consumer's buffer: 8 or 16 (PbGeom::lmax)
#define PB_MAX_PASS 32

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
  uint16_t *slot = nullptr;      // [ccap] consumer order: BYTE offset in the consumer's row-sorted buffer
  int32_t *cons_start = nullptr; // [n_groups * n_prod + 1] start of consumer region (group, s) in `slot`
  int4 *cdesc = nullptr;         // [n_groups * n_prod] the same as one descriptor per consumer unit (PbArgs::cdesc)
  uint32_t *long_bits = nullptr; // [ceil(n_rows / 32)] bit r: row r is long (computed by pb_long_rows_kernel)
  int32_t *long_rows = nullptr;  // [n_long] their indices
  int64_t n_long = 0;
  void *ring = nullptr;          // [depth * (slot_cap + PB_CHUNK)] the product ring
  void *partial = nullptr;       // [n_rows] row sums carried from pass to pass
  uint32_t *sync = nullptr;      // ready[n_groups], consumed[n_groups], exit counter
  int ring_depth_alloc = 0;
  uint64_t device_bytes = 0;
};

#define PB_THREADS 896        // 72 registers per thread (896 x 72 = 64512 of the SM's 65536)
#define PB_PROD_THREADS 384   // twelve producer warps: the role is bound by each warp's serial instruction stream
#define PB_CONS_THREADS (PB_THREADS - PB_PROD_THREADS)
#define PB_PROD_WARPS (PB_PROD_THREADS / 32)
#define PB_CU 5               // quads a consumer thread holds in registers: one round = PB_CU x 512 x 4 = 10240 entries
#define PB_MAX_W 30720
#define PB_MAX_DEPTH 8
#define PB_ARR 16  // arrival counters in shared memory (> PB_MAX_DEPTH + 1 panels can be live in one CTA)
#define PB_SMEM_LIMIT (227 * 1024)
#define PB_MISC_BYTES (PB_ARR * 4 + 8 + 8 + PB_PROD_WARPS * PB_STAGES * 8)  // arrival counters, flag, x mbarrier, slot mbarriers
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
  const int4 *cdesc;  // [n_groups * n_prod] consumer unit: {first entry in `slot`, quads, offset inside the ring slot, 0}
  const uint32_t *long_bits;
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
// relaxed poll (no L1 invalidate, no fence): used where the poller only WRITES afterwards -- a producer waiting for a ring
// slot to be drained: the consumers' reads were complete before they bumped the counter, so a later store cannot reach them
__device__ __forceinline__ uint32_t pb_ld_relaxed(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
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

// ---- the producer's stream pipeline: TMA bulk copies into a per-warp ring of PB_STAGES chunk slots -------------------
// Measured on the way here (profiles/r02_summary.md): (1) register-staged loads do not pipeline -- ptxas has six
// scoreboards per warp, the loads of several chunks in flight end up sharing one, and the wait for the oldest chunk
// then waits for the newest; (2) cp.async (LDGSTS) pipelines, but writes shared memory once per returned 32-byte
// sector: 32 wavefronts for a 1 KB value chunk, 63 % of the producer's shared-memory traffic.  A 1-D bulk copy
// (cp.async.bulk, the TMA unit) lands whole lines, is issued by ONE lane, and completes on an mbarrier per slot.
// Slot layout (PB_STAGE_BYTES): [512 B local columns: lane l owns bytes 16 l ..][512 B values 0-3 of every lane's
// 8 -- i.e. the chunk's values in stream order occupy 1 KB: lane l owns bytes 32 l .. 32 l + 31][48 B descriptor:
// piece-start flags (64 bits), ordinal of the first piece, number of pieces, and the deltas of the first 8 pieces].
#define PB_STAGES 3
#define PB_STAGE_BYTES 1584
#define PB_INLINE_DELTAS 8  // deltas of the first pieces touching a chunk travel in its descriptor (48 bytes)

// the piece of each of the lane's two quads and the pieces' deltas.  The descriptor lists the pieces TOUCHING the chunk in
// order (the first may have started in an earlier chunk: then flag bit 0 is clear); the deltas of the first
// PB_INLINE_DELTAS of them sit in the descriptor itself (shared memory), the rest come from the global table.
template <class T>
__device__ __forceinline__ void pb_stage_deltas(const unsigned char *st, const PbArgs<T> &c, int lane, int &d0, int &d1) {
  const uint4 d = *reinterpret_cast<const uint4 *>(st + 1536);
  const int *inl = reinterpret_cast<const int *>(st + 1552);
  const uint64_t flags = ((uint64_t)d.y << 32) | (uint64_t)d.x;
  const int q0 = lane * 2;
  const int i0 = __popcll(flags & ((2ull << q0) - 1ull)) - (int)(d.x & 1u);  // index among the pieces touching the chunk
  const int i1 = i0 + (int)((flags >> (q0 + 1)) & 1ull);
  const int base = (int)d.z - 1 + (int)(d.x & 1u);  // ordinal of touching piece 0 in the global table
  d0 = i0 < PB_INLINE_DELTAS ? inl[i0] : __ldg(c.delta + base + i0);
  d1 = i1 < PB_INLINE_DELTAS ? inl[i1] : __ldg(c.delta + base + i1);
}

// gather, multiply and store one chunk into ring slot `ring` (consumer order)
template <class S, bool HAS_VAL>
__device__ __forceinline__ void pb_stage_process(const unsigned char *st, const typename S::T *xs, typename S::T *ring, int cc,
                                                 int lane, int d0, int d1, int W, int slot_words) {
  typedef typename S::T T;
  const uint4 lc = *reinterpret_cast<const uint4 *>(st + lane * 16);
  SHB_DEV_CHECK((int)(lc.x & 0xffffu) < W && (int)(lc.x >> 16) < W && (int)(lc.y & 0xffffu) < W && (int)(lc.y >> 16) < W);
  SHB_DEV_CHECK((int)(lc.z & 0xffffu) < W && (int)(lc.z >> 16) < W && (int)(lc.w & 0xffffu) < W && (int)(lc.w >> 16) < W);
  uint4 v0 = make_uint4(0u, 0u, 0u, 0u), v1 = v0;
  if (HAS_VAL) {
    v0 = *reinterpret_cast<const uint4 *>(st + 512 + lane * 32);
    v1 = *reinterpret_cast<const uint4 *>(st + 512 + lane * 32 + 16);
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
  SHB_DEV_CHECK(i0 + d0 >= 0 && i0 + d0 + 4 <= slot_words && i0 + 4 + d1 >= 0 && i0 + 8 + d1 <= slot_words);
  pb_st_v4(ring + (i0 + d0), o0);
  pb_st_v4(ring + (i0 + 4 + d1), o1);
}

// ---------------------------------------------------------------------------------------------
// producer role: PB_PROD_THREADS threads (two warpgroups), named barrier 1 (twice per pass, around the x slice
// load).  Per warp: the streams of chunks k+1 and k+2 are in flight (bulk copies),
// chunk k is gathered, multiplied and stored (its piece deltas come with its descriptor).
// ---------------------------------------------------------------------------------------------
template <class S, bool HAS_VAL>
__device__ __forceinline__ void pb_produce(const PbArgs<typename S::T> &c, const SpmvArgs<typename S::T> &a,
                                           typename S::T *xs, uint32_t *arr, uint64_t *xbar, uint64_t *sbar_all,
                                           unsigned char *pstage, int tid) {
  typedef typename S::T T;
  const int p = blockIdx.x;
  uint32_t xphase = 0;  // completed phases of the x-slice mbarrier
  const int lane = tid & 31, warp = tid >> 5;
  const uint32_t zbits = shb_bits<T>(S::zero());
  const bool x_aligned = ((reinterpret_cast<uintptr_t>(a.x) & 15u) == 0);
  const bool fused = (c.roles & PB_ROLE_CONSUME) != 0;
  const int64_t slot_stride = (int64_t)c.g.slot_cap + PB_CHUNK;
  constexpr int NW = PB_PROD_WARPS;
  unsigned char *const wst = pstage + (size_t)warp * PB_STAGES * PB_STAGE_BYTES;  // this warp's ring of chunk slots
  uint64_t *const sbar = sbar_all + warp * PB_STAGES;                             // ... and their mbarriers
  uint32_t issued = 0, waited = 0;  // chunks issued to / awaited from the ring (slot = n % PB_STAGES, parity = (n / PB_STAGES) & 1)

  // lane 0: arm the slot's mbarrier with the byte count and start the bulk copies of one chunk
  auto issue = [&](int cc) {
    if (lane == 0) {
      const uint32_t st_i = issued % PB_STAGES;
      unsigned char *st = wst + st_i * PB_STAGE_BYTES;
      const int64_t i0 = (int64_t)cc * PB_CHUNK;
      pb_mbar_expect_tx(sbar + st_i, HAS_VAL ? 1584u : 560u);
      pb_bulk_g2s(st, c.lcol + i0, 512u, sbar + st_i);
      if (HAS_VAL) pb_bulk_g2s(st + 512, c.val + i0, 1024u, sbar + st_i);
      pb_bulk_g2s(st + 1536, c.chunk_desc + 3 * (int64_t)cc, 48u, sbar + st_i);
    }
    ++issued;
  };
  auto await = [&]() {  // the oldest chunk not yet awaited has landed (for every lane)
    pb_mbar_wait(sbar + (waited % PB_STAGES), (waited / PB_STAGES) & 1u);
    ++waited;
  };

  // the warp's arrival at the end of its share of a panel; the CTA's last arrival publishes the panel
  // (each warp orders its stores at CTA scope before its count; the last one fences at GPU scope and releases: the same
  //  pattern as a __syncthreads() followed by one thread's __threadfence() + atomic -- fence cumulativity covers the others)
  auto arrive = [&](int g) {
    __syncwarp();
    if (lane == 0) {
      __threadfence_block();
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
      pre = pb_ld_relaxed(c.consumed + (g - c.g.depth));
      pre_g = g;
    }
  };
  auto wait_slot = [&](int g) {
    if (fused && g >= c.g.depth && !(pre_g == g && pre >= (uint32_t)c.g.n_prod)) {
      if (lane == 0)
        while (pb_ld_relaxed(c.consumed + (g - c.g.depth)) < (uint32_t)c.g.n_prod) __nanosleep(100);
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
    // the streams of the first two chunks go out while the slice is still landing
    if (cc < c_end) issue(cc);
    if (cc + NW < c_end) issue(cc + NW);
    if (bulk > 0) {
      pb_mbar_wait(xbar, xphase & 1u);
      ++xphase;
    }
    pb_bar(1, PB_PROD_THREADS);

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
    while (cc < c_end) {
      __syncwarp();  // every lane is past its reads of the slot chunk cc + 2 NW is about to overwrite
      if (cc + 2 * NW < c_end) issue(cc + 2 * NW);
      advance();
      const unsigned char *st = wst + (waited % PB_STAGES) * PB_STAGE_BYTES;
      await();
      int d0, d1;
      pb_stage_deltas(st, c, lane, d0, d1);
      pb_stage_process<S, HAS_VAL>(st, xs, ring, cc, lane, d0, d1, c.g.W, (int)slot_stride);
      cc += NW;
    }
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
// Per unit:  bar | registers -> row buffer (one st.shared per product, the 16-bit slot IS the byte offset) | bar |
//            hand the ring slot back | issue the NEXT unit's ring, slot and carried-sum loads into registers |
//            thread per row: LQ x LDS.128, zeros back, the adds, one store | ...
// so every global load of a unit flies while the previous unit is reduced, and nothing in the loop divides, converts a
// generic address or re-derives a unit's extents: those come as one 16-byte descriptor per unit, read a unit ahead.
// PLAIN = one destination, no y, no convergence test (the plain SpMV call): the epilogue is one expression and the long
// rows need no care -- pb_long_rows_kernel runs after this kernel and overwrites them.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void pb_sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void pb_sts128(uint32_t addr, const uint4 &v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 pb_lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t pb_ld_cg_b32(const void *p) {
  uint32_t v;
  asm volatile("ld.global.cg.b32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void pb_red_relaxed(uint32_t *p) { asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory"); }

__device__ __forceinline__ void pb_scatter4(uint32_t cb, const uint4 &pr, const uint2 &sl, uint32_t dump) {
  SHB_DEV_CHECK((sl.x & 0xffffu) <= dump && (sl.x >> 16) <= dump && (sl.y & 0xffffu) <= dump && (sl.y >> 16) <= dump);
  SHB_DEV_CHECK(((sl.x | sl.y) & 0x00030003u) == 0u);
  pb_sts32(cb + (sl.x & 0xffffu), pr.x);
  pb_sts32(cb + (sl.x >> 16), pr.y);
  pb_sts32(cb + (sl.y & 0xffffu), pr.z);
  pb_sts32(cb + (sl.y >> 16), pr.w);
}

template <class S, int LQ, bool PLAIN>  // LQ = lmax / 4: 128-bit words per row
__device__ __forceinline__ void pb_consume(const PbArgs<typename S::T> &c, const SpmvArgs<typename S::T> &a,
                                           typename S::T *cbuf, uint32_t *flag_s, int tid, bool &differs) {
  typedef typename S::T T;
  constexpr int NT = PB_CONS_THREADS;
  constexpr int U = PB_CU;
  constexpr int RPT = 4;  // rows per thread per sweep (4 x 512 covers the usual unit in one sweep)
  const int s = blockIdx.x;
  const int n_groups = c.g.n_pass * c.g.n_panel;
  const int64_t slot_stride = (int64_t)c.g.slot_cap + PB_CHUNK;
  const bool fused = (c.roles & PB_ROLE_PRODUCE) != 0;
  const uint64_t pol_stream = shb_policy_evict_first();
  const int RS = c.g.RS;
  const uint32_t zb = shb_bits<T>(S::zero());
  const uint4 zero4 = make_uint4(zb, zb, zb, zb);
  const uint32_t cb = pb_smem_addr(cbuf);
  const uint32_t qstride = 16u * (uint32_t)RS;  // bytes between a row's consecutive 128-bit words
  const uint32_t dump = 4u * (uint32_t)c.g.cbuf_cap;  // byte offset of the word padding entries drop their product into
  T *const out0 = a.out[0] + a.row_offset;
  // the sync duties (poll the next panel, hand the slot back) go to the LAST warp: it has the fewest rows to reduce
  const bool syncer = tid == NT - 32;
  // carried sums are fetched a unit ahead unless the unit ahead is the same panel one pass later (a single row panel)
  const bool early_carry = c.g.n_panel > 1;

  // every word of the row buffer starts as the semiring zero; the reduction puts the zeros back as it reads
  for (int i = tid * 4; i < c.g.cbuf_cap + 32; i += NT * 4) pb_sts128(cb + 4u * (uint32_t)i, zero4);

  uint4 pr[U];
  uint2 sl[U];
  T part_nxt[RPT];
#pragma unroll
  for (int k = 0; k < RPT; ++k) part_nxt[k] = S::zero();
  // loads of a unit: its first U quads per thread out of ring slot `rslot`, and (later passes) its first RPT carried sums
  auto issue = [&](const int4 &d, int rslot, int64_t r0n, int nrn, bool carry) {
    const T *ring_u = c.ring + (int64_t)rslot * slot_stride + d.z;
    const uint16_t *slot_u = c.slot + d.x;
    SHB_DEV_CHECK(rslot >= 0 && rslot < c.g.depth && d.z >= 0 && (int64_t)d.z + 4 * (int64_t)d.y <= slot_stride);
    SHB_DEV_CHECK(nrn >= 0 && nrn <= c.g.RS && r0n >= 0);
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const int q = tid + k * NT;
      if (q < d.y) {
        pr[k] = pb_ld_cg_v4(ring_u + 4 * q);
        sl[k] = pb_ld_stream_v2(slot_u + 4 * q, pol_stream);
      }
    }
    if (carry) {
#pragma unroll
      for (int k = 0; k < RPT; ++k) {
        const int i = tid + k * NT;
        if (i < nrn) part_nxt[k] = shb_from_bits<T>(pb_ld_cg_b32(c.partial + r0n + i));
      }
    }
  };

  const int64_t r_step = (int64_t)c.g.n_prod * RS;
  int4 d = __ldg(c.cdesc + s);  // {first entry of the unit in `slot`, quads, offset inside the ring slot, -}
  int64_t r0 = (int64_t)s * RS;
  int rslot = 0;
  bool have = false;  // the registers hold the first round of the current unit
  int pass = 0, panel = 0;
  for (int g = 0; g < n_groups; ++g) {
    const int nr = (int)max((int64_t)0, min((int64_t)RS, a.n_rows - r0));
    // the next unit: descriptor (its latency hides behind this unit), rows, ring slot
    int4 d_n = make_int4(0, 0, 0, 0);
    if (g + 1 < n_groups) d_n = __ldg(c.cdesc + (int64_t)(g + 1) * c.g.n_prod + s);
    const bool wrap = panel + 1 == c.g.n_panel;
    const int64_t r0_n = wrap ? (int64_t)s * RS : r0 + r_step;
    const int nr_n = (int)max((int64_t)0, min((int64_t)RS, a.n_rows - r0_n));
    const int pass_n = pass + (wrap ? 1 : 0);
    const int rslot_n = rslot + 1 == c.g.depth ? 0 : rslot + 1;
    if (!have) {
      if (fused && syncer)
        while (pb_ld_acquire(c.ready + g) < (uint32_t)c.g.n_prod) __nanosleep(20);
      pb_bar(2, NT);  // ring slot g is complete (and everyone is done reducing unit g - 1)
      issue(d, rslot, r0, nr, early_carry && pass > 0);
    } else {
      pb_bar(2, NT);  // everyone is done reducing unit g - 1: the row buffer holds zeros again
    }
#pragma unroll
    for (int k = 0; k < U; ++k)
      if (tid + k * NT < d.y) pb_scatter4(cb, pr[k], sl[k], dump);
    if (d.y > NT * U) {  // units larger than one round (rare geometries): the rest, round by round
      const T *ring_u = c.ring + (int64_t)rslot * slot_stride + d.z;
      const uint16_t *slot_u = c.slot + d.x;
      for (int q = tid + NT * U; q < d.y; q += NT) pb_scatter4(cb, pb_ld_cg_v4(ring_u + 4 * q), pb_ld_stream_v2(slot_u + 4 * q, pol_stream), dump);
    }
    // is the next unit's ring slot complete already?  one thread looks, the barrier spreads the verdict
    if (syncer) *flag_s = (g + 1 < n_groups && (!fused || pb_ld_acquire(c.ready + g + 1) >= (uint32_t)c.g.n_prod)) ? 1u : 0u;
    pb_bar(2, NT);
    // the ring slot is drained (every product went through a register into shared memory): hand it back.  Relaxed: the
    // loads have completed, and the producers only write the slot after seeing the count.
    if (fused && syncer) pb_red_relaxed(c.consumed + g);
    have = *flag_s != 0u;
    const bool last = pass == c.g.n_pass - 1;
    T part[RPT];
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
      part[k] = part_nxt[k];
      part_nxt[k] = S::zero();
    }
    if (!early_carry && pass > 0) {
#pragma unroll
      for (int k = 0; k < RPT; ++k) {
        const int i = tid + k * NT;
        if (i < nr) part[k] = shb_from_bits<T>(pb_ld_cg_b32(c.partial + r0 + i));
      }
    }
    uint32_t lbits[RPT];
    if (!PLAIN) {
#pragma unroll
      for (int k = 0; k < RPT; ++k) {
        const int i = tid + k * NT;
        lbits[k] = i < nr ? __ldg(c.long_bits + ((r0 + i) >> 5)) : 0u;
      }
    }
    if (have) issue(d_n, rslot_n, r0_n, nr_n, early_carry && pass_n > 0);  // in flight during the reduction below
    // the row's lmax words out of the buffer, the semiring zero back in, the sum on top of `acc`
    auto row_sum = [&](int i, T acc) {
      uint4 v[LQ];
      const uint32_t ra = cb + 16u * (uint32_t)i;
#pragma unroll
      for (int q = 0; q < LQ; ++q) {
        v[q] = pb_lds128(ra + q * qstride);
        pb_sts128(ra + q * qstride, zero4);
      }
#pragma unroll
      for (int q = 0; q < LQ; ++q) {
        acc = S::add(acc, shb_from_bits<T>(v[q].x));
        acc = S::add(acc, shb_from_bits<T>(v[q].y));
        acc = S::add(acc, shb_from_bits<T>(v[q].z));
        acc = S::add(acc, shb_from_bits<T>(v[q].w));
      }
      return acc;
    };
    auto finish = [&](int i, T acc, uint32_t lb) {
      if (PLAIN) {
        if (last) out0[r0 + i] = S::epi(acc, a.alpha, S::zero(), a.beta);
        else c.partial[r0 + i] = acc;
      } else if (!((lb >> ((r0 + i) & 31)) & 1u)) {  // long rows belong to pb_long_rows_kernel
        if (last) shb_finish_row<S>(a, r0 + i, acc, differs);
        else c.partial[r0 + i] = acc;
      }
    };
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
      const int i = tid + k * NT;
      if (i < nr) finish(i, row_sum(i, part[k]), PLAIN ? 0u : lbits[k]);
    }
    for (int i = tid + RPT * NT; i < nr; i += NT) {  // units with more than RPT x 512 rows (rare geometries)
      const T carried = pass > 0 ? shb_from_bits<T>(pb_ld_cg_b32(c.partial + r0 + i)) : S::zero();
      finish(i, row_sum(i, carried), PLAIN ? 0u : __ldg(c.long_bits + ((r0 + i) >> 5)));
    }
    d = d_n;
    r0 = r0_n;
    rslot = rslot_n;
    pass = pass_n;
    panel = wrap ? 0 : panel + 1;
  }
}

template <class S, bool HAS_VAL, int LQ, bool PLAIN>
__global__ void __launch_bounds__(PB_THREADS, 1) pb_fused_kernel(const __grid_constant__ PbArgs<typename S::T> c,
                                                                  const __grid_constant__ SpmvArgs<typename S::T> a) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  extern __shared__ __align__(16) unsigned char pb_smem[];
  T *xs = reinterpret_cast<T *>(pb_smem);
  T *cbuf = xs + c.g.W;
  uint32_t *arr = reinterpret_cast<uint32_t *>(cbuf + c.g.cbuf_cap + 32);
  uint32_t *flag_s = arr + PB_ARR;
  uint64_t *xbar = reinterpret_cast<uint64_t *>(arr + PB_ARR + 2);
  uint64_t *sbar = xbar + 1;                                                                   // PB_PROD_WARPS * PB_STAGES slots
  unsigned char *pstage = reinterpret_cast<unsigned char *>(sbar + PB_PROD_WARPS * PB_STAGES);  // 16-byte aligned
  if (threadIdx.x < PB_ARR) arr[threadIdx.x] = 0;
  if (threadIdx.x == 0) {
    pb_mbar_init(xbar, 1);
    for (int i = 0; i < PB_PROD_WARPS * PB_STAGES; ++i) pb_mbar_init(sbar + i, 1);
  }
  __syncthreads();
  bool differs = false;
  if (threadIdx.x < PB_PROD_THREADS) {
    if (c.roles & PB_ROLE_PRODUCE) pb_produce<S, HAS_VAL>(c, a, xs, arr, xbar, sbar, pstage, threadIdx.x);
  } else {
    if (c.roles & PB_ROLE_CONSUME) pb_consume<S, LQ, PLAIN>(c, a, cbuf, flag_s, threadIdx.x - PB_PROD_THREADS, differs);
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

// the rows the blocked copy leaves out (more than lmax entries in some pass): a warp per row over the operator's own
// CSR arrays, lanes striding the row, butterfly reduction, the usual epilogue.  Launched right after the fused kernel.
template <class S, bool HAS_VAL>
__global__ void __launch_bounds__(256) pb_long_rows_kernel(const __grid_constant__ SpmvArgs<typename S::T> a,
                                                          const int32_t *__restrict__ long_rows, int n_long) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  const int lane = threadIdx.x & 31;
  const int n_warps = (gridDim.x * blockDim.x) >> 5;
  bool differs = false;
  for (int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_long; w += n_warps) {
    const int r = __ldg(long_rows + w);
    const int b = __ldg(a.row_ptr + r), e = __ldg(a.row_ptr + r + 1);
    T acc = S::zero();
    for (int k = b + lane; k < e; k += 32) {
      const int32_t col = __ldg(a.col + k);
      const T xv = ((uint32_t)col < a.vlength) ? __ldg(a.x + col) : S::zero();
      acc = S::add(acc, S::mult(xv, HAS_VAL ? __ldg(a.val + k) : S::one()));
    }
    for (int off = 16; off > 0; off >>= 1) acc = S::add(acc, shb_from_bits<T>(__shfl_xor_sync(0xffffffffu, shb_bits<T>(acc), off)));
    if (lane == 0) shb_finish_row<S>(a, r, acc, differs);
  }
  shb_publish_changed(a.ctl, differs);
}

// host side (shb_pb.cu)
size_t shb_pb_smem_bytes(const shb_pb *pb);
int shb_pb_reserve_ring(shb_ctx *c, shb_pb *pb, int depth);
