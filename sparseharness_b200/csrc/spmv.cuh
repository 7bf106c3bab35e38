// spmv.cuh -- the hot path: hand-written sm_100a semiring CSR SpMV kernels.
//
// Replaces the Lift-generated OpenCL "KERNEL" of the reference (example/<alg>/kernel{,2..7}.json:3):
//   out[r] = epi( (+)_k mult( guard(x, col[k]), val[k] ), alpha, y[r], beta )
// Three schedules, all templated on the semiring:
//   spmv_scalar_kernel  thread per row, forward (or reversed) sequential order == the reference's
//                       glb-sdp / glb-sdp-rsa schedules (kernel5/6.json) and the gold loop order
//   spmv_vector_kernel  sub-warp of L lanes per row, R rows in flight per sub-warp, 128-bit
//                       coalesced loads of col/val for the 16-byte aligned body of a row, scalar
//                       predicated loads for the <=3-element head and tail, shuffle reduction,
//                       fused epilogue + convergence test + (multi-GPU) scatter to peer buffers
//   spmv_stream_kernel  nnz-split tiles (perfect load balance for power-law rows): a CTA streams
//                       SHB_TILE_NNZ consecutive non-zeros, multiplies, parks the products in
//                       padded shared memory, then reduces the rows of the tile (thread per short
//                       row piece, warp per long piece); pieces of rows that straddle tiles go to
//                       carry buffers and are combined in tile order by spmv_fixup_kernel
//                       (deterministic: no float atomics)
// Padding entries (col = -1) gather ZERO, and mult(ZERO, v) is the additive identity in all four
// semirings -- the same property the reference's ELL padding relies on (src/sparse_matrix.cpp:318-328).
#pragma once
#include "ldst.cuh"
#include "semiring.cuh"
#include "shb_internal.cuh"

#define SHB_MAX_DST 8

template <class T> struct SpmvArgs {
  const int32_t *row_ptr;
  const int32_t *row_ids;  // nullptr, or: row r of the operator is element row_ids[r] of y / out (compressed row list)
  const int32_t *col;
  const T *val;  // nullptr = pattern
  const T *x;
  const T *y;    // indexed by row_offset + r
  T alpha, beta;
  T *out[SHB_MAX_DST];  // every destination receives out[row_offset + r]
  int n_out;
  int use_y;            // 0: epilogue ignores y, 1: read y[row_offset + r], 2: y is the semiring one everywhere
  int reversed;         // scalar kernel only: accumulate from the row's last entry to its first
  int64_t row_offset;
  int64_t n_rows;
  uint32_t vlength;
  IterCtl ctl;
};

template <class S>
__device__ __forceinline__ void shb_finish_row(const SpmvArgs<typename S::T> &a, int64_t r, typename S::T dp,
                                               bool &differs) {
  typedef typename S::T T;
  const int64_t gr = a.row_offset + (a.row_ids ? (int64_t)a.row_ids[r] : r);
  const T yv = a.use_y == 1 ? a.y[gr] : (a.use_y == 2 ? S::one() : S::zero());
  const T o = S::epi(dp, a.alpha, yv, a.beta);
#pragma unroll 1
  for (int i = 0; i < a.n_out; ++i) a.out[i][gr] = o;
  if (a.ctl.changed) differs |= shb_row_differs(a.x[gr], o, a.ctl.delta, a.ctl.term_mode);
}

__device__ __forceinline__ void shb_publish_changed(const IterCtl &ctl, bool differs) {
  if (ctl.changed) {
    const bool any = __any_sync(0xffffffffu, differs);
    if (any && (threadIdx.x & 31) == 0 && shb_ld_volatile_u32(ctl.changed) == 0) *ctl.changed = 1u;
  }
}

// ---------------------------------------------------------------------------------------------
// scalar: thread per row
// ---------------------------------------------------------------------------------------------
template <class S, bool HAS_VAL>
__global__ void __launch_bounds__(256) spmv_scalar_kernel(const __grid_constant__ SpmvArgs<typename S::T> a) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  bool differs = false;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t n_padded = (a.n_rows + 31) / 32 * 32;  // keep warps converged for the ballot
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_padded; r += stride) {
    if (r < a.n_rows) {
      const int s = a.row_ptr[r], e = a.row_ptr[r + 1];
      T acc = S::zero();
      for (int i = 0; i < e - s; ++i) {
        const int k = a.reversed ? (e - 1 - i) : (s + i);
        const int32_t c = __ldg(a.col + k);
        const T xv = ((uint32_t)c < a.vlength) ? __ldg(a.x + c) : S::zero();
        const T v = HAS_VAL ? __ldg(a.val + k) : S::one();
        acc = S::add(acc, S::mult(xv, v));
      }
      shb_finish_row<S>(a, r, acc, differs);
    }
  }
  shb_publish_changed(a.ctl, differs);
}

// ---------------------------------------------------------------------------------------------
// vector: L lanes per row, R rows in flight per lane group
// ---------------------------------------------------------------------------------------------
template <class S, int L, int R, bool HAS_VAL, bool HINTS>
__global__ void __launch_bounds__(256) spmv_vector_kernel(const __grid_constant__ SpmvArgs<typename S::T> a) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  constexpr int G = 32 / L;  // rows handled side by side in one warp
  constexpr int ROWS_PER_WARP = G * R;
  const int lane = threadIdx.x & 31;
  const int g = lane / L, li = lane % L;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t base = warp * ROWS_PER_WARP;
  const uint64_t pol_stream = HINTS ? shb_policy_evict_first() : 0ull;
  const uint64_t pol_keep = HINTS ? shb_policy_evict_last() : 0ull;
  const uint32_t zbits = shb_bits<T>(S::zero());
  const uint32_t obits = shb_bits<T>(S::one());

  int s[R], e[R], ha[R], tb[R];
  T acc[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int64_t r = base + (int64_t)j * G + g;
    const bool valid = r < a.n_rows;
    s[j] = valid ? __ldg(a.row_ptr + r) : 0;
    e[j] = valid ? __ldg(a.row_ptr + r + 1) : 0;
    ha[j] = min(e[j], (s[j] + 3) & ~3);  // end of the unaligned head
    tb[j] = max(ha[j], e[j] & ~3);       // start of the unaligned tail
    acc[j] = S::zero();
  }

  // head + tail: at most 3 + 3 scalar elements per row, spread over the lanes of the group
  constexpr int EDGE_STEPS = (6 + L - 1) / L;
#pragma unroll
  for (int k = 0; k < EDGE_STEPS; ++k) {
    uint32_t cb[R], vb[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const int idx = k * L + li;
      const int nh = ha[j] - s[j];
      const int pos = idx < nh ? s[j] + idx : tb[j] + (idx - nh);
      const bool p = pos < e[j];
      cb[j] = shb_ld_stream_b32<HINTS>(a.col + pos, p, 0xffffffffu, pol_stream);
      vb[j] = HAS_VAL ? shb_ld_stream_b32<HINTS>(a.val + pos, p, zbits, pol_stream) : obits;
    }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const T xv = shb_from_bits<T>(shb_gather_b32<HINTS>(a.x, (int32_t)cb[j], a.vlength, zbits, pol_keep));
      acc[j] = S::add(acc[j], S::mult(xv, shb_from_bits<T>(vb[j])));
    }
  }

  // aligned body: 128-bit loads, 4 elements per lane per step
  int p[R];
  bool more = false;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    p[j] = ha[j] + 4 * li;
    more |= (p[j] + 4 <= tb[j]);
  }
  while (more) {
    more = false;
    uint4 c4[R], v4[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const bool pr = p[j] + 4 <= tb[j];
      c4[j] = shb_ld_stream_v4<HINTS>(a.col + p[j], pr, 0xffffffffu, pol_stream);
      if (HAS_VAL) v4[j] = shb_ld_stream_v4<HINTS>(a.val + p[j], pr, zbits, pol_stream);
      else v4[j] = make_uint4(obits, obits, obits, obits);
    }
    uint32_t xb[R][4];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      xb[j][0] = shb_gather_b32<HINTS>(a.x, (int32_t)c4[j].x, a.vlength, zbits, pol_keep);
      xb[j][1] = shb_gather_b32<HINTS>(a.x, (int32_t)c4[j].y, a.vlength, zbits, pol_keep);
      xb[j][2] = shb_gather_b32<HINTS>(a.x, (int32_t)c4[j].z, a.vlength, zbits, pol_keep);
      xb[j][3] = shb_gather_b32<HINTS>(a.x, (int32_t)c4[j].w, a.vlength, zbits, pol_keep);
    }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      acc[j] = S::add(acc[j], S::mult(shb_from_bits<T>(xb[j][0]), shb_from_bits<T>(v4[j].x)));
      acc[j] = S::add(acc[j], S::mult(shb_from_bits<T>(xb[j][1]), shb_from_bits<T>(v4[j].y)));
      acc[j] = S::add(acc[j], S::mult(shb_from_bits<T>(xb[j][2]), shb_from_bits<T>(v4[j].z)));
      acc[j] = S::add(acc[j], S::mult(shb_from_bits<T>(xb[j][3]), shb_from_bits<T>(v4[j].w)));
      p[j] += 4 * L;
      more |= (p[j] + 4 <= tb[j]);
    }
  }

  // butterfly reduction inside each lane group: afterwards every lane of the group holds the row sum
#pragma unroll
  for (int j = 0; j < R; ++j) {
#pragma unroll
    for (int off = L / 2; off > 0; off >>= 1) acc[j] = S::add(acc[j], __shfl_xor_sync(0xffffffffu, acc[j], off));
  }

  // epilogue: lane li of group g finishes the rows j with j % L == li, so the warp's stores cover
  // ROWS_PER_WARP consecutive rows
  bool differs = false;
  if (R <= L) {
    T mine = acc[0];
#pragma unroll
    for (int j = 1; j < R; ++j)
      if (li == j) mine = acc[j];
    const int64_t r = base + (int64_t)li * G + g;
    if (li < R && r < a.n_rows) shb_finish_row<S>(a, r, mine, differs);
  } else {
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const int64_t r = base + (int64_t)j * G + g;
      if ((j % L) == li && r < a.n_rows) shb_finish_row<S>(a, r, acc[j], differs);
    }
  }
  shb_publish_changed(a.ctl, differs);
}

// ---------------------------------------------------------------------------------------------
// stream: nnz-split tiles through shared memory
// ---------------------------------------------------------------------------------------------
#define SHB_STREAM_THREADS 256
#define SHB_LONG_PIECE 64      // default: pieces longer than this go to a warp each (phase 3)
#define SHB_LONG_PIECE_MIN 8   // smallest value the launch argument may take (sizes the shared list)
__device__ __forceinline__ int shb_pad(int i) { return i + (i >> 5); }  // 1 pad word per 32: conflict-free strides

// MERGE: the row phase for SHORT, SKEWED rows (column panels of power-law operators: mean 2-4 entries, a few of 30-60 in
// every warp's 32 rows, so the thread-per-row loop below runs at the longest row's length: 8.3 warp instructions per row
// on BASELINE config 5's panels).  Instead every thread walks an equal share of the merged list (row-end markers,
// entries) -- a merge-path partition over the tile's row_ptr segment staged in shared memory -- dropping each finished
// row's sum into shared memory; partial sums of rows that span threads are joined in thread order; a thread-per-row
// pass then runs the epilogue with coalesced y / out accesses.  Same ownership rules for straddling and empty rows as
// the loop it replaces; tiles touching more than SHB_MERGE_ROWS rows fall back to that loop.
#define SHB_MERGE_ROWS 2048

// SCAN (MODE 2): the row phase as a segmented scan.  Row starts inside the tile are flagged in a 4096-bit shared-memory
// mask (thread per row, while the tile's loads are in flight); every thread then takes 16 CONSECUTIVE products into
// registers, scans them with the flags, the open segments are joined across lanes (shuffles) and warps (8 values in
// shared memory), and the inclusive segmented sums go back over the products.  A row's sum is then ONE shared-memory
// read -- the element at its last entry -- whatever its length: no per-row loop, no divergence on skewed rows, no
// separate pass for long rows.  Association order differs from the loop's (deterministic; exact for min-plus / or-and /
// max-min and integer-valued floats, within the stated 1e-5 for general floats).
// TMA (MODE 3): phase 1 through the TMA unit -- one thread starts two 16 KB bulk copies (cp.async.bulk, completion on an
// mbarrier) of the tile's columns and values into shared memory and every thread then takes its entries from there, so
// the operator streams do not go through the LSU / L1TEX request path the gathers saturate.  Row phase = the loop's.
// MODE: 0 = thread-per-row loop, 1 = MERGE, 2 = SCAN, 3 = TMA.
__device__ __forceinline__ uint32_t shb_smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <class S, bool HAS_VAL, bool HINTS, int MODE>
__global__ void __launch_bounds__(SHB_STREAM_THREADS)
    spmv_stream_kernel(const __grid_constant__ SpmvArgs<typename S::T> a, const int32_t *__restrict__ tile_row,
                       int n_tiles, int32_t nnz, typename S::T *__restrict__ carry_tail,
                       typename S::T *__restrict__ carry_head, int long_piece) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  __shared__ __align__(128) T prod[SHB_TILE_NNZ + SHB_TILE_NNZ / 32];
  __shared__ int long_rows[SHB_TILE_NNZ / SHB_LONG_PIECE_MIN + 2];
  __shared__ int n_long;
  constexpr bool MERGE = MODE == 1;
  constexpr bool SCAN = MODE == 2;
  constexpr bool TMA = MODE == 3;
  __shared__ __align__(128) uint32_t tma_col[TMA ? SHB_TILE_NNZ : 4];
  uint32_t *const tma_val = reinterpret_cast<uint32_t *>(prod);  // the values land where the products go later (a barrier between)
  __shared__ __align__(8) uint64_t tma_bar;
  static_assert(SHB_TILE_NNZ == 16 * SHB_STREAM_THREADS, "the scan row phase gives every thread 16 consecutive products");
  __shared__ uint32_t flag_s[SCAN ? SHB_TILE_NNZ / 32 : 1];      // bit k: a row starts at entry lo + k
  __shared__ T wtail_s[SCAN ? SHB_STREAM_THREADS / 32 : 1];      // per warp: sum of the segment open at its end
  __shared__ int wflag_s[SCAN ? SHB_STREAM_THREADS / 32 : 1];    // per warp: a row starts somewhere in it
  __shared__ int rp_s[MERGE ? SHB_MERGE_ROWS + 1 : 1];
  __shared__ T rowsum[MERGE ? SHB_MERGE_ROWS : 1];
  __shared__ T carry_s[MERGE ? SHB_STREAM_THREADS / 32 : 1];   // per warp: scanned trailing sum of its last lane
  __shared__ int first_s[MERGE ? SHB_STREAM_THREADS / 32 : 1];  // per warp: a segment starts somewhere in it
  __shared__ T scan_s[MERGE ? SHB_STREAM_THREADS : 1];
  const int tid = threadIdx.x;
  const int t = blockIdx.x;
  const int lo = t * SHB_TILE_NNZ;
  const int hi = min(nnz, lo + SHB_TILE_NNZ);
  const uint64_t pol_stream = HINTS ? shb_policy_evict_first() : 0ull;
  const uint64_t pol_keep = HINTS ? shb_policy_evict_last() : 0ull;
  const uint32_t zbits = shb_bits<T>(S::zero());
  const uint32_t obits = shb_bits<T>(S::one());
  if (tid == 0) n_long = 0;
  if (SCAN) {
    if (tid < SHB_TILE_NNZ / 32) flag_s[tid] = 0u;
    __syncthreads();
  }

  // phase 1: stream the tile's non-zeros (the arrays are padded to a tile multiple with col = -1)
  constexpr int STEPS = SHB_TILE_NNZ / (SHB_STREAM_THREADS * 4);
  uint4 c4[STEPS], v4[STEPS];
  if (TMA) {
    const uint32_t bar = shb_smem_addr(&tma_bar);
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      const uint32_t bytes = SHB_TILE_NNZ * 4u;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(HAS_VAL ? 2u * bytes : bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(shb_smem_addr(tma_col)),
                   "l"(a.col + lo), "r"(bytes), "r"(bar)
                   : "memory");
      if (HAS_VAL)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(shb_smem_addr(tma_val)),
                     "l"(a.val + lo), "r"(bytes), "r"(bar)
                     : "memory");
    }
    __syncthreads();  // the barrier is initialised before anyone polls it
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok)
                   : "r"(bar), "r"(0u)
                   : "memory");
#pragma unroll
    for (int k = 0; k < STEPS; ++k) {
      const int i = 4 * (tid + SHB_STREAM_THREADS * k);
      c4[k] = *reinterpret_cast<const uint4 *>(tma_col + i);
      if (HAS_VAL) v4[k] = *reinterpret_cast<const uint4 *>(tma_val + i);
      else v4[k] = make_uint4(obits, obits, obits, obits);
    }
  } else {
#pragma unroll
    for (int k = 0; k < STEPS; ++k) {
      const int i = lo + 4 * (tid + SHB_STREAM_THREADS * k);
      c4[k] = shb_ld_stream_v4<HINTS>(a.col + i, true, 0xffffffffu, pol_stream);
      if (HAS_VAL) v4[k] = shb_ld_stream_v4<HINTS>(a.val + i, true, zbits, pol_stream);
      else v4[k] = make_uint4(obits, obits, obits, obits);
    }
  }
  if (SCAN) {  // row starts of this tile -> the mask (the loads above are in flight meanwhile)
    const int rf = tile_row[t];
    const int rl = (t + 1 < n_tiles) ? tile_row[t + 1] : (int)(a.n_rows - 1);
    for (int r = rf + tid; r <= rl; r += SHB_STREAM_THREADS) {
      const int s0 = a.row_ptr[r];
      if (s0 >= lo && s0 < hi) atomicOr(&flag_s[(s0 - lo) >> 5], 1u << ((s0 - lo) & 31));
    }
  }
  uint32_t xb[STEPS][4];
#pragma unroll
  for (int k = 0; k < STEPS; ++k) {
    xb[k][0] = shb_gather_b32<HINTS>(a.x, (int32_t)c4[k].x, a.vlength, zbits, pol_keep);
    xb[k][1] = shb_gather_b32<HINTS>(a.x, (int32_t)c4[k].y, a.vlength, zbits, pol_keep);
    xb[k][2] = shb_gather_b32<HINTS>(a.x, (int32_t)c4[k].z, a.vlength, zbits, pol_keep);
    xb[k][3] = shb_gather_b32<HINTS>(a.x, (int32_t)c4[k].w, a.vlength, zbits, pol_keep);
  }
  if (TMA) __syncthreads();  // every thread holds its values in registers before products overwrite them
#pragma unroll
  for (int k = 0; k < STEPS; ++k) {
    const int i = 4 * (tid + SHB_STREAM_THREADS * k);
    prod[shb_pad(i + 0)] = S::mult(shb_from_bits<T>(xb[k][0]), shb_from_bits<T>(v4[k].x));
    prod[shb_pad(i + 1)] = S::mult(shb_from_bits<T>(xb[k][1]), shb_from_bits<T>(v4[k].y));
    prod[shb_pad(i + 2)] = S::mult(shb_from_bits<T>(xb[k][2]), shb_from_bits<T>(v4[k].z));
    prod[shb_pad(i + 3)] = S::mult(shb_from_bits<T>(xb[k][3]), shb_from_bits<T>(v4[k].w));
  }
  // phase 2: the rows of this tile.  Row r's piece is [max(s,lo), min(e,hi)).
  const int r_first = tile_row[t];
  const int r_last = (t + 1 < n_tiles) ? tile_row[t + 1] : (int)(a.n_rows - 1);
  const int nr = r_last - r_first + 1;
  const bool merge = MERGE && nr <= SHB_MERGE_ROWS;
  if (merge)
    for (int i = tid; i <= nr; i += SHB_STREAM_THREADS) rp_s[i] = a.row_ptr[r_first + i];
  __syncthreads();
  bool differs = false;
  if (SCAN) {
    const int lane = tid & 31, warp = tid >> 5;
    const int base = tid * 16;
    T sv[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) sv[j] = prod[shb_pad(base + j)];
    const uint32_t f = (flag_s[tid >> 1] >> ((tid & 1) * 16)) & 0xffffu;
    T run = S::zero();
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      run = ((f >> j) & 1u) ? sv[j] : S::add(run, sv[j]);
      sv[j] = run;
    }
    // join the open segments across the warp: inclusive segmented scan of (tail, "a row starts in this thread")
    uint32_t v = shb_bits<T>(run);
    int hf = f != 0u;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t pv = __shfl_up_sync(0xffffffffu, v, off);
      const int pf = __shfl_up_sync(0xffffffffu, hf, off);
      if (lane >= off) {
        if (!hf) v = shb_bits<T>(S::add(shb_from_bits<T>(pv), shb_from_bits<T>(v)));
        hf |= pf;
      }
    }
    if (lane == 31) {
      wtail_s[warp] = shb_from_bits<T>(v);
      wflag_s[warp] = hf;
    }
    const uint32_t vprev = __shfl_up_sync(0xffffffffu, v, 1);   // the lanes before this one, scanned
    const int fprev = __shfl_up_sync(0xffffffffu, hf, 1);       // ... does a row start among them?
    __syncthreads();
    T pre = S::zero();  // the segment open at the end of the warps before this one
    for (int w = 0; w < warp; ++w) pre = wflag_s[w] ? wtail_s[w] : S::add(pre, wtail_s[w]);
    const T cin = lane == 0 ? pre : (fprev ? shb_from_bits<T>(vprev) : S::add(pre, shb_from_bits<T>(vprev)));
    const int n_open = f ? (__ffs(f) - 1) : 16;  // leading entries that continue a row begun before this thread
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (j < n_open) sv[j] = S::add(cin, sv[j]);
      prod[shb_pad(base + j)] = sv[j];
    }
    __syncthreads();
    // thread per row: its sum is the scanned element at its last entry in the tile
    for (int r = r_first + tid; r <= r_last; r += SHB_STREAM_THREADS) {
      SHB_DEV_CHECK(r >= 0 && r < a.n_rows);
      const int s0 = a.row_ptr[r], e0 = a.row_ptr[r + 1];
      if (e0 == s0) {
        if ((s0 > lo && s0 <= hi) || (s0 == 0 && t == 0)) shb_finish_row<S>(a, r, S::zero(), differs);
      } else {
        const int ps = max(s0, lo), pe = min(e0, hi);
        if (pe > ps) {
          SHB_DEV_CHECK(pe - 1 - lo >= 0 && pe - 1 - lo < SHB_TILE_NNZ);
          const T sum = prod[shb_pad(pe - 1 - lo)];
          if (s0 >= lo && e0 <= hi) shb_finish_row<S>(a, r, sum, differs);
          else if (e0 > hi) carry_tail[t] = sum;  // continues into the next tile
          else carry_head[t] = sum;                // started in an earlier tile, ends here
        }
      }
    }
    shb_publish_changed(a.ctl, differs);
    return;
  }
  if (merge) {
    const int ne = hi - lo;
    auto rend = [&](int i) { return min(max(rp_s[i + 1] - lo, 0), ne); };  // end of row i inside the tile's entries
    // merge-path split: marker i precedes entry j iff rend(i) <= j; thread tid takes merged items [d0, d1)
    const int total = nr + ne;
    const int ipt = (total + SHB_STREAM_THREADS - 1) / SHB_STREAM_THREADS;
    const int d0 = min(tid * ipt, total), d1 = min(d0 + ipt, total);
    int i_lo = max(0, d0 - ne), i_hi = min(d0, nr);
    while (i_lo < i_hi) {
      const int mid = (i_lo + i_hi) >> 1;
      if (rend(mid) <= d0 - 1 - mid) i_lo = mid + 1;
      else i_hi = mid;
    }
    int i = i_lo, j = d0 - i_lo;
    int cur_end = i < nr ? rend(i) : ne + 1;
    int first = -1;  // first row this thread finishes
    T acc = S::zero();
    for (int step = d0; step < d1; ++step) {
      if (j < cur_end) {
        SHB_DEV_CHECK(j >= 0 && j < ne);
        acc = S::add(acc, prod[shb_pad(j)]);
        ++j;
      } else {
        SHB_DEV_CHECK(i >= 0 && i < nr);
        rowsum[i] = acc;
        if (first < 0) first = i;
        acc = S::zero();
        ++i;
        cur_end = i < nr ? rend(i) : ne + 1;
      }
    }
    // `acc` is what this thread summed of the row it leaves unfinished.  The part of row `first` that earlier threads
    // summed = the trailing sums from the thread that started it up to tid - 1 (the ones in between finished no row):
    // a segmented inclusive scan over the CTA in thread order, a segment starting at every thread that finished a row.
    const int lane = tid & 31, warp = tid >> 5;
    uint32_t v = shb_bits<T>(acc);
    int f = (first >= 0 || tid == 0) ? 1 : 0;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t pv = __shfl_up_sync(0xffffffffu, v, off);
      const int pf = __shfl_up_sync(0xffffffffu, f, off);
      if (lane >= off) {
        if (!f) v = shb_bits<T>(S::add(shb_from_bits<T>(pv), shb_from_bits<T>(v)));
        f |= pf;
      }
    }
    if (lane == 31) {
      carry_s[warp] = shb_from_bits<T>(v);
      first_s[warp] = f;
    }
    __syncthreads();
    if (!f && warp > 0) {  // no segment start in this warp up to here: the sums of the warps before flow in
      int w0 = warp - 1;
      while (w0 > 0 && !first_s[w0]) --w0;
      T pre = carry_s[w0];
      for (int w = w0 + 1; w < warp; ++w) pre = S::add(pre, carry_s[w]);
      v = shb_bits<T>(S::add(pre, shb_from_bits<T>(v)));
    }
    scan_s[tid] = shb_from_bits<T>(v);
    __syncthreads();
    if (first >= 0 && tid > 0) rowsum[first] = S::add(scan_s[tid - 1], rowsum[first]);
    __syncthreads();
    for (int i2 = tid; i2 < nr; i2 += SHB_STREAM_THREADS) {
      const int r = r_first + i2;
      const int s = rp_s[i2], e = rp_s[i2 + 1];
      if (e == s) {
        if ((s > lo && s <= hi) || (s == 0 && t == 0)) shb_finish_row<S>(a, r, S::zero(), differs);
      } else if (min(e, hi) > max(s, lo)) {
        const T sum = rowsum[i2];
        if (s >= lo && e <= hi) shb_finish_row<S>(a, r, sum, differs);
        else if (e > hi) carry_tail[t] = sum;  // continues into the next tile
        else carry_head[t] = sum;               // started in an earlier tile, ends here
      }
    }
    shb_publish_changed(a.ctl, differs);
    return;
  }
  const int n_iter = (r_last - r_first + 1 + SHB_STREAM_THREADS - 1) / SHB_STREAM_THREADS;
  for (int it = 0; it < n_iter; ++it) {
    const int r = r_first + it * SHB_STREAM_THREADS + tid;
    if (r <= r_last) {
      SHB_DEV_CHECK(r >= 0 && r < a.n_rows);
      const int s = a.row_ptr[r], e = a.row_ptr[r + 1];
      SHB_DEV_CHECK(s <= e);
      if (e == s) {
        // empty row: owned by the tile with lo < s <= hi (or tile 0 when s == 0)
        if ((s > lo && s <= hi) || (s == 0 && t == 0)) shb_finish_row<S>(a, r, S::zero(), differs);
      } else {
        const int ps = max(s, lo), pe = min(e, hi);
        if (pe > ps) {
          SHB_DEV_CHECK(ps >= lo && pe <= lo + SHB_TILE_NNZ);
          if (pe - ps > long_piece) {
            const int li = atomicAdd(&n_long, 1);
            SHB_DEV_CHECK(li < SHB_TILE_NNZ / SHB_LONG_PIECE_MIN + 2);
            long_rows[li] = r;
          } else {
            T acc = S::zero();
            for (int k = ps; k < pe; ++k) acc = S::add(acc, prod[shb_pad(k - lo)]);
            if (s >= lo && e <= hi) shb_finish_row<S>(a, r, acc, differs);
            else if (e > hi) carry_tail[t] = acc;  // continues into the next tile
            else carry_head[t] = acc;               // started in an earlier tile, ends here
          }
        }
      }
    }
  }
  __syncthreads();
  // phase 3: long pieces, one warp each (lanes stride, butterfly reduce)
  const int nl = n_long;
  for (int q = tid >> 5; q < nl; q += SHB_STREAM_THREADS / 32) {
    const int r = long_rows[q];
    const int s = a.row_ptr[r], e = a.row_ptr[r + 1];
    const int ps = max(s, lo), pe = min(e, hi);
    T acc = S::zero();
    for (int k = ps + (tid & 31); k < pe; k += 32) acc = S::add(acc, prod[shb_pad(k - lo)]);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc = S::add(acc, __shfl_xor_sync(0xffffffffu, acc, off));
    if ((tid & 31) == 0) {
      if (s >= lo && e <= hi) shb_finish_row<S>(a, r, acc, differs);
      else if (e > hi) carry_tail[t] = acc;
      else carry_head[t] = acc;
    }
  }
  shb_publish_changed(a.ctl, differs);
}

// Rows that straddle tile boundaries: the tile in which such a row ENDS (t >= 1, row = tile_row[t],
// s < lo, e <= hi) combines carry_tail of the tiles it crossed, in tile order, with its own head.
template <class S>
__global__ void __launch_bounds__(256)
    spmv_fixup_kernel(const __grid_constant__ SpmvArgs<typename S::T> a, const int32_t *__restrict__ tile_row,
                      int n_tiles, int32_t nnz, const typename S::T *carry_tail, const typename S::T *carry_head) {
  typedef typename S::T T;
  if (a.ctl.prev_changed && *a.ctl.prev_changed == 0) return;
  bool differs = false;
  const int n_padded = (n_tiles + 31) / 32 * 32;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n_padded; t += gridDim.x * blockDim.x) {
    if (t >= 1 && t < n_tiles) {
      const int lo = t * SHB_TILE_NNZ;
      const int hi = min(nnz, lo + SHB_TILE_NNZ);
      const int r = tile_row[t];
      const int s = a.row_ptr[r], e = a.row_ptr[r + 1];
      if (s < lo && e > lo && e <= hi) {
        const int t0 = s / SHB_TILE_NNZ;
        T acc = carry_tail[t0];
        for (int tt = t0 + 1; tt < t; ++tt) acc = S::add(acc, carry_tail[tt]);
        acc = S::add(acc, carry_head[t]);
        shb_finish_row<S>(a, r, acc, differs);
      }
    }
  }
  shb_publish_changed(a.ctl, differs);
}
