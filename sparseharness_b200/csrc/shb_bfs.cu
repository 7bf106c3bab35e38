// shb_bfs.cu -- BFS levels under the or-and semiring: the row-partitioned, direction-optimising driver
// (shb_bfs_block_*, one process per GPU, frontier bits and counts exchanged by the kernels' own peer stores) and
// shb_bfs, the same kernels on one block that owns every vertex.
//
// Reference counterpart: app/bfs.cpp:94-152 HarnessBFS::executeRun -- a host loop with one launch, one blocking
// read-back of the whole vector and a host-side comparison per step, computing reachability-by-walk-length rather
// than levels.  Here the frontier, the visited set, the levels and the termination state never leave
// the device; the host enqueues batches of levels and polls a 4 KB count ring once per batch.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include <mutex>

#include "shb_internal.cuh"

// The per-operator caches below (row lengths, empty-row bits) are built the first time an operator is traversed.  Two
// contexts or threads may traverse the same operator: the build is serialised, and finished on the device (stream
// synchronise) before the pointer is published, because another context's stream is not ordered after this one's.
static std::mutex g_bfs_cache_mutex;

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
//                       exit; hub rows by the whole warp) and the warp's next-frontier words are stored into every
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
      if (i != next.self) atomicOr_system(&next.bm[i][v >> 5], bit);  // peer mapping: system scope
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

// push: a warp per owned frontier word (32 fetched at a time); lanes stride over each frontier vertex's out-adjacency.  Hub vertices
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
  // Warp i owns words i, i + n_warps, ... (neighbouring words -- the hub region of an RMAT graph -- go to different
  // warps).  It fetches 32 of its words with one load per lane and walks the non-empty ones: a sparse frontier is a
  // scan of the bitmap, and one word per warp step made that scan a chain of dependent loads (30 us on 2^24 vertices).
  for (int64_t base = word_lo + warp; base < word_hi; base += n_warps * 32) {
    const int64_t my_index = base + lane * n_warps;
    const uint32_t my_word = my_index < word_hi ? cur[my_index] : 0u;
    uint32_t live = __ballot_sync(0xffffffffu, my_word != 0);
    while (live) {
      const int src = __ffs(live) - 1;
      live &= live - 1;
      uint32_t word = __shfl_sync(0xffffffffu, my_word, src);
      const int64_t w = base + src * n_warps;
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

// pull, thread per row (SHB_BFS_PULL=rows: kept as the comparison for the compacted mapping below)
template <bool HAS_VAL>
__device__ __forceinline__ void bfs_pull_rows(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                                              const int32_t *__restrict__ val, const int32_t *__restrict__ deg,
                                              const uint32_t *__restrict__ cur, const uint32_t *__restrict__ visited,
                                              const BfsPeers &next, int32_t *levels, long long *acc, uint32_t level,
                                              int64_t n_local, int64_t row_offset) {
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

// Pull, compacted: thread-per-row leaves most lanes idle once a good part of the rows is visited (visited and empty
// rows are skipped; ncu on the first big pull level: 9.7 active threads per instruction).  A warp takes 256 rows,
// compacts the unvisited ones into its own shared-memory list with the 8 visited words, and its lanes then walk that
// list densely (hub rows by the whole warp, as above).  The group's 8 next-frontier words are assembled in shared
// memory and stored once into every rank's bitmap.  No CTA barrier: warps stay independent.  RMAT 24 on one GPU:
// 0.90 -> 0.77 ms for the traversal; groups of 16 or 32 words measured 2-3 % slower than 8.
#define BFS_PULL_GROUP_WORDS 8
template <bool HAS_VAL>
__device__ __forceinline__ void bfs_pull_compact(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                                                 const int32_t *__restrict__ val, const int32_t *__restrict__ deg,
                                                 const uint32_t *__restrict__ cur, const uint32_t *__restrict__ visited,
                                                 const BfsPeers &next, int32_t *levels, long long *acc, uint32_t level,
                                                 int64_t n_local, int64_t row_offset) {
  __shared__ uint8_t s_rows[8][BFS_PULL_GROUP_WORDS * 32];
  __shared__ uint32_t s_bits[8][BFS_PULL_GROUP_WORDS];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t word_offset = row_offset >> 5;
  const int64_t n_words = (n_local + 31) >> 5;
  const int64_t n_groups = (n_words + BFS_PULL_GROUP_WORDS - 1) / BFS_PULL_GROUP_WORDS;
  unsigned long long found_n = 0, push_deg = 0;
  for (int64_t g = warp; g < n_groups; g += n_warps) {
    const int64_t w0 = g * BFS_PULL_GROUP_WORDS;
    const bool word_lane = lane < BFS_PULL_GROUP_WORDS && w0 + lane < n_words;
    const uint32_t seen_mine = word_lane ? visited[word_offset + w0 + lane] : 0xffffffffu;
    if (lane < BFS_PULL_GROUP_WORDS) s_bits[wib][lane] = 0u;
    uint32_t n_act = 0;
#pragma unroll
    for (int j = 0; j < BFS_PULL_GROUP_WORDS; ++j) {
      const uint32_t seen = __shfl_sync(0xffffffffu, seen_mine, j);
      const bool active = (w0 + j) * 32 + lane < n_local && !((seen >> lane) & 1u);
      const uint32_t m = __ballot_sync(0xffffffffu, active);
      SHB_DEV_CHECK(n_act + __popc(m) <= BFS_PULL_GROUP_WORDS * 32);
      if (active) s_rows[wib][n_act + __popc(m & ((1u << lane) - 1u))] = (uint8_t)(j * 32 + lane);
      n_act += __popc(m);
    }
    __syncwarp();
    for (uint32_t i0 = 0; i0 < n_act; i0 += 32) {
      const uint32_t i = i0 + lane;
      bool found = false, big = false;
      int s = 0, e = 0;
      uint32_t rt = 0;
      int64_t r = 0;
      if (i < n_act) {
        rt = s_rows[wib][i];
        r = w0 * 32 + rt;
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
        SHB_DEV_CHECK((rt >> 5) < BFS_PULL_GROUP_WORDS);
        atomicOr(&s_bits[wib][rt >> 5], 1u << (rt & 31));
        ++found_n;
        push_deg += (unsigned long long)(deg ? deg[row_offset + r] : 1);
      }
    }
    __syncwarp();
    if (word_lane) {
      const uint32_t bits = s_bits[wib][lane];
      for (int i = 0; i < next.n; ++i) next.bm[i][word_offset + w0 + lane] = bits;
    }
    __syncwarp();
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

// One launch per level.  by_rows selects the thread-per-row mapping (SHB_BFS_PULL=rows, kept for comparison).
template <bool HAS_VAL>
__global__ void __launch_bounds__(256, 8)
    bfs_block_pull_kernel(const BfsBlockState *st, const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                          const int32_t *__restrict__ val, const int32_t *__restrict__ deg, const uint32_t *__restrict__ cur,
                          const uint32_t *__restrict__ visited, const __grid_constant__ BfsPeers next, int32_t *levels,
                          long long *acc, uint32_t level, int64_t n_local, int64_t row_offset, int by_rows) {
  if (st->done || st->mode != 1) return;
  if (by_rows) bfs_pull_rows<HAS_VAL>(row_ptr, col, val, deg, cur, visited, next, levels, acc, level, n_local, row_offset);
  else bfs_pull_compact<HAS_VAL>(row_ptr, col, val, deg, cur, visited, next, levels, acc, level, n_local, row_offset);
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
  if (A->n_rows > 0) {
    // the empty-row bits depend on the operator only: one pass the first time it is traversed, a 1-bit-per-row copy after
    const size_t iso_bytes = (size_t)((A->n_rows + 31) / 32) * 4;
    {
      std::lock_guard<std::mutex> lk(g_bfs_cache_mutex);
      if (!A->empty_row_bits) {
        uint32_t *bits = nullptr;
        SHB_CUDA(cudaMalloc(&bits, iso_bytes));
        SHB_CUDA(cudaMemsetAsync(bits, 0, iso_bytes, c->stream));
        bfs_block_mark_isolated_kernel<<<grid_for(A->n_rows), 256, 0, c->stream>>>(A->row_ptr, bits, A->n_rows, 0);
        SHB_CUDA(cudaStreamSynchronize(c->stream));
        A->empty_row_bits = bits;
      }
    }
    SHB_CUDA(cudaMemcpyAsync((uint32_t *)visited->ptr + (row_offset >> 5), A->empty_row_bits, iso_bytes,
                             cudaMemcpyDeviceToDevice, c->stream));
  }
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
      const unsigned grid = grid_for((word_hi - word_lo) * 32, 256, c->prop.multiProcessorCount * 8);
      if (AT->val)
        bfs_block_push_kernel<true><<<grid, 256, 0, c->stream>>>(st, AT->row_ptr, AT->col, (const int32_t *)AT->val, dg, cur,
                                                                 (const uint32_t *)visited->ptr, peers, acc, word_lo, word_hi,
                                                                 row_offset, n_local);
      else
        bfs_block_push_kernel<false><<<grid, 256, 0, c->stream>>>(st, AT->row_ptr, AT->col, nullptr, dg, cur,
                                                                  (const uint32_t *)visited->ptr, peers, acc, word_lo, word_hi,
                                                                  row_offset, n_local);
      if (AT->val)
        bfs_block_push_big_kernel<true><<<c->prop.multiProcessorCount * 4, 256, 0, c->stream>>>(st, AT->row_ptr, AT->col, (const int32_t *)AT->val, dg,
                                                                        (const uint32_t *)visited->ptr, peers, acc);
      else
        bfs_block_push_big_kernel<false><<<c->prop.multiProcessorCount * 4, 256, 0, c->stream>>>(st, AT->row_ptr, AT->col, nullptr, dg,
                                                                         (const uint32_t *)visited->ptr, peers, acc);
    }
    static const int by_rows = [] { const char *e = getenv("SHB_BFS_PULL"); return e && !strcmp(e, "rows") ? 1 : 0; }();
    static const int grid_mult = [] { const char *e = getenv("SHB_BFS_PULL_CTAS_PER_SM"); return e ? std::max(1, atoi(e)) : 32; }();
    const unsigned grid = grid_for(n_local, 256, c->prop.multiProcessorCount * grid_mult);
    if (A->val)
      bfs_block_pull_kernel<true><<<grid, 256, 0, c->stream>>>(st, A->row_ptr, A->col, (const int32_t *)A->val, dg, cur,
                                                               (const uint32_t *)visited->ptr, peers, (int32_t *)levels->ptr,
                                                               acc, level, n_local, row_offset, by_rows);
    else
      bfs_block_pull_kernel<false><<<grid, 256, 0, c->stream>>>(st, A->row_ptr, A->col, nullptr, dg, cur,
                                                                (const uint32_t *)visited->ptr, peers, (int32_t *)levels->ptr,
                                                                acc, level, n_local, row_offset, by_rows);
  }
  if (comm.n)
    bfs_block_publish_kernel<<<1, 32, 0, c->stream>>>(st, comm, cnt + BFS_STAGE_WORD, BFS_SLOT(level + 1), BFS_FLAG_WORD,
                                                      self_index, (unsigned long long)barrier_target);
  SHB_CUDA(cudaGetLastError());
  return SHB_OK;
}

// A traversal stopped by its level cap: the vertices of the last frontier built (frontier `level`, by a push level) get
// their level -- a push level leaves that to the NEXT level's prep, which the cap cut off.  (shb_bfs does the same.)
SHB_API int shb_bfs_block_label_frontier(shb_ctx *c, const shb_csr *A, int64_t row_offset, int64_t n_total, uint32_t level,
                                         const shb_buf *frontier, shb_buf *spare, shb_buf *visited, shb_buf *levels,
                                         shb_buf *counters, shb_buf *state) {
  SHB_REQUIRE(c && A && frontier && spare && visited && levels && counters && state, "bad arguments");
  const int64_t n_words = (n_total + 31) / 32;
  const uint64_t bm_bytes = (uint64_t)n_words * 4;
  SHB_REQUIRE(frontier->bytes >= bm_bytes && visited->bytes >= bm_bytes && spare->bytes >= bm_bytes, "bitmap shorter than n_total bits");
  SHB_REQUIRE(spare->ptr != frontier->ptr, "the spare bitmap aliases the current frontier");
  SHB_REQUIRE(levels->bytes >= (uint64_t)A->n_rows * 4 && state->bytes >= sizeof(BfsBlockState), "levels / state too short");
  SHB_CUDA(cudaSetDevice(c->device));
  const int64_t n_local = A->n_rows, word_lo = row_offset >> 5, word_hi = (row_offset + n_local + 31) >> 5;
  bfs_block_prep_kernel<<<grid_for(n_words), 256, 0, c->stream>>>((BfsBlockState *)state->ptr, (long long *)counters->ptr,
                                                                  (const uint32_t *)frontier->ptr, (uint32_t *)visited->ptr,
                                                                  (uint32_t *)spare->ptr, (int32_t *)levels->ptr, level, n_words,
                                                                  word_lo, word_hi, row_offset, n_local);
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
  if (AT && n > 0) {
    std::lock_guard<std::mutex> lk(g_bfs_cache_mutex);
    if (!AT->row_len) {
      int32_t *rl = nullptr;
      SHB_CUDA(cudaMalloc(&rl, (size_t)n * 4));
      csr_row_lengths_kernel<<<grid_for(n), 256, 0, c->stream>>>(AT->row_ptr, rl, n);
      SHB_CUDA(cudaStreamSynchronize(c->stream));
      AT->row_len = rl;
    }
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
  // the traversal ends here: levels are final.  Counting what was reached (the TEPS numerator) is reporting.
  cudaEventRecord(c->ev1, c->stream);
  int fin = shb_bfs_block_finish(c, A, levels, &counters, &state);
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

