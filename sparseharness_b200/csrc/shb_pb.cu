// shb_pb.cu -- builds the BLOCKED (propagation-blocked) copy of an operator on the device (see pb.cuh).
//
// The reference re-encodes every matrix into the layout its kernel wants before the run
// (SparseMatrix::cl_encode / executorEncodeMatrix, src/sparse_matrix.cpp:122-399, inc/kernel_utils.h:35-168);
// this is the same step for the blocked schedule:
//   1. a stable sort of the stored entries by (pass, producer, panel, consumer): the order the producers
//      stream them in -- one contiguous stream per (pass, producer); CSR order is kept inside each piece;
//   2. the host lays the pieces out twice from their lengths: producer order (pieces padded to 4 entries,
//      units to a 256-entry chunk) and consumer order (pass, panel, consumer, producer; pieces padded to 4),
//      which gives every piece its `delta` = ring offset - stream position, every chunk its piece-start
//      flags, every unit its first chunk and every consumer region its start;
//   3. a second stable sort of the entries by (pass, panel, consumer, row): the order of each consumer's
//      row-sorted shared-memory buffer, which gives every entry its slot and every row its offsets.
// cub::DeviceRadixSort does both sorts; they run once per operator, never on the SpMV path.
#include <algorithm>
#include <cub/device/device_radix_sort.cuh>
#include <vector>

#include "pb.cuh"

static const uint32_t PB_INVALID_KEY = 0xffffffffu;

// one thread per row: bucket key, source index and local row of every stored entry
// rows with more than lmax entries in one column pass do not fit the consumer's fixed row slots: they are marked
// "long", left out of the blocked copy and computed by pb_long_rows_kernel
__global__ void pb_mark_long_kernel(PbGeom g, int lmax, const int32_t *row_ptr, const int32_t *col, int64_t n_rows, uint32_t n_cols,
                                    uint32_t *long_bits, int32_t *long_rows, int32_t *n_long) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    const int b = row_ptr[r], e = row_ptr[r + 1];
    bool is_long = false;
    if (e - b > lmax) {
      unsigned char cnt[PB_MAX_PASS];
      for (int p = 0; p < g.n_pass; ++p) cnt[p] = 0;
      const uint32_t pass_cols = (uint32_t)g.W * (uint32_t)g.n_prod;
      for (int k = b; k < e && !is_long; ++k) {
        const uint32_t c = (uint32_t)col[k];
        if (c < n_cols && ++cnt[c / pass_cols] > lmax) is_long = true;
      }
    }
    if (is_long) {
      atomicOr(&long_bits[r >> 5], 1u << (r & 31));
      long_rows[atomicAdd(n_long, 1)] = (int32_t)r;
    }
  }
}

__global__ void pb_keys_kernel(PbGeom g, const int32_t *row_ptr, const int32_t *col, int64_t n_rows, uint32_t n_cols,
                               const uint32_t *long_bits, uint32_t *keys, uint32_t *src, uint16_t *lrow_tmp) {
  const int64_t PR = (int64_t)g.RS * g.n_prod;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    const int panel = (int)(r / PR);
    const int in_panel = (int)(r - (int64_t)panel * PR);
    const int s = in_panel / g.RS;
    const uint16_t lr = (uint16_t)(in_panel - s * g.RS);
    const int b = row_ptr[r], e = row_ptr[r + 1];
    const bool is_long = (long_bits[r >> 5] >> (r & 31)) & 1u;
    for (int k = b; k < e; ++k) {
      const uint32_t c = (uint32_t)col[k];
      uint32_t key = PB_INVALID_KEY;  // out-of-range column: x reads as ZERO, the product is the additive identity
      if (c < n_cols && !is_long) {
        const uint32_t slice = c / (uint32_t)g.W;  // global slice index = pass * n_prod + p
        key = ((slice * (uint32_t)g.n_panel + (uint32_t)panel)) * (uint32_t)g.n_prod + (uint32_t)s;
      }
      keys[k] = key;
      src[k] = (uint32_t)k;
      lrow_tmp[k] = lr;
    }
  }
}

// bound[k] = first sorted position whose key is >= k, for k in [0, n_keys]
__global__ void pb_bounds_kernel(const uint32_t *sorted_keys, int64_t n, uint32_t n_keys, int32_t *bound) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t prev = i > 0 ? (int64_t)min(sorted_keys[i - 1], n_keys) : -1;
    const int64_t cur = i < n ? (int64_t)min(sorted_keys[i], n_keys) : (int64_t)n_keys;
    for (int64_t k = prev + 1; k <= cur; ++k) bound[k] = (int32_t)i;
  }
}

// producer-order scatter of (local column, value); emits the consumer-order key and position of every entry
__global__ void pb_scatter_kernel(PbGeom g, const uint32_t *sorted_keys, const uint32_t *perm, int64_t n_valid,
                                  const int32_t *piece_start, const int32_t *piece_ppos, const int32_t *piece_cpos,
                                  const int32_t *col, const uint32_t *val, const uint16_t *lrow_tmp, uint16_t *lcol,
                                  uint32_t *oval, uint32_t *key2, uint32_t *cdst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_valid; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t key = sorted_keys[i];
    const uint32_t u = key / (uint32_t)g.n_prod, s = key - u * (uint32_t)g.n_prod;  // u = slice * n_panel + panel
    const uint32_t slice = u / (uint32_t)g.n_panel, panel = u - slice * (uint32_t)g.n_panel;
    const uint32_t pass = slice / (uint32_t)g.n_prod;
    const uint32_t grp = pass * (uint32_t)g.n_panel + panel;
    const int64_t in_piece = i - (int64_t)piece_start[key];
    const int64_t dst = (int64_t)piece_ppos[key] + in_piece;
    const uint32_t j = perm[i];
    const uint32_t c = (uint32_t)col[j];
    lcol[dst] = (uint16_t)(c % (uint32_t)g.W);
    if (oval) oval[dst] = val[j];
    key2[i] = (grp * (uint32_t)g.n_prod + s) * (uint32_t)g.RS + (uint32_t)lrow_tmp[j];
    cdst[i] = (uint32_t)((int64_t)piece_cpos[key] + in_piece);
  }
}

// the largest consumer unit (valid entries of one (group, s)): what one round of the consumer's registers must hold
__global__ void pb_max_unit_kernel(PbGeom g, const int32_t *bound2, int64_t n_cons_units, int32_t *max_unit) {
  int m = 0;
  for (int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u < n_cons_units; u += (int64_t)gridDim.x * blockDim.x)
    m = max(m, bound2[(u + 1) * g.RS] - bound2[u * g.RS]);
  if (m) atomicMax(max_unit, m);
}

__global__ void pb_fill16_kernel(uint16_t *p, int64_t n, uint16_t v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

// slot of every entry, written at its consumer-order position: the BYTE offset of word (k / 4) * 4 RS + 4 * local row + k % 4
// buffer, k = rank of the entry among its row's entries of this pass (sorted order = (producer, CSR order): fixed)
__global__ void pb_slots_kernel(PbGeom g, const uint32_t *key2_sorted, const uint32_t *cdst_sorted, const int32_t *bound2,
                                int64_t n_valid, uint16_t *slot) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_valid; j += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t key2 = key2_sorted[j];
    const uint32_t lrow = key2 % (uint32_t)g.RS;
    const uint32_t k = (uint32_t)(j - (int64_t)bound2[key2]);  // < PB_LMAX: longer rows were left out
    slot[cdst_sorted[j]] = (uint16_t)(4u * ((k >> 2) * 4u * (uint32_t)g.RS + 4u * lrow + (k & 3u)));
  }
}

void shb_pb_free(shb_pb *pb) {
  if (!pb) return;
  cudaFree(pb->lcol);
  cudaFree(pb->val);
  cudaFree(pb->chunk_desc);
  cudaFree(pb->delta);
  cudaFree(pb->uchunk);
  cudaFree(pb->slot);
  cudaFree(pb->cons_start);
  cudaFree(pb->cdesc);
  cudaFree(pb->long_bits);
  cudaFree(pb->long_rows);
  cudaFree(pb->ring);
  cudaFree(pb->partial);
  cudaFree(pb->sync);
  delete pb;
}

static bool n_rows_too_many(const PbGeom &g) { return g.RS > 16383; }

static size_t pb_smem_bytes(const PbGeom &g) {
  return (size_t)g.W * 4 + ((size_t)g.cbuf_cap + 32) * 4 + PB_MISC_BYTES + (size_t)PB_PROD_WARPS * PB_STAGES * PB_STAGE_BYTES;
}

// chooses lmax (8 words per row unless that sends more than an eighth of the rows to the long-row kernel; then 16), marks
// the long rows and lists them; needs g.n_prod / g.n_pass / g.W only
static int pb_choose_lmax(shb_ctx *c, const shb_csr *A, PbGeom &g, shb_pb *pb) {
  const size_t words = std::max<size_t>((size_t)(A->n_rows + 31) / 32, 1);
  int32_t *d_n_long = nullptr;
  SHB_CUDA(cudaMalloc(&pb->long_bits, words * 4));
  SHB_CUDA(cudaMalloc(&pb->long_rows, (size_t)A->n_rows * 4));
  SHB_CUDA(cudaMalloc(&d_n_long, 4));
  const unsigned grid = (unsigned)std::min<int64_t>((A->n_rows + 255) / 256, (int64_t)c->prop.multiProcessorCount * 32);
  int32_t h_n_long = 0;
  cudaError_t e1 = cudaSuccess;
  for (int lmax = 8; lmax <= PB_LMAX_MAX; lmax *= 2) {
    g.lmax = lmax;
    if (e1 == cudaSuccess) e1 = cudaMemsetAsync(pb->long_bits, 0, words * 4, c->stream);
    if (e1 == cudaSuccess) e1 = cudaMemsetAsync(d_n_long, 0, 4, c->stream);
    pb_mark_long_kernel<<<grid, 256, 0, c->stream>>>(g, lmax, A->row_ptr, A->col, A->n_rows, (uint32_t)A->n_cols, pb->long_bits,
                                                      pb->long_rows, d_n_long);
    if (e1 == cudaSuccess) e1 = cudaGetLastError();
    if (e1 == cudaSuccess) e1 = cudaMemcpyAsync(&h_n_long, d_n_long, 4, cudaMemcpyDeviceToHost, c->stream);
    if (e1 == cudaSuccess) e1 = cudaStreamSynchronize(c->stream);
    if (e1 != cudaSuccess || (int64_t)h_n_long * 8 <= A->n_rows) break;
  }
  cudaFree(d_n_long);
  SHB_CUDA(e1);
  pb->n_long = h_n_long;
  return SHB_OK;
}

SHB_API int shb_csr_build_blocked(shb_ctx *c, shb_csr *A, int rows_per_consumer, int make_default) {
  SHB_REQUIRE(c && A, "bad arguments");
  SHB_REQUIRE(A->nnz > 0 && A->n_rows > 0 && A->n_cols > 0, "empty operator");
  SHB_CUDA(cudaSetDevice(c->device));
  if (A->pb) {
    A->device_bytes -= A->pb->device_bytes;
    shb_pb_free(A->pb);
    A->pb = nullptr;
    if (A->auto_variant == SHB_VARIANT_BLOCKED) SHB_TRY(shb_csr_analyse(c, A));
  }
  PbGeom g;
  memset(&g, 0, sizeof g);
  g.n_prod = c->prop.multiProcessorCount;
  int max_w = PB_MAX_W;
  if (const char *e = getenv("SHB_PB_MAXW")) max_w = std::max(64, std::min(PB_MAX_W, atoi(e)));
  g.n_pass = (int)((A->n_cols + (int64_t)g.n_prod * max_w - 1) / ((int64_t)g.n_prod * max_w));
  g.W = (int)((A->n_cols + (int64_t)g.n_prod * g.n_pass - 1) / ((int64_t)g.n_prod * g.n_pass));
  g.W = std::max(4, (g.W + 3) & ~3);
  SHB_REQUIRE(g.n_pass <= PB_MAX_PASS, "too many column passes for the blocked layout");
  shb_pb *pb = new shb_pb();
  {
    const int st = pb_choose_lmax(c, A, g, pb);
    if (st != SHB_OK) {
      shb_pb_free(pb);
      return st;
    }
  }
  // rows per consumer: lmax words per row of the shared memory the x slice and the producers' ring leave, and a unit
  // (RS rows x mean row length / n_pass entries) that one round of the consumer's registers holds (checked below)
  const double per_row = std::max(0.25, (double)A->nnz / (double)A->n_rows / g.n_pass);
  const double smem_rows = ((double)PB_SMEM_LIMIT - 1024.0 - (double)PB_PROD_WARPS * PB_STAGES * PB_STAGE_BYTES - (double)g.W * 4) / (4.0 * g.lmax);
  const double reg_rows = (double)(PB_CU * PB_CONS_THREADS * 4) * 0.94 / (per_row * 1.03 + 0.05);
  const double slot_rows = 16383.0 / g.lmax;  // 16-bit byte offsets
  int rs = rows_per_consumer > 0 ? rows_per_consumer
                                 : (int)std::min(8000.0, std::max(32.0, std::min(std::min(smem_rows, slot_rows), reg_rows) - 8.0));
  g.n_panel = (int)((A->n_rows + (int64_t)g.n_prod * rs - 1) / ((int64_t)g.n_prod * rs));
  g.RS = (int)((A->n_rows + (int64_t)g.n_prod * g.n_panel - 1) / ((int64_t)g.n_prod * g.n_panel));
  // (16-bit slots are byte offsets into the row buffer: at most 16383 words, the dump word included)
  if ((int64_t)g.RS * g.lmax > 16383 || n_rows_too_many(g)) {
    shb_pb_free(pb);
    shb_set_error("invalid argument: rows_per_consumer too large (%d rows x %d words)", g.RS, g.lmax);
    return SHB_ERR_INVALID;
  }
  g.cbuf_cap = g.RS * g.lmax;
  const int64_t n_groups = (int64_t)g.n_pass * g.n_panel;
  const int64_t n_units = n_groups * g.n_prod;       // (pass, p, panel) on the producer side, (group, s) on the consumer side
  const int64_t n_buckets = n_units * g.n_prod;      // pieces
  const int64_t n_keys2 = n_units * g.RS;            // (group, s, local row)
  if (!(n_buckets < (int64_t)0x7fffffff && n_keys2 < (int64_t)0x7fffffff)) {
    shb_pb_free(pb);
    shb_set_error("operator too large for the blocked layout");
    return SHB_ERR_TOO_LARGE;
  }
  const int64_t nnz = A->nnz;

  uint32_t *keys = nullptr, *keys_s = nullptr, *src = nullptr, *perm = nullptr;
  uint16_t *lrow_tmp = nullptr;
  int32_t *piece_start = nullptr, *piece_ppos = nullptr, *piece_cpos = nullptr, *bound2 = nullptr, *d_max = nullptr;
  void *cub_tmp = nullptr;
  size_t cub_bytes = 0;
  int status = SHB_OK;
  std::vector<int32_t> h_piece, h_ppos, h_cpos, h_uchunk, h_cons, h_delta;
  std::vector<int4> h_cdesc;   // one per consumer unit
  std::vector<uint4> h_desc;   // one per chunk while the flags are collected
  std::vector<uint32_t> h_desc48;  // 12 words per chunk: flags, first ordinal, piece count, 8 inline deltas
#define PB_CUDA(call)                                                                                   \
  do {                                                                                                  \
    cudaError_t _e = (call);                                                                            \
    if (_e != cudaSuccess) {                                                                            \
      shb_set_error("%s failed: %s (code: %d) in %s at line %d", #call, cudaGetErrorString(_e), (int)_e, \
                    __FILE__, __LINE__);                                                                \
      status = _e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;                            \
      goto done;                                                                                        \
    }                                                                                                   \
  } while (0)
  {
    PB_CUDA(cudaMalloc(&keys, (size_t)nnz * 4));
    PB_CUDA(cudaMalloc(&keys_s, (size_t)nnz * 4));
    PB_CUDA(cudaMalloc(&src, (size_t)nnz * 4));
    PB_CUDA(cudaMalloc(&perm, (size_t)nnz * 4));
    PB_CUDA(cudaMalloc(&lrow_tmp, (size_t)nnz * 2));
    PB_CUDA(cudaMalloc(&piece_start, (size_t)(n_buckets + 1) * 4));
    PB_CUDA(cudaMalloc(&piece_ppos, (size_t)n_buckets * 4));
    PB_CUDA(cudaMalloc(&piece_cpos, (size_t)n_buckets * 4));
    PB_CUDA(cudaMalloc(&bound2, (size_t)(n_keys2 + 1) * 4));
    PB_CUDA(cudaMalloc(&d_max, 4));
    PB_CUDA(cudaMemsetAsync(d_max, 0, 4, c->stream));
    const int cap_blocks = c->prop.multiProcessorCount * 32;
    const unsigned grid = (unsigned)std::min<int64_t>((A->n_rows + 255) / 256, cap_blocks);
    pb_keys_kernel<<<grid, 256, 0, c->stream>>>(g, A->row_ptr, A->col, A->n_rows, (uint32_t)A->n_cols, pb->long_bits, keys, src,
                                                lrow_tmp);
    PB_CUDA(cudaGetLastError());
    // all 32 key bits: the invalid key (all ones) must sort last
    PB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, keys, keys_s, src, perm, (int)nnz, 0, 32, c->stream));
    PB_CUDA(cudaMalloc(&cub_tmp, cub_bytes));
    PB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys_s, src, perm, (int)nnz, 0, 32, c->stream));
    const unsigned grid_n = (unsigned)std::min<int64_t>((nnz + 256) / 256, cap_blocks);
    pb_bounds_kernel<<<grid_n, 256, 0, c->stream>>>(keys_s, nnz, (uint32_t)n_buckets, piece_start);
    PB_CUDA(cudaGetLastError());
    h_piece.resize((size_t)n_buckets + 1);
    PB_CUDA(cudaMemcpyAsync(h_piece.data(), piece_start, (size_t)(n_buckets + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(cudaStreamSynchronize(c->stream));
    pb->n_valid = h_piece[(size_t)n_buckets];

    // ---- consumer order: (group, s, p); piece key = ((pass * n_prod + p) * n_panel + panel) * n_prod + s
    const int P = g.n_prod;
    auto key_of = [&](int64_t pass, int64_t p, int64_t panel, int64_t s) { return ((pass * P + p) * g.n_panel + panel) * P + s; };
    h_cpos.resize((size_t)n_buckets);
    h_cons.resize((size_t)n_units + 1);
    int64_t cpos = 0, max_group = 0;
    for (int64_t gp = 0; gp < n_groups; ++gp) {
      const int64_t pass = gp / g.n_panel, panel = gp % g.n_panel;
      const int64_t gstart = cpos;
      for (int s = 0; s < P; ++s) {
        h_cons[(size_t)(gp * P + s)] = (int32_t)cpos;
        for (int p = 0; p < P; ++p) {
          const int64_t key = key_of(pass, p, panel, s);
          h_cpos[(size_t)key] = (int32_t)cpos;
          cpos += ((int64_t)h_piece[(size_t)key + 1] - h_piece[(size_t)key] + 3) & ~(int64_t)3;
        }
      }
      max_group = std::max(max_group, cpos - gstart);
      if (cpos >= (int64_t)0x7fffffff - 4096) {
        shb_set_error("operator too large for the blocked layout (%lld padded entries)", (long long)cpos);
        status = SHB_ERR_TOO_LARGE;
        goto done;
      }
    }
    h_cons[(size_t)n_units] = (int32_t)cpos;
    pb->ccap = std::max<int64_t>(cpos, 4);
    g.slot_cap = (int)((max_group + 7) & ~(int64_t)7);

    // ---- producer order: (pass, p, panel, s); pieces to 4 entries, units to a chunk; piece ordinals, flags, deltas
    h_ppos.resize((size_t)n_buckets);
    h_uchunk.resize((size_t)n_units + 1);
    h_delta.clear();
    int64_t pos = 0;
    for (int64_t u = 0; u < n_units; ++u) {  // u = (pass * P + p) * n_panel + panel
      h_uchunk[(size_t)u] = (int32_t)(pos / PB_CHUNK);
      for (int s = 0; s < P; ++s) {
        const int64_t key = u * P + s;
        h_ppos[(size_t)key] = (int32_t)pos;
        pos += ((int64_t)h_piece[(size_t)key + 1] - h_piece[(size_t)key] + 3) & ~(int64_t)3;
      }
      pos = (pos + PB_CHUNK - 1) / PB_CHUNK * PB_CHUNK;
      if (pos >= (int64_t)0x7fffffff - 4096) {
        shb_set_error("operator too large for the blocked layout (%lld padded entries)", (long long)pos);
        status = SHB_ERR_TOO_LARGE;
        goto done;
      }
    }
    h_uchunk[(size_t)n_units] = (int32_t)(pos / PB_CHUNK);
    pb->cap = std::max<int64_t>(pos, PB_CHUNK);
    {
      const int64_t n_chunks = pb->cap / PB_CHUNK;
      h_desc.assign((size_t)n_chunks, make_uint4(0u, 0u, 0u, 0u));
      auto start_piece = [&](int64_t at, int32_t delta) {  // a piece (or a padding run) starts at stream position `at`
        const int64_t ch = at / PB_CHUNK;
        const int q = (int)((at % PB_CHUNK) / 4);
        if (q < 32) h_desc[(size_t)ch].x |= 1u << q;
        else h_desc[(size_t)ch].y |= 1u << (q - 32);
        h_delta.push_back(delta);
      };
      for (int64_t u = 0; u < n_units; ++u) {
        const int64_t sl = u / g.n_panel, panel = u % g.n_panel, pass = sl / P;
        const int64_t gp = pass * g.n_panel + panel;
        const int64_t gstart = h_cons[(size_t)(gp * P)];
        int64_t end = (int64_t)h_uchunk[(size_t)u] * PB_CHUNK;
        for (int s = 0; s < P; ++s) {
          const int64_t key = u * P + s;
          const int64_t len = ((int64_t)h_piece[(size_t)key + 1] - h_piece[(size_t)key] + 3) & ~(int64_t)3;
          if (len == 0) continue;
          start_piece(h_ppos[(size_t)key], (int32_t)(((int64_t)h_cpos[(size_t)key] - gstart) - (int64_t)h_ppos[(size_t)key]));
          end = (int64_t)h_ppos[(size_t)key] + len;
        }
        const int64_t uend = (int64_t)h_uchunk[(size_t)u + 1] * PB_CHUNK;
        if (end < uend) start_piece(end, (int32_t)((int64_t)g.slot_cap + (end % PB_CHUNK) - end));  // padding run -> trash words
      }
      // pieces started before each chunk
      uint32_t run = 0;
      for (int64_t ch = 0; ch < n_chunks; ++ch) {
        h_desc[(size_t)ch].z = run;
        run += (uint32_t)__builtin_popcount(h_desc[(size_t)ch].x) + (uint32_t)__builtin_popcount(h_desc[(size_t)ch].y);
      }
      pb->n_pieces = (int64_t)h_delta.size();
      if (h_delta.empty()) h_delta.push_back(0);
      // the descriptor a producer warp reads with the chunk: flags, ordinal of the first piece STARTED in or after the
      // chunk, number of pieces touching it, and the deltas of the first PB_INLINE_DELTAS of those (the first one may
      // have started in an earlier chunk: flag bit 0 clear)
      h_desc48.assign((size_t)n_chunks * 12, 0u);
      for (int64_t ch = 0; ch < n_chunks; ++ch) {
        const uint4 d = h_desc[(size_t)ch];
        uint32_t *o = &h_desc48[(size_t)ch * 12];
        const int starts = __builtin_popcount(d.x) + __builtin_popcount(d.y);
        const int carry = (d.x & 1u) ? 0 : 1;
        const int64_t base = (int64_t)d.z - carry;  // ordinal of touching piece 0
        const int touching = starts + carry;
        o[0] = d.x; o[1] = d.y; o[2] = d.z; o[3] = (uint32_t)touching;
        for (int j = 0; j < PB_INLINE_DELTAS && j < touching; ++j) {
          const int64_t ord = base + j;
          o[4 + j] = (ord >= 0 && ord < (int64_t)h_delta.size()) ? (uint32_t)h_delta[(size_t)ord] : 0u;
        }
      }
    }

    PB_CUDA(cudaMalloc(&pb->lcol, (size_t)pb->cap * 2));
    PB_CUDA(cudaMalloc(&pb->slot, (size_t)pb->ccap * 2));
    if (A->val) PB_CUDA(cudaMalloc(&pb->val, (size_t)pb->cap * 4));
    PB_CUDA(cudaMalloc(&pb->chunk_desc, h_desc48.size() * 4));
    PB_CUDA(cudaMalloc(&pb->delta, h_delta.size() * 4));
    PB_CUDA(cudaMalloc(&pb->uchunk, (size_t)(n_units + 1) * 4));
    PB_CUDA(cudaMalloc(&pb->cons_start, (size_t)(n_units + 1) * 4));
    PB_CUDA(cudaMalloc(&pb->cdesc, (size_t)n_units * sizeof(int4)));
    PB_CUDA(cudaMalloc(&pb->partial, (size_t)A->n_rows * 4));
    PB_CUDA(cudaMalloc(&pb->sync, (size_t)(2 * n_groups + 1) * 4));
    PB_CUDA(cudaMemsetAsync(pb->sync, 0, (size_t)(2 * n_groups + 1) * 4, c->stream));
    PB_CUDA(cudaMemsetAsync(pb->lcol, 0, (size_t)pb->cap * 2, c->stream));
    if (pb->val) PB_CUDA(cudaMemsetAsync(pb->val, 0, (size_t)pb->cap * 4, c->stream));
    PB_CUDA(cudaMemcpyAsync(piece_ppos, h_ppos.data(), (size_t)n_buckets * 4, cudaMemcpyHostToDevice, c->stream));
    PB_CUDA(cudaMemcpyAsync(piece_cpos, h_cpos.data(), (size_t)n_buckets * 4, cudaMemcpyHostToDevice, c->stream));
    PB_CUDA(cudaMemcpyAsync(pb->chunk_desc, h_desc48.data(), h_desc48.size() * 4, cudaMemcpyHostToDevice, c->stream));
    PB_CUDA(cudaMemcpyAsync(pb->delta, h_delta.data(), h_delta.size() * 4, cudaMemcpyHostToDevice, c->stream));
    PB_CUDA(cudaMemcpyAsync(pb->uchunk, h_uchunk.data(), (size_t)(n_units + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    PB_CUDA(cudaMemcpyAsync(pb->cons_start, h_cons.data(), (size_t)(n_units + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    h_cdesc.resize((size_t)n_units);
    for (int64_t gp = 0; gp < n_groups; ++gp)
      for (int s = 0; s < P; ++s) {
        const int64_t u = gp * P + s;
        h_cdesc[(size_t)u] = make_int4(h_cons[(size_t)u], (h_cons[(size_t)u + 1] - h_cons[(size_t)u]) >> 2,
                                       h_cons[(size_t)u] - h_cons[(size_t)(gp * P)], 0);
      }
    PB_CUDA(cudaMemcpyAsync(pb->cdesc, h_cdesc.data(), (size_t)n_units * sizeof(int4), cudaMemcpyHostToDevice, c->stream));
    const unsigned grid_v = (unsigned)std::min<int64_t>((pb->n_valid + 255) / 256 + 1, cap_blocks);
    if (pb->n_valid > 0) {
      // keys / src are free again: they receive the consumer-order keys and positions
      pb_scatter_kernel<<<grid_v, 256, 0, c->stream>>>(g, keys_s, perm, pb->n_valid, piece_start, piece_ppos, piece_cpos, A->col,
                                                       (const uint32_t *)A->val, lrow_tmp, pb->lcol, (uint32_t *)pb->val, keys, src);
      PB_CUDA(cudaGetLastError());
      PB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys_s, src, perm, (int)pb->n_valid, 0, 32, c->stream));
      pb_bounds_kernel<<<grid_v, 256, 0, c->stream>>>(keys_s, pb->n_valid, (uint32_t)n_keys2, bound2);
      PB_CUDA(cudaGetLastError());
    } else {
      PB_CUDA(cudaMemsetAsync(bound2, 0, (size_t)(n_keys2 + 1) * 4, c->stream));
    }
    pb_max_unit_kernel<<<(unsigned)std::min<int64_t>((n_units + 255) / 256, cap_blocks), 256, 0, c->stream>>>(g, bound2, n_units, d_max);
    PB_CUDA(cudaGetLastError());
    PB_CUDA(cudaMemcpyAsync(&pb->max_unit, d_max, 4, cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(cudaStreamSynchronize(c->stream));  // the host staging vectors go out of scope
    g.depth = (int)std::min<int64_t>(n_groups, 4);
    if (const char *e = getenv("SHB_PB_DEPTH")) g.depth = (int)std::min<int64_t>(n_groups, std::max(1, std::min(PB_MAX_DEPTH, atoi(e))));
    // (a unit larger than one round of consumer registers still runs, in extra rounds)
    if (pb_smem_bytes(g) > PB_SMEM_LIMIT) {
      shb_set_error("blocked layout does not fit shared memory: %zu bytes needed (rows_per_consumer %d, %d words per row)",
                    pb_smem_bytes(g), g.RS, g.lmax);
      status = SHB_ERR_UNSUPPORTED;
      goto done;
    }
    // padding entries of the consumer stream drop their (meaningless) product into the dump word
    pb_fill16_kernel<<<(unsigned)std::min<int64_t>((pb->ccap + 255) / 256, cap_blocks), 256, 0, c->stream>>>(pb->slot, pb->ccap,
                                                                                                       (uint16_t)(4 * g.cbuf_cap));
    PB_CUDA(cudaGetLastError());
    if (pb->n_valid > 0) {
      pb_slots_kernel<<<grid_v, 256, 0, c->stream>>>(g, keys_s, perm, bound2, pb->n_valid, pb->slot);
      PB_CUDA(cudaGetLastError());
    }
    PB_CUDA(cudaStreamSynchronize(c->stream));
    PB_CUDA(cudaMalloc(&pb->ring, (size_t)g.depth * ((size_t)g.slot_cap + PB_CHUNK) * 4));
    pb->ring_depth_alloc = g.depth;
    pb->g = g;
    pb->device_bytes = (uint64_t)pb->cap * (2 + (A->val ? 4 : 0)) + (uint64_t)pb->ccap * 2 + h_desc48.size() * 4 + h_delta.size() * 4 +
                       (uint64_t)(n_units + 1) * 8 + (uint64_t)n_units * 16 + (uint64_t)A->n_rows * 4 + (uint64_t)(A->n_rows + 31) / 32 * 4 +
                       (uint64_t)g.depth * ((uint64_t)g.slot_cap + PB_CHUNK) * 4 + (uint64_t)A->n_rows * 4;
  }
done:
#undef PB_CUDA
  cudaFree(keys);
  cudaFree(keys_s);
  cudaFree(src);
  cudaFree(perm);
  cudaFree(lrow_tmp);
  cudaFree(piece_start);
  cudaFree(piece_ppos);
  cudaFree(piece_cpos);
  cudaFree(bound2);
  cudaFree(d_max);
  cudaFree(cub_tmp);
  if (status != SHB_OK) {
    shb_pb_free(pb);
    return status;
  }
  A->pb = pb;
  A->device_bytes += pb->device_bytes;
  if (make_default) A->auto_variant = SHB_VARIANT_BLOCKED;
  return SHB_OK;
}

SHB_API int shb_csr_blocked_info(const shb_csr *A, int32_t *geom, int64_t *entries, uint64_t *device_bytes) {
  SHB_REQUIRE(A, "bad arguments");
  SHB_REQUIRE(A->pb, "the blocked layout has not been built (shb_csr_build_blocked)");
  if (geom) {
    geom[0] = A->pb->g.n_prod;
    geom[1] = A->pb->g.n_pass;
    geom[2] = A->pb->g.n_panel;
    geom[3] = A->pb->g.W;
    geom[4] = A->pb->g.RS;
    geom[5] = A->pb->g.cbuf_cap;
    geom[6] = A->pb->g.slot_cap;
    geom[7] = A->pb->g.depth;
  }
  if (entries) *entries = A->pb->cap;
  if (device_bytes) *device_bytes = A->pb->device_bytes;
  return SHB_OK;
}

// the launch (called from run_spmv in shb_core.cu through shb_launch_blocked<S>)
size_t shb_pb_smem_bytes(const shb_pb *pb) { return pb_smem_bytes(pb->g); }

// two-launch debug mode needs a ring that holds every (pass, panel) group at once
int shb_pb_reserve_ring(shb_ctx *c, shb_pb *pb, int depth) {
  if (pb->ring_depth_alloc >= depth) return SHB_OK;
  SHB_CUDA(cudaStreamSynchronize(c->stream));
  cudaFree(pb->ring);
  pb->ring = nullptr;
  pb->ring_depth_alloc = 0;
  SHB_CUDA(cudaMalloc(&pb->ring, (size_t)depth * ((size_t)pb->g.slot_cap + PB_CHUNK) * 4));
  pb->ring_depth_alloc = depth;
  return SHB_OK;
}
