// shb_pb.cu -- builds the BLOCKED (propagation-blocked) copy of an operator on the device (see pb.cuh).
//
// The reference re-encodes every matrix into the layout its kernel wants before the run
// (SparseMatrix::cl_encode / executorEncodeMatrix, src/sparse_matrix.cpp:122-399, inc/kernel_utils.h:35-168);
// this is the same step for the blocked schedule:
//   1. a stable sort of the stored entries by (pass, panel, producer, consumer): the order the producers
//      stream them in; CSR order -- (row, column) -- is kept inside each piece;
//   2. a second stable sort of those positions by (pass, panel, consumer, row): the order of each
//      consumer's row-sorted shared-memory buffer, which gives every entry its slot and every row its
//      offsets.
// cub::DeviceRadixSort does both sorts; they run once per operator, never on the SpMV path.
#include <algorithm>
#include <cub/device/device_radix_sort.cuh>
#include <vector>

#include "pb.cuh"

static const uint32_t PB_INVALID_KEY = 0xffffffffu;

// one thread per row: bucket key, source index and local row of every stored entry
__global__ void pb_keys_kernel(PbGeom g, const int32_t *row_ptr, const int32_t *col, int64_t n_rows, uint32_t n_cols,
                               uint32_t *keys, uint32_t *src, uint16_t *lrow_tmp) {
  const int64_t PR = (int64_t)g.RS * g.n_prod;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
    const int panel = (int)(r / PR);
    const int in_panel = (int)(r - (int64_t)panel * PR);
    const int s = in_panel / g.RS;
    const uint16_t lr = (uint16_t)(in_panel - s * g.RS);
    const int b = row_ptr[r], e = row_ptr[r + 1];
    for (int k = b; k < e; ++k) {
      const uint32_t c = (uint32_t)col[k];
      uint32_t key = PB_INVALID_KEY;  // out-of-range column: x reads as ZERO, the product is the additive identity
      if (c < n_cols) {
        const uint32_t slice = c / (uint32_t)g.W;  // global slice index = pass * n_prod + p
        const uint32_t pass = slice / (uint32_t)g.n_prod, p = slice - pass * (uint32_t)g.n_prod;
        key = ((pass * (uint32_t)g.n_panel + (uint32_t)panel) * (uint32_t)g.n_prod + p) * (uint32_t)g.n_prod + (uint32_t)s;
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

// producer-order scatter of (local column, value); emits the consumer-order key of every entry
__global__ void pb_scatter_kernel(PbGeom g, const uint32_t *sorted_keys, const uint32_t *perm, int64_t n_valid,
                                  const int32_t *piece_start, const int32_t *piece_pad, const int32_t *col,
                                  const uint32_t *val, const uint16_t *lrow_tmp, uint16_t *lcol, uint32_t *oval,
                                  uint32_t *key2, uint32_t *dst2) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_valid; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t key = sorted_keys[i];
    const uint32_t u = key / (uint32_t)g.n_prod, s = key - u * (uint32_t)g.n_prod;
    const uint32_t grp = u / (uint32_t)g.n_prod;  // (pass, panel)
    const int64_t dst = (int64_t)piece_pad[key] + (i - (int64_t)piece_start[key]);  // pieces start at multiples of 4
    const uint32_t j = perm[i];
    const uint32_t c = (uint32_t)col[j];
    lcol[dst] = (uint16_t)(c % (uint32_t)g.W);
    if (oval) oval[dst] = val[j];
    key2[i] = (grp * (uint32_t)g.n_prod + s) * (uint32_t)g.RS + (uint32_t)lrow_tmp[j];
    dst2[i] = (uint32_t)dst;
  }
}

// consumer order.  The row-sorted buffer of a consumer unit is laid out per group of 32 consecutive rows,
// k-major inside a group: first every row's entry 0 (rows in order, empty rows skipped), then every
// row's entry 1, ...  A warp reducing the group reads consecutive words at every step (no bank
// conflicts) and needs only the row lengths and the group's base.  One warp per (unit, group):
// writes the slot of every entry, the row lengths and the group bases (crp layout: [RS] lengths,
// then [ceil(RS/32)] bases, 16-bit each).
__global__ void pb_slots_kernel(PbGeom g, const uint32_t *dst2_sorted, const int32_t *bound2, int64_t n_cons_units,
                                uint16_t *slot, uint16_t *crp, int32_t *max_unit) {
  const int lane = threadIdx.x & 31;
  const int ngr = (g.RS + 31) / 32;
  const int stride = pb_crp_stride(g.RS);
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_cons_units * ngr; w += n_warps) {
    const int64_t gs = w / ngr;
    const int m = (int)(w - gs * ngr);
    const int i = m * 32 + lane;
    const int64_t kb = gs * g.RS;
    const int ustart = bound2[kb];
    const int rb = i < g.RS ? bound2[kb + i] : 0, re = i < g.RS ? bound2[kb + i + 1] : 0;
    const int len = re - rb;
    const int gbase = __shfl_sync(0xffffffffu, rb, 0) - ustart;
    if (i < g.RS) crp[gs * stride + i] = (uint16_t)min(len, 65535);
    if (lane == 0) {
      crp[gs * stride + g.RS + m] = (uint16_t)min(gbase, 65535);
      if (m == 0) atomicMax(max_unit, bound2[kb + g.RS] - ustart);
    }
    const int maxlen = __reduce_max_sync(0xffffffffu, len);
    int off = gbase;
    for (int k = 0; k < maxlen; ++k) {
      const uint32_t act = __ballot_sync(0xffffffffu, k < len);
      if (k < len) slot[dst2_sorted[rb + k]] = (uint16_t)min(off + __popc(act & ((1u << lane) - 1u)), 65535);
      off += __popc(act);
    }
  }
}

void shb_pb_free(shb_pb *pb) {
  if (!pb) return;
  cudaFree(pb->lcol);
  cudaFree(pb->val);
  cudaFree(pb->slot);
  cudaFree(pb->unit_ptr);
  cudaFree(pb->cons_desc);
  cudaFree(pb->crp);
  cudaFree(pb->ring);
  cudaFree(pb->partial);
  cudaFree(pb->sync);
  delete pb;
}

static size_t pb_smem_bytes(const PbGeom &g) {
  return (size_t)g.W * 4 + (size_t)g.cbuf_cap * 4 + (size_t)pb_crp_stride(g.RS) * 2 * 2;
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
  SHB_REQUIRE(g.n_prod <= 32 * (PB_CONS_THREADS / 32), "more SMs than one descriptor chunk covers");
  int max_w = PB_MAX_W;
  if (const char *e = getenv("SHB_PB_MAXW")) max_w = std::max(64, std::min(PB_MAX_W, atoi(e)));
  g.n_pass = (int)((A->n_cols + (int64_t)g.n_prod * max_w - 1) / ((int64_t)g.n_prod * max_w));
  g.W = (int)((A->n_cols + (int64_t)g.n_prod * g.n_pass - 1) / ((int64_t)g.n_prod * g.n_pass));
  g.W = std::max(4, (g.W + 3) & ~3);
  // rows per consumer: the unit (RS rows x mean row length / n_pass entries) should fill about 44 KB of shared memory
  const double per_row = std::max(1.0, (double)A->nnz / (double)A->n_rows / g.n_pass);
  // automatic: a consumer unit (RS rows x per_row entries, 4 bytes each) takes the shared memory the x slice leaves
  const double free_entries = std::min(30000.0, ((double)PB_SMEM_LIMIT - 6144.0 - (double)g.W * 4) / 4.0);
  int rs = rows_per_consumer > 0 ? rows_per_consumer : (int)std::min(16384.0, std::max(32.0, free_entries / (per_row * 1.06 + 1.1)));
  g.n_panel = (int)((A->n_rows + (int64_t)g.n_prod * rs - 1) / ((int64_t)g.n_prod * rs));
  g.RS = (int)((A->n_rows + (int64_t)g.n_prod * g.n_panel - 1) / ((int64_t)g.n_prod * g.n_panel));
  SHB_REQUIRE(g.RS <= 32767, "rows_per_consumer too large");
  const int64_t n_groups = (int64_t)g.n_pass * g.n_panel;
  const int64_t n_units = n_groups * g.n_prod;
  const int64_t n_buckets = n_units * g.n_prod;
  const int64_t n_keys2 = n_units * g.RS;  // (pass, panel, s, local row)
  SHB_REQUIRE(n_buckets < (int64_t)0x7fffffff && n_keys2 < (int64_t)0x7fffffff, "operator too large for the blocked layout");
  const int64_t nnz = A->nnz;

  uint32_t *keys = nullptr, *keys_s = nullptr, *src = nullptr, *perm = nullptr;
  uint16_t *lrow_tmp = nullptr;
  int32_t *piece_start = nullptr, *piece_pad = nullptr, *bound2 = nullptr, *d_max = nullptr;
  void *cub_tmp = nullptr;
  size_t cub_bytes = 0, cub_bytes2 = 0;
  shb_pb *pb = new shb_pb();
  int status = SHB_OK;
  std::vector<int32_t> h_piece, h_unit, h_ppad;
  std::vector<int2> h_desc;
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
    PB_CUDA(cudaMalloc(&piece_pad, (size_t)n_buckets * 4));
    PB_CUDA(cudaMalloc(&bound2, (size_t)(n_keys2 + 1) * 4));
    PB_CUDA(cudaMalloc(&d_max, 4));
    PB_CUDA(cudaMemsetAsync(d_max, 0, 4, c->stream));
    const unsigned grid = (unsigned)std::min<int64_t>((A->n_rows + 255) / 256, 148 * 32);
    pb_keys_kernel<<<grid, 256, 0, c->stream>>>(g, A->row_ptr, A->col, A->n_rows, (uint32_t)A->n_cols, keys, src, lrow_tmp);
    PB_CUDA(cudaGetLastError());
    // all 32 key bits: the invalid key (all ones) must sort last
    PB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, keys, keys_s, src, perm, (int)nnz, 0, 32, c->stream));
    PB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes2, keys, keys_s, src, perm, (int)nnz, 0, 32, c->stream));
    cub_bytes = std::max(cub_bytes, cub_bytes2);
    PB_CUDA(cudaMalloc(&cub_tmp, cub_bytes));
    PB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys_s, src, perm, (int)nnz, 0, 32, c->stream));
    const unsigned grid_n = (unsigned)std::min<int64_t>((nnz + 256) / 256, 148 * 32);
    pb_bounds_kernel<<<grid_n, 256, 0, c->stream>>>(keys_s, nnz, (uint32_t)n_buckets, piece_start);
    PB_CUDA(cudaGetLastError());
    h_piece.resize((size_t)n_buckets + 1);
    PB_CUDA(cudaMemcpyAsync(h_piece.data(), piece_start, (size_t)(n_buckets + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(cudaStreamSynchronize(c->stream));
    // pieces start at multiples of 4 entries (the consumer's 128-bit ring loads), units at multiples of 8 (the
    // producer's 16-byte groups of local columns); piece descriptors are stored consumer-major
    h_unit.resize((size_t)n_units + 1);
    h_ppad.resize((size_t)n_buckets);
    int64_t pos = 0, max_group = 0;
    for (int64_t u = 0; u < n_units; ++u) {
      if (u % g.n_prod == 0 && u > 0) max_group = std::max(max_group, pos - (int64_t)h_unit[(size_t)(u - g.n_prod)]);
      h_unit[(size_t)u] = (int32_t)pos;
      int64_t in_unit = 0;
      for (int s = 0; s < g.n_prod; ++s) {
        const int64_t key = u * g.n_prod + s;
        h_ppad[(size_t)key] = (int32_t)(pos + in_unit);
        in_unit += ((int64_t)h_piece[(size_t)key + 1] - h_piece[(size_t)key] + 3) & ~(int64_t)3;
      }
      pos += (in_unit + 7) & ~(int64_t)7;
      if (pos >= (int64_t)0x7fffffff - 64) {
        shb_set_error("operator too large for the blocked layout (%lld padded entries)", (long long)pos);
        status = SHB_ERR_TOO_LARGE;
        goto done;
      }
    }
    max_group = std::max(max_group, pos - (int64_t)h_unit[(size_t)(n_units - g.n_prod)]);
    h_unit[(size_t)n_units] = (int32_t)pos;
    pb->cap = std::max<int64_t>(pos, 8);
    pb->n_valid = h_piece[(size_t)n_buckets];
    h_desc.resize((size_t)n_buckets);
    for (int64_t gp = 0; gp < n_groups; ++gp)
      for (int p = 0; p < g.n_prod; ++p) {
        const int64_t u = gp * g.n_prod + p;
        for (int s = 0; s < g.n_prod; ++s) {
          const int64_t key = u * g.n_prod + s;
          int2 d;
          d.x = h_ppad[(size_t)key];
          d.y = (h_piece[(size_t)key + 1] - h_piece[(size_t)key] + 3) & ~3;
          h_desc[(size_t)((gp * g.n_prod + s) * g.n_prod + p)] = d;
        }
      }
    g.slot_cap = (int)((max_group + 7) & ~(int64_t)7);
    PB_CUDA(cudaMalloc(&pb->lcol, (size_t)pb->cap * 2));
    PB_CUDA(cudaMalloc(&pb->slot, (size_t)pb->cap * 2));
    if (A->val) PB_CUDA(cudaMalloc(&pb->val, (size_t)pb->cap * 4));
    PB_CUDA(cudaMalloc(&pb->unit_ptr, (size_t)(n_units + 1) * 4));
    PB_CUDA(cudaMalloc(&pb->cons_desc, (size_t)n_buckets * sizeof(int2)));
    PB_CUDA(cudaMalloc(&pb->crp, (size_t)n_units * pb_crp_stride(g.RS) * 2));
    PB_CUDA(cudaMemsetAsync(pb->crp, 0, (size_t)n_units * pb_crp_stride(g.RS) * 2, c->stream));
    PB_CUDA(cudaMalloc(&pb->partial, (size_t)A->n_rows * 4));
    PB_CUDA(cudaMalloc(&pb->sync, (size_t)(2 * n_groups + 1) * 4));
    PB_CUDA(cudaMemsetAsync(pb->sync, 0, (size_t)(2 * n_groups + 1) * 4, c->stream));
    PB_CUDA(cudaMemsetAsync(pb->lcol, 0, (size_t)pb->cap * 2, c->stream));
    PB_CUDA(cudaMemsetAsync(pb->slot, 0xff, (size_t)pb->cap * 2, c->stream));  // 0xffff = padding
    if (pb->val) PB_CUDA(cudaMemsetAsync(pb->val, 0, (size_t)pb->cap * 4, c->stream));
    PB_CUDA(cudaMemcpyAsync(piece_pad, h_ppad.data(), (size_t)n_buckets * 4, cudaMemcpyHostToDevice, c->stream));
    PB_CUDA(cudaMemcpyAsync(pb->unit_ptr, h_unit.data(), (size_t)(n_units + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    PB_CUDA(cudaMemcpyAsync(pb->cons_desc, h_desc.data(), (size_t)n_buckets * sizeof(int2), cudaMemcpyHostToDevice, c->stream));
    if (pb->n_valid > 0) {
      const unsigned grid_v = (unsigned)std::min<int64_t>((pb->n_valid + 255) / 256, 148 * 32);
      // keys / src are free again: they receive the consumer-order keys and the producer positions
      pb_scatter_kernel<<<grid_v, 256, 0, c->stream>>>(g, keys_s, perm, pb->n_valid, piece_start, piece_pad, A->col,
                                                       (const uint32_t *)A->val, lrow_tmp, pb->lcol, (uint32_t *)pb->val, keys, src);
      PB_CUDA(cudaGetLastError());
      PB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys_s, src, perm, (int)pb->n_valid, 0, 32, c->stream));
      pb_bounds_kernel<<<grid_v, 256, 0, c->stream>>>(keys_s, pb->n_valid, (uint32_t)n_keys2, bound2);
      PB_CUDA(cudaGetLastError());
    } else {
      PB_CUDA(cudaMemsetAsync(bound2, 0, (size_t)(n_keys2 + 1) * 4, c->stream));
    }
    pb_slots_kernel<<<(unsigned)std::min<int64_t>((n_units * ((g.RS + 31) / 32) + 7) / 8, 148 * 32), 256, 0, c->stream>>>(
        g, perm, bound2, n_units, pb->slot, pb->crp, d_max);
    PB_CUDA(cudaGetLastError());
    PB_CUDA(cudaMemcpyAsync(&pb->max_unit, d_max, 4, cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(cudaStreamSynchronize(c->stream));  // the host staging vectors go out of scope
    g.cbuf_cap = std::max(32, (pb->max_unit + 31) & ~31);
    g.depth = (int)std::min<int64_t>(n_groups, 4);
    if (const char *e = getenv("SHB_PB_DEPTH")) g.depth = (int)std::min<int64_t>(n_groups, std::max(1, atoi(e)));
    if (pb->max_unit > 65535 || pb_smem_bytes(g) > PB_SMEM_LIMIT) {
      shb_set_error("blocked layout does not fit shared memory: largest consumer unit %d entries, %zu bytes needed "
                    "(rows_per_consumer %d; the schedule suits operators with balanced rows and columns)",
                    pb->max_unit, pb_smem_bytes(g), g.RS);
      status = SHB_ERR_UNSUPPORTED;
      goto done;
    }
    PB_CUDA(cudaMalloc(&pb->ring, (size_t)g.depth * g.slot_cap * 4));
    pb->ring_depth_alloc = g.depth;
    pb->g = g;
    pb->device_bytes = (uint64_t)pb->cap * (2 + 2 + (A->val ? 4 : 0)) + (uint64_t)(n_units + 1) * 4 + (uint64_t)n_buckets * 8 +
                       (uint64_t)n_units * pb_crp_stride(g.RS) * 2 + (uint64_t)g.depth * g.slot_cap * 4 + (uint64_t)A->n_rows * 4;
  }
done:
#undef PB_CUDA
  cudaFree(keys);
  cudaFree(keys_s);
  cudaFree(src);
  cudaFree(perm);
  cudaFree(lrow_tmp);
  cudaFree(piece_start);
  cudaFree(piece_pad);
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
  SHB_CUDA(cudaMalloc(&pb->ring, (size_t)depth * pb->g.slot_cap * 4));
  pb->ring_depth_alloc = depth;
  return SHB_OK;
}
