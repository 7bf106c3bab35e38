// shb_coo.cu -- COO tuples -> CSR on the device (the encode half of the ingest).
//
// The reference turns the Matrix Market tuples into ragged rows on the host (calculate_ellpack,
// src/sparse_matrix.cpp:72-119) and then into the kernel's byte layout (cl_encode, :122-399).  Its semantics, kept
// here bit for bit:
//   Q1  a tuple (I, J, v) is stored in ROW J, column I;
//   Q2  values pass through `int` on the way into the rows (1474.779 -> 1474; optional here: canonical operators keep them);
//   Q3  rows are NOT sorted: the entries of a row keep the order of the tuple list (file order, each symmetric mirror
//       right after its original), duplicates are kept.
// So the encode is a STABLE sort of the tuples by J: cub::DeviceRadixSort on (J, position) over exactly the key bits
// the row count needs, a gather of (I, v) through the permutation, and the row offsets from the sorted keys.
// The host keeps the parse (host/inc/mm_reader.h); this replaces its counting sort (host/inc/sparse_matrix.h
// csr_encode) and the separate upload of the CSR arrays.
#include <algorithm>
#include <cub/device/device_radix_sort.cuh>

#include "shb_internal.cuh"

__global__ void coo_iota_kernel(uint32_t *p, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = (uint32_t)i;
}

// row_ptr[r] = first sorted position whose row is >= r
__global__ void coo_row_ptr_kernel(const uint32_t *sorted_rows, int64_t n, int64_t n_rows, int32_t *row_ptr) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t prev = i > 0 ? (int64_t)min((int64_t)sorted_rows[i - 1], n_rows) : -1;
    const int64_t cur = i < n ? (int64_t)min((int64_t)sorted_rows[i], n_rows) : n_rows;
    for (int64_t r = prev + 1; r <= cur; ++r) row_ptr[r] = (int32_t)i;
  }
}

// elem_kind 0: float, 1: int32; truncate: float values pass through int (Q2); ints are unchanged
__global__ void coo_gather_kernel(const uint32_t *perm, int64_t n, const int32_t *ci, const uint32_t *v, int elem_kind, int truncate,
                                  int32_t *col, uint32_t *val) {
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t j = perm[k];
    SHB_DEV_CHECK((int64_t)j < n);
    col[k] = ci[j];
    if (val) {
      uint32_t b = v[j];
      if (elem_kind == 0 && truncate) {
        // the host's `(int)value` (Q2) is cvttss2si: NaN and everything outside int range become INT_MIN ("integer
        // indefinite"), where the device conversion would saturate -- x / 0 column sums of pagerank_normalise get here
        const float f = __uint_as_float(b);
        const int t = (f >= -2147483648.0f && f < 2147483648.0f) ? __float2int_rz(f) : (int)0x80000000;
        b = __float_as_uint((float)t);
      }
      val[k] = b;
    }
  }
}

SHB_API int shb_csr_from_coo(shb_ctx *c, int64_t n_rows, int64_t n_cols, int64_t n_entries, const int32_t *col_of_entry,
                             const int32_t *row_of_entry, const void *vals, int elem_kind, int truncate_through_int, shb_csr **out,
                             uint64_t *ns) {
  SHB_REQUIRE(c && out && n_rows >= 0 && n_cols >= 0 && n_entries >= 0, "bad arguments");
  SHB_REQUIRE(n_entries == 0 || (col_of_entry && row_of_entry), "tuple arrays are null");
  SHB_REQUIRE(elem_kind == 0 || elem_kind == 1, "element kind: 0 float, 1 int32");
  SHB_CUDA(cudaSetDevice(c->device));
  shb_csr *A = nullptr;
  SHB_TRY(shb_csr_alloc(c, n_rows, n_cols, n_entries, vals != nullptr, &A));
  uint32_t *d_rows = nullptr, *d_rows_s = nullptr, *d_idx = nullptr, *d_perm = nullptr, *d_v = nullptr;
  int32_t *d_ci = nullptr;
  void *tmp = nullptr;
  size_t tmp_bytes = 0;
  int status = SHB_OK;
  const size_t n = (size_t)std::max<int64_t>(n_entries, 1);
  ScopedTimer t(c, ns);
#define CO_CUDA(call)                                                                                     \
  do {                                                                                                    \
    cudaError_t _e = (call);                                                                              \
    if (_e != cudaSuccess) {                                                                              \
      shb_set_error("%s failed: %s (code: %d) in %s at line %d", #call, cudaGetErrorString(_e), (int)_e, \
                    __FILE__, __LINE__);                                                                  \
      status = _e == cudaErrorMemoryAllocation ? SHB_ERR_OOM : SHB_ERR_CUDA;                              \
      goto done;                                                                                          \
    }                                                                                                     \
  } while (0)
  {
    CO_CUDA(cudaMalloc(&d_rows, n * 4));
    CO_CUDA(cudaMalloc(&d_rows_s, n * 4));
    CO_CUDA(cudaMalloc(&d_idx, n * 4));
    CO_CUDA(cudaMalloc(&d_perm, n * 4));
    CO_CUDA(cudaMalloc(&d_ci, n * 4));
    if (vals) CO_CUDA(cudaMalloc(&d_v, n * 4));
    if (n_entries > 0) {
      CO_CUDA(cudaMemcpyAsync(d_rows, row_of_entry, (size_t)n_entries * 4, cudaMemcpyHostToDevice, c->stream));
      CO_CUDA(cudaMemcpyAsync(d_ci, col_of_entry, (size_t)n_entries * 4, cudaMemcpyHostToDevice, c->stream));
      if (vals) CO_CUDA(cudaMemcpyAsync(d_v, vals, (size_t)n_entries * 4, cudaMemcpyHostToDevice, c->stream));
      const unsigned grid = grid_for(n_entries, 256, c->prop.multiProcessorCount * 16);
      coo_iota_kernel<<<grid, 256, 0, c->stream>>>(d_idx, n_entries);
      CO_CUDA(cudaGetLastError());
      int bits = 1;
      while (bits < 32 && (1ll << bits) < std::max<int64_t>(n_rows, 2)) ++bits;  // stable LSD radix sort on the row bits only
      CO_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_rows, d_rows_s, d_idx, d_perm, (int)n_entries, 0, bits, c->stream));
      CO_CUDA(cudaMalloc(&tmp, tmp_bytes));
      CO_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, d_rows, d_rows_s, d_idx, d_perm, (int)n_entries, 0, bits, c->stream));
      coo_row_ptr_kernel<<<grid_for(n_entries + 1, 256, c->prop.multiProcessorCount * 16), 256, 0, c->stream>>>(d_rows_s, n_entries, n_rows,
                                                                                                             A->row_ptr);
      CO_CUDA(cudaGetLastError());
      coo_gather_kernel<<<grid, 256, 0, c->stream>>>(d_perm, n_entries, d_ci, d_v, elem_kind, truncate_through_int, A->col, (uint32_t *)A->val);
      CO_CUDA(cudaGetLastError());
    } else {
      CO_CUDA(cudaMemsetAsync(A->row_ptr, 0, (size_t)(n_rows + 1) * 4, c->stream));
    }
    CO_CUDA(cudaStreamSynchronize(c->stream));
  }
done:
#undef CO_CUDA
  cudaFree(d_rows);
  cudaFree(d_rows_s);
  cudaFree(d_idx);
  cudaFree(d_perm);
  cudaFree(d_ci);
  cudaFree(d_v);
  cudaFree(tmp);
  if (status == SHB_OK) status = t.finish();
  if (status == SHB_OK) status = shb_csr_analyse(c, A);
  if (status != SHB_OK) {
    shb_csr_destroy(A);
    return status;
  }
  *out = A;
  return SHB_OK;
}
