// shb_noalloc.cu -- shb_spmv_stream_static: the stream-schedule SpMV for callers that cannot allocate.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc
// around a launch: code under CUDA graph capture, tight serving loops, harnesses that forbid it.  The stream schedule
// (spmv.cuh: spmv_stream_kernel + spmv_fixup_kernel) needs scratch memory -- two carry arrays with one element per
// 4096-entry tile, for the partial sums of rows that straddle tiles.  The regular path keeps them in buffers owned by
// the operator (allocated when it is analysed); this variant keeps them in ONE file-scope __device__ array of a fixed
// size, which every GPU that loads the library holds until exit:
//
//   SHB_STATIC_MAX_TILES = 65536 tiles = 2^28 stored entries (BASELINE config 2 exactly), 2 x 65536 x 4 B = 512 KB.
//
// The entry point calls no cudaMalloc / cudaFree and no allocating library routine; operators with more tiles are
// rejected with SHB_ERR_TOO_LARGE.  EVERY call on a device shares the array: calls that overlap on the GPU (two
// streams, two contexts on one device) race on it -- serialise them.  The array keeps the last call's contents; nothing
// has to be cleared between calls, because every carry element a launch reads was written by the same launch (a tile
// that a row crosses writes its carry_tail, the tile a straddling row ends in writes its carry_head; the fix-up reads
// exactly those).
//
// In host code the array's NAME is a host-side stand-in, not a device address: a kernel given it would fault (or, where
// the GPU can reach host memory, silently use the wrong memory).  The device address comes from cudaGetSymbolAddress,
// asked once per device with that device current, below.
#include <algorithm>
#include <mutex>

#include "spmv.cuh"

#define SHB_STATIC_MAX_TILES 65536
#define SHB_STATIC_MAX_DEVICES 64

static __device__ __align__(256) uint32_t shb_static_carry[2 * SHB_STATIC_MAX_TILES];  // [tail | head], 4-byte elements

static std::mutex g_static_mutex;
static uint32_t *g_static_ptr[SHB_STATIC_MAX_DEVICES];  // device address of shb_static_carry on each device, once resolved

// the array's address on device `device` (which must be current)
static int static_carry_address(int device, uint32_t **out) {
  SHB_REQUIRE(device >= 0 && device < SHB_STATIC_MAX_DEVICES, "device ordinal out of range for the static scratch table");
  std::lock_guard<std::mutex> lk(g_static_mutex);
  if (!g_static_ptr[device]) {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, shb_static_carry);
    if (e != cudaSuccess || !p) {
      shb_set_error("cudaGetSymbolAddress(shb_static_carry) failed on device %d: %s (code: %d)", device, cudaGetErrorString(e), (int)e);
      return SHB_ERR_CUDA;
    }
    g_static_ptr[device] = (uint32_t *)p;
  }
  *out = g_static_ptr[device];
  return SHB_OK;
}

template <class S>
static int launch_stream_static(shb_ctx *c, const shb_csr *A, const void *x, const void *y, const void *alpha, const void *beta,
                                void *out, uint32_t *carry) {
  typedef typename S::T T;
  SpmvArgs<T> a;
  memset(&a, 0, sizeof a);
  a.row_ptr = A->row_ptr;
  a.col = A->col;
  a.val = (const T *)A->val;
  a.x = (const T *)x;
  a.y = (const T *)y;
  a.alpha = *(const T *)alpha;
  a.beta = *(const T *)beta;
  a.out[0] = (T *)out;
  a.n_out = 1;
  a.row_offset = 0;
  a.n_rows = A->n_rows;
  a.vlength = (uint32_t)A->n_cols;
  a.ctl = IterCtl{nullptr, nullptr, 0.0, 0};
  a.use_y = S::needs_y(a.beta) ? 1 : 0;
  a.reversed = 0;
  const int nt = (int)A->n_tiles;
  if (A->nnz == 0 || nt == 0) {  // no stored entries: out = epi(ZERO, alpha, y, beta) per row, no tiles, no scratch
    const int64_t blocks = std::min<int64_t>((A->n_rows + 255) / 256, 2147483647LL);
    spmv_scalar_kernel<S, false><<<(unsigned)blocks, 256, 0, c->stream>>>(a);
    SHB_CUDA(cudaGetLastError());
    return SHB_OK;
  }
  T *ct = reinterpret_cast<T *>(carry), *ch = reinterpret_cast<T *>(carry + SHB_STATIC_MAX_TILES);
  if (A->val) spmv_stream_kernel<S, true, true, 0><<<nt, SHB_STREAM_THREADS, 0, c->stream>>>(a, A->tile_row, nt, (int32_t)A->nnz, ct, ch, SHB_LONG_PIECE);
  else spmv_stream_kernel<S, false, true, 0><<<nt, SHB_STREAM_THREADS, 0, c->stream>>>(a, A->tile_row, nt, (int32_t)A->nnz, ct, ch, SHB_LONG_PIECE);
  SHB_CUDA(cudaGetLastError());
  if (A->straddling_rows > 0) {
    const int blocks = std::min((nt + 255) / 256, c->prop.multiProcessorCount * 4);
    spmv_fixup_kernel<S><<<blocks, 256, 0, c->stream>>>(a, A->tile_row, nt, (int32_t)A->nnz, ct, ch);
    SHB_CUDA(cudaGetLastError());
  }
  return SHB_OK;
}

SHB_API int shb_spmv_stream_static_max_nnz(int64_t *max_nnz) {
  SHB_REQUIRE(max_nnz, "bad arguments");
  *max_nnz = (int64_t)SHB_STATIC_MAX_TILES * SHB_TILE_NNZ;
  return SHB_OK;
}

SHB_API int shb_spmv_stream_static(shb_ctx *c, const shb_csr *A, int semiring, const shb_buf *x, const shb_buf *y,
                                   const void *alpha, const void *beta, shb_buf *out, uint64_t *kernel_ns) {
  SHB_REQUIRE(c && A && x && out && alpha && beta, "bad arguments");
  SHB_REQUIRE(x->bytes >= (uint64_t)A->n_cols * 4, "x shorter than n_cols");
  SHB_REQUIRE(out->bytes >= (uint64_t)A->n_rows * 4, "out shorter than n_rows");
  SHB_REQUIRE(!y || y->bytes >= (uint64_t)A->n_rows * 4, "y shorter than n_rows");
  SHB_REQUIRE(out->ptr != x->ptr, "out must not alias x");
  if (A->n_tiles > SHB_STATIC_MAX_TILES) {
    shb_set_error("operator too large for the static scratch: %lld stored entries = %lld tiles, the limit is %d tiles (%lld entries); "
                  "use shb_spmv, whose scratch belongs to the operator",
                  (long long)A->nnz, (long long)A->n_tiles, SHB_STATIC_MAX_TILES, (long long)SHB_STATIC_MAX_TILES * SHB_TILE_NNZ);
    return SHB_ERR_TOO_LARGE;
  }
  if (!y) {  // y omitted: only legal when the semiring's epilogue ignores it for this beta (as shb_spmv)
    bool needs = true;
    switch (semiring) {
      case SHB_PLUS_TIMES_F32: needs = PlusTimesF32::needs_y(*(const float *)beta); break;
      case SHB_MIN_PLUS_F32: needs = true; break;
      case SHB_OR_AND_I32: needs = OrAndI32::needs_y(*(const int32_t *)beta); break;
      case SHB_MAX_MIN_I32: needs = MaxMinI32::needs_y(*(const int32_t *)beta); break;
    }
    SHB_REQUIRE(!needs, "y is required for this semiring / beta");
  }
  SHB_CUDA(cudaSetDevice(c->device));
  if (A->n_rows == 0) return SHB_OK;
  uint32_t *carry = nullptr;
  SHB_TRY(static_carry_address(c->device, &carry));
  const void *yp = y ? y->ptr : x->ptr;  // never dereferenced unless the epilogue needs y
  ScopedTimer t(c, kernel_ns);
  switch (semiring) {
    case SHB_PLUS_TIMES_F32: SHB_TRY(launch_stream_static<PlusTimesF32>(c, A, x->ptr, yp, alpha, beta, out->ptr, carry)); break;
    case SHB_MIN_PLUS_F32: SHB_TRY(launch_stream_static<MinPlusF32>(c, A, x->ptr, yp, alpha, beta, out->ptr, carry)); break;
    case SHB_OR_AND_I32: SHB_TRY(launch_stream_static<OrAndI32>(c, A, x->ptr, yp, alpha, beta, out->ptr, carry)); break;
    case SHB_MAX_MIN_I32: SHB_TRY(launch_stream_static<MaxMinI32>(c, A, x->ptr, yp, alpha, beta, out->ptr, carry)); break;
    default: shb_set_error("unknown semiring %d", semiring); return SHB_ERR_UNSUPPORTED;
  }
  return t.finish();
}
