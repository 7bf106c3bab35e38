/*
 * shb.h -- C ABI of libshb.so: the B200-native replacement for the device side of
 * AdamBrouwersHarries/sparseharness (semiring CSR SpMV  out = alpha*(A (x) x) (+) beta*y  and the
 * iterative BFS / SSSP / PageRank / SCC loops built on it).
 *
 * It replaces, in the reference (paths relative to the reference repo):
 *   inc/opencl_utils.h          (device query, error strings, checkCLError)
 *   inc/cl_memory_manager.h     (the bag of cl_mem handles)
 *   inc/harness.h:13-82         (context / queue / JIT of the Lift "KERNEL")
 *   inc/harness.h:149-411       (executeKernel, allocateBuffers, create/upload/fill/read buffers)
 *   app/{sssp,bfs,pr,scc}.cpp executeRun  (the host-driven fixed-point loops, now device-resident)
 *   example/<alg>/kernel*.json "source"   (the Lift-generated OpenCL kernels)
 *
 * Conventions: plain C types only; every function returns an shb_status (0 = ok) and never exits
 * or throws -- the C++ Harness wrapper (host/inc/harness.h) turns a non-zero status into the
 * reference's "log and exit(1)" behaviour (inc/opencl_utils.h:15-23).  All sizes are 64-bit (the
 * reference's unsigned-int byte sizes cap at 4 GB, inc/kernel_utils.h:27-30).  Functions taking a
 * `uint64_t *ns` time the device work with a cudaEvent pair on the context's stream and
 * synchronise (the reference's event-profiling semantics, inc/harness.h:183-194); pass NULL to
 * enqueue asynchronously.  A context is single-threaded like the reference's in-order queue.
 * There is NO CPU fallback: without a CUDA device shb_ctx_create fails.
 */
#ifndef SHB_H
#define SHB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SHB_ABI_VERSION 1

typedef struct shb_ctx shb_ctx; /* one device + one in-order stream   (inc/harness.h:19-81)        */
typedef struct shb_buf shb_buf; /* one device buffer                   (cl_mem)                      */
typedef struct shb_csr shb_csr; /* device-resident CSR operator + its launch schedule metadata      */

typedef enum {
  SHB_OK = 0,
  SHB_ERR_INVALID = 1,     /* bad argument                                                       */
  SHB_ERR_CUDA = 2,        /* a CUDA runtime call failed; see shb_last_error()                   */
  SHB_ERR_NO_DEVICE = 3,   /* "No OpenCL devices found!" analogue (inc/harness.h:24-27)          */
  SHB_ERR_OOM = 4,         /* allocation exceeds device memory (src/sparse_matrix.cpp:231-233)   */
  SHB_ERR_UNSUPPORTED = 5, /* semiring / type / variant combination not available                */
  SHB_ERR_TOO_LARGE = 6,   /* nnz >= 2^31 in one partition                                       */
  SHB_ERR_NOT_CONVERGED = 7 /* iteration cap reached (the reference loops forever, Q8)           */
} shb_status;

/* The harness's semirings (example/<alg>/kernel*.json user functions) */
typedef enum {
  SHB_PLUS_TIMES_F32 = 0, /* spmv/, pr/ : mult l*r, add x+y, epi dp*alpha + y*beta, ZERO 0.0f              */
  SHB_MIN_PLUS_F32 = 1,   /* sssp/      : |a|+|b|, min(|a|,|b|), min(|dp|+|alpha|,|y|+|beta|), ZERO FLT_MAX */
  SHB_OR_AND_I32 = 2,     /* bfs/       : a&&b, a||b, (dp&&alpha)||(y&&beta), ZERO 0                       */
  SHB_MAX_MIN_I32 = 3     /* scc/       : min, max, max(min(dp,alpha),min(y,beta)), ZERO INT_MIN           */
} shb_semiring;

/* Kernel schedule variants (what the seven Lift schedules, kernel{,2..7}.json, map onto) */
typedef enum {
  SHB_VARIANT_AUTO = 0,   /* chosen from row-length statistics                                            */
  SHB_VARIANT_SCALAR = 1, /* thread per row, forward order (glb-sdp / glb-sdp-rsa, kernel5/6.json)        */
  SHB_VARIANT_VECTOR = 2, /* sub-warp of `lanes` threads per row, 128-bit loads, shuffle reduction        */
  SHB_VARIANT_STREAM = 3, /* nnz-split load-balanced tiles staged through shared memory (power-law rows)   */
  SHB_VARIANT_BLOCKED = 4 /* propagation blocking for gather-bound operators: x slices staged in shared
                             memory, products streamed to the row owners (needs shb_csr_build_blocked)     */
} shb_variant;

enum { SHB_F_NO_CACHE_HINTS = 1u, /* plain ld.global for every stream (A/B experiments)          */
       SHB_F_NO_Y_SKIP = 2u,      /* always read y, even when beta makes it irrelevant            */
       SHB_F_REVERSED_ORDER = 4u, /* SCALAR: accumulate a row from its last entry to its first, the
                                     order of the ELL schedules (kernel{,3,5}.json write products
                                     reversed into the temp buffer)                               */
       SHB_F_Y_IS_ONE = 8u,       /* y is the semiring one in every element (canonical PageRank's
                                     y == 1, app/pr.cpp:183): the y argument is not read          */
       SHB_F_NO_PANELS = 16u,     /* AUTO: ignore the operator's column panels (A/B experiments)   */
       SHB_F_STREAM_MERGE = 32u,  /* STREAM: the merge-path row phase (balanced walk over row ends
                                     and entries; measured slower than the default, kept for A/B)  */
       SHB_F_STREAM_NO_MERGE = 64u, /* STREAM: the thread-per-row row phase                         */
       SHB_F_STREAM_SCAN = 128u,  /* STREAM: the segmented-scan row phase (a row's sum is one shared-
                                     memory read whatever its length)                              */
       SHB_F_STREAM_TMA = 256u    /* STREAM: the tile's columns and values arrive by TMA bulk copies
                                     (cp.async.bulk + mbarrier) instead of through the LSU          */ };

typedef struct {
  int32_t variant;        /* shb_variant                                                          */
  int32_t lanes;          /* VECTOR: threads per row 1,2,4,8,16,32; 0 = auto                      */
  int32_t rows_per_group; /* VECTOR: rows in flight per sub-warp (ILP); 0 = auto                   */
  int32_t block_threads;  /* 0 = auto (reference: Run.local1, src/run.cpp:12)                      */
  int32_t grid_blocks;    /* 0 = auto (reference: Run.global1 / local1)                            */
  uint32_t flags;         /* SHB_F_*                                                               */
} shb_launch_cfg;

typedef struct {
  int64_t n_rows, n_cols, nnz;
  int64_t max_row_len, min_row_len, empty_rows;
  double mean_row_len;
  int32_t has_values;     /* 0 = pattern (every stored value is the semiring one)                  */
  int32_t elem_bytes;     /* 4                                                                    */
  int32_t auto_variant, auto_lanes; /* what SHB_VARIANT_AUTO resolves to                           */
  int64_t stream_tiles, straddling_rows;
  uint64_t device_bytes;  /* bytes held on the device by this operator                            */
  uint64_t algorithmic_bytes_spmv; /* nnz*(4+vb) + (N+1)*4 + 2*N*4, beta == 0        */
} shb_csr_info;

/* ---- library / device (inc/opencl_utils.h:173-226, inc/harness.h:100-108) -------------------- */
int shb_abi_version(void);
const char *shb_status_string(int status);      /* <- getErrorString, inc/opencl_utils.h:25        */
const char *shb_last_error(void);               /* detail of the last failure on this thread       */
int shb_device_count(int *count);
int shb_device_max_alloc_bytes(int device, uint64_t *bytes); /* <- deviceGetMaxAllocSize :216      */
int shb_device_props(int device, char *name, size_t name_cap, int *sm_count, int *cc_major, int *cc_minor,
                     uint64_t *total_mem, uint64_t *l2_bytes);

int shb_ctx_create(int device, shb_ctx **ctx);  /* <- Harness ctor, inc/harness.h:13-82            */
int shb_ctx_destroy(shb_ctx *ctx);
int shb_ctx_set_stream(shb_ctx *ctx, void *cuda_stream); /* run on a caller-owned cudaStream_t      */
int shb_ctx_synchronize(shb_ctx *ctx);
int shb_ctx_device_name(shb_ctx *ctx, char *name, size_t cap); /* <- getDeviceName :100             */
int shb_ctx_device(shb_ctx *ctx, int *device);

/* ---- buffers (inc/harness.h:266-391) ------------------------------------------------------------ */
int shb_buf_create(shb_ctx *ctx, uint64_t bytes, shb_buf **buf);                       /* createGlobalArg  */
int shb_buf_wrap(shb_ctx *ctx, void *device_ptr, uint64_t bytes, shb_buf **buf);       /* caller-owned mem */
int shb_buf_destroy(shb_buf *buf);
int shb_buf_upload(shb_ctx *ctx, shb_buf *buf, uint64_t offset, const void *host, uint64_t bytes,
                   uint64_t *ns);                                                      /* writeToGlobalArg */
int shb_buf_download(shb_ctx *ctx, const shb_buf *buf, uint64_t offset, void *host, uint64_t bytes,
                     uint64_t *ns);                                                    /* readFromGlobalArg */
int shb_buf_fill_zero(shb_ctx *ctx, shb_buf *buf, uint64_t offset, uint64_t bytes, uint64_t *ns); /* fillGlobalArg */
int shb_buf_fill_u32(shb_ctx *ctx, shb_buf *buf, uint64_t offset_elems, uint64_t n_elems, uint32_t pattern);
int shb_buf_copy(shb_ctx *ctx, shb_buf *dst, const shb_buf *src, uint64_t bytes);
int shb_buf_size(const shb_buf *buf, uint64_t *bytes);
int shb_buf_device_ptr(const shb_buf *buf, void **ptr);
int shb_host_alloc_pinned(uint64_t bytes, void **ptr);
int shb_host_free_pinned(void *ptr);

/* ---- the operator: CSR on the device (replaces the ELL / RSA blobs of SparseMatrix::cl_encode,
 *      src/sparse_matrix.cpp:122-399; arguments 0 "idx" and 1 "vals" of the kernel ABI) ----------- */
int shb_csr_create(shb_ctx *ctx, int64_t n_rows, int64_t n_cols, int64_t nnz, const int64_t *row_ptr,
                   const int32_t *col_idx, const void *vals /* NULL = pattern */, int elem_bytes,
                   shb_csr **csr, uint64_t *upload_ns);
/* The same operator from COO tuples, encoded ON THE DEVICE (replaces calculate_ellpack + cl_encode,
 * src/sparse_matrix.cpp:72-399, with the reference's semantics: tuple k lands in row row_of_entry[k] = the Matrix
 * Market entry's SECOND index (Q1), keeps the tuple order inside its row (Q3, stable sort, duplicates kept), and
 * float values pass through int when truncate_through_int (Q2).  elem_kind 0 = float, 1 = int32. */
int shb_csr_from_coo(shb_ctx *ctx, int64_t n_rows, int64_t n_cols, int64_t n_entries, const int32_t *col_of_entry,
                     const int32_t *row_of_entry, const void *vals /* NULL = pattern */, int elem_kind,
                     int truncate_through_int, shb_csr **csr, uint64_t *ns);
int shb_csr_destroy(shb_csr *csr);
int shb_csr_get_info(const shb_csr *csr, shb_csr_info *info);
int shb_csr_download(shb_ctx *ctx, const shb_csr *csr, int64_t row_begin, int64_t row_count,
                     int64_t *row_ptr /* row_count+1, rebased to 0 */, int32_t *col_idx, void *vals);
int shb_csr_transpose(shb_ctx *ctx, const shb_csr *csr, shb_csr **transposed); /* push-side adjacency */
/* 1-D row partition: row_bounds[0..parts] with row_bounds[k] = first row whose prefix nnz
 * reaches k*nnz/parts (contiguous blocks balanced by non-zeros); slice_rows copies one block, on the
 * device, into an operator of its own that keeps GLOBAL column ids. */
/* Builds the column-slice-major copy of the operator that SHB_VARIANT_BLOCKED runs on (the analogue of the
 * reference re-encoding a matrix for its kernel, SparseMatrix::cl_encode, src/sparse_matrix.cpp:122-399).
 * rows_per_consumer: rows of a panel one CTA reduces out of shared memory (0 = sized so that a consumer
 * unit fills about 44 KB).  make_default != 0 makes SHB_VARIANT_AUTO resolve to the blocked schedule for
 * this operator.  SHB_ERR_UNSUPPORTED when a consumer unit cannot fit shared memory (skewed operators). */
int shb_csr_build_blocked(shb_ctx *ctx, shb_csr *csr, int rows_per_consumer, int make_default);
int shb_csr_blocked_info(const shb_csr *csr, int32_t *geom /* 8: n_prod, n_pass, n_panel, W, RS, cbuf, slot_cap, depth */, int64_t *entries,
                         uint64_t *device_bytes);
/* Column panels for operators whose x exceeds the L2 (config 5: 2^26 columns, x = 268 MB): n_panels sub-operators
 * of consecutive column ranges, each a CSR with global column ids over ONLY the rows it has entries in (a compressed
 * row list), so that the gathers of one launch stay inside an L2-resident slice of x and empty (row, panel) pairs
 * cost nothing; SHB_VARIANT_AUTO then runs panel after panel, accumulating the row sums through the epilogue (exact
 * for min-plus / or-and / max-min, summation order for floats).  n_panels 0 = x slices of at most 40 MB;
 * *n_built = 0 when the operator needs no panels.  Doubles the operator's footprint. */
int shb_csr_build_panels(shb_ctx *ctx, shb_csr *csr, int n_panels, int *n_built);
int shb_csr_balance_rows(shb_ctx *ctx, const shb_csr *csr, int parts, int64_t *row_bounds);
int shb_csr_slice_rows(shb_ctx *ctx, const shb_csr *csr, int64_t row_begin, int64_t row_count, shb_csr **slice);

/* ---- the hot path: one semiring SpMV (<- Harness::executeKernel, inc/harness.h:149-195) ---------
 * out[r] = epi( (+)_k mult(x[col k], val k), alpha, y[r], beta ), x <- ZERO for col out of range.
 * alpha/beta point at one element of the semiring's type.  out must not alias x; y may alias x. */
int shb_spmv(shb_ctx *ctx, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const shb_buf *x,
             const shb_buf *y, const void *alpha, const void *beta, shb_buf *out, uint64_t *kernel_ns);

/* The stream-schedule SpMV for callers that cannot allocate.  NO counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * Same arguments and result as shb_spmv with cfg = {SHB_VARIANT_STREAM} (bit for bit: the same kernels), but the
 * schedule's scratch (the carry arrays of rows that straddle tiles) lives in a file-scope __device__ array of the
 * library instead of buffers owned by the operator: 512 KB held by every GPU that loads the library, until exit.
 * The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).  Limit: operators of at most
 * 65536 tiles = 2^28 stored entries (shb_spmv_stream_static_max_nnz reports it); larger ones are rejected with
 * SHB_ERR_TOO_LARGE and a message naming the limit.  ALL calls on a device share the array: calls that overlap on
 * the GPU (two streams, two contexts of one device) race on it -- the caller serialises them.  The array keeps the
 * last call's contents; nothing needs clearing (every carry a launch reads, the same launch wrote).  Column panels
 * and the blocked copy of the operator are ignored. */
int shb_spmv_stream_static(shb_ctx *ctx, const shb_csr *A, int semiring, const shb_buf *x, const shb_buf *y,
                           const void *alpha, const void *beta, shb_buf *out, uint64_t *kernel_ns);
int shb_spmv_stream_static_max_nnz(int64_t *max_nnz);

/* Same call with HOST vectors (the reference-facing plugin boundary: x upload, kernel, out download,
 * all inside *total_ns; A stays resident like the reference's idx/vals buffers). */
int shb_spmv_host(shb_ctx *ctx, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const void *x_host,
                  const void *y_host /* NULL when beta makes y irrelevant */, const void *alpha, const void *beta,
                  void *out_host, uint64_t *total_ns, uint64_t *kernel_ns);

/* A batch of n_vec host vectors through the same resident operator (the trial loop of app/spmv.cpp:52-98 with a
 * fresh x per trial, or several right-hand sides): x_{i+1} is uploaded while SpMV i runs and result i-1 is
 * downloaded (three streams, double-buffered staging).  Every step still uploads its own x (and y) and downloads
 * its own result; *total_ns spans the first upload to the last download.  y_hosts may be NULL when beta makes y
 * irrelevant. */
int shb_spmv_host_batch(shb_ctx *ctx, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, int n_vec,
                        const void *const *x_hosts, const void *const *y_hosts, const void *alpha, const void *beta,
                        void *const *out_hosts, uint64_t *total_ns);

/* ---- device-resident fixed-point iteration (<- app/{sssp,bfs,pr,scc}.cpp executeRun) ------------
 * x_{k+1} = KERNEL(x_k, y_k); y_0 = y0, y_k = x_k afterwards when y_rebind (app/sssp.cpp:150, Q9).
 * term_mode 0: stop when all |x_k[i]-x_{k+1}[i]| < delta (sssp, pr); 1: all equal (bfs, scc).
 * The convergence test runs on the device inside the kernel's epilogue; the host looks at a
 * 4-byte flag once per `check_every` launches.  x and work are ping-pong buffers of n elements;
 * *result_in_work tells which one holds the last output.  per_iter_ns (cap max_iters) optional. */
enum { SHB_TERM_ABSDIFF = 0, SHB_TERM_EQUAL = 1 };
int shb_iterate(shb_ctx *ctx, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, shb_buf *x,
                const shb_buf *y0, const void *alpha, const void *beta, shb_buf *work, double delta,
                int term_mode, int y_rebind, uint32_t max_iters, uint32_t check_every, uint64_t *per_iter_ns,
                uint32_t *n_iters, int *result_in_work, uint64_t *total_ns);

/* ---- graph drivers on the semirings (north_star: frontier / vectors / flag stay on the device) -- */
typedef struct {
  uint32_t iterations;      /* levels (bfs) / relaxation rounds (sssp) / sweeps (pagerank)        */
  uint32_t converged;
  int64_t edges_traversed;  /* TEPS numerator: sum of in-degrees of reached vertices              */
  int64_t reached;
  uint64_t total_ns;        /* device time, first launch to last                                  */
  uint64_t launches;
} shb_graph_stats;

/* BFS levels under or-and: level[root]=0, frontier_{k+1} = (A (x) frontier_k) AND NOT visited,
 * level[v] = first k with v in frontier, -1 when unreached (SURVEY.md Q8 build decision).
 * AT (optional transpose) enables push steps for small frontiers. */
int shb_bfs(shb_ctx *ctx, const shb_csr *A, const shb_csr *AT, int64_t root, shb_buf *levels,
            uint32_t max_levels, shb_graph_stats *stats);
/* SSSP under min-plus to the exact fixed point (delta -> no change), frontier-aware. */
int shb_sssp(shb_ctx *ctx, const shb_csr *A, const shb_csr *AT, int64_t source, shb_buf *dist,
             uint32_t max_rounds, shb_graph_stats *stats);
/* Canonical PageRank (Q9): x_{k+1} = A x_k + beta (y == 1 not re-bound), stop on |dx| < delta. */
int shb_pagerank(shb_ctx *ctx, const shb_csr *A, const shb_launch_cfg *cfg, shb_buf *x, shb_buf *work,
                 float x_init, float beta, double delta, uint32_t max_iters, uint32_t check_every,
                 int *result_in_work, shb_graph_stats *stats);

/* Power iteration under plus-times (the eigenvector app; app/eigenvector.cpp:42-137 is dead code in the reference:
 * x <- A x, at most 10 launches, no normalisation -- that recurrence is shb_iterate with max_iters = 10).  This is
 * the canonical driver: out = A x_k, s_k = max |out_i|, x_{k+1} = out / s_k, until all |x_{k+1} - x_k| < delta.
 * x holds x_0 on entry; *eigenvalue = the last s_k (the dominant eigenvalue's magnitude on convergence). */
int shb_power_iteration(shb_ctx *ctx, const shb_csr *A, const shb_launch_cfg *cfg, shb_buf *x, shb_buf *work, double delta,
                        uint32_t max_iters, float *eigenvalue, int *result_in_work, shb_graph_stats *stats);

/* ---- synthetic workloads generated on the device (bit-identical to the oracle's
 *      generators so results can be cross-checked) ---------------------------------------------- */
int shb_synth_uniform(shb_ctx *ctx, int64_t n_rows, int64_t n_cols, int k, uint64_t seed, int val_mode,
                      int64_t row_begin, int64_t row_count, shb_csr **csr);
int shb_synth_vector(shb_ctx *ctx, shb_buf *x, int64_t n, uint64_t seed, int mode);
int shb_synth_rmat(shb_ctx *ctx, int scale, int64_t n_edges, uint64_t seed, int symmetrise, int value_mode,
                   float damping, shb_csr **csr);
int shb_synth_grid2d(shb_ctx *ctx, int64_t w, int64_t h, uint64_t seed, shb_csr **csr);

/* ---- multi-GPU: 1-D row partition, x exchanged every iteration (new capability) ----
 * One process per GPU.  Each rank exports its x buffers once with CUDA IPC; the fused kernel
 * writes every finished out[r] straight into all peers' next-x buffers over NVLink. */
#define SHB_IPC_HANDLE_BYTES 64
int shb_ipc_export(shb_buf *buf, void *handle /* SHB_IPC_HANDLE_BYTES */);
int shb_ipc_open(shb_ctx *ctx, const void *handle, uint64_t bytes, shb_buf **buf);
int shb_ipc_close(shb_buf *buf);
/* out slice [row_offset, row_offset + A.n_rows) of every dst[i] (i < n_dst) receives the result.
 * y == NULL means y is x (the re-bound y of the iterative apps) unless cfg->flags has SHB_F_Y_IS_ONE.
 * changed_flag (optional, >= 4 bytes): set to 1 when a row of this block fails the termination test
 * against x[row_offset + r].  prev_changed (optional): when its first word is 0 the launch returns at
 * once -- the fixed point was reached in the previous iteration (all ranks see the all-reduced flag). */
int shb_spmv_scatter(shb_ctx *ctx, const shb_csr *A, int semiring, const shb_launch_cfg *cfg, const shb_buf *x,
                     const shb_buf *y, const void *alpha, const void *beta, shb_buf *const *dst, int n_dst,
                     int64_t row_offset, double delta, int term_mode, shb_buf *changed_flag,
                     const shb_buf *prev_changed, uint64_t *kernel_ns);


/* Row-partitioned, direction-optimising BFS (new capability; the reference's bfs app is single-device,
 * app/bfs.cpp:94-152).  This rank owns the vertices [row_offset, row_offset + A.n_rows) of the n_total-vertex
 * graph (row_offset a multiple of 32): their rows of A (in-edges, pull) and optionally of A^T (AT, out-edges,
 * push; NULL = pull only), both with GLOBAL column ids; its slice of `levels`; full-length frontier / visited
 * bitmaps (n_total bits) and, optionally, `deg` = push degree of every vertex (int32 x n_total, steers Beamer's
 * heuristic).  counters: int64 words: 0,1 = reached vertices / TEPS numerator of this rank (after finish),
 * 2 = arrival word of the peer barrier (monotonic: zero it once, when the buffer is created), 4,5 = staging,
 * slot F = words 6+2(F mod ring), +1 = (vertices, push degrees) of frontier F summed over ranks (a ring of slot_ring
 * frontiers, shb_bfs_block_counter_bytes: read a slot before the traversal is slot_ring - 2 levels further).
 *   init     levels = -1, frontier0 = {root}, frontier1 = frontier2 = {}, counters (but the arrival word) = 0; the
 *            three frontier bitmaps rotate: level L reads bitmap L % 3, builds (L+1) % 3 and zeroes (L+2) % 3 for
 *            the level after (so no separate clearing pass and no store can race a clear); the root's owner reports
 *            frontier 0 -- into the staging words when `staged` (publish follows), else into slot 0.
 *   publish  adds the staged pair to `slot` of EVERY rank's counters (peer_counters[i], CUDA-IPC mappings) with
 *            system-scope atomics over NVLink and waits until barrier_target arrivals have been counted here
 *            (world x number of publishes so far; 0 = do not wait).  Needs all ranks past their init.
 *   level    enqueues prep (direction decision + visited update + zeroing of `spare`) / push / pull for `level`
 *            (frontier = bitmap level % 3, next[i] = every rank's bitmap (level+1) % 3, next[self_index] the local
 *            one, spare = the local bitmap (level+2) % 3) and, when peer_counters is given,
 *            the publish of slot level+1 -- the whole level is kernels of this stream, the frontier and the counts
 *            travel as the kernels' own peer stores, no collective library on the path.  peer_counters == NULL:
 *            the counts land in the local slot level+1 and the caller all-reduces it (NCCL baseline).
 *            A level launched after the frontier emptied does nothing but keep the barrier count in step.
 *   finish   per-rank reached / edge counts from the levels; fails if a peer barrier timed out (~4 s). */
int shb_csr_row_lengths(shb_ctx *ctx, const shb_csr *A, shb_buf *out /* int32 x n_rows */);
int shb_bfs_block_state_bytes(uint64_t *bytes);
int shb_bfs_block_counter_bytes(uint64_t *bytes, uint32_t *slot_ring); /* size of `counters`; slots form a ring */
int shb_bfs_block_init(shb_ctx *ctx, const shb_csr *A, const shb_csr *AT, int64_t row_offset, int64_t n_total,
                       int64_t nnz_total, int64_t root, const shb_buf *deg, shb_buf *levels, shb_buf *frontier0,
                       shb_buf *frontier1, shb_buf *frontier2, shb_buf *visited, shb_buf *counters, shb_buf *state,
                       int staged);
int shb_bfs_block_publish(shb_ctx *ctx, uint32_t slot, shb_buf *counters, shb_buf *state, shb_buf *const *peer_counters,
                          int n_ranks, int self_index, uint64_t barrier_target);
int shb_bfs_block_level(shb_ctx *ctx, const shb_csr *A, const shb_csr *AT, int64_t row_offset, int64_t n_total,
                        uint32_t level, const shb_buf *deg, const shb_buf *frontier, shb_buf *const *next, int n_next,
                        int self_index, shb_buf *spare, shb_buf *visited, shb_buf *levels, shb_buf *counters,
                        shb_buf *state, shb_buf *const *peer_counters, uint64_t barrier_target);
/* after a traversal stopped by its level cap: label the vertices of the last frontier built (frontier `level`) */
int shb_bfs_block_label_frontier(shb_ctx *ctx, const shb_csr *A, int64_t row_offset, int64_t n_total, uint32_t level,
                                 const shb_buf *frontier, shb_buf *spare, shb_buf *visited, shb_buf *levels,
                                 shb_buf *counters, shb_buf *state);
int shb_bfs_block_finish(shb_ctx *ctx, const shb_csr *A, const shb_buf *levels, shb_buf *counters, const shb_buf *state);

/* Row-partitioned SSSP (new capability; the reference's sssp app is single-device, app/sssp.cpp:97-155):
 * block-asynchronous near-far.  This rank owns the vertices [row_offset, row_offset + AT.n_rows): AT holds their
 * OUT-edges (rows of A^T, GLOBAL target ids; for a symmetric graph the block's rows of A), dist is a full-length
 * array on every rank whose owned slice is authoritative, prev (n_local) the distance each owned vertex was last
 * expanded with, work an opaque scratch of shb_sssp_block_work_bytes(n_local) bytes.
 *   init    dist = FLT_MAX, dist[source] = 0 on every rank, prev = FLT_MAX, bucket width from the block's weights.
 *   seed    first half of a round: the owned v with dist[v] < prev[v] (lowered by a peer since their last
 *           expansion, or the source) enter the work pile; their number is ADDED to int64 word `word` of seed_count;
 *           the smallest of their distances (float bits) that lies at or beyond the previous round's window is
 *           MIN-ed into word `word` of seed_min (optional, preset large): the window follows the front, late
 *           corrections below it are expanded along the way.
 *   solve   second half: the near-far solve of the block: improved owned targets are queued, improved targets of
 *           other ranks are sent to their owner as red.min on peer_dist[owner] (CUDA-IPC mappings, NVLink).
 *           bounds[0..n_ranks] = vertex ownership.  window_buckets > 0: only distances below seed_min[word] (all-
 *           reduced MIN by the caller) + window_buckets * bucket width are expanded this round; the rest stays
 *           pending and is seeded again, so all ranks sweep the distance range together instead of one block
 *           running ahead on distances a neighbour will still lower.
 *           The caller separates rounds with a barrier (the all-reduce of the seed count) and stops after a round
 *           whose seed count is 0 on every rank.
 *   finish  status of the last solve + (reached, sum of A_rows' row lengths over reached) into stats_out (2 x int64).
 * Same least fixed point as shb_sssp and the reference's dense sweeps, bit for bit (monotone float adds). */
int shb_sssp_block_work_bytes(int64_t n_local, uint64_t *bytes);
int shb_sssp_block_init(shb_ctx *ctx, const shb_csr *AT, int64_t row_offset, int64_t n_total, int64_t source,
                        shb_buf *dist, shb_buf *prev, shb_buf *work);
int shb_sssp_block_seed(shb_ctx *ctx, const shb_csr *AT, int64_t row_offset, int64_t n_total, const shb_buf *dist,
                        const shb_buf *prev, shb_buf *work, shb_buf *seed_count, shb_buf *seed_min, uint64_t word);
int shb_sssp_block_solve(shb_ctx *ctx, const shb_csr *AT, int64_t row_offset, int64_t n_total, shb_buf *dist,
                         shb_buf *const *peer_dist, int n_ranks, int self_index, const int64_t *bounds, shb_buf *prev,
                         shb_buf *work, const shb_buf *seed_min, uint64_t word, float window_buckets, uint32_t max_phases);
int shb_sssp_block_finish(shb_ctx *ctx, const shb_csr *A_rows, int64_t row_offset, const shb_buf *dist,
                          const shb_buf *work, shb_buf *stats_out);

/* ---- multi-GPU in ONE process: a group of devices driven by one caller ----------
 * Generalises the reference's single queue on a single device (inc/harness.h:78-81): shb_group owns one context
 * + stream per device with peer access enabled between all of them; shb_gpart is an operator row-partitioned over
 * the group (contiguous blocks balanced by non-zeros, bounds multiples of 32, GLOBAL column ids, block i resident
 * on device i; `AT` blocks = the same vertices' out-edges, needed by BFS push levels and by SSSP -- for a
 * symmetric operator the row blocks serve as both).  The drivers below are the round / level loops of the
 * reference's iterative apps (app/{pr,bfs,sssp}.cpp executeRun) on that partition: the per-block kernels are the
 * shb_spmv_scatter / shb_bfs_block_* / shb_sssp_block_* primitives above; every iteration ends with one publish
 * kernel per device that all-reduces a few words over NVLink with system-scope atomics and is the inter-device
 * barrier.  The host polls device 0's copy once per `check_every` iterations.  No collective library, no
 * PyTorch, no second process.  Times are device times (events), the maximum over the devices. */
typedef struct shb_group shb_group;
typedef struct shb_gpart shb_gpart;
int shb_group_create(int n_dev /* 1..8 */, const int *devices /* NULL: 0..n_dev-1 */, shb_group **grp);
int shb_group_destroy(shb_group *grp);
int shb_group_size(const shb_group *grp, int *n_dev);
int shb_group_ctx(shb_group *grp, int i, shb_ctx **ctx); /* to build blocks on device i (synthetic generators, slices) */
/* a HOST CSR, split and uploaded; symmetric != 0: the operator equals its transpose (no AT copy is made) */
int shb_group_partition_host(shb_group *grp, int64_t n_rows, int64_t n_cols, const int64_t *row_ptr,
                             const int32_t *col_idx, const void *vals /* NULL = pattern */, int symmetric,
                             shb_gpart **part);
/* blocks the caller built on the group's devices (blocks[i] on device i, rows [bounds[i], bounds[i+1]));
 * at_blocks == NULL: symmetric.  The partition takes ownership of the operators. */
int shb_group_adopt(shb_group *grp, shb_csr *const *blocks, shb_csr *const *at_blocks, const int64_t *bounds,
                    int64_t n_cols, shb_gpart **part);
int shb_gpart_destroy(shb_gpart *part);
int shb_gpart_info(const shb_gpart *part, int64_t *n_rows, int64_t *n_cols, int64_t *nnz, int64_t *bounds /* n_dev+1 */);
int shb_gpart_block(const shb_gpart *part, int i, shb_csr **A, shb_csr **AT);
/* x_{k+1} = KERNEL(x_k): shb_iterate on the partition (same arguments; x0_host == NULL: x0 filled with the 32-bit
 * pattern x0_fill_bits; y_mode 0: y0 every launch, 1: y0 for the first launch then y = x, 2: y is the semiring one).
 * result_host / prev_host (optional, n_rows elements): x_k and x_{k-1}.  SHB_ERR_NOT_CONVERGED at the cap. */
int shb_group_iterate(shb_group *grp, const shb_gpart *part, int semiring, const shb_launch_cfg *cfg, const void *x0_host,
                      uint32_t x0_fill_bits, const void *y0_host, const void *alpha, const void *beta, double delta,
                      int term_mode, int y_mode, uint32_t max_iters, uint32_t check_every, void *result_host,
                      void *prev_host, uint32_t *n_iters, uint64_t *total_ns);
int shb_group_pagerank(shb_group *grp, const shb_gpart *part, const shb_launch_cfg *cfg, float x_init, float beta,
                       double delta, uint32_t max_iters, uint32_t check_every, void *result_host, void *prev_host,
                       shb_graph_stats *stats);
int shb_group_bfs(shb_group *grp, const shb_gpart *part, int64_t root, int32_t *levels_host /* n_rows, optional */,
                  uint32_t max_levels, uint32_t check_every, shb_graph_stats *stats);
int shb_group_sssp(shb_group *grp, const shb_gpart *part, int64_t source, float *dist_host /* n_rows, optional */,
                   float window_buckets /* < 0: default (8 for n_dev > 1, none for 1) */, uint32_t max_rounds,
                   uint32_t check_every, shb_graph_stats *stats);

#ifdef __cplusplus
}
#endif
#endif /* SHB_H */
