/*
 * shb_oracle.c -- CPU ORACLE for the semiring CSR SpMV hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product path
 * (sparseharness_b200/csrc + host/) never links, loads or calls anything in oracle/.
 *
 * It is a plain-C restatement of the reference algorithm (AdamBrouwersHarries/sparseharness):
 *   - Matrix Market ingest with the reference's semantics (Q1-Q4 of SURVEY.md section 8a):
 *       src/sparse_matrix.cpp:11-70   (load_from_file)
 *       src/sparse_matrix.cpp:72-119  (calculate_ellpack: row = J, col = I, int truncation)
 *   - the five semiring kernels, following the simplest Lift kernels
 *       example/{spmv,sssp,bfs,pr,scc}/kernel5.json:3 (ELL, reversed accumulation order)
 *       example/spmv/kernel6.json:3                   (RSA, forward accumulation order)
 *   - the CPU gold loop inc/spmv_gold.h:9-28 (verbatim arithmetic, including its beta quirk Q5)
 *   - the iterative drivers app/{sssp,bfs,pr,scc}.cpp executeRun + should_terminate_iteration
 *   - the normalisations src/sparse_matrix.cpp:409-456
 *   - canonical graph answers (queue BFS levels, Dijkstra with the kernel's float adds, PageRank)
 *   - the synthetic workload generators of SURVEY.md section 8d (uniform / RMAT / 2-D grid)
 *
 * Parity pinning: tests/test_oracle_pinned.py checks this file against (a) the reference's own
 * loader + Gold<float>::spmv compiled from /root/reference into oracle/_ref (oracle/Makefile) and
 * (b) the golden vectors under tests/golden/ which were produced by that binary.
 * The min-plus / or-and / max-min kernels have NO CPU implementation in the reference
 * (app/sssp.cpp:243 passes an empty gold, results are "notchecked"): for those the restatement
 * of the kernel source is the only oracle -> "parity unpinned by the reference" for those rows.
 */
#include <float.h>
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ctype.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

enum { ORC_PLUS_TIMES = 0, ORC_MIN_PLUS = 1, ORC_OR_AND = 2, ORC_MAX_MIN = 3 };
enum { ORC_TERM_ABSDIFF = 0, ORC_TERM_EQUAL = 1 };

/* Writes a CSR block as a Matrix Market "coordinate real general" file (1-based, %.9g values): the only
 * input format the reference's loader accepts (src/sparse_matrix.cpp:11-70).  Used by bench.py to hand a
 * bounded sample of a synthetic operator to oracle/_ref/ref_driver.  Returns 0 on success. */
ORC_API int orc_write_mtx(const char *path, int64_t n_rows, int64_t n_cols, const int64_t *row_ptr, const int32_t *col,
                          const float *val, int second_index_is_row) {
  FILE *f = fopen(path, "w");
  if (!f) return 1;
  static char buf[1 << 20];
  setvbuf(f, buf, _IOFBF, sizeof buf);
  fprintf(f, "%%%%MatrixMarket matrix coordinate real general\n%lld %lld %lld\n", (long long)n_rows, (long long)n_cols,
          (long long)row_ptr[n_rows]);
  for (int64_t r = 0; r < n_rows; ++r)
    for (int64_t k = row_ptr[r]; k < row_ptr[r + 1]; ++k)
      if (second_index_is_row)  /* the reference keeps (I, J) as (x, y) and uses y as the row (src/sparse_matrix.cpp:59,86) */
        fprintf(f, "%d %lld %.9g\n", col[k] + 1, (long long)(r + 1), val ? (double)val[k] : 1.0);
      else
        fprintf(f, "%lld %d %.9g\n", (long long)(r + 1), col[k] + 1, val ? (double)val[k] : 1.0);
  return fclose(f) != 0;
}

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* ------------------------------------------------------------------------------------------
 * Semiring user functions, restated from the kernel sources (example/<alg>/kernel5.json:3).
 * Float ops are written as separate statements through volatile-free temporaries and this file
 * is compiled with -ffp-contract=off so that no FMA contraction changes the rounding.
 * ---------------------------------------------------------------------------------------- */
static inline float pt_mult(float xv, float a) { return xv * a; }            /* mult           */
static inline float pt_add(float acc, float p) { return acc + p; }           /* add            */
static inline float pt_epi(float dp, float alpha, float y, float beta) {     /* doubleMultiplyAdd */
  float l = dp * alpha;
  float r = y * beta;
  return l + r;
}
static inline float mp_mult(float xv, float a) { return fabsf(xv) + fabsf(a); }          /* absadd */
static inline float mp_add(float a, float b) { return fabsf(a) < fabsf(b) ? fabsf(a) : fabsf(b); } /* clmin */
static inline float mp_epi(float dp, float alpha, float y, float beta) {
  float a = fabsf(dp) + fabsf(alpha);
  float b = fabsf(y) + fabsf(beta);
  return fabsf(a) < fabsf(b) ? fabsf(a) : fabsf(b);
}
static inline int oa_mult(int a, int b) { return (a != 0) && (b != 0) ? 1 : 0; }  /* bool_and */
static inline int oa_add(int a, int b) { return (a != 0) || (b != 0) ? 1 : 0; }   /* bool_or  */
static inline int oa_epi(int dp, int alpha, int y, int beta) {                    /* doubleAndOr */
  int r1 = ((dp != 0) && (alpha != 0)) ? 1 : 0;
  int r2 = ((y != 0) && (beta != 0)) ? 1 : 0;
  return (r1 || r2);
}
static inline int mm_mult(int a, int b) { return a < b ? a : b; }  /* int_min */
static inline int mm_add(int a, int b) { return a > b ? a : b; }   /* int_max */
static inline int mm_epi(int dp, int alpha, int y, int beta) {     /* doubleMinMax */
  int m1 = dp < alpha ? dp : alpha;
  int m2 = y < beta ? y : beta;
  return m1 > m2 ? m1 : m2;
}

/* ZERO used for accumulator init and for out-of-range x (kernel5.json: v_tmp_* constants) */
ORC_API float orc_zero_f32(int semiring) { return semiring == ORC_MIN_PLUS ? 3.4028235E38f : 0.0f; }
ORC_API int orc_zero_i32(int semiring) { return semiring == ORC_MAX_MIN ? INT_MIN : 0; }

/* ------------------------------------------------------------------------------------------
 * One semiring SpMV on CSR:  out[r] = epi( ADD_k mult(x[col[k]], val[k]), alpha, y[r], beta )
 * with x <- ZERO when col < 0 or col >= vlength (kernel5.json:3 index guards).
 * order: 0 = forward (RSA kernels 6/7 and the gold loop), 1 = reversed (ELL kernels 1/3/5 write
 * products to the temp buffer at index W-1-i and then reduce ascending).
 * val == NULL means "pattern": every stored value is the semiring one (1.0f / 1).
 * threads: 1 = the sequential order-defining loop; >1 = OpenMP over rows (same per-row order).
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_spmv_f32(int semiring, int64_t n_rows, int64_t vlength, const int64_t *row_ptr,
                          const int32_t *col, const float *val, const float *x, const float *y,
                          float alpha, float beta, float *out, int order, int threads) {
  const float zero = orc_zero_f32(semiring);
  (void)threads;
#pragma omp parallel for schedule(static) num_threads(threads) if (threads > 1)
  for (int64_t r = 0; r < n_rows; ++r) {
    const int64_t s = row_ptr[r], e = row_ptr[r + 1];
    float acc = zero;
    for (int64_t i = 0; i < e - s; ++i) {
      const int64_t k = order ? (e - 1 - i) : (s + i);
      const int32_t c = col[k];
      const float xv = (c < 0 || c >= vlength) ? zero : x[c];
      const float a = val ? val[k] : 1.0f;
      if (semiring == ORC_MIN_PLUS)
        acc = mp_add(acc, mp_mult(xv, a));
      else
        acc = pt_add(acc, pt_mult(xv, a));
    }
    out[r] = (semiring == ORC_MIN_PLUS) ? mp_epi(acc, alpha, y[r], beta) : pt_epi(acc, alpha, y[r], beta);
  }
}

ORC_API void orc_spmv_i32(int semiring, int64_t n_rows, int64_t vlength, const int64_t *row_ptr,
                          const int32_t *col, const int32_t *val, const int32_t *x, const int32_t *y,
                          int32_t alpha, int32_t beta, int32_t *out, int order, int threads) {
  const int zero = orc_zero_i32(semiring);
  (void)threads;
#pragma omp parallel for schedule(static) num_threads(threads) if (threads > 1)
  for (int64_t r = 0; r < n_rows; ++r) {
    const int64_t s = row_ptr[r], e = row_ptr[r + 1];
    int acc = zero;
    for (int64_t i = 0; i < e - s; ++i) {
      const int64_t k = order ? (e - 1 - i) : (s + i);
      const int32_t c = col[k];
      const int xv = (c < 0 || c >= vlength) ? zero : x[c];
      const int a = val ? val[k] : 1;
      if (semiring == ORC_MAX_MIN)
        acc = mm_add(acc, mm_mult(xv, a));
      else
        acc = oa_add(acc, oa_mult(xv, a));
    }
    out[r] = (semiring == ORC_MAX_MIN) ? mm_epi(acc, alpha, y[r], beta) : oa_epi(acc, alpha, y[r], beta);
  }
}

/* ------------------------------------------------------------------------------------------
 * The reference's CPU gold, arithmetic verbatim (inc/spmv_gold.h:17-26):
 *   acc = zero; for each stored (c, v) of the row in file order:
 *       acc += (alpha * (x[c] * v)) + (beta * y[(int)v]);      <- Q5: beta term per non-zero,
 *   result[row] = acc;                                            indexed by the VALUE
 * Only meaningful for beta == 0 (the spmv app's setting); kept verbatim for the cross-check.
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_gold_spmv_f32(int64_t n_rows, const int64_t *row_ptr, const int32_t *col, const float *val,
                               const float *x, const float *y, int64_t ylen, float alpha, float beta,
                               float zero, float *out) {
  for (int64_t r = 0; r < n_rows; ++r) {
    float acc = zero;
    for (int64_t k = row_ptr[r]; k < row_ptr[r + 1]; ++k) {
      const float v = val[k];
      float t1 = x[col[k]] * v;
      t1 = alpha * t1;
      float t2 = 0.0f;
      if (beta != 0.0f) {
        int64_t yi = (int64_t)(int)v;
        if (yi < 0) yi = 0;
        if (yi >= ylen) yi = ylen - 1;
        t2 = beta * y[yi];
      } else {
        t2 = beta * 0.0f; /* y is finite in every app: beta*y == +-0 */
      }
      const float t3 = t1 + t2;
      acc = acc + t3;
    }
    out[r] = acc;
  }
}

/* ------------------------------------------------------------------------------------------
 * Matrix Market ingest with the reference's semantics.
 *   banner / size line:  src/mmio.cpp:92-199 (NIST mmio: first line banner, '%' comments, M N nz)
 *   entries:             src/sparse_matrix.cpp:43-63
 *     pattern -> val = 1.0;  1-based -> 0-based;  value parsed as double then static_cast<T>;
 *     symmetric -> mirror pushed right after its original when I != J; skew/hermitian = general.
 * elem_type: 0 = float (SparseMatrix<float>), 1 = int (SparseMatrix<int>).
 * Output COO is in the loader's tuple order: (I, J, val) with val already cast to T, stored in a
 * double (exact for both float and int32).
 * Returns 0 ok, <0 on the conditions for which the reference exits (-1).
 * ---------------------------------------------------------------------------------------- */
typedef struct {
  int32_t rows, cols;
  int64_t file_nnz;  /* nz of the size line */
  int64_t n;         /* tuples stored (after symmetric expansion) */
  int32_t *I, *J;
  double *V;
  int is_pattern, is_symmetric;
} orc_coo;

static void lower(char *p) {
  for (; *p; ++p) *p = (char)tolower((unsigned char)*p);
}

ORC_API int orc_mm_read(const char *path, int elem_type, orc_coo *out) {
  memset(out, 0, sizeof(*out));
  FILE *f = fopen(path, "r");
  if (!f) return -1;
  char line[1100], banner[80], mtx[80], crd[80], dtype[80], scheme[80];
  if (!fgets(line, sizeof line, f)) { fclose(f); return -2; }
  if (sscanf(line, "%64s %64s %64s %64s %64s", banner, mtx, crd, dtype, scheme) != 5) { fclose(f); return -2; }
  lower(mtx); lower(crd); lower(dtype); lower(scheme);
  if (strncmp(banner, "%%MatrixMarket", 14) != 0) { fclose(f); return -2; }
  if (strcmp(mtx, "matrix") != 0) { fclose(f); return -2; }
  const int coordinate = strcmp(crd, "coordinate") == 0;
  const int real = strcmp(dtype, "real") == 0, integer = strcmp(dtype, "integer") == 0;
  const int pattern = strcmp(dtype, "pattern") == 0;
  if (!coordinate || !(real || integer || pattern)) { fclose(f); return -3; } /* "Cannot process this matrix type" */
  const int symmetric = strcmp(scheme, "symmetric") == 0;
  /* skip comments, read sizes */
  int M = 0, N = 0, nz = 0;
  do {
    if (!fgets(line, sizeof line, f)) { fclose(f); return -4; }
  } while (line[0] == '%');
  if (sscanf(line, "%d %d %d", &M, &N, &nz) != 3) {
    int got;
    do {
      got = fscanf(f, "%d %d %d", &M, &N, &nz);
      if (got == EOF) { fclose(f); return -4; }
    } while (got != 3);
  }
  out->rows = M; out->cols = N; out->file_nnz = nz;
  out->is_pattern = pattern; out->is_symmetric = symmetric;
  int64_t cap = (int64_t)nz * 2 + 1;
  out->I = (int32_t *)malloc(sizeof(int32_t) * cap);
  out->J = (int32_t *)malloc(sizeof(int32_t) * cap);
  out->V = (double *)malloc(sizeof(double) * cap);
  int64_t n = 0;
  for (int i = 0; i < nz; ++i) {
    int I = 0, J = 0;
    double v = 1.0;
    if (pattern) {
      if (fscanf(f, "%d %d\n", &I, &J) != 2) break;
    } else {
      if (fscanf(f, "%d %d %lg\n", &I, &J, &v) != 3) break;
    }
    I--; J--;
    double tv = elem_type == 0 ? (double)(float)v : (double)(int)v; /* static_cast<T>(val) */
    out->I[n] = I; out->J[n] = J; out->V[n] = tv; ++n;
    if (symmetric && I != J) { out->I[n] = J; out->J[n] = I; out->V[n] = tv; ++n; }
  }
  out->n = n;
  fclose(f);
  return 0;
}

ORC_API void orc_coo_free(orc_coo *c) {
  free(c->I); free(c->J); free(c->V);
  memset(c, 0, sizeof(*c));
}

/* src/sparse_matrix.cpp:409-431 pagerank_normalise(dampingFactor, zero) for T = float.
 * column_sums[I] accumulated in float, in tuple order.  new = (fabs(v) / colsum[I]) * d.
 * The reference calls the unqualified ::fabs on a float with only <cmath> included, i.e. the
 * double overload, so the quotient and the product are evaluated in double and rounded to float
 * once on assignment.  (Not observable through the reference binary because of Q2 truncation.) */
ORC_API void orc_pagerank_normalise_f32(orc_coo *c, float damping, float zero) {
  float *colsum = (float *)malloc(sizeof(float) * (size_t)(c->cols > 0 ? c->cols : 1));
  for (int i = 0; i < c->cols; ++i) colsum[i] = zero;
  for (int64_t i = 0; i < c->n; ++i) colsum[c->I[i]] = colsum[c->I[i]] + (float)c->V[i];
  for (int64_t i = 0; i < c->n; ++i) {
    const float v = (float)c->V[i];
    const double q = (fabs((double)v) / (double)colsum[c->I[i]]) * (double)damping;
    c->V[i] = (double)(float)q;
  }
  free(colsum);
}

/* src/sparse_matrix.cpp:433-456 scc_normalise for T = int: diag -> numeric_limits<int>::min(),
 * off-diag -> (int)J  (the tuple's second index, i.e. the row the entry lands in). */
ORC_API void orc_scc_normalise_i32(orc_coo *c) {
  for (int64_t i = 0; i < c->n; ++i) c->V[i] = (c->I[i] == c->J[i]) ? (double)INT_MIN : (double)c->J[i];
}

/* src/sparse_matrix.cpp:72-119 calculate_ellpack restated as a stable counting sort into CSR:
 * tuple (I, J, v) goes to ROW J with COLUMN I (Q1), value truncated through int (Q2: "int val =
 * std::get<2>(...)") when truncate != 0, in-row order = tuple order (Q3: the sort there sorts a copy).
 * row_ptr has rows+1 entries; col/val have c->n entries.  val is written as float or int32
 * according to elem_type. */
ORC_API int orc_coo_to_csr(const orc_coo *c, int elem_type, int truncate, int64_t *row_ptr, int32_t *col,
                           void *val) {
  const int64_t H = c->rows;
  for (int64_t r = 0; r <= H; ++r) row_ptr[r] = 0;
  for (int64_t i = 0; i < c->n; ++i) {
    if (c->J[i] < 0 || c->J[i] >= H) return -1;
    row_ptr[c->J[i] + 1]++;
  }
  for (int64_t r = 0; r < H; ++r) row_ptr[r + 1] += row_ptr[r];
  int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * (size_t)(H > 0 ? H : 1));
  for (int64_t r = 0; r < H; ++r) cur[r] = row_ptr[r];
  for (int64_t i = 0; i < c->n; ++i) {
    const int64_t p = cur[c->J[i]]++;
    col[p] = c->I[i];
    double v = c->V[i];
    if (truncate) {
      /* float -> int conversion of an out-of-range value is UB in C; every fixture is in range */
      v = (double)(int)v;
    }
    if (elem_type == 0) ((float *)val)[p] = (float)v;
    else ((int32_t *)val)[p] = (int32_t)v;
  }
  free(cur);
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * Iterative drivers, FAITHFUL to app/{sssp,bfs,pr,scc}.cpp executeRun:
 *   in = x0 ; out = zero-filled ; y = y0 for iteration 0, then y is re-bound to the current
 *   input every iteration (app/sssp.cpp:150 "also set the y vector!")  [Q9]
 *   do { out = KERNEL(in, y) ; stop = should_terminate(in, out) ; swap(in,out) ; y = in } while(!stop)
 * term: ORC_TERM_ABSDIFF  all i: fabs(in[i]-out[i]) < delta   (app/sssp.cpp:170, app/pr.cpp:170;
 *                          float subtract, compare as double)
 *       ORC_TERM_EQUAL    all i: in[i] == out[i]              (app/bfs.cpp:167, app/scc.cpp:169)
 * The reference has NO iteration cap; max_iters guards the known non-terminating cases (Q8).
 * y_rebind = 1 reproduces the reference; y_rebind = 0 keeps y = y0 (canonical PageRank, Q9).
 * Returns the number of kernel launches performed; result (the last kernel output) in `result`.
 * ---------------------------------------------------------------------------------------- */
ORC_API int64_t orc_iterate_f32(int semiring, int64_t n, const int64_t *row_ptr, const int32_t *col,
                                const float *val, const float *x0, const float *y0, float alpha, float beta,
                                double delta, int term, int y_rebind, int64_t max_iters, int order,
                                int threads, float *result) {
  float *a = (float *)malloc(sizeof(float) * (size_t)n), *b = (float *)malloc(sizeof(float) * (size_t)n);
  memcpy(a, x0, sizeof(float) * (size_t)n);
  memset(b, 0, sizeof(float) * (size_t)n);
  float *in = a, *out = b;
  const float *y = y0;
  int64_t it = 0;
  int stop = 0;
  do {
    orc_spmv_f32(semiring, n, n, row_ptr, col, val, in, y, alpha, beta, out, order, threads);
    stop = 1;
    for (int64_t i = 0; i < n; ++i) {
      int ok;
      if (term == ORC_TERM_ABSDIFF) {
        const float d = in[i] - out[i];
        ok = fabs((double)d) < delta;
      } else {
        ok = in[i] == out[i];
      }
      if (!ok) { stop = 0; break; }
    }
    float *t = in; in = out; out = t;
    if (y_rebind) y = in;
    ++it;
  } while (!stop && it < max_iters);
  memcpy(result, in, sizeof(float) * (size_t)n);
  free(a); free(b);
  return it;
}

ORC_API int64_t orc_iterate_i32(int semiring, int64_t n, const int64_t *row_ptr, const int32_t *col,
                                const int32_t *val, const int32_t *x0, const int32_t *y0, int32_t alpha,
                                int32_t beta, int term, int y_rebind, int64_t max_iters, int order,
                                int threads, int32_t *result) {
  int32_t *a = (int32_t *)malloc(sizeof(int32_t) * (size_t)n), *b = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
  memcpy(a, x0, sizeof(int32_t) * (size_t)n);
  memset(b, 0, sizeof(int32_t) * (size_t)n);
  int32_t *in = a, *out = b;
  const int32_t *y = y0;
  int64_t it = 0;
  int stop = 0;
  (void)term;
  do {
    orc_spmv_i32(semiring, n, n, row_ptr, col, val, in, y, alpha, beta, out, order, threads);
    stop = 1;
    for (int64_t i = 0; i < n; ++i)
      if (in[i] != out[i]) { stop = 0; break; }
    int32_t *t = in; in = out; out = t;
    if (y_rebind) y = in;
    ++it;
  } while (!stop && it < max_iters);
  memcpy(result, in, sizeof(int32_t) * (size_t)n);
  free(a); free(b);
  return it;
}

/* ------------------------------------------------------------------------------------------
 * Canonical graph answers on the PULL CSR (row r lists the in-neighbours c of r: edge c -> r).
 * ---------------------------------------------------------------------------------------- */

/* transpose structure (+ optional float values): out-adjacency of every column vertex */
static void transpose_csr(int64_t n, const int64_t *row_ptr, const int32_t *col, const float *val,
                          int64_t **t_ptr, int32_t **t_idx, float **t_val) {
  const int64_t nnz = row_ptr[n];
  int64_t *p = (int64_t *)calloc((size_t)n + 1, sizeof(int64_t));
  int32_t *idx = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nnz > 0 ? nnz : 1));
  float *v = val ? (float *)malloc(sizeof(float) * (size_t)(nnz > 0 ? nnz : 1)) : NULL;
  for (int64_t k = 0; k < nnz; ++k) p[col[k] + 1]++;
  for (int64_t i = 0; i < n; ++i) p[i + 1] += p[i];
  int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
  memcpy(cur, p, sizeof(int64_t) * (size_t)n);
  for (int64_t r = 0; r < n; ++r)
    for (int64_t k = row_ptr[r]; k < row_ptr[r + 1]; ++k) {
      const int64_t q = cur[col[k]]++;
      idx[q] = (int32_t)r;
      if (v) v[q] = val[k];
    }
  free(cur);
  *t_ptr = p; *t_idx = idx; if (t_val) *t_val = v; else free(v);
}

/* BFS levels (Q8 build decision): level[root] = 0, level[v] = first k with v in
 * frontier_k where frontier_{k+1} = (A (x) frontier_k) AND NOT visited under the or-and semiring;
 * an edge counts when its stored value != 0 (bool_and's "b != 0"); val == NULL -> all ones.
 * Unreached vertices get -1.  Classic queue BFS over the transposed structure.
 * Returns the number of levels (max level + 1); *edges_traversed = sum of in-degrees of reached
 * vertices (the TEPS numerator). */
ORC_API int32_t orc_bfs_levels(int64_t n, const int64_t *row_ptr, const int32_t *col, const int32_t *val,
                               int64_t root, int32_t *level, int64_t *edges_traversed) {
  int64_t *tp; int32_t *ti;
  /* build transpose keeping only edges with val != 0 */
  const int64_t nnz = row_ptr[n];
  tp = (int64_t *)calloc((size_t)n + 1, sizeof(int64_t));
  ti = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nnz > 0 ? nnz : 1));
  for (int64_t k = 0; k < nnz; ++k)
    if (!val || val[k] != 0) tp[col[k] + 1]++;
  for (int64_t i = 0; i < n; ++i) tp[i + 1] += tp[i];
  int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
  memcpy(cur, tp, sizeof(int64_t) * (size_t)n);
  for (int64_t r = 0; r < n; ++r)
    for (int64_t k = row_ptr[r]; k < row_ptr[r + 1]; ++k)
      if (!val || val[k] != 0) ti[cur[col[k]]++] = (int32_t)r;
  free(cur);
  for (int64_t i = 0; i < n; ++i) level[i] = -1;
  int32_t *queue = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  int64_t head = 0, tail = 0;
  level[root] = 0; queue[tail++] = (int32_t)root;
  int32_t maxl = 0;
  while (head < tail) {
    const int32_t u = queue[head++];
    for (int64_t k = tp[u]; k < tp[u + 1]; ++k) {
      const int32_t v = ti[k];
      if (level[v] < 0) { level[v] = level[u] + 1; if (level[v] > maxl) maxl = level[v]; queue[tail++] = v; }
    }
  }
  if (edges_traversed) {
    int64_t e = 0;
    for (int64_t r = 0; r < n; ++r) if (level[r] >= 0) e += row_ptr[r + 1] - row_ptr[r];
    *edges_traversed = e;
  }
  free(queue); free(tp); free(ti);
  return maxl + 1;
}

/* SSSP fixed point of the min-plus recurrence  d[r] = min(d[r], min_k(|d[col k]| + |val k|))
 * from source (d = 0 there, FLT_MAX elsewhere): Dijkstra using exactly the kernel's float add
 * (absadd).  |val| >= 0 so Dijkstra is valid; fl(a + w) is monotone in a, so the label-setting
 * answer equals the Bellman-Ford fixed point bit for bit (SURVEY.md 8a determinism note).
 * Unreachable vertices keep 3.4028235E38f. */
typedef struct { float d; int32_t v; } heap_item;
static void heap_push(heap_item *h, int64_t *n, heap_item it) {
  int64_t i = (*n)++;
  while (i > 0) {
    int64_t p = (i - 1) / 2;
    if (h[p].d <= it.d) break;
    h[i] = h[p]; i = p;
  }
  h[i] = it;
}
static heap_item heap_pop(heap_item *h, int64_t *n) {
  heap_item top = h[0];
  heap_item last = h[--(*n)];
  int64_t i = 0;
  for (;;) {
    int64_t c = 2 * i + 1;
    if (c >= *n) break;
    if (c + 1 < *n && h[c + 1].d < h[c].d) c++;
    if (last.d <= h[c].d) break;
    h[i] = h[c]; i = c;
  }
  h[i] = last;
  return top;
}
ORC_API void orc_sssp_dijkstra(int64_t n, const int64_t *row_ptr, const int32_t *col, const float *val,
                               int64_t source, float *dist) {
  int64_t *tp; int32_t *ti; float *tv;
  transpose_csr(n, row_ptr, col, val, &tp, &ti, &tv);
  const int64_t nnz = row_ptr[n];
  for (int64_t i = 0; i < n; ++i) dist[i] = 3.4028235E38f;
  dist[source] = 0.0f;
  heap_item *heap = (heap_item *)malloc(sizeof(heap_item) * (size_t)(nnz + n + 1));
  int64_t hn = 0;
  heap_item s = {0.0f, (int32_t)source};
  heap_push(heap, &hn, s);
  while (hn > 0) {
    heap_item it = heap_pop(heap, &hn);
    if (it.d > dist[it.v]) continue;
    for (int64_t k = tp[it.v]; k < tp[it.v + 1]; ++k) {
      const float w = tv ? tv[k] : 1.0f;
      const float cand = mp_mult(it.d, w);
      const int32_t v = ti[k];
      /* clmin on non-negative operands is plain min; relax when strictly smaller */
      if (cand < dist[v]) { dist[v] = cand; heap_item ni = {cand, v}; heap_push(heap, &hn, ni); }
    }
  }
  free(heap); free(tp); free(ti); free(tv);
}

/* Canonical PageRank (Q9): x_{k+1}[r] = (sum_k x_k[col k] * val k) * 1 + 1 * beta, i.e. the plus-times
 * kernel with alpha = 1, y == 1, beta = (1-d)/N, NOT re-binding y; stop when all |x_{k+1}-x_k| < delta
 * or after max_iters.  acc_double = 1 gives the double-accumulated reference used for the tolerance
 * test, 0 the float forward-order kernel restatement.  Returns iterations run. */
ORC_API int64_t orc_pagerank(int64_t n, const int64_t *row_ptr, const int32_t *col, const float *val,
                             float x_init, float beta, double delta, int64_t max_iters, int acc_double,
                             int threads, float *result) {
  float *a = (float *)malloc(sizeof(float) * (size_t)n), *b = (float *)malloc(sizeof(float) * (size_t)n);
  for (int64_t i = 0; i < n; ++i) a[i] = x_init;
  float *in = a, *out = b;
  int64_t it = 0;
  int stop = 0;
  (void)threads;
  do {
    int changed = 0;
#pragma omp parallel for schedule(static) num_threads(threads) reduction(| : changed) if (threads > 1)
    for (int64_t r = 0; r < n; ++r) {
      float res;
      if (acc_double) {
        double acc = 0.0;
        for (int64_t k = row_ptr[r]; k < row_ptr[r + 1]; ++k) acc += (double)in[col[k]] * (double)(val ? val[k] : 1.0f);
        res = (float)(acc + (double)beta);
      } else {
        float acc = 0.0f;
        for (int64_t k = row_ptr[r]; k < row_ptr[r + 1]; ++k) acc = pt_add(acc, pt_mult(in[col[k]], val ? val[k] : 1.0f));
        res = pt_epi(acc, 1.0f, 1.0f, beta);
      }
      out[r] = res;
      const float d = in[r] - res;
      if (!(fabs((double)d) < delta)) changed |= 1;
    }
    stop = !changed;
    float *t = in; in = out; out = t;
    ++it;
  } while (!stop && it < max_iters);
  memcpy(result, in, sizeof(float) * (size_t)n);
  free(a); free(b);
  return it;
}

/* Power iteration (canonical mode of the eigenvector app; the reference's own driver, app/eigenvector.cpp:42-137, is
 * dead code: x <- A x, at most 10 launches, no normalisation -- that is orc_iterate with max_iters = 10).
 *   out = A x_k   (plus-times kernel loop, alpha 1, beta 0; `order` as orc_spmv_f32)
 *   s_k = max_i |out_i| ;  x_{k+1} = out / s_k  (left as it is when s_k == 0)
 * until all |x_{k+1} - x_k| < delta (float subtract, double compare, as app/sssp.cpp:170) or max_iters.
 * Returns the iterations run; *eigen = the last s_k. */
ORC_API int64_t orc_power_iteration(int64_t n, const int64_t *row_ptr, const int32_t *col, const float *val, const float *x0,
                                    double delta, int64_t max_iters, int order, float *result, float *eigen) {
  float *a = (float *)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1)), *b = (float *)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
  memcpy(a, x0, sizeof(float) * (size_t)n);
  float *in = a, *out = b, s = 0.0f;
  int64_t it = 0;
  int stop = 0;
  do {
    orc_spmv_f32(ORC_PLUS_TIMES, n, n, row_ptr, col, val, in, in, 1.0f, 0.0f, out, order, 1);
    s = 0.0f;
    for (int64_t i = 0; i < n; ++i) {
      const float m = fabsf(out[i]);
      if (m > s) s = m;
    }
    int changed = 0;
    for (int64_t i = 0; i < n; ++i) {
      const float o = s != 0.0f ? out[i] / s : out[i];
      out[i] = o;
      const float d = in[i] - o;
      if (!(fabs((double)d) < delta)) changed = 1;
    }
    stop = !changed;
    float *t = in; in = out; out = t;
    ++it;
  } while (!stop && it < max_iters);
  memcpy(result, in, sizeof(float) * (size_t)n);
  if (eigen) *eigen = s;
  free(a); free(b);
  return it;
}

/* ------------------------------------------------------------------------------------------
 * Synthetic workloads.  The same integer hash arithmetic is used by the device
 * generators in sparseharness_b200/csrc/synth.cu so that tests can compare the two bit for bit.
 * ---------------------------------------------------------------------------------------- */
static inline uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
ORC_API uint64_t orc_splitmix64(uint64_t z) { return splitmix64(z); }

static inline float u24_to_unit(uint32_t u) { return (float)u * (1.0f / 8388608.0f) - 1.0f; } /* [-1,1) */

/* C2 family: n_rows x n_cols, exactly k (<= 32) stored entries per row.
 *   h = splitmix64(seed ^ (row*k + j));  col = (hi32(h) * n_cols) >> 32;  u = h & 0xFFFFFF
 *   val_mode 0: val = u*2^-23 - 1 in [-1,1);  1: val = (u % 17) - 8 (small ints);  2: val = 1
 *   entries of a row sorted by (col, j) ascending; duplicates kept.                           */
ORC_API void orc_gen_uniform(int64_t n_rows, int64_t n_cols, int k, uint64_t seed, int val_mode,
                             int64_t row_begin, int64_t row_count, int64_t *row_ptr, int32_t *col, float *val) {
  for (int64_t i = 0; i <= row_count; ++i) row_ptr[i] = i * k;
  (void)n_rows;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < row_count; ++i) {
    const int64_t r = row_begin + i;
    uint64_t key[32];
    uint32_t uu[32];
    for (int j = 0; j < k; ++j) {
      const uint64_t h = splitmix64(seed ^ (uint64_t)(r * k + j));
      const uint64_t c = ((h >> 32) * (uint64_t)n_cols) >> 32;
      key[j] = (c << 5) | (uint64_t)j;
      uu[j] = (uint32_t)(h & 0xFFFFFFu);
    }
    for (int a = 1; a < k; ++a) { /* insertion sort on (col, j) */
      uint64_t kk = key[a]; int b = a - 1;
      while (b >= 0 && key[b] > kk) { key[b + 1] = key[b]; --b; }
      key[b + 1] = kk;
    }
    for (int j = 0; j < k; ++j) {
      const int src = (int)(key[j] & 31);
      col[i * k + j] = (int32_t)(key[j] >> 5);
      const uint32_t u = uu[src];
      val[i * k + j] = val_mode == 0 ? u24_to_unit(u) : val_mode == 1 ? (float)((int)(u % 17u) - 8) : 1.0f;
    }
  }
}

/* x vector: mode 0 uniform [-1,1) from splitmix64(seed ^ i); mode 1 all ones */
ORC_API void orc_gen_vector(int64_t n, uint64_t seed, int mode, float *x) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i)
    x[i] = mode == 0 ? u24_to_unit((uint32_t)(splitmix64(seed ^ (uint64_t)i) & 0xFFFFFFu)) : 1.0f;
}

/* RMAT edge (Graph500 style, a,b,c,d = .57,.19,.19,.05 as 24-bit integer thresholds), followed
 * by a bijective label scramble on `scale` bits. */
static inline uint32_t rmat_scramble(uint32_t v, int scale, uint64_t seed) {
  const uint32_t mask = scale >= 32 ? 0xFFFFFFFFu : ((1u << scale) - 1u);
  const uint32_t m1 = (uint32_t)(splitmix64(seed ^ 0xA5A5A5A5ULL) | 1ULL);
  const uint32_t m2 = (uint32_t)(splitmix64(seed ^ 0x5A5A5A5AULL) | 1ULL);
  const uint32_t add = (uint32_t)splitmix64(seed ^ 0x1234567ULL);
  v = (v * m1 + add) & mask;
  v ^= v >> ((scale + 1) / 2);
  v = (v * m2) & mask;
  v ^= v >> ((scale + 1) / 2);
  return v & mask;
}
ORC_API void orc_rmat_edge(uint64_t e, int scale, uint64_t seed, uint32_t *src, uint32_t *dst) {
  uint32_t s = 0, d = 0;
  const uint64_t base = splitmix64(seed ^ e);
  for (int l = 0; l < scale; ++l) {
    const uint32_t r = (uint32_t)(splitmix64(base + (uint64_t)l) >> 40); /* 24 bits */
    uint32_t bi, bj;
    if (r < 9563013u) { bi = 0; bj = 0; }
    else if (r < 12750684u) { bi = 0; bj = 1; }
    else if (r < 15938355u) { bi = 1; bj = 0; }
    else { bi = 1; bj = 1; }
    s = (s << 1) | bi; d = (d << 1) | bj;
  }
  *src = rmat_scramble(s, scale, seed);
  *dst = rmat_scramble(d, scale, seed);
}

static int cmp_u64(const void *a, const void *b) {
  const uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
  return x < y ? -1 : x > y;
}

/* Build the pull CSR of an RMAT graph: edge src -> dst is stored in ROW dst, COLUMN src.
 * symmetrise != 0: self loops dropped, both directions stored, (row,col) de-duplicated  (C3, BFS)
 * symmetrise == 0: every generated edge kept, duplicates and self loops included          (C5, PR)
 * Entries sorted by (row, col).  Call with col == NULL to obtain only *nnz_out (row_ptr filled).
 * Returns 0; needs up to 2*n_edges*8 bytes of scratch. */
ORC_API int orc_gen_rmat_csr(int scale, int64_t n_edges, uint64_t seed, int symmetrise, int64_t *row_ptr,
                             int32_t *col, int64_t *nnz_out) {
  const int64_t n = (int64_t)1 << scale;
  const int64_t cap = symmetrise ? 2 * n_edges : n_edges;
  uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(cap > 0 ? cap : 1));
  int64_t m = 0;
  for (int64_t e = 0; e < n_edges; ++e) {
    uint32_t s, d;
    orc_rmat_edge((uint64_t)e, scale, seed, &s, &d);
    if (symmetrise) {
      if (s == d) continue;
      keys[m++] = ((uint64_t)d << 32) | s;
      keys[m++] = ((uint64_t)s << 32) | d;
    } else {
      keys[m++] = ((uint64_t)d << 32) | s;
    }
  }
  qsort(keys, (size_t)m, sizeof(uint64_t), cmp_u64);
  if (symmetrise) {
    int64_t w = 0;
    for (int64_t i = 0; i < m; ++i)
      if (i == 0 || keys[i] != keys[i - 1]) keys[w++] = keys[i];
    m = w;
  }
  for (int64_t r = 0; r <= n; ++r) row_ptr[r] = 0;
  for (int64_t i = 0; i < m; ++i) row_ptr[(keys[i] >> 32) + 1]++;
  for (int64_t r = 0; r < n; ++r) row_ptr[r + 1] += row_ptr[r];
  if (col)
    for (int64_t i = 0; i < m; ++i) col[i] = (int32_t)(keys[i] & 0xFFFFFFFFu);
  *nnz_out = m;
  free(keys);
  return 0;
}

/* PageRank values for a pattern pull-CSR (canonical mode, float values kept):
 * pagerank_normalise applied to all-ones entries: val = (|1| / outdeg[col]) * d, where
 * outdeg[col] = number of stored entries with that column (float accumulation of ones). */
ORC_API void orc_pagerank_values(int64_t n, const int64_t *row_ptr, const int32_t *col, float damping,
                                 float *val) {
  const int64_t nnz = row_ptr[n];
  float *colsum = (float *)calloc((size_t)(n > 0 ? n : 1), sizeof(float));
  for (int64_t k = 0; k < nnz; ++k) colsum[col[k]] = colsum[col[k]] + 1.0f;
  for (int64_t k = 0; k < nnz; ++k) val[k] = (float)((1.0 / (double)colsum[col[k]]) * (double)damping);
  free(colsum);
}

/* C4: W x H 4-neighbour grid, vertex id = y*W + x, both directions stored, neighbour order
 * (y-1,x), (y,x-1), (y,x+1), (y+1,x) = ascending column.  Edge weight symmetric:
 *   w(u,v) = 1 + splitmix64(seed ^ (min(u,v)*2 + dir)) % 255,  dir = 0 horizontal, 1 vertical. */
static inline float grid_weight(uint64_t lo, int dir, uint64_t seed) {
  return (float)(1u + (uint32_t)(splitmix64(seed ^ (lo * 2u + (uint64_t)dir)) % 255u));
}
ORC_API int64_t orc_gen_grid2d(int64_t W, int64_t H, uint64_t seed, int64_t *row_ptr, int32_t *col, float *val) {
  int64_t p = 0;
  for (int64_t y = 0; y < H; ++y)
    for (int64_t x = 0; x < W; ++x) {
      const int64_t u = y * W + x;
      row_ptr[u] = p;
      if (y > 0) { if (col) { col[p] = (int32_t)(u - W); val[p] = grid_weight((uint64_t)(u - W), 1, seed); } ++p; }
      if (x > 0) { if (col) { col[p] = (int32_t)(u - 1); val[p] = grid_weight((uint64_t)(u - 1), 0, seed); } ++p; }
      if (x + 1 < W) { if (col) { col[p] = (int32_t)(u + 1); val[p] = grid_weight((uint64_t)u, 0, seed); } ++p; }
      if (y + 1 < H) { if (col) { col[p] = (int32_t)(u + W); val[p] = grid_weight((uint64_t)u, 1, seed); } ++p; }
    }
  row_ptr[W * H] = p;
  return p;
}
