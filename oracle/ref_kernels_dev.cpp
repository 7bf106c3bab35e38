// ref_kernels_dev.cpp -- runs the reference's OWN device kernel text on the CPU.
//
// TEST INFRASTRUCTURE (oracle/).  This translation unit contains no reference code: it defines the
// handful of OpenCL C built-ins the Lift-generated sources use, then #includes
// oracle/_ref/gen/ref_kernels.inc, which oracle/gen_ref_kernels.py writes at build time from
//   /root/reference/example/{spmv,sssp,bfs,pr,scc}/kernel5.json:3   and   .../kernel6.json:3
// (the unmodified "source" strings, one namespace each).  It deliberately includes NO system header,
// so that `fabs`, `min` and `max` below are the only candidates the kernel text can bind to:
//   OpenCL C  fabs(float) -> float            (6.12.2; NOT C's double fabs)
//             min/max(int, int) -> int        (6.12.3)
//             get_global_id(0) / get_global_size(0): one work-item at a time, chosen by refk_ndrange
// A 1-D NDRange of G work-items is emulated by running the kernel G times with ids 0..G-1; kernels 5
// and 6 have no barrier and no cross-work-item communication (each row is written by one work-item).
#define kernel
#define global
#define restrict __restrict__

static int refk_gid = 0, refk_gsz = 1;
static inline int get_global_id(int) { return refk_gid; }
static inline int get_global_size(int) { return refk_gsz; }
static inline float fabs(float v) { return __builtin_fabsf(v); }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }

#include "_ref/gen/ref_kernels.inc"

// argument order = the harness's fixed kernel-arg ABI (inc/harness.h:199-249, inc/kernel_utils.h:145-160):
//   idx, vals, x, y, alpha, beta, out, temp global, then the int sizes in paramVars order
//   kernel5: MHeight, MWidthC, VLength          kernel6: MHeight, MWidthC
#define RUN5(NS) NS::KERNEL(idx, (const T *)vals, x, y, alpha, beta, out, temp, mheight, mwidth, vlength)
#define RUN6(NS) NS::KERNEL(idx, (const T *)vals, x, y, alpha, beta, out, temp, mheight, mwidth)

template <typename T, typename F> static void ndrange(int gsz, F f) {
  refk_gsz = gsz < 1 ? 1 : gsz;
  for (refk_gid = 0; refk_gid < refk_gsz; ++refk_gid) f();
  refk_gid = 0;
  refk_gsz = 1;
}

// alg: 0 spmv, 1 sssp, 3 pr (float) -- variant 5 or 6
int refk_run_f32(int alg, int variant, int gsz, const int *idx, const void *vals, const float *x, const float *y,
                 float alpha, float beta, float *out, float *temp, int mheight, int mwidth, int vlength) {
  typedef float T;
  if (variant == 5) {
    if (alg == 0) ndrange<T>(gsz, [&] { RUN5(k_spmv_5); });
    else if (alg == 1) ndrange<T>(gsz, [&] { RUN5(k_sssp_5); });
    else if (alg == 3) ndrange<T>(gsz, [&] { RUN5(k_pr_5); });
    else return 1;
  } else if (variant == 6) {
    if (alg == 0) ndrange<T>(gsz, [&] { RUN6(k_spmv_6); });
    else if (alg == 1) ndrange<T>(gsz, [&] { RUN6(k_sssp_6); });
    else if (alg == 3) ndrange<T>(gsz, [&] { RUN6(k_pr_6); });
    else return 1;
  } else
    return 1;
  return 0;
}

// alg: 2 bfs, 4 scc (int)
int refk_run_i32(int alg, int variant, int gsz, const int *idx, const void *vals, const int *x, const int *y,
                 int alpha, int beta, int *out, int *temp, int mheight, int mwidth, int vlength) {
  typedef int T;
  if (variant == 5) {
    if (alg == 2) ndrange<T>(gsz, [&] { RUN5(k_bfs_5); });
    else if (alg == 4) ndrange<T>(gsz, [&] { RUN5(k_scc_5); });
    else return 1;
  } else if (variant == 6) {
    if (alg == 2) ndrange<T>(gsz, [&] { RUN6(k_bfs_6); });
    else if (alg == 4) ndrange<T>(gsz, [&] { RUN6(k_scc_6); });
    else return 1;
  } else
    return 1;
  return 0;
}
