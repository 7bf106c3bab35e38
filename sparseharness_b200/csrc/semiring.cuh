// semiring.cuh -- the harness's four semirings as device functors.
//
// Each struct restates the three user functions at the top of the corresponding Lift kernel
// (reference: example/<alg>/kernel5.json:3) plus the ZERO constant used there for accumulator
// initialisation and for out-of-range x:
//   PlusTimesF32  example/spmv/kernel5.json:3, example/pr/kernel5.json:3   mult / add / doubleMultiplyAdd
//   MinPlusF32    example/sssp/kernel5.json:3                              absadd / clmin / doubleMultiplyAdd
//   OrAndI32      example/bfs/kernel5.json:3                               bool_and / bool_or / doubleAndOr
//   MaxMinI32     example/scc/kernel5.json:3                               int_min / int_max / doubleMinMax
// Float multiply/add use the _rn intrinsics so nvcc never contracts them into an FMA: with the same
// accumulation order the result is bit-identical to the CPU loop (inc/spmv_gold.h:17-26).
#pragma once
#include <cuda_runtime.h>
#include <limits.h>
#include <stdint.h>

#include "../../include/shb.h"

struct PlusTimesF32 {
  typedef float T;
  static constexpr int id = SHB_PLUS_TIMES_F32;
  __device__ __forceinline__ static T zero() { return 0.0f; }
  __device__ __forceinline__ static T one() { return 1.0f; }
  __device__ __forceinline__ static T mult(T xv, T a) { return __fmul_rn(xv, a); }
  __device__ __forceinline__ static T add(T acc, T p) { return __fadd_rn(acc, p); }
  __device__ __forceinline__ static T epi(T dp, T alpha, T y, T beta) {
    return __fadd_rn(__fmul_rn(dp, alpha), __fmul_rn(y, beta));
  }
  __host__ __device__ static bool needs_y(T beta) { return beta != 0.0f; }
};

struct MinPlusF32 {
  typedef float T;
  static constexpr int id = SHB_MIN_PLUS_F32;
  __device__ __forceinline__ static T zero() { return 3.4028235E38f; }
  __device__ __forceinline__ static T one() { return 1.0f; }  // pattern edges weigh 1 (loader: pattern -> 1.0)
  __device__ __forceinline__ static T mult(T xv, T a) { return __fadd_rn(fabsf(xv), fabsf(a)); }
  __device__ __forceinline__ static T add(T a, T b) { return fabsf(a) < fabsf(b) ? fabsf(a) : fabsf(b); }
  __device__ __forceinline__ static T epi(T dp, T alpha, T y, T beta) {
    const T a = __fadd_rn(fabsf(dp), fabsf(alpha));
    const T b = __fadd_rn(fabsf(y), fabsf(beta));
    return fabsf(a) < fabsf(b) ? fabsf(a) : fabsf(b);
  }
  __host__ __device__ static bool needs_y(T) { return true; }
};

struct OrAndI32 {
  typedef int32_t T;
  static constexpr int id = SHB_OR_AND_I32;
  __device__ __forceinline__ static T zero() { return 0; }
  __device__ __forceinline__ static T one() { return 1; }
  __device__ __forceinline__ static T mult(T a, T b) { return ((a != 0) && (b != 0)) ? 1 : 0; }
  __device__ __forceinline__ static T add(T a, T b) { return ((a != 0) || (b != 0)) ? 1 : 0; }
  __device__ __forceinline__ static T epi(T dp, T alpha, T y, T beta) {
    const int r1 = ((dp != 0) && (alpha != 0)) ? 1 : 0;
    const int r2 = ((y != 0) && (beta != 0)) ? 1 : 0;
    return (r1 || r2) ? 1 : 0;
  }
  __host__ __device__ static bool needs_y(T beta) { return beta != 0; }
};

struct MaxMinI32 {
  typedef int32_t T;
  static constexpr int id = SHB_MAX_MIN_I32;
  __device__ __forceinline__ static T zero() { return INT_MIN; }
  __device__ __forceinline__ static T one() { return INT_MAX; }  // multiplicative identity of min
  __device__ __forceinline__ static T mult(T a, T b) { return a < b ? a : b; }
  __device__ __forceinline__ static T add(T a, T b) { return a > b ? a : b; }
  __device__ __forceinline__ static T epi(T dp, T alpha, T y, T beta) {
    const T m1 = dp < alpha ? dp : alpha;
    const T m2 = y < beta ? y : beta;
    return m1 > m2 ? m1 : m2;
  }
  __host__ __device__ static bool needs_y(T beta) { return beta != INT_MIN; }
};

// Convergence predicate of the iterative apps, evaluated per row on the device:
//   term 0: fabs(in - out) < delta, float subtract then double compare (app/sssp.cpp:170, app/pr.cpp:170)
//   term 1: in == out                                                  (app/bfs.cpp:167, app/scc.cpp:169)
// returns true when the row has NOT converged.
__device__ __forceinline__ bool shb_row_differs(float in, float out, double delta, int term) {
  if (term == SHB_TERM_EQUAL) return !(in == out);
  const float d = __fsub_rn(in, out);
  return !(fabs((double)d) < delta);
}
__device__ __forceinline__ bool shb_row_differs(int32_t in, int32_t out, double delta, int term) {
  if (term == SHB_TERM_EQUAL) return in != out;
  const double d = (double)in - (double)out;
  return !(fabs(d) < delta);
}
