// ldst.cuh -- cache-hinted, predicated global loads for the CSR streams and the x gathers.
//
// Streams (col_idx / vals) are read exactly once: ld.global.nc, no L1 allocation, L2 evict-first
// policy so that they do not displace x from the 126 MB L2.  x is gathered many times: normal L1
// allocation, L2 evict-last policy.  Predicates live INSIDE the asm so that every load of an
// unrolled group is emitted as straight-line "@p ld" code and ptxas can issue them back to back
// (memory-level parallelism) instead of branching around each one.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ uint64_t shb_policy_evict_first() {
  uint64_t p;
  asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t shb_policy_evict_last() {
  uint64_t p;
  asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// 128-bit streaming load of 4 consecutive 32-bit words; lanes with pred == false get `dflt` in all four.
template <bool HINTS>
__device__ __forceinline__ uint4 shb_ld_stream_v4(const void *ptr, bool pred, uint32_t dflt, uint64_t pol) {
  uint4 r;
  if (HINTS) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b32 %0, %6;\n\tmov.b32 %1, %6;\n\tmov.b32 %2, %6;\n\tmov.b32 %3, %6;\n\t"
        "@p ld.global.nc.L1::no_allocate.L2::cache_hint.v4.b32 {%0, %1, %2, %3}, [%4], %7;\n\t}"
        : "=&r"(r.x), "=&r"(r.y), "=&r"(r.z), "=&r"(r.w)
        : "l"(ptr), "r"((int)pred), "r"(dflt), "l"(pol));
  } else {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b32 %0, %6;\n\tmov.b32 %1, %6;\n\tmov.b32 %2, %6;\n\tmov.b32 %3, %6;\n\t"
        "@p ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];\n\t}"
        : "=&r"(r.x), "=&r"(r.y), "=&r"(r.z), "=&r"(r.w)
        : "l"(ptr), "r"((int)pred), "r"(dflt));
  }
  return r;
}

// 32-bit streaming load (row heads / tails that are not 16-byte aligned)
template <bool HINTS>
__device__ __forceinline__ uint32_t shb_ld_stream_b32(const void *ptr, bool pred, uint32_t dflt, uint64_t pol) {
  uint32_t r;
  if (HINTS) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.b32 %0, %3;\n\t"
        "@p ld.global.nc.L1::no_allocate.L2::cache_hint.b32 %0, [%1], %4;\n\t}"
        : "=&r"(r)
        : "l"(ptr), "r"((int)pred), "r"(dflt), "l"(pol));
  } else {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.b32 %0, %3;\n\t"
        "@p ld.global.nc.b32 %0, [%1];\n\t}"
        : "=&r"(r)
        : "l"(ptr), "r"((int)pred), "r"(dflt));
  }
  return r;
}

// x gather with the reference kernels' index guard (example/spmv/kernel5.json:3:
// "if (idx < 0) ZERO else if (idx >= VLength) ZERO else x[idx]"): one unsigned compare.
template <bool HINTS>
__device__ __forceinline__ uint32_t shb_gather_b32(const void *x, int32_t idx, uint32_t vlength, uint32_t zero_bits,
                                                   uint64_t pol) {
  uint32_t r;
  const char *ptr = (const char *)x + ((uint64_t)(uint32_t)idx << 2);
  const bool pred = (uint32_t)idx < vlength;
  if (HINTS) {
    asm("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.b32 %0, %3;\n\t"
        "@p ld.global.nc.L2::cache_hint.b32 %0, [%1], %4;\n\t}"
        : "=&r"(r)
        : "l"(ptr), "r"((int)pred), "r"(zero_bits), "l"(pol));
  } else {
    asm("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.b32 %0, %3;\n\t"
        "@p ld.global.nc.b32 %0, [%1];\n\t}"
        : "=&r"(r)
        : "l"(ptr), "r"((int)pred), "r"(zero_bits));
  }
  return r;
}

__device__ __forceinline__ uint32_t shb_ld_volatile_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

template <class T> __device__ __forceinline__ uint32_t shb_bits(T v);
template <> __device__ __forceinline__ uint32_t shb_bits<float>(float v) { return __float_as_uint(v); }
template <> __device__ __forceinline__ uint32_t shb_bits<int32_t>(int32_t v) { return (uint32_t)v; }
template <class T> __device__ __forceinline__ T shb_from_bits(uint32_t b);
template <> __device__ __forceinline__ float shb_from_bits<float>(uint32_t b) { return __uint_as_float(b); }
template <> __device__ __forceinline__ int32_t shb_from_bits<int32_t>(uint32_t b) { return (int32_t)b; }
