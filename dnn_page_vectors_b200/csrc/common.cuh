// Shared helpers for libclsm_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <float.h>
#include "../../include/clsm_b200.h"

// ---- error plumbing ---------------------------------------------------------------------
void clsm_set_error(const char* fmt, ...);

#define CLSM_REQUIRE(cond, ...)                                   \
  do {                                                            \
    if (!(cond)) {                                                \
      clsm_set_error(__VA_ARGS__);                                \
      return CLSM_EINVAL;                                         \
    }                                                             \
  } while (0)

#define CLSM_CUDA(expr)                                                           \
  do {                                                                            \
    cudaError_t e__ = (expr);                                                     \
    if (e__ != cudaSuccess) {                                                     \
      clsm_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,                 \
                     cudaGetErrorString(e__));                                    \
      return CLSM_ECUDA;                                                          \
    }                                                                             \
  } while (0)

extern thread_local void* g_clsm_marks[2];       // measurement marks (api.cu), consumed by clsm_conv_pool_bwd
extern unsigned long long g_clsm_launches;      // kernels of this library launched so far (api.cu)
#define CLSM_LAUNCH_CHECK()            \
  do {                                 \
    ++g_clsm_launches;                 \
    CLSM_CUDA(cudaGetLastError());     \
  } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int clsm_sm_count();   // cached per device

// ---- order-preserving float <-> uint32 ------------------------------------------------------
// larger float  <=>  larger unsigned.  -0.0f and +0.0f map to different keys; accumulators
// never produce -0.0 from a sum that starts at +0, and the finalize step decodes exactly.
__host__ __device__ __forceinline__ uint32_t f32_to_ordered(float f) {
#ifdef __CUDA_ARCH__
  uint32_t u = __float_as_uint(f);
#else
  union { float f; uint32_t u; } c; c.f = f; uint32_t u = c.u;
#endif
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ordered_to_f32(uint32_t k) {
  uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
// key = (ordered(value) << 32) | (0xFFFFFFFF - index): max over keys = max value, lowest index
__device__ __forceinline__ unsigned long long pool_key(float v, uint32_t t) {
  return (static_cast<unsigned long long>(f32_to_ordered(v)) << 32) |
         static_cast<unsigned long long>(0xFFFFFFFFu - t);
}

// ---- activations ----------------------------------------------------------------------------
__device__ __forceinline__ float act_apply(float x, int act) {
  if (act == CLSM_ACT_RELU) return fmaxf(x, 0.0f);
  if (act == CLSM_ACT_TANH) return tanhf(x);
  return x;
}
// derivative expressed through the activation OUTPUT y
__device__ __forceinline__ float act_grad_out(float y, int act) {
  if (act == CLSM_ACT_RELU) return y > 0.0f ? 1.0f : 0.0f;
  if (act == CLSM_ACT_TANH) return 1.0f - y * y;
  return 1.0f;
}

// ---- counter-based dropout bits (Philox4x32-10) -------------------------------------------------
// One call yields 128 random bits = 16 keep/drop decisions (one byte each: keep iff byte < thr).
// Counter = (row, 16-column block); key = seed.  A pure function, so the forward gather and the
// backward scatter regenerate the same mask without storing it.
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                       uint32_t k0, uint32_t k1, uint32_t out[4]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = static_cast<uint64_t>(M0) * c0;
    uint64_t p1 = static_cast<uint64_t>(M1) * c2;
    uint32_t hi0 = static_cast<uint32_t>(p0 >> 32), lo0 = static_cast<uint32_t>(p0);
    uint32_t hi1 = static_cast<uint32_t>(p1 >> 32), lo1 = static_cast<uint32_t>(p1);
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// threshold byte for keep probability (quantised to 1/256)
__host__ __device__ __forceinline__ uint32_t keep_threshold(float keep) {
  int t = static_cast<int>(keep * 256.0f + 0.5f);
  return t < 0 ? 0u : (t > 256 ? 256u : static_cast<uint32_t>(t));
}
// One element of Keras-1 Adam.  Every operation carries its rounding (.rn intrinsics are never contracted, by nvcc or
// by ptxas), so the dense pass, the data-parallel pass and the deferred-row replay (adam.cu) produce the same bits
// from the same inputs whatever code surrounds them.  sqrt.approx and the rcp-based divide are <= 2 ulp each: the
// dense pass is HBM-bound only if the ~600 M sqrt+divide pairs per step stay off the issue limit (the IEEE
// sequences cost ~40 instructions per element), and 2 ulp on a 1e-3 step is far inside the parity tolerance.
__device__ __forceinline__ void adam_step1(float& p, float& m, float& v, float g, float b1, float b2, float omb1, float omb2,
                                           float lr_t, float eps) {
  m = __fmaf_rn(b1, m, __fmul_rn(omb1, g));
  v = __fmaf_rn(b2, v, __fmul_rn(__fmul_rn(omb2, g), g));
  float s;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(s) : "f"(v));
  p = __fsub_rn(p, __fdividef(__fmul_rn(lr_t, m), __fadd_rn(s, eps)));
}
// inverted-dropout scale = 1 / the realised keep probability (thr / 256), so E[X] is unbiased for every keep
__host__ __device__ __forceinline__ float dropout_scale(float keep) {
  const uint32_t t = keep_threshold(keep);
  return t == 0 ? 0.0f : 256.0f / static_cast<float>(t);
}
// 16 decision bytes for (row, block of 16 columns)
__host__ __device__ __forceinline__ void dropout_bytes16(uint64_t seed, uint64_t row, uint32_t cblk,
                                                         uint32_t out[4]) {
  philox4x32_10(static_cast<uint32_t>(row), static_cast<uint32_t>(row >> 32), cblk, 0x434c534du,
                static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32), out);
}
__host__ __device__ __forceinline__ bool dropout_keep(const uint32_t bytes[4], int c_in_blk, uint32_t thr) {
  uint32_t b = (bytes[c_in_blk >> 2] >> ((c_in_blk & 3) * 8)) & 0xffu;
  return b < thr;
}
