// Keras-1 Adam (optimizer="adam", dssm_cnn_v2/cnn_dssm_tf.py:182) as fused HBM-bound passes.
//   t = step (1-based); lr_t = lr*sqrt(1-b2^t)/(1-b1^t)
//   m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; p -= lr_t m / (sqrt(v) + eps)
// adam_table walks the WHOLE embedding table every step (Keras updates every row; rows without
// gradient still move by momentum) but takes the gradient row-sparsely, so no (V, D) gradient
// is ever materialised: 24 B/param of p,m,v traffic (+2 B for the bf16 compute copy).
#include "common.cuh"
#include <math.h>

namespace {

struct AdamC { float b1, b2, omb1, omb2, lr_t, eps; };

// The step uses the hardware approximations (sqrt.approx, rcp-based divide: <= 2 ulp each).  The pass is
// HBM-bound only if the ~600 M sqrt+divide pairs per step do not become the issue bottleneck; the IEEE
// sequences cost ~40 instructions per element.  2 ulp on a 1e-3 step is far inside the parity tolerance.
__device__ __forceinline__ void adam1(float& p, float& m, float& v, float g, const AdamC& c) {
  m = c.b1 * m + c.omb1 * g;
  v = c.b2 * v + c.omb2 * g * g;
  float s;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(s) : "f"(v));
  p = p - __fdividef(c.lr_t * m, s + c.eps);
}

__global__ void __launch_bounds__(256)
adam_dense_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                  const float* __restrict__ g, int64_t n, AdamC c) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float pi = p[i], mi = m[i], vi = v[i];
  adam1(pi, mi, vi, g[i], c);
  p[i] = pi; m[i] = mi; v[i] = vi;
}

// D % 4 == 0: one thread = one float4 of one row.  rows_sel == nullptr: all rows (dense Keras
// semantics, gradient looked up through row_of_id); else only rows uniq[0:n_uniq] (lazy Adam).
template <bool VEC4>
__global__ void __launch_bounds__(256)
adam_table_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                  int64_t n_rows, int D, const int32_t* __restrict__ row_of_id,
                  const int32_t* __restrict__ uniq, const int32_t* __restrict__ n_uniq,
                  const float* __restrict__ dE, int ldg, AdamC c,
                  __nv_bfloat16* __restrict__ copy, int ld_copy) {
  constexpr int W = VEC4 ? 4 : 1;
  const int per_row = D / W;
  const int64_t item = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  int64_t r;
  if (item < 0xffffffffll) r = static_cast<uint32_t>(item) / static_cast<uint32_t>(per_row);   // 32-bit divide: ~4x cheaper
  else r = item / per_row;
  const int q = static_cast<int>(item - r * per_row);
  int64_t grow;
  if (uniq != nullptr) {
    if (r >= n_uniq[0]) return;
    grow = r;
    r = uniq[r];
  } else {
    if (r >= n_rows) return;
    grow = row_of_id[r];
  }
  const int64_t off = r * D + q * W;
  if (VEC4) {
    // (streaming ld.cs / st.cs hints on p/m/v were measured: 1.23 -> 1.34 ms per table, so plain accesses stay)
    float4 pi = *reinterpret_cast<float4*>(p + off);
    float4 mi = *reinterpret_cast<float4*>(m + off);
    float4 vi = *reinterpret_cast<float4*>(v + off);
    float4 gi = make_float4(0.f, 0.f, 0.f, 0.f);
    if (grow >= 0) gi = __ldg(reinterpret_cast<const float4*>(dE + grow * ldg + q * 4));
    adam1(pi.x, mi.x, vi.x, gi.x, c); adam1(pi.y, mi.y, vi.y, gi.y, c);
    adam1(pi.z, mi.z, vi.z, gi.z, c); adam1(pi.w, mi.w, vi.w, gi.w, c);
    *reinterpret_cast<float4*>(p + off) = pi;
    *reinterpret_cast<float4*>(m + off) = mi;
    *reinterpret_cast<float4*>(v + off) = vi;
    if (copy) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pi.x, pi.y), hi = __floats2bfloat162_rn(pi.z, pi.w);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(copy + r * ld_copy + q * 4) = pk;
    }
  } else {
    float pi = p[off], mi = m[off], vi = v[off];
    const float gi = grow >= 0 ? dE[grow * ldg + q] : 0.0f;
    adam1(pi, mi, vi, gi, c);
    p[off] = pi; m[off] = mi; v[off] = vi;
    if (copy) copy[r * ld_copy + q] = __float2bfloat16_rn(pi);
  }
}

AdamC make_c(int step, float lr, float b1, float b2, float eps) {
  AdamC c;
  c.b1 = b1; c.b2 = b2; c.omb1 = 1.0f - b1; c.omb2 = 1.0f - b2; c.eps = eps;
  const double t = static_cast<double>(step);
  c.lr_t = static_cast<float>(static_cast<double>(lr) * sqrt(1.0 - pow(static_cast<double>(b2), t)) /
                              (1.0 - pow(static_cast<double>(b1), t)));
  return c;
}

}  // namespace

extern "C" {

int clsm_adam_dense(float* p, float* m, float* v, const float* g, int64_t n, int step,
                    float lr, float b1, float b2, float eps, void* stream) {
  CLSM_REQUIRE(n >= 0 && step >= 1, "adam_dense: bad n/step");
  if (n == 0) return CLSM_OK;
  adam_dense_kernel<<<static_cast<unsigned>(ceil_div64(n, 256)), 256, 0, as_stream(stream)>>>(p, m, v, g, n, make_c(step, lr, b1, b2, eps));
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_adam_table(float* p, float* m, float* v, int64_t n_rows, int D,
                    const int32_t* row_of_id, const float* dE_rows, int ldg, int step,
                    float lr, float b1, float b2, float eps,
                    void* bf16_copy, int ld_copy, void* stream) {
  CLSM_REQUIRE(n_rows > 0 && D > 0 && step >= 1 && ldg >= D, "adam_table: bad sizes");
  CLSM_REQUIRE(bf16_copy == nullptr || ld_copy >= D, "adam_table: ld_copy < D");
  const AdamC c = make_c(step, lr, b1, b2, eps);
  cudaStream_t st = as_stream(stream);
  __nv_bfloat16* cp = static_cast<__nv_bfloat16*>(bf16_copy);
  if (D % 4 == 0 && ldg % 4 == 0 && (cp == nullptr || ld_copy % 4 == 0)) {
    const int64_t items = n_rows * (D / 4);
    adam_table_kernel<true><<<static_cast<unsigned>(ceil_div64(items, 256)), 256, 0, st>>>(p, m, v, n_rows, D, row_of_id, nullptr, nullptr, dE_rows, ldg, c, cp, ld_copy);
  } else {
    const int64_t items = n_rows * D;
    adam_table_kernel<false><<<static_cast<unsigned>(ceil_div64(items, 256)), 256, 0, st>>>(p, m, v, n_rows, D, row_of_id, nullptr, nullptr, dE_rows, ldg, c, cp, ld_copy);
  }
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_adam_rows(float* p, float* m, float* v, int D, const int32_t* uniq_ids,
                   const int32_t* n_uniq, int64_t max_uniq, const float* dE_rows, int ldg, int step,
                   float lr, float b1, float b2, float eps, void* bf16_copy, int ld_copy, void* stream) {
  CLSM_REQUIRE(D > 0 && step >= 1 && max_uniq > 0 && ldg >= D, "adam_rows: bad sizes");
  const AdamC c = make_c(step, lr, b1, b2, eps);
  cudaStream_t st = as_stream(stream);
  __nv_bfloat16* cp = static_cast<__nv_bfloat16*>(bf16_copy);
  if (D % 4 == 0 && ldg % 4 == 0 && (cp == nullptr || ld_copy % 4 == 0)) {
    const int64_t items = max_uniq * (D / 4);
    adam_table_kernel<true><<<static_cast<unsigned>(ceil_div64(items, 256)), 256, 0, st>>>(p, m, v, 0, D, nullptr, uniq_ids, n_uniq, dE_rows, ldg, c, cp, ld_copy);
  } else {
    const int64_t items = max_uniq * D;
    adam_table_kernel<false><<<static_cast<unsigned>(ceil_div64(items, 256)), 256, 0, st>>>(p, m, v, 0, D, nullptr, uniq_ids, n_uniq, dE_rows, ldg, c, cp, ld_copy);
  }
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

}  // extern "C"
