// n-gram window convolution + max over time, CUDA-core implementation.
//   replaces Convolution1D(valid) + MaxPooling1D + Flatten, dssm_cnn_v2/cnn_dssm_tf.py:84-100.
// This is the fp32 (exact-parity) path and the cross-check for the tcgen05 kernel in conv_tc.cu.
// Implicit GEMM over the flat token matrix X (S*L, Dpad): output row r = (s, t) uses input rows
// r .. r+k-1, so the im2col matrix is never built.  Tile 64 rows x 64 filters, K step 16.
#include "common.cuh"

namespace {

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

constexpr int BM = 64, BN = 64, BK = 16;

template <typename T>
__global__ void __launch_bounds__(256)
conv_pool_simt_kernel(const T* __restrict__ X, int64_t n_rows, int L, int Dpad,
                      const T* __restrict__ Wt, int k, int F, int Fpad,
                      unsigned long long* __restrict__ keys, int ldk, int col0) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * BM;
  const int n0 = blockIdx.y * BN;
  const int K = k * Dpad;
  const int ty = tid / 16, tx = tid % 16;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  // loader mapping: 64 rows x 16 k-elements, 4 consecutive k per thread
  const int lm = tid / 4, lk = (tid % 4) * 4;
  for (int kk = 0; kk < K; kk += BK) {
    const int j = kk / Dpad, c = kk - j * Dpad;
    {
      const int64_t row = r0 + lm + j;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (row < n_rows) {
        const T* src = X + row * Dpad + c + lk;
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = to_f32(src[e]);
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) As[lk + e][lm] = v[e];
    }
    {
      const int f = n0 + lm;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (f < Fpad) {
        const T* src = Wt + static_cast<int64_t>(f) * K + kk + lk;
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = to_f32(src[e]);
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) Bs[lk + e][lm] = v[e];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < BK; ++q) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[q][ty * 4 + i];
#pragma unroll
      for (int jn = 0; jn < 4; ++jn) b[jn] = Bs[q][tx * 4 + jn];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int jn = 0; jn < 4; ++jn) acc[i][jn] = fmaf(a[i], b[jn], acc[i][jn]);
    }
    __syncthreads();
  }

  // epilogue: per column, fold this thread's 4 consecutive rows per sequence, then atomicMax.
  const int Tn = L - k + 1;
#pragma unroll
  for (int jn = 0; jn < 4; ++jn) {
    const int f = n0 + tx * 4 + jn;
    if (f >= F) continue;
    unsigned long long best = 0ull;
    int64_t best_s = -1;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t r = r0 + ty * 4 + i;
      if (r >= n_rows) break;
      const int64_t s = r / L;
      const int t = static_cast<int>(r - s * L);
      if (t >= Tn) continue;
      const unsigned long long key = pool_key(acc[i][jn] + 0.0f, static_cast<uint32_t>(t));
      if (s != best_s) {
        if (best_s >= 0) atomicMax(keys + best_s * ldk + col0 + f, best);
        best_s = s;
        best = key;
      } else if (key > best) {
        best = key;
      }
    }
    if (best_s >= 0) atomicMax(keys + best_s * ldk + col0 + f, best);
  }
}

__global__ void clear_keys_kernel(unsigned long long* keys, int64_t S, int F, int ldk, int col0) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= S * F) return;
  const int64_t s = i / F;
  const int f = static_cast<int>(i - s * F);
  keys[s * ldk + col0 + f] = 0ull;
}

__global__ void pool_finalize_kernel(const unsigned long long* __restrict__ keys, int64_t S, int P, int ldk,
                                     const float* __restrict__ bias, int act,
                                     float* __restrict__ pooled, int32_t* __restrict__ argmax, int ldp) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= S * P) return;
  const int64_t s = i / P;
  const int p = static_cast<int>(i - s * P);
  const unsigned long long key = keys[s * ldk + p];
  const float mx = ordered_to_f32(static_cast<uint32_t>(key >> 32));
  int32_t t = static_cast<int32_t>(0xFFFFFFFFu - static_cast<uint32_t>(key & 0xFFFFFFFFull));
  const float y = act_apply(mx + bias[p], act);
  if (act == CLSM_ACT_RELU && !(y > 0.0f)) t = 0;
  pooled[s * ldp + p] = y;
  if (argmax) argmax[s * ldp + p] = t;
}

}  // namespace

int clsm_conv_pool_fwd_tc(const void* X, int64_t S, int L, int D, int Dpad, const void* Wt, int k, int F, int Fpad,
                          uint64_t* keys, int ldk, int col0, cudaStream_t st);   // conv_tc.cu
int clsm_conv_pool_fwd_tok_tc(const int32_t* tok, const void* table, int64_t table_rows, int64_t S, int L, int D, int Dpad,
                              const void* Wt, int k, int F, int Fpad, uint64_t* keys, int ldk, int col0, cudaStream_t st);

extern "C" {

int clsm_conv_pool_fwd(const void* X, int dtype, int64_t S, int L, int Dpad,
                       const void* Wt, int k, int F, int Fpad,
                       uint64_t* keys, int ldk, int col0, int impl, void* stream) {
  return clsm_conv_pool_fwd_d(X, dtype, S, L, 0, Dpad, Wt, k, F, Fpad, keys, ldk, col0, impl, stream);
}

int clsm_conv_pool_fwd_d(const void* X, int dtype, int64_t S, int L, int D, int Dpad,
                         const void* Wt, int k, int F, int Fpad,
                         uint64_t* keys, int ldk, int col0, int impl, void* stream) {
  CLSM_REQUIRE(D >= 0 && D <= Dpad, "conv_pool_fwd: D must be 0 (unknown) or in [1, Dpad]");
  CLSM_REQUIRE(dtype == CLSM_BF16 || dtype == CLSM_F32, "conv_pool_fwd: bad dtype %d", dtype);
  CLSM_REQUIRE(S >= 0 && L >= k && k >= 1 && F >= 1 && Fpad >= F && Dpad % 16 == 0 && Dpad > 0,
               "conv_pool_fwd: bad sizes (S=%lld L=%d k=%d F=%d Fpad=%d Dpad=%d)", (long long)S, L, k, F, Fpad, Dpad);
  CLSM_REQUIRE(ldk >= col0 + F, "conv_pool_fwd: key row too short");
  if (S == 0) return CLSM_OK;
  cudaStream_t st = as_stream(stream);
  const int64_t n_rows = S * L;
  clear_keys_kernel<<<static_cast<unsigned>(ceil_div64(S * F, 256)), 256, 0, st>>>(
      reinterpret_cast<unsigned long long*>(keys), S, F, ldk, col0);
  CLSM_LAUNCH_CHECK();
  const bool tc_ok = dtype == CLSM_BF16 && Dpad % 64 == 0 && Fpad % 16 == 0 && Fpad <= 512;
  if (impl == CLSM_IMPL_TCGEN05 && !tc_ok) {
    clsm_set_error("conv_pool_fwd: tcgen05 path needs bf16, Dpad%%64==0, Fpad%%16==0, Fpad<=512");
    return CLSM_EUNSUPPORTED;
  }
  if (impl == CLSM_IMPL_TCGEN05 || (impl == CLSM_IMPL_AUTO && tc_ok))
    return clsm_conv_pool_fwd_tc(X, S, L, D, Dpad, Wt, k, F, Fpad, keys, ldk, col0, st);
  dim3 grid(static_cast<unsigned>(ceil_div64(n_rows, BM)), static_cast<unsigned>((F + BN - 1) / BN));
  if (dtype == CLSM_BF16)
    conv_pool_simt_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(X), n_rows, L, Dpad,
        static_cast<const __nv_bfloat16*>(Wt), k, F, Fpad, reinterpret_cast<unsigned long long*>(keys), ldk, col0);
  else
    conv_pool_simt_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(X), n_rows, L, Dpad,
        static_cast<const float*>(Wt), k, F, Fpad, reinterpret_cast<unsigned long long*>(keys), ldk, col0);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_conv_pool_fwd_tok(const int32_t* tok, const void* table, int64_t n_rows, int ld_table, int64_t S, int L, int D, int Dpad,
                           const void* Wt, int k, int F, int Fpad, uint64_t* keys, int ldk, int col0, void* stream) {
  CLSM_REQUIRE(D >= 0 && D <= Dpad, "conv_pool_fwd_tok: D must be 0 (unknown) or in [1, Dpad]");
  CLSM_REQUIRE(tok != nullptr && table != nullptr && n_rows > 0, "conv_pool_fwd_tok: null tokens / table");
  CLSM_REQUIRE(S >= 0 && L >= k && k >= 1 && F >= 1 && Fpad >= F, "conv_pool_fwd_tok: bad sizes");
  CLSM_REQUIRE(Dpad % 64 == 0 && ld_table == Dpad && Fpad % 16 == 0 && Fpad <= 512,
               "conv_pool_fwd_tok: needs a bf16 table with row pitch Dpad (a multiple of 64), Fpad %% 16 == 0, Fpad <= 512");
  CLSM_REQUIRE(ldk >= col0 + F, "conv_pool_fwd_tok: key row too short");
  if (S == 0) return CLSM_OK;
  cudaStream_t st = as_stream(stream);
  clear_keys_kernel<<<static_cast<unsigned>(ceil_div64(S * F, 256)), 256, 0, st>>>(
      reinterpret_cast<unsigned long long*>(keys), S, F, ldk, col0);
  CLSM_LAUNCH_CHECK();
  return clsm_conv_pool_fwd_tok_tc(tok, table, n_rows, S, L, D, Dpad, Wt, k, F, Fpad, keys, ldk, col0, st);
}

int clsm_pool_finalize(const uint64_t* keys, int64_t S, int P, int ldk, const float* bias,
                       int act, float* pooled, int32_t* argmax, int ldp, void* stream) {
  CLSM_REQUIRE(S >= 0 && P > 0 && ldk >= P && ldp >= P, "pool_finalize: bad sizes");
  if (S == 0) return CLSM_OK;
  pool_finalize_kernel<<<static_cast<unsigned>(ceil_div64(S * P, 256)), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const unsigned long long*>(keys), S, P, ldk, bias, act, pooled, argmax, ldp);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

}  // extern "C"
