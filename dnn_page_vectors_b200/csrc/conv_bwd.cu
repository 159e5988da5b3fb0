// Backward of Convolution1D + global max-pool, arg-max sparse form.
//   (gradient of dssm_cnn_v2/cnn_dssm_tf.py:84-100)
// Only the arg-max window of each (sequence, filter) receives gradient, so instead of the two
// dense-equivalent GEMMs (2x the forward FLOPs) this does S*F*k*D multiply-adds for each of dW
// and dX -- ~T (= L-k+1) times fewer.  Both kernels are bound by L1/L2 row reads, not FLOPs.
//   dW[j,c,f] = sum_s G[s,f] X[s, a(s,f)+j, c]      (dw kernel: lanes over c, registers over f,j)
//   dX[s,t,c] = sum_{f,j: a(s,f)+j = t} G[s,f] W[j,c,f]   (dx kernel: one warp per output row)
// Every sum is taken in a fixed order (s ascending / f ascending), no atomics: deterministic.
#include "common.cuh"

int clsm_colsum(const float* A, int64_t rows, int ld, int col0, int n, float* out, float* part,
                int n_chunks, cudaStream_t st);   // dense.cu

int clsm_conv_bwd_dw_fast(const void* X, int dtype, int L, int Dpad, int64_t S, int n_groups, const float* G,
                          const int32_t* A, int ldp, int col0, int F, int k, float* part, cudaStream_t st);   // conv_bwd_fast.cu
size_t clsm_conv_bwd_dx_fast_bytes(int64_t S, int L, int F);
int clsm_conv_bwd_dx_fast(const void* Wt, int dtype, int k, int Dpad, int L, int64_t S, const float* G, const int32_t* A,
                          int ldp, int col0, int F, float* dX, int accumulate, void* scratch, cudaStream_t st);

int clsm_conv_bwd_dw_tc(const void* X, int dtype, int L, int Dpad, int64_t S, int max_splits, const float* G,
                        const int32_t* A, int ldp, int col0, int F, int k, float* part, cudaStream_t st);          // conv_bwd_tc.cu
size_t clsm_conv_bwd_dx_tc_bytes(int k, int Dpad, int F);                                               // conv_bwd_tc.cu
int clsm_conv_bwd_dx_tc_tiling2_ok(int k, int Dpad, int F, int64_t n_rows);
int clsm_conv_bwd_dx_multi_tc_ok(int n_windows, const int* ks, int Dpad, int F, int64_t n_rows);
size_t clsm_conv_bwd_dx_multi_tc_bytes(int n_windows, const int* ks, int Dpad, int F);
int clsm_conv_bwd_dx_multi_tc(const void* const* Wt, const int* ks, int n_windows, int Dpad, int L, int64_t S, const float* G,
                              const int32_t* A, int ldp, int col0, int F, void* dX, int dx_bf16, void* scratch, cudaStream_t st);
int clsm_conv_bwd_dx_tc(const void* Wt, int dtype, int k, int Dpad, int L, int64_t S, const float* G, const int32_t* A,
                        int ldp, int col0, int F, float* dX, int accumulate, int dx_bf16, void* scratch, cudaStream_t st);

namespace {

constexpr int FPW = 4;            // filters per warp
constexpr int DW_WARPS = 32;      // warps per CTA -> 128 filters per CTA
constexpr int FT = FPW * DW_WARPS;

template <typename T> struct Pair;
template <> struct Pair<float> {
  static __device__ __forceinline__ float2 ld(const float* p) { return *reinterpret_cast<const float2*>(p); }
};
template <> struct Pair<__nv_bfloat16> {
  static __device__ __forceinline__ float2 ld(const __nv_bfloat16* p) {
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  }
};

// grid (n_groups, Dpad/64, ceil(F/FT)); part[g][f][j][c]
// Each warp owns FPW filters and one 64-column slice; lane owns 2 columns.  The (G, argmax) entries
// of 32 sequences are fetched in one go (lane l holds sequence s0+l), so the only dependent global
// loads left in the inner loop are the row reads themselves, which are independent of each other.
template <typename T, int KW>
__global__ void __launch_bounds__(DW_WARPS * 32, 1)
conv_bwd_dw_kernel(const T* __restrict__ X, int L, int Dpad, int64_t S, int64_t seq_per_group,
                   const float* __restrict__ G, const int32_t* __restrict__ A, int ldp, int col0, int F,
                   float* __restrict__ part) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int cbase = blockIdx.y * 64 + lane * 2;
  const int f0 = blockIdx.z * FT + warp * FPW;
  const int64_t s0 = static_cast<int64_t>(blockIdx.x) * seq_per_group;
  const int64_t s1 = (s0 + seq_per_group < S) ? s0 + seq_per_group : S;
  float2 acc[FPW][KW];
#pragma unroll
  for (int i = 0; i < FPW; ++i)
#pragma unroll
    for (int j = 0; j < KW; ++j) acc[i][j] = make_float2(0.f, 0.f);

  for (int64_t sb = s0; sb < s1; sb += 32) {
    float gv[FPW]; int av[FPW];
    const int64_t sl = sb + lane;
#pragma unroll
    for (int i = 0; i < FPW; ++i) {
      const bool ok = sl < s1 && f0 + i < F;
      gv[i] = ok ? __ldg(G + sl * ldp + col0 + f0 + i) : 0.0f;
      av[i] = ok ? __ldg(A + sl * ldp + col0 + f0 + i) : 0;
    }
    const int cnt = static_cast<int>((s1 - sb < 32) ? s1 - sb : 32);
    for (int q = 0; q < cnt; ++q) {
      const T* xs = X + (sb + q) * L * static_cast<int64_t>(Dpad) + cbase;
#pragma unroll
      for (int i = 0; i < FPW; ++i) {
        const float gi = __shfl_sync(0xffffffffu, gv[i], q);
        const int ai = __shfl_sync(0xffffffffu, av[i], q);
        if (gi != 0.0f) {
#pragma unroll
          for (int j = 0; j < KW; ++j) {
            const float2 x = Pair<T>::ld(xs + static_cast<int64_t>(ai + j) * Dpad);
            acc[i][j].x = fmaf(gi, x.x, acc[i][j].x);
            acc[i][j].y = fmaf(gi, x.y, acc[i][j].y);
          }
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < FPW; ++i) {
    const int f = f0 + i;
    if (f >= F) continue;
#pragma unroll
    for (int j = 0; j < KW; ++j) {
      float* dst = part + ((static_cast<int64_t>(blockIdx.x) * F + f) * KW + j) * Dpad + cbase;
      *reinterpret_cast<float2*>(dst) = acc[i][j];
    }
  }
}

// dW_ref[j][c][f] = sum_g part[g][f][j][c]
__global__ void conv_bwd_dw_reduce_kernel(const float* __restrict__ part, int n_groups, int k, int D, int Dpad, int F,
                                          float* __restrict__ dW) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t n = static_cast<int64_t>(k) * D * F;
  if (i >= n) return;
  // consecutive threads walk c (contiguous in part); dW writes are strided by F but tiny.
  const int c = static_cast<int>(i % D);
  const int64_t r = i / D;
  const int j = static_cast<int>(r % k);
  const int f = static_cast<int>(r / k);
  const int64_t stride = static_cast<int64_t>(F) * k * Dpad;
  const float* src = part + (static_cast<int64_t>(f) * k + j) * Dpad + c;
  float s = 0.0f;
  for (int g = 0; g < n_groups; ++g) s += src[g * stride];
  dW[(static_cast<int64_t>(j) * D + c) * F + f] = s;
}

// dX: one CTA per (sequence, chunk of DX_ROWS output rows).  The CTA first turns the sequence's F
// (arg-max, gradient) pairs into a bitmap  position -> set of filters  in shared memory (atomicOr:
// the RESULT is order independent), then each warp produces whole output rows: for row t and window
// offset j it walks the set bits of position t-j in ascending filter order and accumulates
// g * Wt[f, j, :] -- a fixed order, no floating-point atomics, every row written exactly once.
constexpr int DX_ROWS = 128;
constexpr int DX_MAXF = 512;                    // filters per window supported by the bitmap (16 words)
constexpr int DX_WORDS = DX_MAXF / 32;

template <typename T>
__global__ void __launch_bounds__(256)
conv_bwd_dx_kernel(const T* __restrict__ Wt, int k, int Dpad, int L, int chunks_per_seq,
                   const float* __restrict__ G, const int32_t* __restrict__ A, int ldp, int col0, int F,
                   float* __restrict__ dX, int accumulate) {
  __shared__ float s_g[DX_MAXF];
  __shared__ unsigned s_bits[(DX_ROWS + 8) * DX_WORDS];     // positions t0-(k-1) .. t0+DX_ROWS-1, k <= 8
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t s = blockIdx.x / chunks_per_seq;
  const int t0 = (blockIdx.x % chunks_per_seq) * DX_ROWS;
  const int nwords = (F + 31) / 32;
  const int pos_lo = t0 - (k - 1);                           // first window start that can touch this chunk
  const int npos = DX_ROWS + k - 1;
  for (int i = threadIdx.x; i < npos * nwords; i += blockDim.x) s_bits[i] = 0u;
  __syncthreads();
  for (int f = threadIdx.x; f < F; f += blockDim.x) {
    const float g = __ldg(G + s * ldp + col0 + f);
    const int a = __ldg(A + s * ldp + col0 + f);
    s_g[f] = g;
    const int p = a - pos_lo;
    if (g != 0.0f && p >= 0 && p < npos) atomicOr(&s_bits[p * nwords + (f >> 5)], 1u << (f & 31));
  }
  __syncthreads();
  const int np = Dpad / 64;
  const int64_t K = static_cast<int64_t>(k) * Dpad;
  const int rows = (L - t0 < DX_ROWS) ? L - t0 : DX_ROWS;
  for (int r = warp; r < rows; r += 8) {
    float2 acc[8];
#pragma unroll
    for (int p = 0; p < 8; ++p) acc[p] = make_float2(0.f, 0.f);
    for (int j = 0; j < k; ++j) {
      const unsigned* bits = s_bits + (r + (k - 1) - j) * nwords;      // position t - j
      for (int w = 0; w < nwords; ++w) {
        unsigned m = bits[w];
        while (m) {
          const int f = w * 32 + __ffs(m) - 1;
          m &= m - 1;
          const float g = s_g[f];
          const T* wrow = Wt + static_cast<int64_t>(f) * K + static_cast<int64_t>(j) * Dpad + lane * 2;
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            if (p < np) {
              const float2 x = Pair<T>::ld(wrow + p * 64);
              acc[p].x = fmaf(g, x.x, acc[p].x);
              acc[p].y = fmaf(g, x.y, acc[p].y);
            }
          }
        }
      }
    }
    float* dst = dX + (s * L + t0 + r) * Dpad + lane * 2;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      if (p < np) {
        float2 v = acc[p];
        if (accumulate) { const float2 o = *reinterpret_cast<const float2*>(dst + p * 64); v.x += o.x; v.y += o.y; }
        *reinterpret_cast<float2*>(dst + p * 64) = v;
      }
    }
  }
}

int dw_groups(int64_t S, int Dpad, int F) {
  // sized for the coarser of the two dW tilings (generic: 128 filters / CTA, staged: 160) so that
  // both fit the same partial-sum workspace
  const int per = (Dpad / 64) * ((F + 160 - 1) / 160);
  int64_t g = (4 * static_cast<int64_t>(clsm_sm_count()) + per - 1) / per;
  if (g > S) g = S;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}
const int kColsumChunks = 64;

template <typename T>
int launch_dw(const T* X, int L, int Dpad, int64_t S, int64_t spg, const float* G, const int32_t* A, int ldp,
              int col0, int F, int k, float* part, dim3 grid, cudaStream_t st) {
  switch (k) {
    case 1: conv_bwd_dw_kernel<T, 1><<<grid, DW_WARPS * 32, 0, st>>>(X, L, Dpad, S, spg, G, A, ldp, col0, F, part); break;
    case 2: conv_bwd_dw_kernel<T, 2><<<grid, DW_WARPS * 32, 0, st>>>(X, L, Dpad, S, spg, G, A, ldp, col0, F, part); break;
    case 3: conv_bwd_dw_kernel<T, 3><<<grid, DW_WARPS * 32, 0, st>>>(X, L, Dpad, S, spg, G, A, ldp, col0, F, part); break;
    case 4: conv_bwd_dw_kernel<T, 4><<<grid, DW_WARPS * 32, 0, st>>>(X, L, Dpad, S, spg, G, A, ldp, col0, F, part); break;
    case 5: conv_bwd_dw_kernel<T, 5><<<grid, DW_WARPS * 32, 0, st>>>(X, L, Dpad, S, spg, G, A, ldp, col0, F, part); break;
    case 6: conv_bwd_dw_kernel<T, 6><<<grid, DW_WARPS * 32, 0, st>>>(X, L, Dpad, S, spg, G, A, ldp, col0, F, part); break;
    case 7: conv_bwd_dw_kernel<T, 7><<<grid, DW_WARPS * 32, 0, st>>>(X, L, Dpad, S, spg, G, A, ldp, col0, F, part); break;
    default: return CLSM_EUNSUPPORTED;
  }
  return CLSM_OK;
}

}  // namespace

extern "C" {

size_t clsm_conv_bwd_workspace_bytes(int64_t S, int L, int k, int D, int F) {
  const int Dpad = (D + 63) / 64 * 64;
  const size_t part = align_up(static_cast<size_t>(dw_groups(S, Dpad, F)) * F * k * Dpad * 4, 256);
  const size_t cs = align_up(static_cast<size_t>(kColsumChunks) * F * 4, 256);
  return part + cs + clsm_conv_bwd_dx_fast_bytes(S, L, F) + clsm_conv_bwd_dx_tc_bytes(k, Dpad, F);
}

int clsm_conv_bwd_dx_bf16_ok(int64_t S, int L, int k, int D, int F) {
  const int Dpad = (D + 63) / 64 * 64;
  return (k >= 1 && k <= 7 && Dpad <= 512) ? clsm_conv_bwd_dx_tc_tiling2_ok(k, Dpad, F, S * L) : 0;
}

int clsm_conv_pool_bwd(const void* X, int dtype, int64_t S, int L, int D, int Dpad,
                       const void* Wt, int k, int F, int Fpad,
                       const float* G, const int32_t* argmax, int ldp, int col0,
                       float* dW_ref, float* db, float* dX, int first_window,
                       void* ws, size_t ws_bytes, void* stream) {
  return clsm_conv_pool_bwd_ex(X, dtype, S, L, D, Dpad, Wt, k, F, Fpad, G, argmax, ldp, col0, dW_ref, db, dX, CLSM_F32,
                               first_window, ws, ws_bytes, stream);
}

int clsm_conv_pool_bwd_ex(const void* X, int dtype, int64_t S, int L, int D, int Dpad,
                          const void* Wt, int k, int F, int Fpad,
                          const float* G, const int32_t* argmax, int ldp, int col0,
                          float* dW_ref, float* db, void* dX_any, int dx_dtype, int first_window,
                          void* ws, size_t ws_bytes, void* stream) {
  return clsm_conv_pool_bwd_parts(X, dtype, S, L, D, Dpad, Wt, k, F, Fpad, G, argmax, ldp, col0, dW_ref, db, dX_any, dx_dtype,
                                  first_window, CLSM_BWD_DW | CLSM_BWD_DX, ws, ws_bytes, stream);
}

// The two gradients read the same inputs, write different outputs and use disjoint workspace regions, so a caller may
// issue them as two calls on two streams (parts = CLSM_BWD_DX on one, CLSM_BWD_DW on the other).
int clsm_conv_pool_bwd_parts(const void* X, int dtype, int64_t S, int L, int D, int Dpad,
                             const void* Wt, int k, int F, int Fpad,
                             const float* G, const int32_t* argmax, int ldp, int col0,
                             float* dW_ref, float* db, void* dX_any, int dx_dtype, int first_window, int parts,
                             void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE((parts & ~(CLSM_BWD_DW | CLSM_BWD_DX)) == 0 && parts != 0, "conv_pool_bwd: parts must be CLSM_BWD_DW, CLSM_BWD_DX or both");
  CLSM_REQUIRE(dtype == CLSM_BF16 || dtype == CLSM_F32, "conv_pool_bwd: bad dtype");
  CLSM_REQUIRE(dx_dtype == CLSM_F32 || dx_dtype == CLSM_BF16, "conv_pool_bwd: dX must be fp32 or bf16");
  const int dx_bf16 = dx_dtype == CLSM_BF16 ? 1 : 0;
  if (dx_bf16 && !(dtype == CLSM_BF16 && first_window != 0 && clsm_conv_bwd_dx_bf16_ok(S, L, k, D, F))) {
    clsm_set_error("conv_pool_bwd: bf16 dX rows need the bf16 tensor-core path, first_window = 1 and a shape "
                   "clsm_conv_bwd_dx_bf16_ok() accepts");
    return CLSM_EUNSUPPORTED;
  }
  float* dX = static_cast<float*>(dX_any);
  CLSM_REQUIRE(S > 0 && L >= k && k >= 1 && k <= 7 && F >= 1 && Fpad >= F && F <= DX_MAXF, "conv_pool_bwd: bad sizes (k must be 1..7, F <= 512)");
  CLSM_REQUIRE(S * ((L + DX_ROWS - 1) / DX_ROWS) < (1ll << 31), "conv_pool_bwd: too many sequences for one launch");
  CLSM_REQUIRE(Dpad == (D + 63) / 64 * 64 && Dpad <= 512, "conv_pool_bwd: Dpad must be D rounded up to 64 and <= 512");
  CLSM_REQUIRE(ldp >= col0 + F, "conv_pool_bwd: G row too short");
  if (ws_bytes < clsm_conv_bwd_workspace_bytes(S, L, k, D, F)) {
    clsm_set_error("conv_pool_bwd: workspace too small (%zu < %zu)", ws_bytes, clsm_conv_bwd_workspace_bytes(S, L, k, D, F));
    return CLSM_EWORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  const int n_groups = dw_groups(S, Dpad, F);
  const int64_t spg = ceil_div64(S, n_groups);
  char* w = static_cast<char*>(ws);
  float* part = reinterpret_cast<float*>(w);
  w += align_up(static_cast<size_t>(n_groups) * F * k * Dpad * 4, 256);
  float* cs = reinterpret_cast<float*>(w);
  w += align_up(static_cast<size_t>(kColsumChunks) * F * 4, 256);
  void* sort_scratch = w;

  int rc = CLSM_OK;
  if (parts & CLSM_BWD_DW) {
  // dW: tensor cores (bf16 mode) -> shared-memory staged CUDA cores -> generic
  int n_parts = n_groups;
  rc = clsm_conv_bwd_dw_tc(X, dtype, L, Dpad, S, n_groups, G, argmax, ldp, col0, F, k, part, st);
  if (rc > 0) { n_parts = rc; rc = CLSM_OK; }
  else if (rc == CLSM_EUNSUPPORTED) rc = clsm_conv_bwd_dw_fast(X, dtype, L, Dpad, S, n_groups, G, argmax, ldp, col0, F, k, part, st);
  if (rc == CLSM_EUNSUPPORTED) {            // shape outside the shared-memory plan: generic kernel
    dim3 grid(n_groups, Dpad / 64, (F + FT - 1) / FT);
    rc = dtype == CLSM_BF16
        ? launch_dw<__nv_bfloat16>(static_cast<const __nv_bfloat16*>(X), L, Dpad, S, spg, G, argmax, ldp, col0, F, k, part, grid, st)
        : launch_dw<float>(static_cast<const float*>(X), L, Dpad, S, spg, G, argmax, ldp, col0, F, k, part, grid, st);
    if (rc != CLSM_OK) { clsm_set_error("conv_pool_bwd: unsupported window %d", k); return rc; }
    CLSM_LAUNCH_CHECK();
  } else if (rc != CLSM_OK) {
    return rc;
  }
  if (g_clsm_marks[0] != nullptr) { CLSM_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(g_clsm_marks[0]), st)); g_clsm_marks[0] = nullptr; }
  const int64_t ne = static_cast<int64_t>(k) * D * F;
  conv_bwd_dw_reduce_kernel<<<static_cast<unsigned>(ceil_div64(ne, 256)), 256, 0, st>>>(part, n_parts, k, D, Dpad, F, dW_ref);
  CLSM_LAUNCH_CHECK();
  if (clsm_colsum(G, S, ldp, col0, F, db, cs, kColsumChunks, st) != CLSM_OK) { clsm_set_error("conv_pool_bwd: colsum launch failed"); return CLSM_ECUDA; }
  }
  if (!(parts & CLSM_BWD_DW)) g_clsm_marks[0] = nullptr;      // diagnostics marks of a part that does not run are dropped, not kept
  if (!(parts & CLSM_BWD_DX)) { g_clsm_marks[1] = nullptr; return CLSM_OK; }

  // dX: tensor cores (bf16 mode) -> shared-memory staged CUDA cores -> generic
  void* wb_scratch = static_cast<char*>(sort_scratch) + clsm_conv_bwd_dx_fast_bytes(S, L, F);
  if (g_clsm_marks[1] != nullptr) { CLSM_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(g_clsm_marks[1]), st)); g_clsm_marks[1] = nullptr; }
  rc = clsm_conv_bwd_dx_tc(Wt, dtype, k, Dpad, L, S, G, argmax, ldp, col0, F, dX, first_window ? 0 : 1, dx_bf16, wb_scratch, st);
  if (dx_bf16 && rc == CLSM_EUNSUPPORTED) { clsm_set_error("conv_pool_bwd: bf16 dX rows requested but the tensor-core dX kernel refused the shape"); return rc; }
  if (rc == CLSM_OK) return CLSM_OK;
  if (rc != CLSM_EUNSUPPORTED) return rc;
  rc = clsm_conv_bwd_dx_fast(Wt, dtype, k, Dpad, L, S, G, argmax, ldp, col0, F, dX, first_window ? 0 : 1, sort_scratch, st);
  if (rc == CLSM_OK) return CLSM_OK;
  if (rc != CLSM_EUNSUPPORTED) return rc;
  const int cps = (L + DX_ROWS - 1) / DX_ROWS;
  const unsigned gx = static_cast<unsigned>(S * cps);
  if (dtype == CLSM_BF16)
    conv_bwd_dx_kernel<__nv_bfloat16><<<gx, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(Wt), k, Dpad, L, cps, G, argmax, ldp, col0, F, dX, first_window ? 0 : 1);
  else
    conv_bwd_dx_kernel<float><<<gx, 256, 0, st>>>(static_cast<const float*>(Wt), k, Dpad, L, cps, G, argmax, ldp, col0, F, dX, first_window ? 0 : 1);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

/* ---- dX of several window sizes in one pass (bf16 tensor-core path; conv_bwd_tc.cu) ---- */
int clsm_conv_bwd_dx_multi_ok(int64_t S, int L, int n_windows, const int* ks, int D, int F) {
  if (n_windows < 2 || ks == nullptr || S <= 0 || D <= 0) return 0;
  for (int w = 0; w < n_windows; ++w) if (ks[w] < 1 || L < ks[w]) return 0;
  return clsm_conv_bwd_dx_multi_tc_ok(n_windows, ks, (D + 63) / 64 * 64, F, S * L);
}

size_t clsm_conv_bwd_dx_multi_workspace_bytes(int n_windows, const int* ks, int D, int F) {
  if (n_windows < 1 || ks == nullptr) return 0;
  return clsm_conv_bwd_dx_multi_tc_bytes(n_windows, ks, (D + 63) / 64 * 64, F);
}

int clsm_conv_pool_bwd_dx_multi(int64_t S, int L, int D, int Dpad, const void* const* Wt, const int* ks, int n_windows, int F,
                                const float* G, const int32_t* argmax, int ldp, int col0, void* dX, int dx_dtype,
                                void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(Wt != nullptr && ks != nullptr && n_windows >= 2 && n_windows <= 4, "conv_pool_bwd_dx_multi: 2..4 windows");
  CLSM_REQUIRE(dx_dtype == CLSM_F32 || dx_dtype == CLSM_BF16, "conv_pool_bwd_dx_multi: dX must be fp32 or bf16");
  CLSM_REQUIRE(S > 0 && Dpad == (D + 63) / 64 * 64 && ldp >= col0 + n_windows * F, "conv_pool_bwd_dx_multi: bad sizes");
  for (int w = 0; w < n_windows; ++w) CLSM_REQUIRE(ks[w] >= 1 && ks[w] <= 7 && L >= ks[w], "conv_pool_bwd_dx_multi: windows must be 1..7 and <= L");
  if (ws_bytes < clsm_conv_bwd_dx_multi_workspace_bytes(n_windows, ks, D, F)) {
    clsm_set_error("conv_pool_bwd_dx_multi: workspace too small");
    return CLSM_EWORKSPACE;
  }
  const int rc = clsm_conv_bwd_dx_multi_tc(Wt, ks, n_windows, Dpad, L, S, G, argmax, ldp, col0, F, dX, dx_dtype == CLSM_BF16 ? 1 : 0, ws, as_stream(stream));
  if (rc == CLSM_EUNSUPPORTED) clsm_set_error("conv_pool_bwd_dx_multi: shape outside the kernel's plan (ask clsm_conv_bwd_dx_multi_ok first)");
  return rc;
}

}  // extern "C"
