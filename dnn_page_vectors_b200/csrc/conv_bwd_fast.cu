// Shared-memory staged variants of the arg-max sparse conv backward (see conv_bwd.cu for the math).
//
//  dW : the CTA's 64-column slice of the CURRENT sequence's token rows is prefetched into shared
//       memory with cp.async (4-deep ring), so the inner loop has no global-latency dependence: per
//       live (sequence, filter) pair a warp does k 128-byte shared loads and 2k FMAs per lane.
//  dX : a pre-pass sorts each sequence's live filters by arg-max position (counting sort, ascending
//       filter inside a position -> a fixed order).  The main kernel keeps its 64-column slice of
//       ALL filter weights resident in shared memory (F*k*128 B) for the CTA's lifetime and walks,
//       for output row t, the buckets of positions t, t-1, .., t-k+1.  One global read of the sorted
//       lists and one write of dX per row; no atomics, deterministic.
// Both return CLSM_EUNSUPPORTED when a shape does not fit their shared-memory plan; conv_bwd.cu then
// uses its generic kernels.
#include "common.cuh"

namespace {

constexpr int FPW = 5;                       // filters per warp
constexpr int DW_WARPS = 32;
constexpr int FT = FPW * DW_WARPS;            // 160 filters per CTA
constexpr int DW_STAGES = 4;
constexpr int SMEM_BUDGET = 200 * 1024;

template <typename T> struct Pair;
template <> struct Pair<float> {
  static __device__ __forceinline__ float2 ld(const float* p) { return *reinterpret_cast<const float2*>(p); }
};
template <> struct Pair<__nv_bfloat16> {
  static __device__ __forceinline__ float2 ld(const __nv_bfloat16* p) {
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  }
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ------------------------------------------------------------------------------------------ dW
// grid (n_groups, Dpad/64, ceil(F/FT)); part[g][f][j][c].  Stage = X slice tile (L x 64) | g[FT] | a[FT]
template <typename T, int KW>
__global__ void __launch_bounds__(DW_WARPS * 32, 1)
dw_fast_kernel(const T* __restrict__ X, int L, int Dpad, int64_t S, int64_t seq_per_group,
               const float* __restrict__ G, const int32_t* __restrict__ A, int ldp, int col0, int F,
               float* __restrict__ part) {
  extern __shared__ __align__(16) uint8_t dw_smem[];
  constexpr int ROWB = 64 * sizeof(T);            // bytes per row slice
  constexpr int CPR = ROWB / 16;                  // 16-byte chunks per row
  const int tile_bytes = L * ROWB;
  const int stage_bytes = tile_bytes + FT * 8;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int c0 = blockIdx.y * 64;
  const int ft0 = blockIdx.z * FT;
  const int nf = (F - ft0 < FT) ? F - ft0 : FT;   // filters of this CTA
  const int64_t s0 = static_cast<int64_t>(blockIdx.x) * seq_per_group;
  const int64_t s1 = (s0 + seq_per_group < S) ? s0 + seq_per_group : S;
  const int n_units = static_cast<int>(s1 > s0 ? s1 - s0 : 0);

  auto issue = [&](int u) {
    if (u < n_units) {
      uint8_t* dst = dw_smem + (u % DW_STAGES) * stage_bytes;
      const uint8_t* src = reinterpret_cast<const uint8_t*>(X + ((s0 + u) * L) * static_cast<int64_t>(Dpad) + c0);
      const int chunks = L * CPR;
      for (int i = threadIdx.x; i < chunks; i += DW_WARPS * 32) {
        const int row = i / CPR, pc = i - row * CPR;
        cp_async16(dst + row * ROWB + pc * 16, src + static_cast<int64_t>(row) * Dpad * sizeof(T) + pc * 16);
      }
      if (threadIdx.x < nf) {
        cp_async4(dst + tile_bytes + threadIdx.x * 4, G + (s0 + u) * ldp + col0 + ft0 + threadIdx.x);
        cp_async4(dst + tile_bytes + FT * 4 + threadIdx.x * 4, A + (s0 + u) * ldp + col0 + ft0 + threadIdx.x);
      }
    }
    cp_async_commit();
  };

  float2 acc[FPW][KW];
#pragma unroll
  for (int i = 0; i < FPW; ++i)
#pragma unroll
    for (int j = 0; j < KW; ++j) acc[i][j] = make_float2(0.f, 0.f);

#pragma unroll
  for (int u = 0; u < DW_STAGES - 1; ++u) issue(u);
  const int fw0 = warp * FPW;                      // first filter of this warp inside the CTA tile
  for (int u = 0; u < n_units; ++u) {
    cp_async_wait<DW_STAGES - 2>();
    __syncthreads();                               // stage u landed; everyone is done with stage u-1
    issue(u + DW_STAGES - 1);                      // refills the slot of stage u-1
    const uint8_t* st = dw_smem + (u % DW_STAGES) * stage_bytes;
    const T* tile = reinterpret_cast<const T*>(st) + lane * 2;
    const float* sg = reinterpret_cast<const float*>(st + tile_bytes);
    const int* sa = reinterpret_cast<const int*>(st + tile_bytes + FT * 4);
#pragma unroll
    for (int i = 0; i < FPW; ++i) {
      if (fw0 + i < nf) {
        const float gi = sg[fw0 + i];
        if (gi != 0.0f) {
          const T* row = tile + sa[fw0 + i] * 64;
#pragma unroll
          for (int j = 0; j < KW; ++j) {
            const float2 x = Pair<T>::ld(row + j * 64);
            acc[i][j].x = fmaf(gi, x.x, acc[i][j].x);
            acc[i][j].y = fmaf(gi, x.y, acc[i][j].y);
          }
        }
      }
    }
  }
  cp_async_wait<0>();
#pragma unroll
  for (int i = 0; i < FPW; ++i) {
    const int f = ft0 + fw0 + i;
    if (fw0 + i >= nf) continue;
#pragma unroll
    for (int j = 0; j < KW; ++j) {
      float* dst = part + ((static_cast<int64_t>(blockIdx.x) * F + f) * KW + j) * Dpad + c0 + lane * 2;
      *reinterpret_cast<float2*>(dst) = acc[i][j];
    }
  }
}

// ------------------------------------------------------------------------------------------ sort pre-pass
// one warp per sequence: ord[s][0:n_live] = {(f*k - p)*128, g} of the live filters sorted by (arg-max position p,
// filter): output row t reads weight row f*k + (t - p), i.e. slice byte offset entry.x + t*128;
// bstart[s][p] = first entry of position p, p in [0, L]  (bstart[s][L] = n_live)
constexpr int SORT_WARPS = 4;
constexpr int SORT_MAXL = 1024;

__global__ void __launch_bounds__(SORT_WARPS * 32)
pool_sort_kernel(const float* __restrict__ G, const int32_t* __restrict__ A, int64_t S, int ldp, int col0, int F, int L, int k,
                 uint2* __restrict__ ord, int ldo, uint16_t* __restrict__ bstart, int ldb) {
  __shared__ int s_cnt[SORT_WARPS][SORT_MAXL + 1];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t s = static_cast<int64_t>(blockIdx.x) * SORT_WARPS + warp;
  if (s >= S) return;
  int* cnt = s_cnt[warp];
  for (int p = lane; p <= L; p += 32) cnt[p] = 0;
  __syncwarp();
  const float* g = G + s * ldp + col0;
  const int32_t* a = A + s * ldp + col0;
  for (int fb = 0; fb < F; fb += 32) {
    const int f = fb + lane;
    if (f < F && __ldg(g + f) != 0.0f) atomicAdd(&cnt[__ldg(a + f)], 1);
  }
  __syncwarp();
  // exclusive scan over positions; cnt[p] becomes the running cursor of bucket p
  int carry = 0;
  for (int pb = 0; pb <= L; pb += 32) {
    const int p = pb + lane;
    const int c = p <= L ? cnt[p] : 0;
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += n; }
    const int excl = carry + incl - c;
    if (p <= L) { cnt[p] = excl; bstart[s * ldb + p] = static_cast<uint16_t>(excl); }
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  __syncwarp();
  for (int fb = 0; fb < F; fb += 32) {
    const int f = fb + lane;
    const bool live = f < F && __ldg(g + f) != 0.0f;
    const int pos = live ? __ldg(a + f) : -1 - lane;               // dead lanes get unique keys
    const unsigned peers = __match_any_sync(0xffffffffu, pos);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    int base = 0;
    if (live) base = cnt[pos];
    __syncwarp();
    if (live && rank == 0) cnt[pos] = base + __popc(peers);
    __syncwarp();
    if (live) ord[s * ldo + base + rank] = make_uint2(static_cast<uint32_t>((f * k - pos) * 128), __float_as_uint(__ldg(g + f)));
  }
}

// ------------------------------------------------------------------------------------------ dX
// grid (n_groups, Dpad/64), 1024 threads.  smem: W slice [F][k][64] bf16 | 2 x {entries (8 B), bstart}
constexpr int DX_WARPS = 32;

__device__ __forceinline__ void dx_hit(const uint8_t* s_wt, uint2 ent, float2& acc) {
  const uint32_t w = *reinterpret_cast<const uint32_t*>(s_wt + static_cast<int>(ent.x));
  const float g = __uint_as_float(ent.y);
  acc.x = fmaf(g, __uint_as_float(w << 16), acc.x);
  acc.y = fmaf(g, __uint_as_float(w & 0xffff0000u), acc.y);
}

__global__ void __launch_bounds__(DX_WARPS * 32, 1)
dx_fast_kernel(const __nv_bfloat16* __restrict__ Wt, int k, int Dpad, int L, int64_t S, int64_t seq_per_group, int F,
               const uint2* __restrict__ ord, int ldo, const uint16_t* __restrict__ bstart, int ldb,
               float* __restrict__ dX, int accumulate) {
  extern __shared__ __align__(16) uint8_t dx_smem[];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int c0 = blockIdx.y * 64;
  const int64_t s0 = static_cast<int64_t>(blockIdx.x) * seq_per_group;
  const int64_t s1 = (s0 + seq_per_group < S) ? s0 + seq_per_group : S;
  const int n_units = static_cast<int>(s1 > s0 ? s1 - s0 : 0);
  const int w_bytes = F * k * 128;
  const int ord_bytes = ldo * 8, bs_bytes = ldb * 2;
  const int stage_bytes = ord_bytes + bs_bytes;
  uint8_t* s_w = dx_smem;
  uint8_t* s_stage = dx_smem + w_bytes;

  {   // resident weight slice: byte (f*k + j)*128 + c*2
    const int K = k * Dpad;
    const int chunks = F * k * 8;
    for (int i = threadIdx.x; i < chunks; i += DX_WARPS * 32) {
      const int row = i >> 3, pc = i & 7;
      const int f = row / k, j = row - f * k;
      cp_async16(s_w + row * 128 + pc * 16,
                 reinterpret_cast<const uint8_t*>(Wt + static_cast<int64_t>(f) * K + j * Dpad + c0) + pc * 16);
    }
  }
  auto issue = [&](int u) {
    if (u < n_units) {
      uint8_t* st = s_stage + (u & 1) * stage_bytes;
      const int64_t s = s0 + u;
      const uint8_t* so = reinterpret_cast<const uint8_t*>(ord + s * ldo);
      const uint8_t* sb = reinterpret_cast<const uint8_t*>(bstart + s * ldb);
      for (int i = threadIdx.x; i < ord_bytes / 16; i += DX_WARPS * 32) cp_async16(st + i * 16, so + i * 16);
      for (int i = threadIdx.x; i < bs_bytes / 16; i += DX_WARPS * 32) cp_async16(st + ord_bytes + i * 16, sb + i * 16);
    }
    cp_async_commit();
  };
  issue(0);
  const int Tn = L - k + 1;
  const uint8_t* s_wl = s_w + lane * 4;
  for (int u = 0; u < n_units; ++u) {
    issue(u + 1);                                   // slot (u+1)&1 was last read in iteration u-1 (sync below)
    cp_async_wait<1>();
    __syncthreads();
    const uint8_t* st = s_stage + (u & 1) * stage_bytes;
    const uint2* s_ord = reinterpret_cast<const uint2*>(st);
    const uint16_t* s_bs = reinterpret_cast<const uint16_t*>(st + ord_bytes);
    const int64_t rowbase = (s0 + u) * L;
#pragma unroll 1
    for (int t = warp; t < L; t += DX_WARPS) {
      // positions t-k+1 .. t are one contiguous run of the position-sorted list
      const int p_lo = (t - k + 1 > 0) ? t - k + 1 : 0;
      const int p_hi = (t < Tn - 1) ? t : Tn - 1;
      int e = s_bs[p_lo];
      const int e1 = (p_hi >= p_lo) ? s_bs[p_hi + 1] : e;
      const uint8_t* s_wt = s_wl + t * 128;
      float2 acc = make_float2(0.f, 0.f);
#pragma unroll 1
      for (; e + 4 <= e1; e += 4) {                 // four independent shared-memory chains in flight
        const uint2 a0 = s_ord[e], a1 = s_ord[e + 1], a2 = s_ord[e + 2], a3 = s_ord[e + 3];
        dx_hit(s_wt, a0, acc);
        dx_hit(s_wt, a1, acc);
        dx_hit(s_wt, a2, acc);
        dx_hit(s_wt, a3, acc);
      }
#pragma unroll 1
      for (; e < e1; ++e) dx_hit(s_wt, s_ord[e], acc);
      float* dst = dX + (rowbase + t) * Dpad + c0 + lane * 2;
      if (accumulate) { const float2 o = *reinterpret_cast<const float2*>(dst); acc.x += o.x; acc.y += o.y; }
      *reinterpret_cast<float2*>(dst) = acc;
    }
    __syncthreads();                                // all warps done with slot u&1 before it is refilled
  }
  cp_async_wait<0>();
}

int groups_for(int64_t S, int per_group_ctas, int waves) {
  int64_t g = (static_cast<int64_t>(waves) * clsm_sm_count() + per_group_ctas - 1) / per_group_ctas;
  if (g > S) g = S;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

}  // namespace

// ---- host entry points used by conv_bwd.cu --------------------------------------------------------
int clsm_dw_fast_groups(int64_t S, int Dpad, int F) { return groups_for(S, (Dpad / 64) * ((F + FT - 1) / FT), 4); }

template <typename T>
static int launch_dw_fast(const T* X, int L, int Dpad, int64_t S, int n_groups, const float* G, const int32_t* A,
                          int ldp, int col0, int F, int k, float* part, cudaStream_t st) {
  const int smem = DW_STAGES * (L * 64 * static_cast<int>(sizeof(T)) + FT * 8);
  if (smem > SMEM_BUDGET || k * FPW * 2 > 40) return CLSM_EUNSUPPORTED;   // accumulators must stay in registers (64/thread)
  const int64_t spg = ceil_div64(S, n_groups);
  dim3 grid(n_groups, Dpad / 64, (F + FT - 1) / FT);
#define DWF_CASE(KW)                                                                                             \
  case KW:                                                                                                       \
    CLSM_CUDA(cudaFuncSetAttribute(dw_fast_kernel<T, KW>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));  \
    dw_fast_kernel<T, KW><<<grid, DW_WARPS * 32, smem, st>>>(X, L, Dpad, S, spg, G, A, ldp, col0, F, part);      \
    break;
  switch (k) {
    DWF_CASE(1) DWF_CASE(2) DWF_CASE(3) DWF_CASE(4) DWF_CASE(5) DWF_CASE(6) DWF_CASE(7)
    default: return CLSM_EUNSUPPORTED;
  }
#undef DWF_CASE
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_conv_bwd_dw_fast(const void* X, int dtype, int L, int Dpad, int64_t S, int n_groups, const float* G,
                          const int32_t* A, int ldp, int col0, int F, int k, float* part, cudaStream_t st) {
  if (dtype == CLSM_BF16)
    return launch_dw_fast<__nv_bfloat16>(static_cast<const __nv_bfloat16*>(X), L, Dpad, S, n_groups, G, A, ldp, col0, F, k, part, st);
  return launch_dw_fast<float>(static_cast<const float*>(X), L, Dpad, S, n_groups, G, A, ldp, col0, F, k, part, st);
}

// scratch for the sorted lists: ord (S, ldo) {u32, f32} + bstart (S, ldb) u16
size_t clsm_conv_bwd_dx_fast_bytes(int64_t S, int L, int F) {
  const size_t ldo = (F + 1) / 2 * 2, ldb = (L + 1 + 7) / 8 * 8;
  return align_up(static_cast<size_t>(S) * ldo * 8, 256) + align_up(static_cast<size_t>(S) * ldb * 2, 256);
}

int clsm_conv_bwd_dx_fast(const void* Wt, int dtype, int k, int Dpad, int L, int64_t S, const float* G, const int32_t* A,
                          int ldp, int col0, int F, float* dX, int accumulate, void* scratch, cudaStream_t st) {
  if (dtype != CLSM_BF16 || L > SORT_MAXL || F > 65535) return CLSM_EUNSUPPORTED;
  const int ldo = (F + 1) / 2 * 2, ldb = (L + 1 + 7) / 8 * 8;
  const int smem = F * k * 128 + 2 * (ldo * 8 + ldb * 2);
  if (smem > SMEM_BUDGET) return CLSM_EUNSUPPORTED;
  uint2* ord = static_cast<uint2*>(scratch);
  uint16_t* bstart = reinterpret_cast<uint16_t*>(static_cast<char*>(scratch) + align_up(static_cast<size_t>(S) * ldo * 8, 256));
  pool_sort_kernel<<<static_cast<unsigned>(ceil_div64(S, SORT_WARPS)), SORT_WARPS * 32, 0, st>>>(G, A, S, ldp, col0, F, L, k, ord, ldo, bstart, ldb);
  CLSM_LAUNCH_CHECK();
  const int n_groups = groups_for(S, Dpad / 64, 2);
  const int64_t spg = ceil_div64(S, n_groups);
  CLSM_CUDA(cudaFuncSetAttribute(dx_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dim3 grid(n_groups, Dpad / 64);
  dx_fast_kernel<<<grid, DX_WARPS * 32, smem, st>>>(static_cast<const __nv_bfloat16*>(Wt), k, Dpad, L, S, spg, F,
                                                    ord, ldo, bstart, ldb, dX, accumulate);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}
