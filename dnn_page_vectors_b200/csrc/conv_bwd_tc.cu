// dX of Convolution1D + global max-pool on the tensor cores (bf16 mode).
//   (gradient of dssm_cnn_v2/cnn_dssm_tf.py:84-100 with respect to the gathered token rows)
//
//   dX[r, c] = sum_{j, f}  Gt[r, (j,f)] * W[j, c, f],     Gt[r = (s,t), (j,f)] = G[s,f] if argmax[s,f] + j == t else 0
//
// i.e. a GEMM with K = k * F whose left operand has exactly k non-zeros per (sequence, filter).  Computed
// TRANSPOSED like the forward kernel, D[c, token] = Wb[c, :] . Gt[token, :], so that
//   A operand = Wb (Dpad, k*Fk) bf16, K-major, M tiles of 128 columns c, streamed by TMA;
//   B operand = the Gt tile (NT token rows x 64 K values per chunk), never materialised in HBM: four builder
//               warps zero a shared-memory tile and scatter the <= (sequences in tile) x 64 non-zeros of the
//               chunk straight into the 128-byte-swizzled K-major layout the UMMA descriptor expects
//               (16-byte chunk index XOR row%8), then fence.proxy.async + mbarrier arrive;
//   D in TMEM (c on lanes, tokens on columns), so the epilogue's stores are 128-byte coalesced per token row.
// G enters the tensor core as bf16 (8-bit mantissa), the usual mixed-precision treatment of activation
// gradients; the fp32 parity mode keeps using the exact CUDA-core kernels in conv_bwd.cu / conv_bwd_fast.cu.
//
// 448 threads: warp 0 TMA producer (weights), warp 1 MMA issuer + TMEM owner, warps 2-9 epilogue,
// warps 10-13 Gt builders.  One CTA per SM, persistent over 160-token tiles.
#include "common.cuh"
#include "tc_ptx.cuh"
#include <stdlib.h>

using namespace tcptx;

namespace {

constexpr int MT = 128;
constexpr int KCHUNK = 64;
constexpr int MT_BYTES = MT * 128;
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_BLD_WARPS = 4;
constexpr int NUM_THREADS = 64 + (NUM_EPI_WARPS + NUM_BLD_WARPS) * 32;
constexpr int SMEM_LIMIT = 227 * 1024;

struct DxParams {
  int64_t n_rows;
  int L, k, F, Dpad;
  int n_mt, NT, n_tiles;
  int fblocks;                 // 64-filter blocks per window offset
  const float* G; const int32_t* A; int ldp, col0;
  float* dX; int accumulate;
  int stages;
};

// Wb[c, j*Fk + f] = Wt[f, j*Dpad + c]   (Fk = fblocks*64, zero padded)
__global__ void pack_wb_kernel(const __nv_bfloat16* __restrict__ Wt, int k, int Dpad, int F, int Fk,
                               __nv_bfloat16* __restrict__ Wb) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t K = static_cast<int64_t>(k) * Fk;
  if (i >= K * Dpad) return;
  const int c = static_cast<int>(i / K);
  const int kk = static_cast<int>(i - c * K);
  const int j = kk / Fk, f = kk - j * Fk;
  Wb[i] = f < F ? Wt[static_cast<int64_t>(f) * k * Dpad + static_cast<int64_t>(j) * Dpad + c] : __float2bfloat16_rn(0.0f);
}

template <int NCH>
__global__ void __launch_bounds__(NUM_THREADS, 1)
dx_tc_kernel(const __grid_constant__ CUtensorMap tmWb, const DxParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_bytes = static_cast<uint32_t>(p.n_mt) * MT_BYTES;
  const uint32_t b_bytes = static_cast<uint32_t>(p.NT) * 128u;
  const uint32_t stage_bytes = a_bytes + ((b_bytes + 1023u) & ~1023u);
  const uint32_t bars = base + static_cast<uint32_t>(p.stages) * stage_bytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (p.stages + s); };
  const uint32_t tfull_bar = bars + 8u * (2 * p.stages);
  const uint32_t tempty_bar = bars + 8u * (2 * p.stages + 1);
  const uint32_t tmem_slot = bars + 8u * (2 * p.stages + 2);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar(s), 1 + NUM_BLD_WARPS); mbar_init(empty_bar(s), 1); }
    mbar_init(tfull_bar, 1);
    mbar_init(tempty_bar, NUM_EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (warp == 0 && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmWb) : "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int nk = p.k * p.fblocks;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        for (int kc = 0; kc < nk; ++kc) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          mbar_expect_tx(full_bar(stage), a_bytes);
          const uint32_t a_dst = base + stage * stage_bytes;
          for (int m = 0; m < p.n_mt; ++m) tma_load_2d(a_dst + m * MT_BYTES, &tmWb, kc * KCHUNK, m * MT, full_bar(stage));
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc(p.NT);
      int stage = 0; uint32_t phase = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        mbar_wait(tempty_bar, acc_phase ^ 1u);
        tc_fence_after();
        for (int kc = 0; kc < nk; ++kc) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t a_src = base + stage * stage_bytes;
          const uint32_t b_src = a_src + a_bytes;
#pragma unroll
          for (int ks = 0; ks < KCHUNK / 16; ++ks) {
            const uint32_t accum = (kc | ks) != 0 ? 1u : 0u;
            const uint64_t bdesc = make_sw128_desc(b_src + ks * 32);
            for (int m = 0; m < p.n_mt; ++m)
              tc_mma_bf16(tmem_base + static_cast<uint32_t>(m * p.NT), make_sw128_desc(a_src + m * MT_BYTES + ks * 32), bdesc, idesc, accum);
          }
          tc_commit(empty_bar(stage));
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
        tc_commit(tfull_bar);
        acc_phase ^= 1u;
      }
    }
  } else if (warp < 2 + NUM_EPI_WARPS) {
    // ---- epilogue: D[c, token] -> dX[token, c]; a warp's 32 lanes are 32 consecutive c of one token row
    const int lq = warp & 3;
    const int half = (warp - 2) >> 2;
    const int ncol = p.NT / 2;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t rbeg = static_cast<int64_t>(tile) * p.NT + half * ncol;
      mbar_wait(tfull_bar, acc_phase);
      tc_fence_after();
      for (int m = 0; m < p.n_mt; ++m) {
        const int c = m * MT + lq * 32 + lane;
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + static_cast<uint32_t>(m * p.NT + half * ncol);
        uint32_t v[NCH][16];
#pragma unroll
        for (int q = 0; q < NCH; ++q) tc_ld16_async(taddr + q * 16, v[q]);
        tc_wait_ld();
        if (m == p.n_mt - 1) { tc_fence_before(); if (lane == 0) mbar_arrive(tempty_bar); }
        if (c < p.Dpad) {
          float* dst = p.dX + rbeg * p.Dpad + c;
#pragma unroll
          for (int q = 0; q < NCH; ++q) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int n = q * 16 + i;
              if (rbeg + n < p.n_rows) {
                float x = __uint_as_float(v[q][i]);
                if (p.accumulate) x += dst[static_cast<int64_t>(n) * p.Dpad];
                dst[static_cast<int64_t>(n) * p.Dpad] = x;
              }
            }
          }
        }
      }
      acc_phase ^= 1u;
    }
  } else {
    // ---- Gt builders: 4 warps fill the B tile of every stage
    const int bt = threadIdx.x - (2 + NUM_EPI_WARPS) * 32;        // 0..127
    constexpr int NB = NUM_BLD_WARPS * 32;
    int stage = 0; uint32_t phase = 0;
    const uint32_t zero_chunks = b_bytes / 16;                     // 16-byte chunks of the tile
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t r0 = static_cast<int64_t>(tile) * p.NT;
      const int64_t r_last = (r0 + p.NT < p.n_rows ? r0 + p.NT : p.n_rows) - 1;
      const int64_t s_first = r0 / p.L;
      const int nseq = r_last >= r0 ? static_cast<int>(r_last / p.L - s_first) + 1 : 0;
      for (int kc = 0; kc < nk; ++kc) {
        const int j = kc / p.fblocks, fb = kc - j * p.fblocks;
        mbar_wait(empty_bar(stage), phase ^ 1u);
        uint8_t* btile = base_ptr + stage * stage_bytes + a_bytes;
        for (uint32_t i = bt; i < zero_chunks; i += NB) reinterpret_cast<uint4*>(btile)[i] = make_uint4(0u, 0u, 0u, 0u);
        asm volatile("bar.sync 1, %0;" ::"n"(NB) : "memory");        // zeros land before anyone scatters
        // candidates: (sequence in tile, filter in this 64-block)
        for (int cidx = bt; cidx < nseq * 64; cidx += NB) {
          const int sq = cidx >> 6, fl = cidx & 63;
          const int f = fb * 64 + fl;
          if (f < p.F) {
            const int64_t s = s_first + sq;
            const float g = __ldg(p.G + s * p.ldp + p.col0 + f);
            if (g != 0.0f) {
              const int a = __ldg(p.A + s * p.ldp + p.col0 + f);
              const int64_t n = s * p.L + a + j - r0;                  // token column of this hit inside the tile
              if (n >= 0 && n < p.NT) {
                const uint32_t off = static_cast<uint32_t>(n) * 128u + ((static_cast<uint32_t>(fl >> 3) ^ (static_cast<uint32_t>(n) & 7u)) << 4) +
                                     static_cast<uint32_t>(fl & 7) * 2u;
                *reinterpret_cast<__nv_bfloat16*>(btile + off) = __float2bfloat16_rn(g);
              }
            }
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA's async reads
        __syncwarp();
        if (lane == 0) mbar_arrive(full_bar(stage));
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

}  // namespace

// scratch: Wb (Dpad x k*Fk) bf16
size_t clsm_conv_bwd_dx_tc_bytes(int k, int Dpad, int F) {
  const size_t Fk = static_cast<size_t>((F + 63) / 64) * 64;
  return align_up(static_cast<size_t>(Dpad) * k * Fk * 2, 256);
}

int clsm_conv_bwd_dx_tc(const void* Wt, int dtype, int k, int Dpad, int L, int64_t S, const float* G, const int32_t* A,
                        int ldp, int col0, int F, float* dX, int accumulate, void* scratch, cudaStream_t st) {
  if (dtype != CLSM_BF16) return CLSM_EUNSUPPORTED;
  const char* env = getenv("CLSM_CONV_BWD_DX");
  if (env != nullptr && env[0] == 's') return CLSM_EUNSUPPORTED;       // force the CUDA-core path
  DxParams p{};
  p.n_rows = S * L; p.L = L; p.k = k; p.F = F; p.Dpad = Dpad;
  p.n_mt = (Dpad + MT - 1) / MT;
  if (p.n_mt > 4 || p.n_rows + 512 >= (1ll << 31)) return CLSM_EUNSUPPORTED;
  p.NT = (512 / p.n_mt) / 32 * 32;
  if (p.NT > 256) p.NT = 256;
  p.n_tiles = static_cast<int>(ceil_div64(p.n_rows, p.NT));
  p.fblocks = (F + 63) / 64;
  p.G = G; p.A = A; p.ldp = ldp; p.col0 = col0; p.dX = dX; p.accumulate = accumulate;
  const int Fk = p.fblocks * 64;
  const int a_rows = p.n_mt * MT;
  const int stage_bytes = a_rows * 128 + ((p.NT * 128 + 1023) / 1024) * 1024;
  int stages = (SMEM_LIMIT - 2048) / stage_bytes;
  if (stages > 6) stages = 6;
  if (stages < 2) return CLSM_EUNSUPPORTED;
  p.stages = stages;
  const int smem_bytes = stages * stage_bytes + 1024 + 8 * (2 * stages + 4);

  __nv_bfloat16* Wb = static_cast<__nv_bfloat16*>(scratch);
  const int64_t n_wb = static_cast<int64_t>(Dpad) * k * Fk;
  pack_wb_kernel<<<static_cast<unsigned>(ceil_div64(n_wb, 256)), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(Wt), k, Dpad, F, Fk, Wb);
  CLSM_LAUNCH_CHECK();

  EncodeTiledFn enc = get_encode();
  if (!enc) { clsm_set_error("conv_pool_bwd(tcgen05): cuTensorMapEncodeTiled unavailable"); return CLSM_ECUDA; }
  CUtensorMap tmWb;
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(k) * Fk, static_cast<cuuint64_t>(Dpad)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(k) * Fk * 2};
  cuuint32_t box[2] = {KCHUNK, MT};
  cuuint32_t estr[2] = {1, 1};
  if (enc(&tmWb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Wb, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
    clsm_set_error("conv_pool_bwd(tcgen05): cuTensorMapEncodeTiled failed");
    return CLSM_ECUDA;
  }
  int grid = clsm_sm_count();
  if (grid > p.n_tiles) grid = p.n_tiles;
  const int nch = p.NT / 32;
#define DX_LAUNCH(N)                                                                                          \
  do {                                                                                                        \
    CLSM_CUDA(cudaFuncSetAttribute(dx_tc_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)); \
    dx_tc_kernel<N><<<grid, NUM_THREADS, smem_bytes, st>>>(tmWb, p);                                          \
  } while (0)
  if (nch == 4) DX_LAUNCH(4); else if (nch == 5) DX_LAUNCH(5); else DX_LAUNCH(8);
#undef DX_LAUNCH
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}
