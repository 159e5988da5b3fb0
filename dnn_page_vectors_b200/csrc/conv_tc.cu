// n-gram window convolution + max over time on the 5th-generation tensor cores (sm_100a).
//   replaces Convolution1D(valid) + MaxPooling1D + Flatten, dssm_cnn_v2/cnn_dssm_tf.py:84-100.
//
// Implicit GEMM, no im2col: the gathered token matrix X (S*L, Dpad) bf16 is tiled in 128 flat rows;
// output row r = (s, t) needs input rows r .. r+k-1, so K-chunk (j, c) of the A operand is simply the
// TMA box at row coordinate r0 + j, column c*64 of the SAME 2-D tensor map.  B = Wt (Fpad, k*Dpad)
// K-major.  Accumulators (128 x Fpad fp32) live in TMEM; the epilogue reads them back 32 columns at a
// time (tcgen05.ld 32x32b), takes the per-sequence max over the 32 time steps a warp owns with
// CREDUX.MAX.F32 + a ballot for the first arg-max lane, and folds the result into the (S, P) key
// array with one 64-bit atomicMax per (warp, sequence, filter).  Pre-activations never reach HBM.
//
// Warp roles (192 threads, 1 CTA / SM, persistent over tiles):
//   warp 0  lane 0 : TMA producer   (smem ring: full/empty mbarriers)
//   warp 1  lane 0 : tcgen05.mma issuer; warp 1 also owns the TMEM allocation
//   warps 2..5     : epilogue, TMEM lane quarter = warp % 4
#include "common.cuh"
#include <cuda.h>

namespace {

constexpr int TILE_M = 128;
constexpr int KCHUNK = 64;                    // bf16 elements per 128-byte swizzle row
constexpr int A_BYTES = TILE_M * 128;         // 16 KB
constexpr int NUM_THREADS = 192;
constexpr int SMEM_LIMIT = 227 * 1024;

struct TcParams {
  int64_t n_rows;        // S * L
  int L, k, F, Fpad;
  int cpj;               // 64-element chunks per window offset (Dpad / 64)
  int n_tiles;
  int ldk, col0;
  unsigned long long* keys;
  int n1, n2;            // MMA N split (n2 == 0: single MMA)
  int wbox, n_wbox;      // TMA box rows for W and how many boxes
  int stages;
  int acc_stages, acc_cols;
};

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ float credux_max(float v) {
  float m;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(m) : "f"(v));
  return m;
}

// K-major operand, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart (SBO), version 1.
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t saddr) {
  return static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4) | (static_cast<uint64_t>(1) << 16) |
         (static_cast<uint64_t>(1024 >> 4) << 32) | (static_cast<uint64_t>(1) << 46) |
         (static_cast<uint64_t>(2) << 61);
}
// kind::f16, A/B = bf16 (1), D = fp32 (1), both K-major, M = 128
__device__ __forceinline__ uint32_t make_idesc(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(n >> 3) << 17) |
         (static_cast<uint32_t>(TILE_M >> 4) << 24);
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_pool_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, const TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stage_bytes = A_BYTES + static_cast<uint32_t>(p.Fpad) * 128u;
  const uint32_t bars = base + static_cast<uint32_t>(p.stages) * stage_bytes;
  // barrier layout: full[stages] | empty[stages] | tfull[2] | tempty[2] | tmem ptr
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (p.stages + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * p.stages + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * p.stages + 2 + a); };
  const uint32_t tmem_slot = bars + 8u * (2 * p.stages + 4);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmX) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int nk = p.k * p.cpj;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        const int r0 = tile * TILE_M;
        for (int kc = 0; kc < nk; ++kc) {
          const int j = kc / p.cpj, cc = kc - j * p.cpj;
          mbar_wait(empty_bar(stage), phase ^ 1u);
          mbar_expect_tx(full_bar(stage), stage_bytes);
          const uint32_t a_dst = base + stage * stage_bytes;
          tma_load_2d(a_dst, &tmX, cc * KCHUNK, r0 + j, full_bar(stage));
          for (int b = 0; b < p.n_wbox; ++b)
            tma_load_2d(a_dst + A_BYTES + static_cast<uint32_t>(b * p.wbox) * 128u, &tmW, kc * KCHUNK, b * p.wbox, full_bar(stage));
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc1 = make_idesc(p.n1);
      const uint32_t idesc2 = make_idesc(p.n2 > 0 ? p.n2 : 16);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * p.acc_cols);
        for (int kc = 0; kc < nk; ++kc) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t a_src = base + stage * stage_bytes;
          const uint32_t b_src = a_src + A_BYTES;
#pragma unroll
          for (int ks = 0; ks < KCHUNK / 16; ++ks) {
            const uint32_t accum = (kc | ks) != 0 ? 1u : 0u;
            const uint64_t adesc = make_sw128_desc(a_src + ks * 32);
            tc_mma_bf16(d_tmem, adesc, make_sw128_desc(b_src + ks * 32), idesc1, accum);
            if (p.n2 > 0)
              tc_mma_bf16(d_tmem + p.n1, adesc, make_sw128_desc(b_src + static_cast<uint32_t>(p.n1) * 128u + ks * 32), idesc2, accum);
          }
          tc_commit(empty_bar(stage));                 // smem slot free once these MMAs retire
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
        tc_commit(tfull_bar(acc));                     // accumulator complete
        if (++acc == p.acc_stages) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    const int lq = warp & 3;                           // TMEM lane quarter this warp may read
    const int Tn = p.L - p.k + 1;
    int acc = 0; uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t rowbase = static_cast<int64_t>(tile) * TILE_M + lq * 32;
      const int64_t r = rowbase + lane;
      const bool rv = r < p.n_rows;
      const int64_t s = rv ? r / p.L : -1;
      const int t = static_cast<int>(r - s * p.L);
      const bool valid = rv && t < Tn;
      int64_t s_lo = 0, s_hi = -1;
      if (rowbase < p.n_rows) {
        s_lo = rowbase / p.L;
        const int64_t rl = (rowbase + 31 < p.n_rows) ? rowbase + 31 : p.n_rows - 1;
        s_hi = rl / p.L;
      }
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + static_cast<uint32_t>(acc * p.acc_cols);
      const int n_chunks = (p.F + 31) / 32;
      for (int ch = 0; ch < n_chunks; ++ch) {
        uint32_t v[32];
        tc_ld32(taddr0 + ch * 32, v);
        if (ch == n_chunks - 1) {                      // everything is in registers: release TMEM
          tc_fence_before();
          if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
        for (int64_t sg = s_lo; sg <= s_hi; ++sg) {
          const bool in = valid && s == sg;
          if (!__any_sync(0xffffffffu, in)) continue;
          float mym = 0.0f; unsigned myb = 0u;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float x = in ? __uint_as_float(v[i]) + 0.0f : -INFINITY;
            const float m = credux_max(x);
            const unsigned b = __ballot_sync(0xffffffffu, x == m);
            if (lane == i) { mym = m; myb = b; }
          }
          const int f = ch * 32 + lane;
          if (f < p.F && myb != 0u) {
            const int lstar = __ffs(myb) - 1;
            const uint32_t tstar = static_cast<uint32_t>(rowbase + lstar - sg * p.L);
            atomicMax(p.keys + sg * p.ldk + p.col0 + f, pool_key(mym, tstar));
          }
        }
      }
      if (++acc == p.acc_stages) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
  }
}

// ---------------------------------------------------------------- host side
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

bool make_map_2d(CUtensorMap* map, const void* ptr, uint64_t cols, uint64_t rows, uint32_t box_cols, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

int clsm_conv_pool_fwd_tc(const void* X, int64_t S, int L, int Dpad, const void* Wt, int k, int F, int Fpad,
                          uint64_t* keys, int ldk, int col0, cudaStream_t st) {
  const int64_t n_rows = S * L;
  CLSM_REQUIRE(n_rows + TILE_M < (1ll << 31), "conv_pool_fwd(tcgen05): S*L too large for 32-bit TMA coordinates");
  CLSM_REQUIRE(Dpad % 64 == 0 && Fpad % 16 == 0 && Fpad >= 16 && Fpad <= 512, "conv_pool_fwd(tcgen05): bad Dpad/Fpad");
  TcParams p{};
  p.n_rows = n_rows; p.L = L; p.k = k; p.F = F; p.Fpad = Fpad; p.cpj = Dpad / 64;
  p.n_tiles = static_cast<int>(ceil_div64(n_rows, TILE_M));
  p.ldk = ldk; p.col0 = col0; p.keys = reinterpret_cast<unsigned long long*>(keys);
  if (Fpad <= 256) { p.n1 = Fpad; p.n2 = 0; p.wbox = Fpad; p.n_wbox = 1; }
  else {
    p.n1 = ((Fpad / 2) + 15) / 16 * 16; p.n2 = Fpad - p.n1;
    p.wbox = Fpad / 2; p.n_wbox = 2;
  }
  p.acc_cols = Fpad <= 256 ? 256 : 512;
  p.acc_stages = Fpad <= 256 ? 2 : 1;
  const int stage_bytes = A_BYTES + Fpad * 128;
  int stages = (SMEM_LIMIT - 2048) / stage_bytes;
  if (stages > 8) stages = 8;
  CLSM_REQUIRE(stages >= 2, "conv_pool_fwd(tcgen05): Fpad too large for a 2-stage pipeline");
  p.stages = stages;
  const int smem_bytes = stages * stage_bytes + 1024 /*alignment slack*/ + 8 * (2 * stages + 5);

  CUtensorMap tmX, tmW;
  if (!make_map_2d(&tmX, X, static_cast<uint64_t>(Dpad), static_cast<uint64_t>(n_rows), KCHUNK, TILE_M) ||
      !make_map_2d(&tmW, Wt, static_cast<uint64_t>(k) * Dpad, static_cast<uint64_t>(Fpad), KCHUNK, static_cast<uint32_t>(p.wbox))) {
    clsm_set_error("conv_pool_fwd(tcgen05): cuTensorMapEncodeTiled failed");
    return CLSM_ECUDA;
  }
  CLSM_CUDA(cudaFuncSetAttribute(conv_pool_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  int grid = clsm_sm_count();
  if (grid > p.n_tiles) grid = p.n_tiles;
  conv_pool_tc_kernel<<<grid, NUM_THREADS, smem_bytes, st>>>(tmX, tmW, p);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}
