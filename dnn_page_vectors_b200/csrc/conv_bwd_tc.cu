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
// This file holds two tilings of each gradient: dx_tc / dw_tc (first: per-(window offset, filter block) operand
// tiles) and dx2_tc / dw2_tc (default where they fit: the one-hot tile is built once and shared by all window
// offsets through row-shifted descriptors).  DESIGN.md section 4b has the measurements.
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
  long long* prof;             // optional cycle counters (CLSM_CONV_BWD_PROF=1), else nullptr
};

__device__ long long g_conv_bwd_prof[8];

// Wb[c, j*Fk + f] = Wt[f, j*Dpad + c]   (Fk = fblocks*64, zero padded)
// (ldw = row length of Wb in elements, col_off = first column written: several windows side by side in one matrix)
__global__ void pack_wb_kernel(const __nv_bfloat16* __restrict__ Wt, int k, int Dpad, int F, int Fk,
                               __nv_bfloat16* __restrict__ Wb, int64_t ldw, int64_t col_off) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t K = static_cast<int64_t>(k) * Fk;
  if (i >= K * Dpad) return;
  const int c = static_cast<int>(i / K);
  const int kk = static_cast<int>(i - c * K);
  const int j = kk / Fk, f = kk - j * Fk;
  Wb[c * ldw + col_off + kk] = f < F ? Wt[static_cast<int64_t>(f) * k * Dpad + static_cast<int64_t>(j) * Dpad + c] : __float2bfloat16_rn(0.0f);
}

template <int NCH, int MAXC>      // MAXC = candidate slots per builder thread per chunk: ceil(sequences per tile * 64 / 128)
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

  const int warp = warp_id_uniform(), lane = threadIdx.x % 32;

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
          const int fbq = kc / p.k, jq = kc - fbq * p.k;               // chunk order: filter block outer, window offset inner
          for (int m = 0; m < p.n_mt; ++m) tma_load_2d(a_dst + m * MT_BYTES, &tmWb, (jq * p.fblocks + fbq) * KCHUNK, m * MT, full_bar(stage));
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // whole warp, warp-uniform issue loop (tc_ptx.cuh)
    {
      const uint32_t idesc = make_idesc(p.NT);
      int stage = 0; uint32_t phase = 0, acc_phase = 0;
      long long w_tmem = 0, w_full = 0;
      const long long t_begin = clock64();
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        long long c0 = clock64();
        mbar_wait(tempty_bar, acc_phase ^ 1u);
        w_tmem += clock64() - c0;
        tc_fence_after();
        for (int kc = 0; kc < nk; ++kc) {
          c0 = clock64();
          mbar_wait(full_bar(stage), phase);
          w_full += clock64() - c0;
          tc_fence_after();
          const uint32_t a_lo = sw128_desc_lo(base + stage * stage_bytes);
          const uint32_t b_lo = sw128_desc_lo(base + stage * stage_bytes + a_bytes);
#pragma unroll
          for (int ks = 0; ks < KCHUNK / 16; ++ks) {
            const uint32_t accum = (kc | ks) != 0 ? 1u : 0u;
            for (int m = 0; m < p.n_mt; ++m)
              tc_mma_bf16_u(tmem_base + static_cast<uint32_t>(m * p.NT), a_lo + static_cast<uint32_t>(m * (MT_BYTES / 16) + ks * 2),
                            b_lo + ks * 2, idesc, accum);
          }
          tc_commit_u(empty_bar(stage));
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
        tc_commit_u(tfull_bar);
        acc_phase ^= 1u;
      }
      if (p.prof != nullptr && lane == 0) {
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 4), static_cast<unsigned long long>(clock64() - t_begin));
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 5), static_cast<unsigned long long>(w_tmem));
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 6), static_cast<unsigned long long>(w_full));
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
          // Later windows add into dX.  Each element has exactly one writer, so a fire-and-forget reduction
          // (RED.ADD.F32, result unused; denormal sums flush to zero) equals load + add + store without putting a global-load
          // latency on every element (that chain made the 4-window configuration 20x slower) and without registers.
          if (p.accumulate) {
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int n = q * 16 + i;
                if (rbeg + n < p.n_rows) atomicAdd(dst + static_cast<int64_t>(n) * p.Dpad, __uint_as_float(v[q][i]));
              }
            }
          } else {
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int n = q * 16 + i;
                if (rbeg + n < p.n_rows) dst[static_cast<int64_t>(n) * p.Dpad] = __uint_as_float(v[q][i]);
              }
            }
          }
        }
      }
      acc_phase ^= 1u;
    }
  } else {
    // ---- Gt builders: 4 warps fill the B tile of every stage.  Chunks run filter-block outer / window offset
    // inner, so a thread's (G, arg-max) candidates are fetched once per filter block (prefetched one block
    // ahead, in registers) and reused for the k offsets; tiles start all-zero and each thread un-writes what it
    // wrote the last time the stage was used.
    const int bt = threadIdx.x - (2 + NUM_EPI_WARPS) * 32;        // 0..127
    constexpr int NB = NUM_BLD_WARPS * 32;
    constexpr uint32_t NONE = 0xFFFFFFFFu;
    const uint32_t zero_chunks = (stage_bytes - a_bytes) / 16;
    for (int st = 0; st < p.stages; ++st) {
      uint8_t* btile = base_ptr + st * stage_bytes + a_bytes;
      for (uint32_t i = bt; i < zero_chunks; i += NB) reinterpret_cast<uint4*>(btile)[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(NB) : "memory");
    uint32_t wrote[3][MAXC];
#pragma unroll
    for (int st = 0; st < 3; ++st)
#pragma unroll
      for (int c = 0; c < MAXC; ++c) wrote[st][c] = NONE;
    float g_nxt[MAXC]; int a_nxt[MAXC];
    // state of the block being prefetched: (tile, fb); of the block being built: tile geometry
    int ptile = blockIdx.x, pfb = 0;
    auto tile_geom = [&](int tile, int64_t& s_first, int& t0, int& nseq) {
      const int64_t r0 = static_cast<int64_t>(tile) * p.NT;
      const int64_t r_last = (r0 + p.NT < p.n_rows ? r0 + p.NT : p.n_rows) - 1;
      s_first = r0 / p.L;
      t0 = static_cast<int>(r0 - s_first * p.L);
      nseq = r_last >= r0 ? static_cast<int>(r_last / p.L - s_first) + 1 : 0;
    };
    int64_t ps_first = 0; int pt0 = 0, pnseq = 0;
    if (ptile < p.n_tiles) tile_geom(ptile, ps_first, pt0, pnseq);
    auto prefetch = [&]() {
#pragma unroll
      for (int c = 0; c < MAXC; ++c) {
        const int cidx = bt + c * NB;
        const int sq = cidx >> 6, f = pfb * 64 + (cidx & 63);
        const bool ok = ptile < p.n_tiles && sq < pnseq && f < p.F;
        g_nxt[c] = ok ? __ldg(p.G + (ps_first + sq) * p.ldp + p.col0 + f) : 0.0f;
        a_nxt[c] = ok ? __ldg(p.A + (ps_first + sq) * p.ldp + p.col0 + f) : 0;
      }
      if (++pfb == p.fblocks) {
        pfb = 0;
        ptile += gridDim.x;
        if (ptile < p.n_tiles) tile_geom(ptile, ps_first, pt0, pnseq);
      }
    };
    prefetch();
    int st = 0; uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      int64_t s_first; int t0, nseq;
      tile_geom(tile, s_first, t0, nseq);
      for (int fb = 0; fb < p.fblocks; ++fb) {
        float g_cur[MAXC]; int a_cur[MAXC];
#pragma unroll
        for (int c = 0; c < MAXC; ++c) { g_cur[c] = g_nxt[c]; a_cur[c] = a_nxt[c]; }
        prefetch();
        for (int j = 0; j < p.k; ++j) {
          mbar_wait(empty_bar(st), phase ^ 1u);
          uint8_t* btile = base_ptr + st * stage_bytes + a_bytes;
          // the stage index is a runtime value: select the bookkeeping row with predicated copies
          uint32_t w_old[MAXC];
#pragma unroll
          for (int c = 0; c < MAXC; ++c) w_old[c] = st == 0 ? wrote[0][c] : st == 1 ? wrote[1][c] : wrote[2][c];
#pragma unroll
          for (int c = 0; c < MAXC; ++c)
            if (w_old[c] != NONE) *reinterpret_cast<uint16_t*>(btile + w_old[c]) = 0;
#pragma unroll
          for (int c = 0; c < MAXC; ++c) {
            const int cidx = bt + c * NB;
            const int sq = cidx >> 6, fl = cidx & 63;
            const int n = sq * p.L + a_cur[c] + j - t0;               // token column of this hit inside the tile
            uint32_t off = NONE;
            if (g_cur[c] != 0.0f && n >= 0 && n < p.NT) {
              off = static_cast<uint32_t>(n) * 128u + ((static_cast<uint32_t>(fl >> 3) ^ (static_cast<uint32_t>(n) & 7u)) << 4) +
                    static_cast<uint32_t>(fl & 7) * 2u;
              *reinterpret_cast<__nv_bfloat16*>(btile + off) = __float2bfloat16_rn(g_cur[c]);
            }
            if (st == 0) wrote[0][c] = off; else if (st == 1) wrote[1][c] = off; else wrote[2][c] = off;
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(full_bar(st));
          if (++st == p.stages) { st = 0; phase ^= 1u; }
        }
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

// =====================================================================================================
// dX, second tiling (the default when it fits shared memory): the structure of the forward kernel's resident-tile
// mode.  dX[r, c] = sum_j sum_f dY[r - j, f] * W[j, c, f] with dY[(s,t), f] = G[s,f] at the arg-max row of (s,f):
// the one-hot dY tile (NT + k - 1 token rows x F filters, 64 filters per shared-memory slot) is built ONCE per
// tile and serves every window offset j through a B descriptor whose start is advanced by (k-1-j) rows, so the
// builders do a k-th of the work of the per-(j, filter block) tiling above.  Weights Wb stream one (M tile, j,
// filter block) chunk of 16 KB per slot; the (tile, M tile) units rotate through 512 / NT accumulator regions
// so that the store epilogue of one unit overlaps the MMAs of the next.
struct Dx2Params {
  int64_t n_rows;
  int L, k, F, Dpad;
  int n_mt, NT, n_tiles, fblocks;
  const float* G; const int32_t* A; int ldp, col0;
  float* dX; int accumulate;
  int last_ks;                 // K = 16 steps of the last 64-filter block that hold real filters (F = 300: 3 of 4)
  int dx_bf16;                 // dX rows are bf16 (overwrite mode only: the single-window case); halves the largest stream of the step
  int wstages, xrows, nreg;
  long long* prof;
};

template <int NCH>
__global__ void __launch_bounds__(NUM_THREADS, 1)
dx2_tc_kernel(const __grid_constant__ CUtensorMap tmWb, const Dx2Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t x_slot_bytes = ((static_cast<uint32_t>(p.xrows) * 128u) + 1023u) & ~1023u;
  const uint32_t w_ring = base;
  const uint32_t x_ring = base + static_cast<uint32_t>(p.wstages) * MT_BYTES;
  const int stages = p.wstages + p.fblocks;                   // barrier pairs: weight slots, then dY slots
  const uint32_t bars = x_ring + static_cast<uint32_t>(p.fblocks) * x_slot_bytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (stages + s); };
  auto tfull_bar = [&](int r) { return bars + 8u * (2 * stages + r); };
  auto tempty_bar = [&](int r) { return bars + 8u * (2 * stages + 4 + r); };
  const uint32_t tmem_slot = bars + 8u * (2 * stages + 8);

  const int warp = warp_id_uniform(), lane = threadIdx.x % 32;

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(full_bar(s), s < p.wstages ? 1 : NUM_BLD_WARPS); mbar_init(empty_bar(s), 1); }
    for (int r = 0; r < 4; ++r) { mbar_init(tfull_bar(r), 1); mbar_init(tempty_bar(r), NUM_EPI_WARPS); }
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

  if (warp == 0) {
    // ---- weights: one (M tile, filter block, window offset) chunk per slot, warp-uniform loop
    int ws = 0; uint32_t wphase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      for (int m = 0; m < p.n_mt; ++m)
        for (int fb = 0; fb < p.fblocks; ++fb)
          for (int j = 0; j < p.k; ++j) {
            mbar_wait(empty_bar(ws), wphase ^ 1u);
            if (elect_one()) {
              mbar_expect_tx(full_bar(ws), MT_BYTES);
              tma_load_2d(w_ring + ws * MT_BYTES, &tmWb, (j * p.fblocks + fb) * KCHUNK, m * MT, full_bar(ws));
            }
            if (++ws == p.wstages) { ws = 0; wphase ^= 1u; }
          }
    }
  } else if (warp == 1) {
    // ---- MMA issue, warp-uniform loop (tc_ptx.cuh)
    const uint32_t idesc = make_idesc(p.NT);
    int ws = 0; uint32_t wphase = 0, xphase = 0;
    int reg = 0; uint32_t rphase = 0;
    long long w_tmem = 0, w_full = 0;
    const long long t_begin = clock64();
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      for (int m = 0; m < p.n_mt; ++m) {
        long long c0 = clock64();
        mbar_wait(tempty_bar(reg), rphase ^ 1u);
        w_tmem += clock64() - c0;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(reg * p.NT);
        for (int fb = 0; fb < p.fblocks; ++fb) {
          if (m == 0) { c0 = clock64(); mbar_wait(full_bar(p.wstages + fb), xphase); w_full += clock64() - c0; }
          const uint32_t x_lo = sw128_desc_lo(x_ring + fb * x_slot_bytes);
          for (int j = 0; j < p.k; ++j) {
            c0 = clock64();
            mbar_wait(full_bar(ws), wphase);
            w_full += clock64() - c0;
            tc_fence_after();
            const uint32_t a_lo = sw128_desc_lo(w_ring + ws * MT_BYTES);
            const int nks = fb == p.fblocks - 1 ? p.last_ks : KCHUNK / 16;      // filters past F are zero in Wb and never built into dY
#pragma unroll
            for (int ks = 0; ks < KCHUNK / 16; ++ks)      // output token n reads tile row n + (k-1-j): 8 units of 16 B per row
              if (ks < nks)
                tc_mma_bf16_u(d_tmem, a_lo + ks * 2, x_lo + static_cast<uint32_t>(p.k - 1 - j) * 8u + ks * 2, idesc, (fb | j | ks) != 0 ? 1u : 0u);
            tc_commit_u(empty_bar(ws));
            if (++ws == p.wstages) { ws = 0; wphase ^= 1u; }
          }
          if (m == p.n_mt - 1) tc_commit_u(empty_bar(p.wstages + fb));     // the builders may rebuild this slot
        }
        tc_commit_u(tfull_bar(reg));
        if (++reg == p.nreg) { reg = 0; rphase ^= 1u; }
      }
      xphase ^= 1u;
    }
    if (p.prof != nullptr && lane == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 4), static_cast<unsigned long long>(clock64() - t_begin));
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 5), static_cast<unsigned long long>(w_tmem));
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 6), static_cast<unsigned long long>(w_full));
    }
  } else if (warp < 2 + NUM_EPI_WARPS) {
    // ---- epilogue: D[c, token] -> dX[token, c]; a warp's 32 lanes are 32 consecutive c of one token row
    const int lq = warp & 3;
    const int half = (warp - 2) >> 2;
    const int ncol = p.NT / 2;
    int reg = 0; uint32_t rphase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t rbeg = static_cast<int64_t>(tile) * p.NT + half * ncol;
      for (int m = 0; m < p.n_mt; ++m) {
        mbar_wait(tfull_bar(reg), rphase);
        tc_fence_after();
        const int c = m * MT + lq * 32 + lane;
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + static_cast<uint32_t>(reg * p.NT + half * ncol);
        uint32_t v[NCH][16];
#pragma unroll
        for (int q = 0; q < NCH; ++q) tc_ld16_async(taddr + q * 16, v[q]);
        tc_wait_ld();
        tc_fence_before();
        if (lane == 0) mbar_arrive(tempty_bar(reg));
        if (++reg == p.nreg) { reg = 0; rphase ^= 1u; }
        if (c < p.Dpad) {
          float* dst = p.dX + rbeg * p.Dpad + c;
          // Later windows add into dX.  Each element has exactly one writer, so a fire-and-forget reduction
          // (RED.ADD.F32, result unused; denormal sums flush to zero) equals load + add + store without putting a global-load
          // latency on every element (that chain made the 4-window configuration 20x slower) and without registers.
          if (p.accumulate) {
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int n = q * 16 + i;
                if (rbeg + n < p.n_rows) atomicAdd(dst + static_cast<int64_t>(n) * p.Dpad, __uint_as_float(v[q][i]));
              }
            }
          } else if (p.dx_bf16) {
            __nv_bfloat16* dst16 = reinterpret_cast<__nv_bfloat16*>(p.dX) + rbeg * p.Dpad + c;
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int n = q * 16 + i;
                if (rbeg + n < p.n_rows) dst16[static_cast<int64_t>(n) * p.Dpad] = __float2bfloat16_rn(__uint_as_float(v[q][i]));
              }
            }
          } else {
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int n = q * 16 + i;
                if (rbeg + n < p.n_rows) dst[static_cast<int64_t>(n) * p.Dpad] = __uint_as_float(v[q][i]);
              }
            }
          }
        }
      }
    }
  } else {
    // ---- dY builders: 4 warps, thread = (filter of the block, parity of the sequence slot).  Each slot starts
    // all-zero; before a tile's hits are written, the previous tile's hits are cleared by recomputing their
    // positions from (G, arg-max) (L2-resident), so nothing is remembered in registers and the sequence count
    // per tile is unbounded.  A named barrier separates clearing from writing: a thread's new hit may land
    // where another thread's old hit was.
    const int bt = threadIdx.x - (2 + NUM_EPI_WARPS) * 32;        // 0..127
    constexpr int NB = NUM_BLD_WARPS * 32;
    const int fl = bt & 63, par = bt >> 6;
    const uint32_t zero_chunks = static_cast<uint32_t>(p.fblocks) * x_slot_bytes / 16;
    uint8_t* xbase = base_ptr + (x_ring - base);
    for (uint32_t i = bt; i < zero_chunks; i += NB) reinterpret_cast<uint4*>(xbase)[i] = make_uint4(0u, 0u, 0u, 0u);
    asm volatile("bar.sync 1, %0;" ::"n"(NB) : "memory");
    const int rows_used = p.NT + p.k - 1;
    auto tile_geom = [&](int tile, int64_t& r_lo, int64_t& s_first, int& nseq) {
      r_lo = static_cast<int64_t>(tile) * p.NT - (p.k - 1);                 // token row of slot row 0 (may be < 0)
      const int64_t r_first = r_lo < 0 ? 0 : r_lo;
      int64_t r_last = static_cast<int64_t>(tile) * p.NT + p.NT - 1;
      if (r_last > p.n_rows - 1) r_last = p.n_rows - 1;
      s_first = r_first / p.L;
      nseq = static_cast<int>(r_last / p.L - s_first) + 1;
    };
    auto scatter = [&](uint8_t* slot, int f, int64_t r_lo, int64_t s_first, int nseq, bool clear) {
      for (int sq = par; sq < nseq; sq += 2) {
        const int64_t sidx = s_first + sq;
        const float g = __ldg(p.G + sidx * p.ldp + p.col0 + f);
        const int a = __ldg(p.A + sidx * p.ldp + p.col0 + f);
        const int64_t row = sidx * p.L + a - r_lo;
        if (g != 0.0f && row >= 0 && row < rows_used) {
          const uint32_t r = static_cast<uint32_t>(row);
          const uint32_t off = r * 128u + ((static_cast<uint32_t>(fl >> 3) ^ (r & 7u)) << 4) + static_cast<uint32_t>(fl & 7) * 2u;
          *reinterpret_cast<__nv_bfloat16*>(slot + off) = clear ? __float2bfloat16_rn(0.0f) : __float2bfloat16_rn(g);
        }
      }
    };
    uint32_t xphase = 0;
    bool has_prev = false;
    int64_t pr_lo = 0, ps_first = 0; int pnseq = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      int64_t r_lo, s_first; int nseq;
      tile_geom(tile, r_lo, s_first, nseq);
      for (int fb = 0; fb < p.fblocks; ++fb) {
        const int f = fb * 64 + fl;
        uint8_t* slot = xbase + static_cast<size_t>(fb) * x_slot_bytes;
        mbar_wait(empty_bar(p.wstages + fb), xphase ^ 1u);
        if (has_prev && f < p.F) scatter(slot, f, pr_lo, ps_first, pnseq, true);
        asm volatile("bar.sync 2, %0;" ::"n"(NB) : "memory");
        if (f < p.F) scatter(slot, f, r_lo, s_first, nseq, false);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(full_bar(p.wstages + fb));
      }
      xphase ^= 1u;
      has_prev = true; pr_lo = r_lo; ps_first = s_first; pnseq = nseq;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
  }
}

// =====================================================================================================
// dX of SEVERAL window sizes in one pass (the reference's default model has windows 3 and 4; BASELINE cfg4 has 1, 2, 3, 5):
//   dX[r, c] = sum_w sum_{j < k_w} sum_f dY_w[r - j, f] * W_w[j, c, f]
// One launch per window makes every window after the first add into fp32 dX rows in HBM (a 5.2 GB read-modify-write per
// window on cfg4).  Here the accumulator of a (token tile, M tile) stays in TMEM across the windows: the n_mt regions
// of a tile are all live until the last window has added its part, the one-hot dY slots are rebuilt per (tile, window),
// and dX is written once -- as bf16 when asked.  Same roles, barriers and descriptors as dx2_tc_kernel above; the dY
// tile carries a halo of kmax - 1 rows, so window offset j reads it advanced by (kmax - 1 - j) rows for every window.
// The weights of all windows sit side by side in ONE matrix Wb (Dpad x sum_w k_w * Fk), window w at column wcol[w].
struct DxmParams {
  int64_t n_rows;
  int L, F, Dpad;
  int n_mt, NT, n_tiles, fblocks;
  int nwin, kmax;
  int k[4], wcol[4], gcol[4];
  const float* G; const int32_t* A; int ldp;
  void* dX; int dx_bf16;
  int last_ks, wstages, xrows;
};

template <int NCH>
__global__ void __launch_bounds__(NUM_THREADS, 1)
dxm_tc_kernel(const __grid_constant__ CUtensorMap tmWb, const DxmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t x_slot_bytes = ((static_cast<uint32_t>(p.xrows) * 128u) + 1023u) & ~1023u;
  const uint32_t w_ring = base;
  const uint32_t x_ring = base + static_cast<uint32_t>(p.wstages) * MT_BYTES;
  const int stages = p.wstages + p.fblocks;                   // barrier pairs: weight slots, then dY slots
  const uint32_t bars = x_ring + static_cast<uint32_t>(p.fblocks) * x_slot_bytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (stages + s); };
  auto tfull_bar = [&](int r) { return bars + 8u * (2 * stages + r); };
  auto tempty_bar = [&](int r) { return bars + 8u * (2 * stages + 4 + r); };
  const uint32_t tmem_slot = bars + 8u * (2 * stages + 8);

  const int warp = warp_id_uniform(), lane = threadIdx.x % 32;

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(full_bar(s), s < p.wstages ? 1 : NUM_BLD_WARPS); mbar_init(empty_bar(s), 1); }
    for (int r = 0; r < 4; ++r) { mbar_init(tfull_bar(r), 1); mbar_init(tempty_bar(r), NUM_EPI_WARPS); }
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

  if (warp == 0) {
    // ---- weights: one (window, M tile, filter block, window offset) chunk per slot, warp-uniform loop
    int ws = 0; uint32_t wphase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x)
      for (int w = 0; w < p.nwin; ++w)
        for (int m = 0; m < p.n_mt; ++m)
          for (int fb = 0; fb < p.fblocks; ++fb)
            for (int j = 0; j < p.k[w]; ++j) {
              mbar_wait(empty_bar(ws), wphase ^ 1u);
              if (elect_one()) {
                mbar_expect_tx(full_bar(ws), MT_BYTES);
                tma_load_2d(w_ring + ws * MT_BYTES, &tmWb, p.wcol[w] + (j * p.fblocks + fb) * KCHUNK, m * MT, full_bar(ws));
              }
              if (++ws == p.wstages) { ws = 0; wphase ^= 1u; }
            }
  } else if (warp == 1) {
    // ---- MMA issue, warp-uniform loop.  Region m of TMEM = M tile m of the current token tile, for all windows.
    const uint32_t idesc = make_idesc(p.NT);
    int ws = 0; uint32_t wphase = 0, xphase = 0, tphase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      for (int w = 0; w < p.nwin; ++w) {
        for (int m = 0; m < p.n_mt; ++m) {
          if (w == 0) { mbar_wait(tempty_bar(m), tphase ^ 1u); tc_fence_after(); }      // the previous tile's rows of this region are stored
          const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(m * p.NT);
          for (int fb = 0; fb < p.fblocks; ++fb) {
            if (m == 0) mbar_wait(full_bar(p.wstages + fb), xphase);
            const uint32_t x_lo = sw128_desc_lo(x_ring + fb * x_slot_bytes);
            for (int j = 0; j < p.k[w]; ++j) {
              mbar_wait(full_bar(ws), wphase);
              tc_fence_after();
              const uint32_t a_lo = sw128_desc_lo(w_ring + ws * MT_BYTES);
              const int nks = fb == p.fblocks - 1 ? p.last_ks : KCHUNK / 16;
#pragma unroll
              for (int ks = 0; ks < KCHUNK / 16; ++ks)
                if (ks < nks)
                  tc_mma_bf16_u(d_tmem, a_lo + ks * 2, x_lo + static_cast<uint32_t>(p.kmax - 1 - j) * 8u + ks * 2, idesc, (w | fb | j | ks) != 0 ? 1u : 0u);
              tc_commit_u(empty_bar(ws));
              if (++ws == p.wstages) { ws = 0; wphase ^= 1u; }
            }
            if (m == p.n_mt - 1) tc_commit_u(empty_bar(p.wstages + fb));     // the builders may rebuild this slot
          }
          if (w == p.nwin - 1) tc_commit_u(tfull_bar(m));
        }
        xphase ^= 1u;
      }
      tphase ^= 1u;
    }
  } else if (warp < 2 + NUM_EPI_WARPS) {
    // ---- epilogue: D[c, token] -> dX[token, c], once per (tile, M tile), after the last window
    const int lq = warp & 3;
    const int half = (warp - 2) >> 2;
    const int ncol = p.NT / 2;
    uint32_t tphase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t rbeg = static_cast<int64_t>(tile) * p.NT + half * ncol;
      for (int m = 0; m < p.n_mt; ++m) {
        mbar_wait(tfull_bar(m), tphase);
        tc_fence_after();
        const int c = m * MT + lq * 32 + lane;
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + static_cast<uint32_t>(m * p.NT + half * ncol);
        uint32_t v[NCH][16];
#pragma unroll
        for (int q = 0; q < NCH; ++q) tc_ld16_async(taddr + q * 16, v[q]);
        tc_wait_ld();
        tc_fence_before();
        if (lane == 0) mbar_arrive(tempty_bar(m));
        if (c < p.Dpad) {
          if (p.dx_bf16) {
            __nv_bfloat16* dst16 = static_cast<__nv_bfloat16*>(p.dX) + rbeg * p.Dpad + c;
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int n = q * 16 + i;
                if (rbeg + n < p.n_rows) dst16[static_cast<int64_t>(n) * p.Dpad] = __float2bfloat16_rn(__uint_as_float(v[q][i]));
              }
            }
          } else {
            float* dst = static_cast<float*>(p.dX) + rbeg * p.Dpad + c;
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int n = q * 16 + i;
                if (rbeg + n < p.n_rows) dst[static_cast<int64_t>(n) * p.Dpad] = __uint_as_float(v[q][i]);
              }
            }
          }
        }
      }
      tphase ^= 1u;
    }
  } else {
    // ---- dY builders (see dx2_tc_kernel): the slots are rebuilt for every (tile, window); what a slot held before is
    // cleared by recomputing the previous (tile, window)'s hit positions
    const int bt = threadIdx.x - (2 + NUM_EPI_WARPS) * 32;        // 0..127
    constexpr int NB = NUM_BLD_WARPS * 32;
    const int fl = bt & 63, par = bt >> 6;
    const uint32_t zero_chunks = static_cast<uint32_t>(p.fblocks) * x_slot_bytes / 16;
    uint8_t* xbase = base_ptr + (x_ring - base);
    for (uint32_t i = bt; i < zero_chunks; i += NB) reinterpret_cast<uint4*>(xbase)[i] = make_uint4(0u, 0u, 0u, 0u);
    asm volatile("bar.sync 1, %0;" ::"n"(NB) : "memory");
    const int rows_used = p.NT + p.kmax - 1;
    auto tile_geom = [&](int tile, int64_t& r_lo, int64_t& s_first, int& nseq) {
      r_lo = static_cast<int64_t>(tile) * p.NT - (p.kmax - 1);               // token row of slot row 0 (may be < 0)
      const int64_t r_first = r_lo < 0 ? 0 : r_lo;
      int64_t r_last = static_cast<int64_t>(tile) * p.NT + p.NT - 1;
      if (r_last > p.n_rows - 1) r_last = p.n_rows - 1;
      s_first = r_first / p.L;
      nseq = static_cast<int>(r_last / p.L - s_first) + 1;
    };
    auto scatter = [&](uint8_t* slot, int col, int64_t r_lo, int64_t s_first, int nseq, bool clear) {
      for (int sq = par; sq < nseq; sq += 2) {
        const int64_t sidx = s_first + sq;
        const float g = __ldg(p.G + sidx * p.ldp + col);
        const int a = __ldg(p.A + sidx * p.ldp + col);
        const int64_t row = sidx * p.L + a - r_lo;
        if (g != 0.0f && row >= 0 && row < rows_used) {
          const uint32_t r = static_cast<uint32_t>(row);
          const uint32_t off = r * 128u + ((static_cast<uint32_t>(fl >> 3) ^ (r & 7u)) << 4) + static_cast<uint32_t>(fl & 7) * 2u;
          *reinterpret_cast<__nv_bfloat16*>(slot + off) = clear ? __float2bfloat16_rn(0.0f) : __float2bfloat16_rn(g);
        }
      }
    };
    uint32_t xphase = 0;
    bool has_prev = false;
    int64_t pr_lo = 0, ps_first = 0; int pnseq = 0, pgcol = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      int64_t r_lo, s_first; int nseq;
      tile_geom(tile, r_lo, s_first, nseq);
      for (int w = 0; w < p.nwin; ++w) {
        for (int fb = 0; fb < p.fblocks; ++fb) {
          const int f = fb * 64 + fl;
          uint8_t* slot = xbase + static_cast<size_t>(fb) * x_slot_bytes;
          mbar_wait(empty_bar(p.wstages + fb), xphase ^ 1u);
          if (has_prev && f < p.F) scatter(slot, pgcol + f, pr_lo, ps_first, pnseq, true);
          asm volatile("bar.sync 2, %0;" ::"n"(NB) : "memory");
          if (f < p.F) scatter(slot, p.gcol[w] + f, r_lo, s_first, nseq, false);
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(full_bar(p.wstages + fb));
        }
        xphase ^= 1u;
        has_prev = true; pr_lo = r_lo; ps_first = s_first; pnseq = nseq; pgcol = p.gcol[w];
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

// =====================================================================================================
// dW on the tensor cores:  dWt[f, (j,c)] = sum_r dY[r, f] * X[r + j, c]   (r = flat token row; dY[r, f] =
// G[s,f] at the arg-max row of (s,f), zero elsewhere).  The reduction runs over TOKENS, so
//   A operand = dY^T tile (128 filters x 64 tokens per M tile, K-major), built in swizzled shared memory by
//               the builder warps exactly like the Gt tile above (one non-zero per (sequence, filter));
//   B operand = the token tile itself, X[r0+j .. r0+j+63, c0 .. c0+127], whose K dimension (tokens) is the
//               SLOW dimension of X: an MN-major operand.  A TMA box of 64 token rows x 64 columns with the
//               128-byte swizzle is precisely the canonical MN-major SW128 layout (8-row K atoms 1024 B apart
//               = SBO, 64-column blocks one box apart = LBO), so the rows of X feed the tensor core untouched;
//               the window offset j is again just the TMA row coordinate.
// Each CTA owns one (j, 128-column range) slice of dW for all filters (3 M tiles x 128 TMEM columns) and one
// split of the token range; partial sums go to the same part[split][f][j][c] buffer the CUDA-core kernels use
// and are reduced in fixed order by conv_bwd_dw_reduce_kernel.
struct DwParams {
  int64_t n_rows;
  int L, k, F, Dpad;
  int n_mt;                    // filter M tiles
  int n_cr;                    // 128-column ranges of Dpad
  int n_slices;                // k * n_cr
  int64_t chunks_total, chunks_per_split;
  const float* G; const int32_t* A; int ldp, col0;
  float* part;
  int stages;
  long long* prof;             // optional cycle counters (CLSM_CONV_BWD_PROF=1), else nullptr
};



constexpr int DW_NT = 128;          // (j, c) columns per CTA = UMMA N
constexpr int DW_KT = 64;           // tokens per chunk
constexpr int DW_BLD_WARPS = 8;     // warps 2..9: builders during the main loop, epilogue at the end
constexpr int DW_THREADS = 64 + DW_BLD_WARPS * 32;

template <int MAXSEQ>          // sequences a 64-token chunk can touch: 2 when L >= 64, else up to 5 (L >= 16)
__global__ void __launch_bounds__(DW_THREADS, 1)
dw_tc_kernel(const __grid_constant__ CUtensorMap tmX, const DwParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_bytes = static_cast<uint32_t>(p.n_mt) * MT_BYTES;        // dY^T: n_mt x (128 f x 64 tokens)
  const uint32_t b_bytes = 2u * DW_KT * 128u;                               // X: two boxes of 64 tokens x 64 columns
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t bars = base + static_cast<uint32_t>(p.stages) * stage_bytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (p.stages + s); };
  const uint32_t tfull_bar = bars + 8u * (2 * p.stages);
  const uint32_t tmem_slot = bars + 8u * (2 * p.stages + 1);

  const int warp = warp_id_uniform(), lane = threadIdx.x % 32;
  const int split = blockIdx.x / p.n_slices;
  const int slice = blockIdx.x - split * p.n_slices;
  const int j = slice / p.n_cr, c0 = (slice - j * p.n_cr) * DW_NT;
  const int64_t ch_begin = static_cast<int64_t>(split) * p.chunks_per_split;
  const int64_t ch_end = (ch_begin + p.chunks_per_split < p.chunks_total) ? ch_begin + p.chunks_per_split : p.chunks_total;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar(s), 1 + DW_BLD_WARPS); mbar_init(empty_bar(s), 1); }
    mbar_init(tfull_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (warp == 0 && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmX) : "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int64_t ch = ch_begin; ch < ch_end; ++ch) {
        const int r0 = static_cast<int>(ch * DW_KT);
        mbar_wait(empty_bar(stage), phase ^ 1u);
        mbar_expect_tx(full_bar(stage), b_bytes);
        const uint32_t b_dst = base + stage * stage_bytes + a_bytes;
        tma_load_2d(b_dst, &tmX, c0, r0 + j, full_bar(stage));
        tma_load_2d(b_dst + DW_KT * 128u, &tmX, c0 + 64, r0 + j, full_bar(stage));
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    {
      const uint32_t idesc = make_idesc(DW_NT) | (1u << 16);                 // B operand MN-major
      int stage = 0; uint32_t phase = 0;
      bool first = true;
      for (int64_t ch = ch_begin; ch < ch_end; ++ch) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t a_lo = sw128_desc_lo(base + stage * stage_bytes);
        const uint32_t b_lo = sw128_mn_desc_lo(base + stage * stage_bytes + a_bytes, DW_KT * 128u);
#pragma unroll
        for (int ks = 0; ks < DW_KT / 16; ++ks) {
          const uint32_t accum = (first && ks == 0) ? 0u : 1u;            // 16 tokens = two 8-row K atoms = 2048 B
          for (int m = 0; m < p.n_mt; ++m)
            tc_mma_bf16_u(tmem_base + static_cast<uint32_t>(m * DW_NT), a_lo + static_cast<uint32_t>(m * (MT_BYTES / 16) + ks * 2),
                          b_lo + ks * 128, idesc, accum);
        }
        first = false;
        tc_commit_u(empty_bar(stage));
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
      tc_commit_u(tfull_bar);
    }
  } else {
    // ---- builders: dY^T tile of every chunk.  The (G, arg-max) pairs a thread needs for chunk ch+1 are
    // fetched into registers while it scatters chunk ch, so no global-load latency sits on the critical path.
    const int bt = threadIdx.x - 64;
    constexpr int NB = DW_BLD_WARPS * 32;
    constexpr int MAXFI = 2;                           // F <= 512
    uint32_t phase = 0;
    const uint32_t zero_chunks = a_bytes / 16;
    float g_nxt[MAXSEQ * MAXFI]; int a_nxt[MAXSEQ * MAXFI];
    // (first sequence, offset of the chunk's first token inside it) of the chunk being PREFETCHED and of the
    // chunk being BUILT, advanced by 64 tokens per chunk -- no division on the per-chunk path
    int64_t ps = (ch_begin * DW_KT) / p.L;
    int pt = static_cast<int>(ch_begin * DW_KT - ps * p.L);
    int64_t cs = ps; int ct = pt;
    const int64_t S_total = p.n_rows / p.L;
    auto prefetch = [&](int64_t ch) {
      // sequences touched: tokens [pt, pt+63] of sequence ps onward
      const int nseq = (pt + DW_KT - 1) / p.L + 1;
      const float* gp = p.G + ps * p.ldp + p.col0 + bt;
      const int32_t* ap = p.A + ps * p.ldp + p.col0 + bt;
#pragma unroll
      for (int sq = 0; sq < MAXSEQ; ++sq)
#pragma unroll
        for (int fi = 0; fi < MAXFI; ++fi) {
          const bool ok = ch < ch_end && sq < nseq && ps + sq < S_total && bt + fi * NB < p.F;
          g_nxt[sq * MAXFI + fi] = ok ? __ldg(gp + static_cast<int64_t>(sq) * p.ldp + fi * NB) : 0.0f;
          a_nxt[sq * MAXFI + fi] = ok ? __ldg(ap + static_cast<int64_t>(sq) * p.ldp + fi * NB) : 0;
        }
      pt += DW_KT;
      while (pt >= p.L) { pt -= p.L; ++ps; }
    };
    // The tiles start all-zero and every thread un-writes exactly what IT wrote the last time the stage was
    // used (its candidates never collide with another thread's), so no 48 KB re-zeroing and no barrier per chunk.
    for (int st = 0; st < p.stages; ++st) {
      uint8_t* atile = base_ptr + st * stage_bytes;
      for (uint32_t i = bt; i < zero_chunks; i += NB) reinterpret_cast<uint4*>(atile)[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(NB) : "memory");
    constexpr uint32_t NONE = 0xFFFFFFFFu;
    uint32_t wrote[3][MAXSEQ * MAXFI];                 // byte offsets written per stage (p.stages == 3)
#pragma unroll
    for (int st = 0; st < 3; ++st)
#pragma unroll
      for (int i = 0; i < MAXSEQ * MAXFI; ++i) wrote[st][i] = NONE;
    prefetch(ch_begin);
    for (int64_t chb = ch_begin; chb < ch_end; chb += 3) {
#pragma unroll
      for (int st = 0; st < 3; ++st) {
        const int64_t ch = chb + st;
        if (ch < ch_end) {
          float g_cur[MAXSEQ * MAXFI]; int a_cur[MAXSEQ * MAXFI];
#pragma unroll
          for (int i = 0; i < MAXSEQ * MAXFI; ++i) { g_cur[i] = g_nxt[i]; a_cur[i] = a_nxt[i]; }
          prefetch(ch + 1);
          mbar_wait(empty_bar(st), phase ^ 1u);
          uint8_t* atile = base_ptr + st * stage_bytes;
#pragma unroll
          for (int i = 0; i < MAXSEQ * MAXFI; ++i)
            if (wrote[st][i] != NONE) *reinterpret_cast<uint16_t*>(atile + wrote[st][i]) = 0;
#pragma unroll
          for (int sq = 0; sq < MAXSEQ; ++sq)
#pragma unroll
            for (int fi = 0; fi < MAXFI; ++fi) {
              const float g = g_cur[sq * MAXFI + fi];
              const int f = bt + fi * NB;
              const int n = sq * p.L + a_cur[sq * MAXFI + fi] - ct;          // token column inside the chunk
              uint32_t off = NONE;
              if (g != 0.0f && n >= 0 && n < DW_KT) {
                const uint32_t fl = static_cast<uint32_t>(f) & 127u;
                off = static_cast<uint32_t>(f >> 7) * MT_BYTES + fl * 128u +
                      (((static_cast<uint32_t>(n) >> 3) ^ (fl & 7u)) << 4) + (static_cast<uint32_t>(n) & 7u) * 2u;
                *reinterpret_cast<__nv_bfloat16*>(atile + off) = __float2bfloat16_rn(g);
              }
              wrote[st][sq * MAXFI + fi] = off;
            }
          ct += DW_KT;
          while (ct >= p.L) { ct -= p.L; ++cs; }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(full_bar(st));
        }
      }
      phase ^= 1u;
    }
    // ---- epilogue (same warps): D[f, n] -> part[split][f][j][c0 + n]
    const int lq = warp & 3;
    const int half = (warp - 2) >> 2;                      // 64 of the 128 columns each
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    for (int m = 0; m < p.n_mt; ++m) {
      const int f = m * MT + lq * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + static_cast<uint32_t>(m * DW_NT + half * 64);
      uint32_t v[4][16];
#pragma unroll
      for (int q = 0; q < 4; ++q) tc_ld16_async(taddr + q * 16, v[q]);
      tc_wait_ld();
      if (f < p.F) {
        float* dst = p.part + ((static_cast<int64_t>(split) * p.F + f) * p.k + j) * p.Dpad + c0 + half * 64;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
#pragma unroll
          for (int i = 0; i < 16; i += 4) {
            const int n = q * 16 + i;
            if (c0 + half * 64 + n < p.Dpad)
              *reinterpret_cast<float4*>(dst + n) = make_float4(__uint_as_float(v[q][i]), __uint_as_float(v[q][i + 1]),
                                                                 __uint_as_float(v[q][i + 2]), __uint_as_float(v[q][i + 3]));
          }
        }
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

// ---- dW, second tiling (k * 128 <= 512 accumulator columns, i.e. k <= 4): a CTA owns ONE filter M tile and one
// 128-column slice of the token rows for ALL k window offsets.  The k accumulators D_j[f, c] share the dY^T tile
// (built once per chunk, for 128 filters only -- a third of the builder work of the per-offset tiling above, and
// each CTA's builders touch only their own filters) and the token box: 64 + 8 rows are loaded once and offset j
// is the same MN-major descriptor advanced by j rows (the swizzle is a function of absolute address bits).
// The k accumulators are independent, so consecutive MMAs never wait on each other's result.
constexpr int DW2_XROWS = DW_KT + 8;                        // token rows per box: the chunk plus the window halo
constexpr uint32_t DW2_BOX_BYTES = DW2_XROWS * 128u;        // 9216 = 9 swizzle atoms

template <int MAXSEQ, int NW>        // NW = token-row columns per CTA (UMMA N): 128, or 160 when Dpad is a multiple of 160
__global__ void __launch_bounds__(DW_THREADS, 1)
dw2_tc_kernel(const __grid_constant__ CUtensorMap tmX, const DwParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_bytes = MT_BYTES;                                          // dY^T: 128 f x 64 tokens
  constexpr int NBOX = (NW + 63) / 64;
  const uint32_t b_bytes = NBOX * DW2_BOX_BYTES;                              // X: boxes of 72 tokens x 64 columns
  const uint32_t stage_bytes = a_bytes + ((b_bytes + 1023u) & ~1023u);
  const uint32_t bars = base + static_cast<uint32_t>(p.stages) * stage_bytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (p.stages + s); };
  const uint32_t tfull_bar = bars + 8u * (2 * p.stages);
  const uint32_t tmem_slot = bars + 8u * (2 * p.stages + 1);

  const int warp = warp_id_uniform(), lane = threadIdx.x % 32;
  const int split = blockIdx.x / p.n_slices;
  const int slice = blockIdx.x - split * p.n_slices;
  const int mt = slice / p.n_cr, c0 = (slice - mt * p.n_cr) * NW;           // filter M tile, column slice
  const int64_t ch_begin = static_cast<int64_t>(split) * p.chunks_per_split;
  const int64_t ch_end = (ch_begin + p.chunks_per_split < p.chunks_total) ? ch_begin + p.chunks_per_split : p.chunks_total;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar(s), 1 + DW_BLD_WARPS / 2); mbar_init(empty_bar(s), 1); }
    mbar_init(tfull_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (warp == 0 && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmX) : "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int64_t ch = ch_begin; ch < ch_end; ++ch) {
        const int r0 = static_cast<int>(ch * DW_KT);
        mbar_wait(empty_bar(stage), phase ^ 1u);
        mbar_expect_tx(full_bar(stage), b_bytes);
        const uint32_t b_dst = base + stage * stage_bytes + a_bytes;
#pragma unroll
        for (int b = 0; b < NBOX; ++b) tma_load_2d(b_dst + b * DW2_BOX_BYTES, &tmX, c0 + 64 * b, r0, full_bar(stage));
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    {
      const uint32_t idesc = make_idesc(NW) | (1u << 16);                 // B operand MN-major
      int stage = 0; uint32_t phase = 0;
      bool first = true;
      long long w_full = 0;
      const long long t_begin = clock64();
      for (int64_t ch = ch_begin; ch < ch_end; ++ch) {
        const long long c0 = clock64();
        mbar_wait(full_bar(stage), phase);
        w_full += clock64() - c0;
        tc_fence_after();
        const uint32_t a_lo = sw128_desc_lo(base + stage * stage_bytes);
        const uint32_t b_lo = sw128_mn_desc_lo(base + stage * stage_bytes + a_bytes, DW2_BOX_BYTES);
#pragma unroll
        for (int ks = 0; ks < DW_KT / 16; ++ks) {
          const uint32_t accum = (first && ks == 0) ? 0u : 1u;
          for (int j = 0; j < p.k; ++j)                                        // token rows 16 ks + j .. of the box
            tc_mma_bf16_u(tmem_base + static_cast<uint32_t>(j * NW), a_lo + ks * 2, b_lo + ks * 128 + static_cast<uint32_t>(j) * 8u, idesc, accum);
        }
        first = false;
        tc_commit_u(empty_bar(stage));
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
      tc_commit_u(tfull_bar);
      if (p.prof != nullptr && lane == 0) {
        mbar_wait(tfull_bar, 0);
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 0), static_cast<unsigned long long>(clock64() - t_begin));
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 1), static_cast<unsigned long long>(w_full));
      }
    }
  } else {
    // ---- builders: dY^T of the CTA's 128 filters.  The eight warps form two groups of 128 threads (thread =
    // filter); group g builds the chunks of parity g, so a group has two MMA periods per tile.  With FOUR stages
    // group g only ever touches stages g and g + 2: a thread owns the locations it writes and "un-write what I
    // wrote the last time this stage was used, then write" needs no barrier between builder threads.
    const int bt = threadIdx.x - 64;
    const int fl = bt & 127, grp = bt >> 7;
    const int f = mt * MT + fl;
    constexpr int NS = MAXSEQ;
    const uint32_t zero_chunks = a_bytes / 16;
    constexpr int NB = DW_BLD_WARPS * 32;
    float g_nxt[NS]; int a_nxt[NS];
    const int64_t tok0 = (ch_begin + grp) * DW_KT;            // first token of this group's first chunk
    int64_t ps = tok0 / p.L;
    int pt = static_cast<int>(tok0 - ps * p.L);
    int ct = pt;
    const int64_t S_total = p.n_rows / p.L;
    auto prefetch = [&](int64_t ch) {
      const int nseq = (pt + DW_KT - 1) / p.L + 1;
      const float* gp = p.G + ps * p.ldp + p.col0 + f;
      const int32_t* ap = p.A + ps * p.ldp + p.col0 + f;
#pragma unroll
      for (int sq = 0; sq < NS; ++sq) {
        const bool ok = ch < ch_end && sq < nseq && ps + sq < S_total && f < p.F;
        g_nxt[sq] = ok ? __ldg(gp + static_cast<int64_t>(sq) * p.ldp) : 0.0f;
        a_nxt[sq] = ok ? __ldg(ap + static_cast<int64_t>(sq) * p.ldp) : 0;
      }
      pt += 2 * DW_KT;
      while (pt >= p.L) { pt -= p.L; ++ps; }
    };
    for (int st = 0; st < p.stages; ++st) {
      uint8_t* atile = base_ptr + st * stage_bytes;
      for (uint32_t i = bt; i < zero_chunks; i += NB) reinterpret_cast<uint4*>(atile)[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(NB) : "memory");
    constexpr uint32_t NONE = 0xFFFFFFFFu;
    uint32_t wrote[2][NS];                               // per owned stage (g, g + 2): byte offsets written last time
#pragma unroll
    for (int sl = 0; sl < 2; ++sl)
#pragma unroll
      for (int i = 0; i < NS; ++i) wrote[sl][i] = NONE;
    prefetch(ch_begin + grp);
    long long w_empty = 0;
    const long long tb_begin = clock64();
    uint32_t phase = 0;
    for (int64_t chb = ch_begin + grp; chb < ch_end; chb += 4) {
#pragma unroll
      for (int sl = 0; sl < 2; ++sl) {
        const int64_t ch = chb + 2 * sl;
        const int st = grp + 2 * sl;
        if (ch < ch_end) {
          float g_cur[NS]; int a_cur[NS];
#pragma unroll
          for (int i = 0; i < NS; ++i) { g_cur[i] = g_nxt[i]; a_cur[i] = a_nxt[i]; }
          prefetch(ch + 2);
          const long long c0 = clock64();
          mbar_wait(empty_bar(st), phase ^ 1u);
          w_empty += clock64() - c0;
          uint8_t* atile = base_ptr + st * stage_bytes;
#pragma unroll
          for (int i = 0; i < NS; ++i)
            if (wrote[sl][i] != NONE) *reinterpret_cast<uint16_t*>(atile + wrote[sl][i]) = 0;
#pragma unroll
          for (int sq = 0; sq < NS; ++sq) {
            const int n = sq * p.L + a_cur[sq] - ct;                         // token column inside the chunk
            uint32_t off = NONE;
            if (g_cur[sq] != 0.0f && n >= 0 && n < DW_KT) {
              off = static_cast<uint32_t>(fl) * 128u + (((static_cast<uint32_t>(n) >> 3) ^ (static_cast<uint32_t>(fl) & 7u)) << 4) +
                    (static_cast<uint32_t>(n) & 7u) * 2u;
              *reinterpret_cast<__nv_bfloat16*>(atile + off) = __float2bfloat16_rn(g_cur[sq]);
            }
            wrote[sl][sq] = off;
          }
          ct += 2 * DW_KT;
          while (ct >= p.L) ct -= p.L;
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(full_bar(st));
        }
      }
      phase ^= 1u;
    }
    if (p.prof != nullptr && bt == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 2), static_cast<unsigned long long>(clock64() - tb_begin));
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 3), static_cast<unsigned long long>(w_empty));
    }
    // ---- epilogue (same warps): D_j[f, n] -> part[split][f][j][c0 + n]
    const int lq = warp & 3;
    const int half = (warp - 2) >> 2;                      // half of the NW columns each
    constexpr int HW = NW / 2, NQ = HW / 16;
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    const int fo = mt * MT + lq * 32 + lane;
    for (int j = 0; j < p.k; ++j) {
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + static_cast<uint32_t>(j * NW + half * HW);
      uint32_t v[NQ][16];
#pragma unroll
      for (int q = 0; q < NQ; ++q) tc_ld16_async(taddr + q * 16, v[q]);
      tc_wait_ld();
      if (fo < p.F) {
        float* dst = p.part + ((static_cast<int64_t>(split) * p.F + fo) * p.k + j) * p.Dpad + c0 + half * HW;
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
#pragma unroll
          for (int i = 0; i < 16; i += 4) {
            const int n = q * 16 + i;
            if (c0 + half * HW + n < p.Dpad)
              *reinterpret_cast<float4*>(dst + n) = make_float4(__uint_as_float(v[q][i]), __uint_as_float(v[q][i + 1]),
                                                                 __uint_as_float(v[q][i + 2]), __uint_as_float(v[q][i + 3]));
          }
        }
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

// 1 when the resident-tile dX kernel (the only one that can write bf16 dX rows) takes this shape
int clsm_conv_bwd_dx_tc_tiling2_ok(int k, int Dpad, int F, int64_t n_rows) {
  const char* env = getenv("CLSM_CONV_BWD_DX");
  if (env != nullptr && (env[0] == 's' || env[0] == '1')) return 0;
  if ((Dpad + MT - 1) / MT > 4 || n_rows + 512 >= (1ll << 31)) return 0;
  int nt = 224;
  while (nt > 64 && nt + k - 1 > 256) nt -= 32;
  const int xrows = (nt + k - 1 + 7) / 8 * 8;
  const int x_slot = (xrows * 128 + 1023) / 1024 * 1024;
  const int wstages = (SMEM_LIMIT - 2048 - ((F + 63) / 64) * x_slot) / MT_BYTES;
  return wstages >= 3 ? 1 : 0;
}

int clsm_conv_bwd_dx_tc(const void* Wt, int dtype, int k, int Dpad, int L, int64_t S, const float* G, const int32_t* A,
                        int ldp, int col0, int F, float* dX, int accumulate, int dx_bf16, void* scratch, cudaStream_t st) {
  if (dtype != CLSM_BF16) return CLSM_EUNSUPPORTED;
  const char* env = getenv("CLSM_CONV_BWD_DX");
  if (env != nullptr && env[0] == 's') return CLSM_EUNSUPPORTED;       // force the CUDA-core path
  if (dx_bf16 && (accumulate || !clsm_conv_bwd_dx_tc_tiling2_ok(k, Dpad, F, S * L))) return CLSM_EUNSUPPORTED;
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
  if (stages < 3) return CLSM_EUNSUPPORTED;
  stages = 3;                                   // the builders' un-write bookkeeping is laid out for 3 stages
  p.stages = stages;
  const int smem_bytes = stages * stage_bytes + 1024 + 8 * (2 * stages + 4);
  // sequences one tile can touch -> candidate slots per builder thread (128 threads, 64 filters per block)
  const int max_seq = (p.NT + L - 1) / L + 1;
  const int maxc = (max_seq * 64 + 127) / 128;

  // tiling 2 (resident dY tile) when its shared-memory plan fits; CLSM_CONV_BWD_DX=1 forces tiling 1
  Dx2Params q{};
  bool tiling2 = !(env != nullptr && env[0] == '1');
  if (tiling2) {
    int nt = 224;
    while (nt > 64 && nt + k - 1 > 256) nt -= 32;
    q.n_rows = p.n_rows; q.L = L; q.k = k; q.F = F; q.Dpad = Dpad; q.n_mt = p.n_mt; q.NT = nt; q.fblocks = p.fblocks;
    q.n_tiles = static_cast<int>(ceil_div64(p.n_rows, nt));
    q.G = G; q.A = A; q.ldp = ldp; q.col0 = col0; q.dX = dX; q.accumulate = accumulate; q.dx_bf16 = dx_bf16;
    q.last_ks = (F - (q.fblocks - 1) * 64 + 15) / 16;
    q.xrows = (nt + k - 1 + 7) / 8 * 8;
    q.nreg = 512 / nt; if (q.nreg > 4) q.nreg = 4;
    const int x_slot = (q.xrows * 128 + 1023) / 1024 * 1024;
    q.wstages = (SMEM_LIMIT - 2048 - q.fblocks * x_slot) / MT_BYTES;
    if (q.wstages > 8) q.wstages = 8;
    if (q.wstages < 3) tiling2 = false;
  }
  if (!tiling2 && maxc > 5) return CLSM_EUNSUPPORTED;

  __nv_bfloat16* Wb = static_cast<__nv_bfloat16*>(scratch);
  const int64_t n_wb = static_cast<int64_t>(Dpad) * k * Fk;
  pack_wb_kernel<<<static_cast<unsigned>(ceil_div64(n_wb, 256)), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(Wt), k, Dpad, F, Fk, Wb, static_cast<int64_t>(k) * Fk, 0);
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
  if (grid > (tiling2 ? q.n_tiles : p.n_tiles)) grid = tiling2 ? q.n_tiles : p.n_tiles;
  static int prof_on = -1;
  if (prof_on < 0) { const char* e = getenv("CLSM_CONV_BWD_PROF"); prof_on = (e != nullptr && e[0] == '1') ? 1 : 0; }
  if (prof_on) {
    void* sym = nullptr;
    CLSM_CUDA(cudaGetSymbolAddress(&sym, g_conv_bwd_prof));
    p.prof = static_cast<long long*>(sym);
    CLSM_CUDA(cudaMemsetAsync(sym, 0, sizeof(long long) * 8, st));
  }
  if (tiling2) {
    q.prof = p.prof;
    const int x_slot = (q.xrows * 128 + 1023) / 1024 * 1024;
    const int stages2 = q.wstages + q.fblocks;
    const int smem2 = q.wstages * MT_BYTES + q.fblocks * x_slot + 1024 + 8 * (2 * stages2 + 9);
#define DX2_LAUNCH(N)                                                                                            \
  do {                                                                                                           \
    CLSM_CUDA(cudaFuncSetAttribute(dx2_tc_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));  \
    dx2_tc_kernel<N><<<grid, NUM_THREADS, smem2, st>>>(tmWb, q);                                                 \
  } while (0)
    switch (q.NT / 32) {
      case 2: DX2_LAUNCH(2); break;
      case 3: DX2_LAUNCH(3); break;
      case 4: DX2_LAUNCH(4); break;
      case 5: DX2_LAUNCH(5); break;
      case 6: DX2_LAUNCH(6); break;
      default: DX2_LAUNCH(7); break;
    }
#undef DX2_LAUNCH
    CLSM_LAUNCH_CHECK();
    if (prof_on) {
      long long h[8];
      CLSM_CUDA(cudaStreamSynchronize(st));
      CLSM_CUDA(cudaMemcpy(h, q.prof, sizeof(h), cudaMemcpyDeviceToHost));
      fprintf(stderr, "[dx_tc prof] tiling 2 NT %d tiles %d grid %d wstages %d | MMA thread: total %.0f, wait TMEM %.0f, wait operands %.0f (cycles per CTA)\n",
              q.NT, q.n_tiles, grid, q.wstages, h[4] / static_cast<double>(grid), h[5] / static_cast<double>(grid), h[6] / static_cast<double>(grid));
    }
    return CLSM_OK;
  }
  const int nch = p.NT / 32;
#define DX_LAUNCH(N, C)                                                                                          \
  do {                                                                                                           \
    CLSM_CUDA(cudaFuncSetAttribute(dx_tc_kernel<N, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)); \
    dx_tc_kernel<N, C><<<grid, NUM_THREADS, smem_bytes, st>>>(tmWb, p);                                          \
  } while (0)
  if (maxc <= 2) { if (nch == 4) DX_LAUNCH(4, 2); else if (nch == 5) DX_LAUNCH(5, 2); else DX_LAUNCH(8, 2); }
  else           { if (nch == 4) DX_LAUNCH(4, 5); else if (nch == 5) DX_LAUNCH(5, 5); else DX_LAUNCH(8, 5); }
#undef DX_LAUNCH
  CLSM_LAUNCH_CHECK();
  if (prof_on) {           // diagnostics only: synchronises and prints average cycles per CTA
    long long h[8];
    CLSM_CUDA(cudaStreamSynchronize(st));
    CLSM_CUDA(cudaMemcpy(h, p.prof, sizeof(h), cudaMemcpyDeviceToHost));
    fprintf(stderr, "[dx_tc prof] NT %d tiles %d grid %d maxc %d | MMA thread: total %.0f, wait TMEM %.0f, wait operands %.0f (cycles per CTA)\n",
            p.NT, p.n_tiles, grid, maxc, h[4] / static_cast<double>(grid), h[5] / static_cast<double>(grid), h[6] / static_cast<double>(grid));
  }
  return CLSM_OK;
}

// ---- dX of several windows in one pass (dxm_tc_kernel) -------------------------------------------------------
static bool dxm_plan(int n_windows, const int* ks, int Dpad, int F, int64_t n_rows, DxmParams* out) {
  const char* env = getenv("CLSM_CONV_BWD_DX");
  if (env != nullptr && (env[0] == 's' || env[0] == '1')) return false;
  if (n_windows < 2 || n_windows > 4 || Dpad % 64 != 0 || F < 1 || F > 512) return false;
  DxmParams q{};
  q.n_mt = (Dpad + MT - 1) / MT;
  if (q.n_mt > 4 || n_rows + 512 >= (1ll << 31)) return false;
  q.kmax = 0;
  int ktot = 0;
  for (int w = 0; w < n_windows; ++w) { if (ks[w] < 1 || ks[w] > 7) return false; if (ks[w] > q.kmax) q.kmax = ks[w]; ktot += ks[w]; }
  q.fblocks = (F + 63) / 64;
  int nt = (512 / q.n_mt) / 32 * 32;                 // all n_mt regions of a tile are live at once
  if (nt > 224) nt = 224;
  q.NT = nt; q.nwin = n_windows; q.F = F; q.Dpad = Dpad; q.n_rows = n_rows;
  q.n_tiles = static_cast<int>(ceil_div64(n_rows, nt));
  q.last_ks = (F - (q.fblocks - 1) * 64 + 15) / 16;
  q.xrows = (nt + q.kmax - 1 + 7) / 8 * 8;
  const int x_slot = (q.xrows * 128 + 1023) / 1024 * 1024;
  q.wstages = (SMEM_LIMIT - 2048 - q.fblocks * x_slot) / MT_BYTES;
  if (q.wstages > 8) q.wstages = 8;
  if (q.wstages < 3) return false;
  int col = 0;
  for (int w = 0; w < n_windows; ++w) { q.k[w] = ks[w]; q.wcol[w] = col; q.gcol[w] = w * F; col += ks[w] * q.fblocks * 64; }
  (void)ktot;
  *out = q;
  return true;
}

int clsm_conv_bwd_dx_multi_tc_ok(int n_windows, const int* ks, int Dpad, int F, int64_t n_rows) {
  DxmParams q;
  return dxm_plan(n_windows, ks, Dpad, F, n_rows, &q) ? 1 : 0;
}

size_t clsm_conv_bwd_dx_multi_tc_bytes(int n_windows, const int* ks, int Dpad, int F) {
  size_t ktot = 0;
  for (int w = 0; w < n_windows; ++w) ktot += static_cast<size_t>(ks[w]);
  return align_up(static_cast<size_t>(Dpad) * ktot * (static_cast<size_t>((F + 63) / 64) * 64) * 2, 256);
}

// Wt[w]: window w's compute weights (Fpad, k_w * Dpad) bf16; G / A: (S, ldp) with window w at columns col0 + w * F.
int clsm_conv_bwd_dx_multi_tc(const void* const* Wt, const int* ks, int n_windows, int Dpad, int L, int64_t S, const float* G,
                              const int32_t* A, int ldp, int col0, int F, void* dX, int dx_bf16, void* scratch, cudaStream_t st) {
  DxmParams q;
  if (!dxm_plan(n_windows, ks, Dpad, F, S * L, &q)) return CLSM_EUNSUPPORTED;
  q.L = L; q.G = G; q.A = A; q.ldp = ldp; q.dX = dX; q.dx_bf16 = dx_bf16;
  for (int w = 0; w < n_windows; ++w) q.gcol[w] += col0;
  const int Fk = q.fblocks * 64;
  int64_t ldw = 0;
  for (int w = 0; w < n_windows; ++w) ldw += static_cast<int64_t>(ks[w]) * Fk;
  __nv_bfloat16* Wb = static_cast<__nv_bfloat16*>(scratch);
  for (int w = 0; w < n_windows; ++w) {
    const int64_t n_wb = static_cast<int64_t>(Dpad) * ks[w] * Fk;
    pack_wb_kernel<<<static_cast<unsigned>(ceil_div64(n_wb, 256)), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(Wt[w]), ks[w], Dpad, F, Fk, Wb, ldw, q.wcol[w]);
    CLSM_LAUNCH_CHECK();
  }
  EncodeTiledFn enc = get_encode();
  if (!enc) { clsm_set_error("conv_pool_bwd(tcgen05): cuTensorMapEncodeTiled unavailable"); return CLSM_ECUDA; }
  CUtensorMap tmWb;
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(ldw), static_cast<cuuint64_t>(Dpad)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ldw) * 2};
  cuuint32_t box[2] = {KCHUNK, MT};
  cuuint32_t estr[2] = {1, 1};
  if (enc(&tmWb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Wb, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
    clsm_set_error("conv_pool_bwd(tcgen05): cuTensorMapEncodeTiled failed");
    return CLSM_ECUDA;
  }
  int grid = clsm_sm_count();
  if (grid > q.n_tiles) grid = q.n_tiles;
  const int x_slot = (q.xrows * 128 + 1023) / 1024 * 1024;
  const int stages = q.wstages + q.fblocks;
  const int smem = q.wstages * MT_BYTES + q.fblocks * x_slot + 1024 + 8 * (2 * stages + 9);
#define DXM_LAUNCH(N)                                                                                            \
  do {                                                                                                           \
    CLSM_CUDA(cudaFuncSetAttribute(dxm_tc_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));  \
    dxm_tc_kernel<N><<<grid, NUM_THREADS, smem, st>>>(tmWb, q);                                                  \
  } while (0)
  switch (q.NT / 32) {
    case 4: DXM_LAUNCH(4); break;
    case 5: DXM_LAUNCH(5); break;
    default: DXM_LAUNCH(7); break;
  }
#undef DXM_LAUNCH
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

// dW on the tensor cores.  Returns the number of token splits written to `part` (>= 1), or a negative code.
int clsm_conv_bwd_dw_tc(const void* X, int dtype, int L, int Dpad, int64_t S, int max_splits, const float* G,
                        const int32_t* A, int ldp, int col0, int F, int k, float* part, cudaStream_t st) {
  if (dtype != CLSM_BF16) return CLSM_EUNSUPPORTED;
  const char* env = getenv("CLSM_CONV_BWD_DW");
  if (env != nullptr && env[0] == 's') return CLSM_EUNSUPPORTED;       // force the CUDA-core path
  DwParams p{};
  p.n_rows = S * L; p.L = L; p.k = k; p.F = F; p.Dpad = Dpad;
  p.n_mt = (F + MT - 1) / MT;
  // tiling 2 (CTA = filter M tile x column slice, all window offsets) when the k accumulators fit TMEM; else
  // tiling 1 (CTA = window offset x column slice, all filters).  CLSM_CONV_BWD_DW=1 forces tiling 1.
  // Slice width of tiling 2: 160 columns when that divides Dpad (D = 300 -> Dpad = 320 = 2 x 160: no padded
  // columns, against 3 x 128 = 384), else 128.
  const bool force1 = env != nullptr && env[0] == '1';
  // Wider windows take narrower slices (k = 5: 5 x 96 = 480 accumulator columns; k = 6, 7: 64) rather than tiling 1,
  // which was 2x slower per FLOP at k = 5 (cfg4: 6.1 ms against 1.8 ms for k = 3).
  int nw = DW_NT;
  if (!force1) {
    if (Dpad % 160 == 0 && k * 160 <= 512) nw = 160;
    else if (k * 128 <= 512) nw = 128;
    else if (k * 96 <= 512) nw = 96;
    else nw = 64;
    const char* e = getenv("CLSM_CONV_BWD_DW_NW");             // experiments: 64 / 96 / 128 / 160 where k * nw fits
    if (e != nullptr) { const int v = atoi(e); if ((v == 64 || v == 96 || v == 128 || v == 160) && k * v <= 512) nw = v; }
  }
  const bool tiling2 = k * nw <= 512 && !force1;
  p.n_cr = tiling2 ? (Dpad + nw - 1) / nw : (Dpad + DW_NT - 1) / DW_NT;
  p.n_slices = (tiling2 ? p.n_mt : k) * p.n_cr;
  if ((!tiling2 && p.n_mt * DW_NT > 512) || p.n_rows + 512 >= (1ll << 31) || L < 16 || F > 512) return CLSM_EUNSUPPORTED;
  int splits = clsm_sm_count() / p.n_slices;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) return CLSM_EUNSUPPORTED;
  p.chunks_total = ceil_div64(p.n_rows, DW_KT);
  if (splits > p.chunks_total) splits = static_cast<int>(p.chunks_total);
  p.chunks_per_split = ceil_div64(p.chunks_total, splits);
  // no empty split: a CTA without chunks would issue no MMA and flush an unwritten accumulator
  splits = static_cast<int>(ceil_div64(p.chunks_total, p.chunks_per_split));
  p.G = G; p.A = A; p.ldp = ldp; p.col0 = col0; p.part = part;
  const int stage_bytes = tiling2 ? MT_BYTES + ((((nw + 63) / 64) * static_cast<int>(DW2_BOX_BYTES) + 1023) / 1024) * 1024
                                  : p.n_mt * MT_BYTES + 2 * DW_KT * 128;
  int stages = (SMEM_LIMIT - 2048) / stage_bytes;
  if (stages < (tiling2 ? 4 : 3)) return CLSM_EUNSUPPORTED;
  stages = tiling2 ? 4 : 3;                     // the builders' un-write bookkeeping is laid out for exactly this many
  p.stages = stages;
  const int smem_bytes = stages * stage_bytes + 1024 + 8 * (2 * stages + 3);
  EncodeTiledFn enc = get_encode();
  if (!enc) { clsm_set_error("conv_pool_bwd(tcgen05): cuTensorMapEncodeTiled unavailable"); return CLSM_ECUDA; }
  CUtensorMap tmX;
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(Dpad), static_cast<cuuint64_t>(p.n_rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(Dpad) * 2};
  cuuint32_t box[2] = {64, static_cast<cuuint32_t>(tiling2 ? DW2_XROWS : DW_KT)};
  cuuint32_t estr[2] = {1, 1};
  if (enc(&tmX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(X), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
    clsm_set_error("conv_pool_bwd(tcgen05): cuTensorMapEncodeTiled failed");
    return CLSM_ECUDA;
  }
  static int prof_on = -1;
  if (prof_on < 0) { const char* e = getenv("CLSM_CONV_BWD_PROF"); prof_on = (e != nullptr && e[0] == '1') ? 1 : 0; }
  if (prof_on) {
    void* sym = nullptr;
    CLSM_CUDA(cudaGetSymbolAddress(&sym, g_conv_bwd_prof));
    p.prof = static_cast<long long*>(sym);
    CLSM_CUDA(cudaMemsetAsync(sym, 0, sizeof(long long) * 8, st));
  }
#define DW2_LAUNCH(Q, W)                                                                                          \
  do {                                                                                                            \
    CLSM_CUDA(cudaFuncSetAttribute(dw2_tc_kernel<Q, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)); \
    dw2_tc_kernel<Q, W><<<splits * p.n_slices, DW_THREADS, smem_bytes, st>>>(tmX, p);                             \
  } while (0)
  if (tiling2) {
    if (L >= DW_KT) { if (nw == 160) DW2_LAUNCH(2, 160); else if (nw == 128) DW2_LAUNCH(2, 128); else if (nw == 96) DW2_LAUNCH(2, 96); else DW2_LAUNCH(2, 64); }
    else            { if (nw == 160) DW2_LAUNCH(5, 160); else if (nw == 128) DW2_LAUNCH(5, 128); else if (nw == 96) DW2_LAUNCH(5, 96); else DW2_LAUNCH(5, 64); }
  } else if (L >= DW_KT) {
    CLSM_CUDA(cudaFuncSetAttribute(dw_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    dw_tc_kernel<2><<<splits * p.n_slices, DW_THREADS, smem_bytes, st>>>(tmX, p);
  } else {
    CLSM_CUDA(cudaFuncSetAttribute(dw_tc_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    dw_tc_kernel<5><<<splits * p.n_slices, DW_THREADS, smem_bytes, st>>>(tmX, p);
  }
  CLSM_LAUNCH_CHECK();
  if (prof_on && tiling2) {           // diagnostics only: synchronises and prints average cycles per CTA
    long long h[8];
    CLSM_CUDA(cudaStreamSynchronize(st));
    CLSM_CUDA(cudaMemcpy(h, p.prof, sizeof(h), cudaMemcpyDeviceToHost));
    const double g = static_cast<double>(splits) * p.n_slices;
    fprintf(stderr, "[dw_tc prof] chunks/CTA %lld grid %d | MMA thread: total %.0f, wait operands %.0f | builder: total %.0f, "
            "wait free stage %.0f (cycles per CTA)\n", static_cast<long long>(p.chunks_per_split), splits * p.n_slices,
            h[0] / g, h[1] / g, h[2] / g, h[3] / g);
  }
  return splits;
}
