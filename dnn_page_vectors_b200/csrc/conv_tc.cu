// n-gram window convolution + max over time on the 5th-generation tensor cores (sm_100a).
//   replaces Convolution1D(valid) + MaxPooling1D + Flatten, dssm_cnn_v2/cnn_dssm_tf.py:84-100.
//
// Implicit GEMM, no im2col, computed TRANSPOSED:  D[f, t] = sum_{j,c} Wt[f, j*Dpad + c] * X[r0 + t + j, c].
//   A operand = filter weights Wt (Fpad, k*Dpad) K-major, in M tiles of 128 filters, streamed one
//               (M tile, j, c) chunk of 16 KB per shared-memory slot;
//   B operand = the gathered token matrix X (S*L, Dpad) bf16.  Default (resident-tile) mode: one TMA box of
//               NT + k - 1 token rows per 64-column chunk stays in shared memory for the whole tile; output
//               column t needs input rows t .. t+k-1, so window offset j is the SAME box read through a B
//               descriptor advanced by j rows (the UMMA swizzle works on absolute address bits).
// Filters sit on TMEM lanes and time on TMEM columns, so the max over time is a per-thread running
// maximum over the columns a thread reads back (tcgen05.ld 32x32b) -- no cross-lane reduction at all
// (a 32x32 cross-lane max costs ~700-850 cycles per warp whichever way it is done: measured
// CREDUX+ballot 852, 64-bit butterfly 688; this orientation needs ~100).  Sequence boundaries inside
// a tile are handled by flushing the running (max, first arg-max) at each boundary with one 64-bit
// atomicMax per (thread, sequence piece) into the (S, P) key array.  Pre-activations never reach HBM.
//
// Warp roles (320 threads, 1 CTA / SM, persistent over tiles):
//   warp 0 : TMA producer   (weight ring + token slots: full/empty mbarriers)
//   warp 1 : tcgen05.mma issuer, also owns the TMEM allocation.  Both run their loops with the WHOLE warp
//            (warp-uniform addresses -> uniform registers, see tc_ptx.cuh); one elected lane issues.
//   warps 2..9 : epilogue; TMEM lane quarter = warp % 4, column half = (warp - 2) / 4
// The (tile, M tile) units rotate through 512 / NT accumulator regions, so the epilogue of one unit overlaps
// the MMAs of the next.  Switches (A/B measurements): CLSM_CONV_TC_SHIFT=0 = first tiling (K-outer loop, all M
// tiles per stage, one accumulator set); CLSM_CONV_TC_VARIANT=1 = CTA pairs with multicast weights (first tiling).
#include "common.cuh"
#include "tc_ptx.cuh"
#include <stdlib.h>

namespace {

constexpr int MT = 128;                       // filters per M tile (UMMA M)
constexpr int KCHUNK = 64;                    // bf16 elements per 128-byte swizzle row
constexpr int MT_BYTES = MT * 128;            // one M tile of one K chunk: 16 KB
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_THREADS = 64 + NUM_EPI_WARPS * 32;
constexpr int SMEM_LIMIT = 227 * 1024;

struct TcParams {
  int64_t n_rows;        // S * L
  int L, k, F;
  int cpj;               // 64-element chunks per window offset (Dpad / 64)
  int n_mt;              // M tiles (ceil(F / 128))
  int NT;                // token rows per tile (UMMA N), multiple of 32, n_mt * NT <= 512
  int n_tiles;
  int ldk, col0;
  unsigned long long* keys;
  int wbox, n_wbox;      // TMA box rows for W and how many boxes per stage
  int stages;
  int shift_mode;        // 0: K-outer loop, one X box per (j, c) chunk, all M tiles per stage, one accumulator barrier.
                         // 1 (default): M-tile-outer loop over a RESIDENT token tile: one X box (NT + halo rows) per c chunk stays in
                         //    shared memory for the whole tile and serves every window offset j through a B descriptor whose
                         //    start is advanced by j rows (the UMMA swizzle works on absolute address bits, base_offset 0);
                         //    weights stream one (M tile, j, c) chunk per slot; each M tile's TMEM region has its own
                         //    full/empty barrier, so the epilogue of region m overlaps the MMAs of region m+1.
  int xrows;             // rows of the X box in mode 1 (NT + k - 1 rounded up to 8)
  int xstages, wstages;  // mode 1: xstages = cpj token slots, wstages weight slots of one M tile chunk (16 KB)
  int nreg;              // mode 1: TMEM accumulator regions of NT columns used round-robin by the (tile, M tile) units
  long long* prof;       // optional cycle counters (CLSM_CONV_TC_PROF=1), else nullptr
  int last_ks;           // K = 16 steps of the LAST 64-column chunk that hold real columns (D = 300: 304 = 19 x 16, so 3 of
                         // 4): the zero pad columns beyond them are skipped -- exact, and 5 % fewer MMAs
  const int32_t* tok;    // GATHER kernels: token ids (n_rows); tmX then maps the embedding TABLE and the token tile is
                         // fetched row by row with TMA gather4 (ids as row coordinates) -- X is never materialised
};

using namespace tcptx;

__device__ long long g_conv_tc_prof[8];      // CLSM_CONV_TC_PROF=1 diagnostics

template <bool CLUSTER, int NCH, bool GATHER = false>     // NCH = 16-column chunks per epilogue warp (NT / 32)
__global__ void __launch_bounds__(NUM_THREADS + 32, 1)    // GATHER adds an 11th warp: the token-row gather producer
conv_pool_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, const TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_bytes = static_cast<uint32_t>(p.n_mt) * MT_BYTES;           // weights of one K chunk
  const uint32_t b_bytes = static_cast<uint32_t>(p.NT) * 128u;                 // token rows of one K chunk
  const uint32_t stage_bytes = a_bytes + ((b_bytes + 1023u) & ~1023u);
  const uint32_t data_bytes = p.shift_mode == 0
      ? static_cast<uint32_t>(p.stages) * stage_bytes
      : static_cast<uint32_t>(p.wstages) * MT_BYTES + static_cast<uint32_t>(p.xstages) * (((static_cast<uint32_t>(p.xrows) * 128u) + 1023u) & ~1023u);
  const uint32_t bars = base + data_bytes;
  // barrier layout: full[stages] | empty[stages] | tfull[4] | tempty[4] | tmem ptr   (mode 0 uses tfull[0]/tempty[0])
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (p.stages + s); };
  auto tfull_bar = [&](int m) { return bars + 8u * (2 * p.stages + m); };
  auto tempty_bar = [&](int m) { return bars + 8u * (2 * p.stages + 4 + m); };
  const uint32_t tmem_slot = bars + 8u * (2 * p.stages + 8);

  const int warp = warp_id_uniform(), lane = threadIdx.x % 32;   // warp index as a provably warp-uniform value

  if (threadIdx.x == 0) {
    // weight slots of a CTA pair are freed by BOTH CTAs' MMAs; token slots (mode 1: s >= wstages) are CTA-local
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), (CLUSTER && (p.shift_mode == 0 || s < p.wstages)) ? 2 : 1);
    }
    for (int m = 0; m < 4; ++m) { mbar_init(tfull_bar(m), 1); mbar_init(tempty_bar(m), NUM_EPI_WARPS); }
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
  if (CLUSTER) cluster_sync_all();                  // the peer's barriers exist before anything is multicast to them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int nk = p.k * p.cpj;
  // tile schedule: alone, CTA b takes tiles b, b+grid, ..; in a cluster, pair q = clusterid + i*nclusters
  // and this CTA takes tile 2q + rank (a tile index past the end is a dummy: all rows out of range)
  const int crank = CLUSTER ? static_cast<int>(cluster_ctarank()) : 0;
  const int tile_first = CLUSTER ? 2 * static_cast<int>(cluster_id_x()) + crank : static_cast<int>(blockIdx.x);
  const int tile_step = CLUSTER ? 2 * static_cast<int>(n_clusters_x()) : static_cast<int>(gridDim.x);
  const int tile_end = CLUSTER ? ((p.n_tiles + 1) / 2) * 2 : p.n_tiles;     // both CTAs run the same count

  // ---- shift mode: separate rings for weights (one (j, c) chunk per slot) and tokens (one c chunk per slot)
  const uint32_t x_slot_bytes = ((static_cast<uint32_t>(p.xrows) * 128u) + 1023u) & ~1023u;
  const uint32_t w_ring = base;                                             // wstages * MT_BYTES
  const uint32_t x_ring = base + static_cast<uint32_t>(p.wstages) * MT_BYTES;  // xstages * x_slot_bytes
  // barriers in shift mode reuse the same array: full[i]/empty[i] for i < wstages are the weight ring,
  // i in [wstages, wstages + xstages) the token ring (p.stages == wstages + xstages)

  if (warp == 0) {
    if (lane == 0 && p.shift_mode == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = tile_first; tile < tile_end; tile += tile_step) {
        const int r0 = tile * p.NT;
        for (int kc0 = 0; kc0 < nk; ++kc0) {
          const int kc = kc0;
          const int j = kc / p.cpj, cc = kc - j * p.cpj;
          mbar_wait(empty_bar(stage), phase ^ 1u);
          mbar_expect_tx(full_bar(stage), a_bytes + b_bytes);
          const uint32_t a_dst = base + stage * stage_bytes;
          if (CLUSTER) {      // my half of the weight chunk, written into both CTAs of the pair
            tma_load_2d_mc(a_dst + static_cast<uint32_t>(crank * p.wbox) * 128u, &tmW, kc * KCHUNK, crank * p.wbox,
                           full_bar(stage), static_cast<uint16_t>(3));
          } else {
            for (int b = 0; b < p.n_wbox; ++b)
              tma_load_2d(a_dst + static_cast<uint32_t>(b * p.wbox) * 128u, &tmW, kc * KCHUNK, b * p.wbox, full_bar(stage));
          }
          tma_load_2d(a_dst + a_bytes, &tmX, cc * KCHUNK, r0 + j, full_bar(stage));
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    } else if (p.shift_mode != 0) {
      // whole warp, warp-uniform loop: TMA descriptors, coordinates and barrier addresses live in uniform registers
      int ws = 0; uint32_t wphase = 0, xphase = 0;
      for (int tile = tile_first; tile < tile_end; tile += tile_step) {
        const int r0 = tile * p.NT;
        for (int m = 0; m < p.n_mt; ++m) {
          for (int cc = 0; cc < p.cpj; ++cc) {
            if (m == 0 && !GATHER) {                 // the token tile: one slot per c chunk, loaded once per tile
              mbar_wait(empty_bar(p.wstages + cc), xphase ^ 1u);
              if (elect_one()) {
                mbar_expect_tx(full_bar(p.wstages + cc), static_cast<uint32_t>(p.xrows) * 128u);
                tma_load_2d(x_ring + cc * x_slot_bytes, &tmX, cc * KCHUNK, r0, full_bar(p.wstages + cc));
              }
            }
            for (int j = 0; j < p.k; ++j) {
              mbar_wait(empty_bar(ws), wphase ^ 1u);
              if (elect_one()) {
                mbar_expect_tx(full_bar(ws), MT_BYTES);
                if (CLUSTER)      // my 64-filter half of the chunk, written into both CTAs of the pair
                  tma_load_2d_mc(w_ring + ws * MT_BYTES + static_cast<uint32_t>(crank) * (MT_BYTES / 2), &tmW, (j * p.cpj + cc) * KCHUNK,
                                 m * MT + crank * (MT / 2), full_bar(ws), static_cast<uint16_t>(3));
                else
                  tma_load_2d(w_ring + ws * MT_BYTES, &tmW, (j * p.cpj + cc) * KCHUNK, m * MT, full_bar(ws));
              }
              if (++ws == p.wstages) { ws = 0; wphase ^= 1u; }
            }
          }
        }
        xphase ^= 1u;
      }
    }
  } else if (warp == 1) {
    // the whole warp runs the issue loop (uniform addresses -> uniform registers); see tc_ptx.cuh
    if (p.shift_mode == 0) {
      const uint32_t idesc = make_idesc(p.NT);
      int stage = 0; uint32_t phase = 0;
      uint32_t acc_phase = 0;
      long long w_tmem = 0, w_feed = 0;
      const long long t_begin = clock64();
      for (int tile = tile_first; tile < tile_end; tile += tile_step) {
        long long c0 = clock64();
        mbar_wait(tempty_bar(0), acc_phase ^ 1u);
        w_tmem += clock64() - c0;
        tc_fence_after();
        for (int kc = 0; kc < nk; ++kc) {
          c0 = clock64();
          mbar_wait(full_bar(stage), phase);
          w_feed += clock64() - c0;
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
          if (CLUSTER) { if (elect_one()) tc_commit_mc(empty_bar(stage), static_cast<uint16_t>(3)); }   // frees the slot in BOTH CTAs
          else tc_commit_u(empty_bar(stage));          // smem slot free once these MMAs retire
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
        tc_commit_u(tfull_bar(0));                     // accumulators complete
        acc_phase ^= 1u;
      }
      if (p.prof != nullptr && lane == 0) {
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 0), static_cast<unsigned long long>(clock64() - t_begin));
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 1), static_cast<unsigned long long>(w_tmem));
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 2), static_cast<unsigned long long>(w_feed));
      }
    } else {
      const uint32_t idesc = make_idesc(p.NT);
      int ws = 0; uint32_t wphase = 0, xphase = 0;
      int reg = 0; uint32_t rphase = 0;                // accumulator region of the current unit, and its use parity
      long long w_tmem = 0, w_feed = 0;
      const long long t_begin = clock64();
      for (int tile = tile_first; tile < tile_end; tile += tile_step) {
        for (int m = 0; m < p.n_mt; ++m) {
          long long c0 = clock64();
          mbar_wait(tempty_bar(reg), rphase ^ 1u);     // region drained by the epilogue of the unit that used it last
          w_tmem += clock64() - c0;
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(reg * p.NT);
          for (int cc = 0; cc < p.cpj; ++cc) {
            if (m == 0) { c0 = clock64(); mbar_wait(full_bar(p.wstages + cc), xphase); w_feed += clock64() - c0; }
            const uint32_t x_lo = sw128_desc_lo(x_ring + cc * x_slot_bytes);
            for (int j = 0; j < p.k; ++j) {
              c0 = clock64();
              mbar_wait(full_bar(ws), wphase);
              w_feed += clock64() - c0;
              tc_fence_after();
              const uint32_t a_lo = sw128_desc_lo(w_ring + ws * MT_BYTES);
              const int nks = cc == p.cpj - 1 ? p.last_ks : KCHUNK / 16;
#pragma unroll
              for (int ks = 0; ks < KCHUNK / 16; ++ks)   // token rows j .. j+NT-1 of the resident box (128 B = 8 units per row)
                if (ks < nks)
                  tc_mma_bf16_u(d_tmem, a_lo + ks * 2, x_lo + static_cast<uint32_t>(j) * 8u + ks * 2, idesc, (cc | j | ks) != 0 ? 1u : 0u);
              if (CLUSTER) { if (elect_one()) tc_commit_mc(empty_bar(ws), static_cast<uint16_t>(3)); }   // frees the slot in BOTH CTAs
              else tc_commit_u(empty_bar(ws));
              if (++ws == p.wstages) { ws = 0; wphase ^= 1u; }
            }
            if (m == p.n_mt - 1) tc_commit_u(empty_bar(p.wstages + cc));   // last M tile is done with this token slot
          }
          tc_commit_u(tfull_bar(reg));                 // unit complete: its epilogue overlaps the next unit's MMAs
          if (++reg == p.nreg) { reg = 0; rphase ^= 1u; }
        }
        xphase ^= 1u;
      }
      if (p.prof != nullptr && lane == 0) {
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 0), static_cast<unsigned long long>(clock64() - t_begin));
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 1), static_cast<unsigned long long>(w_tmem));
        atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 2), static_cast<unsigned long long>(w_feed));
      }
    }
  } else if (GATHER && warp == 2 + NUM_EPI_WARPS) {
    // ---- token-row gather (inference): the tile's NT + k - 1 token rows come straight from the embedding table, four
    // rows per TMA gather4 with the token ids as row coordinates, 64 columns (one swizzle row) per c chunk.  Lane l
    // owns row groups l and l + 32 of the tile and issues them for every chunk; one lane arms the slot's barrier.
    const int ngroups = p.xrows / 4;
    uint32_t xphase = 0;
    for (int tile = tile_first; tile < tile_end; tile += tile_step) {
      const int64_t r0 = static_cast<int64_t>(tile) * p.NT;
      int ids[2][4];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int64_t r = r0 + (lane + 32 * h) * 4 + i;
          ids[h][i] = (lane + 32 * h < ngroups && r < p.n_rows) ? __ldg(p.tok + r) : 0;       // rows past the end: any valid row (masked later)
        }
      }
      for (int cc = 0; cc < p.cpj; ++cc) {
        mbar_wait(empty_bar(p.wstages + cc), xphase ^ 1u);
        if (lane == 0) mbar_expect_tx(full_bar(p.wstages + cc), static_cast<uint32_t>(p.xrows) * 128u);
        __syncwarp();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int g = lane + 32 * h;
          if (g < ngroups)
            tma_gather4_2d(x_ring + cc * x_slot_bytes + static_cast<uint32_t>(g) * 512u, &tmX, cc * KCHUNK,
                           ids[h][0], ids[h][1], ids[h][2], ids[h][3], full_bar(p.wstages + cc));
        }
      }
      xphase ^= 1u;
    }
  } else {
    const int lq = warp & 3;                           // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;                  // which half of the tile's columns
    const int Tn = p.L - p.k + 1;
    const int ncol = p.NT / 2;                         // columns this warp walks, multiple of 16
    uint32_t acc_phase = 0;
    int reg = 0; uint32_t rphase = 0;                  // mode 1: accumulator region of the current unit
    long long w_full = 0, w_work = 0;
    for (int tile = tile_first; tile < tile_end; tile += tile_step) {
      const int64_t rbeg = static_cast<int64_t>(tile) * p.NT + half * ncol;       // flat row of my first column
      if (p.shift_mode == 0) {
        const long long c0 = clock64();
        mbar_wait(tfull_bar(0), acc_phase);
        tc_fence_after();
        w_full += clock64() - c0;
      }
      for (int m = 0; m < p.n_mt; ++m) {
        const long long c0 = clock64();
        if (p.shift_mode != 0) { mbar_wait(tfull_bar(reg), rphase); tc_fence_after(); }
        const long long c1 = clock64();
        w_full += c1 - c0;
        const int f = m * MT + lq * 32 + lane;
        unsigned long long* kcol = p.keys + p.col0 + f;
        // running (max, first arg-max) of the current sequence piece
        int64_t s = rbeg / p.L;
        int t = static_cast<int>(rbeg - s * p.L);
        float best = -INFINITY; int bt = -1;
        const bool fok = f < p.F;
        auto flush = [&](int64_t last_row) {
          if (bt >= 0 && fok && last_row < p.n_rows) atomicMax(kcol + s * p.ldk, pool_key(best + 0.0f, static_cast<uint32_t>(bt)));
          best = -INFINITY; bt = -1;
        };
        // 16 columns held in registers.  Common case: no sequence boundary strictly inside the chunk ->
        // tree max over the 16 values, then (only if it beats the running max) a backward scan for the
        // FIRST column that attains it; everything independent, no long select chain.
        // columns [lo, hi) of the chunk (relative indices) belong to the current sequence piece and start
        // at time step t0: fold them into the running (best, bt).  Tree max, then a backward scan for the
        // FIRST column that attains it -- all independent instructions, static register indices.
        auto fold = [&](const uint32_t (&v)[16], int lo, int hi, int t0) {
          const int vhi = min(hi, lo + (Tn - t0));           // columns >= vhi are not valid windows
          float x[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) x[i] = (i >= lo && i < vhi) ? __uint_as_float(v[i]) : -INFINITY;
          const float m01 = fmaxf(x[0], x[1]), m23 = fmaxf(x[2], x[3]), m45 = fmaxf(x[4], x[5]), m67 = fmaxf(x[6], x[7]);
          const float m89 = fmaxf(x[8], x[9]), mab = fmaxf(x[10], x[11]), mcd = fmaxf(x[12], x[13]), mef = fmaxf(x[14], x[15]);
          const float m16 = fmaxf(fmaxf(fmaxf(m01, m23), fmaxf(m45, m67)), fmaxf(fmaxf(m89, mab), fmaxf(mcd, mef)));
          if (m16 > best) {                                  // strict >: an earlier chunk keeps a tie
            int idx = 15;
#pragma unroll
            for (int i = 14; i >= 0; --i) idx = (x[i] == m16) ? i : idx;
            best = m16; bt = t0 + idx - lo;
          }
        };
        // same, for a chunk that lies wholly inside one sequence and holds only valid windows (no masks)
        auto fold_all = [&](const uint32_t (&v)[16], int t0) {
          float x[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) x[i] = __uint_as_float(v[i]);
          const float m01 = fmaxf(x[0], x[1]), m23 = fmaxf(x[2], x[3]), m45 = fmaxf(x[4], x[5]), m67 = fmaxf(x[6], x[7]);
          const float m89 = fmaxf(x[8], x[9]), mab = fmaxf(x[10], x[11]), mcd = fmaxf(x[12], x[13]), mef = fmaxf(x[14], x[15]);
          const float m16 = fmaxf(fmaxf(fmaxf(m01, m23), fmaxf(m45, m67)), fmaxf(fmaxf(m89, mab), fmaxf(mcd, mef)));
          if (m16 > best) {
            int idx = 15;
#pragma unroll
            for (int i = 14; i >= 0; --i) idx = (x[i] == m16) ? i : idx;
            best = m16; bt = t0 + idx;
          }
        };
        auto process = [&](const uint32_t (&v)[16], int cb) {
          const int nb = p.L - t;                            // columns left in the current sequence
          if (nb > 16) {
            if (t + 16 <= Tn) fold_all(v, t); else fold(v, 0, 16, t);
            t += 16;
          } else if (p.L >= 16) {                            // exactly one boundary at relative column nb
            fold(v, 0, nb, t);
            flush(rbeg + cb + nb - 1);
            ++s;
            if (nb < 16) fold(v, nb, 16, 0);
            t = 16 - nb;
          } else {                                           // very short sequences: several boundaries per chunk
#pragma unroll                     // (static register indices: a rolled loop would push v[] to local memory)
            for (int i = 0; i < 16; ++i) {
              const float x = __uint_as_float(v[i]);
              if (t < Tn && x > best) { best = x; bt = t; }
              if (++t == p.L) { flush(rbeg + cb + i); t = 0; ++s; }
            }
          }
        };
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) +
                               static_cast<uint32_t>((p.shift_mode != 0 ? reg : m) * p.NT + half * ncol);
        // all of this warp's columns of region m in one go: up to 8 x16 loads in flight, one wait
        uint32_t v[NCH][16];
#pragma unroll
        for (int c = 0; c < NCH; ++c) tc_ld16_async(taddr + c * 16, v[c]);
        tc_wait_ld();
        if (p.shift_mode != 0) {                                                                        // the unit is in registers
          tc_fence_before();
          if (lane == 0) mbar_arrive(tempty_bar(reg));
          if (++reg == p.nreg) { reg = 0; rphase ^= 1u; }
        }
        else if (m == p.n_mt - 1) { tc_fence_before(); if (lane == 0) mbar_arrive(tempty_bar(0)); }    // last read: release TMEM
#pragma unroll
        for (int c = 0; c < NCH; ++c) process(v[c], c * 16);
        flush(s * p.L + (bt >= 0 ? bt : 0));
        w_work += clock64() - c1;
      }
      acc_phase ^= 1u;
    }
    if (p.prof != nullptr && warp == 2 && lane == 0) {
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 3), static_cast<unsigned long long>(w_full));
      atomicAdd(reinterpret_cast<unsigned long long*>(p.prof + 4), static_cast<unsigned long long>(w_work));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CLUSTER) cluster_sync_all();                  // nobody leaves while the peer can still write here
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

static int conv_fwd_tc_impl(const void* X, const int32_t* tok, int64_t table_rows, int64_t S, int L, int D, int Dpad, const void* Wt,
                            int k, int F, int Fpad, uint64_t* keys, int ldk, int col0, cudaStream_t st);

// D: the number of real columns (0 = unknown: all Dpad columns are multiplied); columns D..Dpad-1 of X and Wt are zero
int clsm_conv_pool_fwd_tc(const void* X, int64_t S, int L, int D, int Dpad, const void* Wt, int k, int F, int Fpad,
                          uint64_t* keys, int ldk, int col0, cudaStream_t st) {
  return conv_fwd_tc_impl(X, nullptr, 0, S, L, D, Dpad, Wt, k, F, Fpad, keys, ldk, col0, st);
}

// Same kernel, token tile fetched from the embedding table by TMA gather4 (tok != nullptr, X = the (table_rows, Dpad)
// bf16 table): the inference path, no dropout, X never written.
int clsm_conv_pool_fwd_tok_tc(const int32_t* tok, const void* table, int64_t table_rows, int64_t S, int L, int D, int Dpad,
                              const void* Wt, int k, int F, int Fpad, uint64_t* keys, int ldk, int col0, cudaStream_t st) {
  return conv_fwd_tc_impl(table, tok, table_rows, S, L, D, Dpad, Wt, k, F, Fpad, keys, ldk, col0, st);
}

static int conv_fwd_tc_impl(const void* X, const int32_t* tok, int64_t table_rows, int64_t S, int L, int D, int Dpad, const void* Wt,
                            int k, int F, int Fpad, uint64_t* keys, int ldk, int col0, cudaStream_t st) {
  const bool gather = tok != nullptr;
  const int64_t n_rows = S * L;
  CLSM_REQUIRE(n_rows + 512 < (1ll << 31), "conv_pool_fwd(tcgen05): S*L too large for 32-bit TMA coordinates");
  CLSM_REQUIRE(Dpad % 64 == 0 && Fpad % 16 == 0 && Fpad >= 16 && Fpad <= 512, "conv_pool_fwd(tcgen05): bad Dpad/Fpad");
  TcParams p{};
  p.n_rows = n_rows; p.L = L; p.k = k; p.F = F; p.cpj = Dpad / 64;
  p.n_mt = (F + MT - 1) / MT;
  p.NT = (512 / p.n_mt) / 32 * 32;
  if (p.NT > 256) p.NT = 256;
  p.n_tiles = static_cast<int>(ceil_div64(n_rows, p.NT));
  p.ldk = ldk; p.col0 = col0; p.keys = reinterpret_cast<unsigned long long*>(keys);
  p.tok = tok;
  p.last_ks = KCHUNK / 16;
  if (D > 0 && D <= Dpad && D > Dpad - KCHUNK) p.last_ks = (D - (Dpad - KCHUNK) + 15) / 16;
  // variant: 0 (default) = independent CTAs, 1 = CTA pairs with multicast weights.  CLSM_CONV_TC_VARIANT overrides.
  static int variant = -1;
  if (variant < 0) {
    const char* e = getenv("CLSM_CONV_TC_VARIANT");
    variant = (e != nullptr && e[0] == '1') ? 1 : 0;
  }
  const bool cluster = variant == 1 && clsm_sm_count() >= 2 && !gather;
  const int a_rows = p.n_mt * MT;
  if (cluster) { p.wbox = a_rows / 2; p.n_wbox = 2; }
  else { p.wbox = MT; p.n_wbox = p.n_mt; }
  // mode 1 (default) = M-tile-outer loop over a resident token tile; CLSM_CONV_TC_SHIFT=0 selects mode 0 (K-outer,
  // all M tiles per stage).  Measured on cfg2 documents (cycles per CTA, CLSM_CONV_TC_PROF=1): mode 0 spends 21 % of
  // the MMA thread's time waiting for the epilogue to drain TMEM (n_mt * NT = 480 columns leave no second
  // accumulator set) and runs at 112 cycles per MMA against a floor of 80, which is the shared-memory bound
  // (4 KB weights written + 4 KB read + 5 KB tokens read + 1.7 KB tokens written per MMA at 128 B/clk).  Mode 1
  // with 224-token tiles needs 37 % fewer bytes per flop, never waits for TMEM, and is 14 % faster end to end.
  { const char* e = getenv("CLSM_CONV_TC_SHIFT"); p.shift_mode = (e != nullptr && e[0] == '0') ? 0 : 1; }
  int smem_bytes = 0;
  if (p.shift_mode != 0) {
    // mode 1 is not tied to n_mt * NT <= 512: the (tile, M tile) units rotate through 512 / NT regions.
    // NT = 224: the widest tile whose NT + k - 1 token rows still fit one TMA box (256 rows) and whose UMMA
    // floor (N / 2 = 112 cycles) covers the ~105-cycle latency of back-to-back MMAs into one accumulator.
    int nt = 224;
    { const char* e = getenv("CLSM_CONV_TC_NT"); if (e != nullptr && atoi(e) >= 64 && atoi(e) <= 256 && atoi(e) % 32 == 0) nt = atoi(e); }
    while (nt > 64 && nt + k - 1 > 256) nt -= 32;
    p.NT = nt;
    p.n_tiles = static_cast<int>(ceil_div64(n_rows, p.NT));
    p.nreg = 512 / p.NT;
    if (p.nreg > 4) p.nreg = 4;
  }
  p.xrows = (p.NT + k - 1 + 7) / 8 * 8;
  if (p.shift_mode != 0 && p.xrows <= 256) {
    const int x_slot = (p.xrows * 128 + 1023) / 1024 * 1024;
    p.xstages = p.cpj;
    p.wstages = (SMEM_LIMIT - 2048 - p.xstages * x_slot) / MT_BYTES;
    if (p.wstages > 8) p.wstages = 8;
    if (p.wstages < 3) p.shift_mode = 0;
    else {
      p.stages = p.wstages + p.xstages;
      smem_bytes = p.wstages * MT_BYTES + p.xstages * x_slot + 1024 + 8 * (2 * p.stages + 9);
    }
  } else {
    p.shift_mode = 0;
  }
  if (p.shift_mode == 0) {
    p.NT = (512 / p.n_mt) / 32 * 32;           // (a fallback from mode 1 returns to the n_mt-region tile)
    if (p.NT > 256) p.NT = 256;
    p.n_tiles = static_cast<int>(ceil_div64(n_rows, p.NT));
    const int stage_bytes = a_rows * 128 + ((p.NT * 128 + 1023) / 1024) * 1024;
    int stages = (SMEM_LIMIT - 2048) / stage_bytes;
    if (stages > 8) stages = 8;
    CLSM_REQUIRE(stages >= 2, "conv_pool_fwd(tcgen05): tile too large for a 2-stage pipeline");
    p.stages = stages;
    smem_bytes = stages * stage_bytes + 1024 /*alignment slack*/ + 8 * (2 * stages + 9);
  }

  if (gather && p.shift_mode == 0) {
    clsm_set_error("conv_pool_fwd(tcgen05, token gather): shape outside the resident-tile plan");
    return CLSM_EUNSUPPORTED;
  }
  CUtensorMap tmX, tmW;
  // rows of Wt past Fpad (the last M tile is usually partial) are zero-filled by TMA
  // gather: the map covers the TABLE, box = one row of 64 columns (tile::gather4 fetches four such rows per instruction)
  if (!(gather ? make_map_2d(&tmX, X, static_cast<uint64_t>(Dpad), static_cast<uint64_t>(table_rows), KCHUNK, 1)
               : make_map_2d(&tmX, X, static_cast<uint64_t>(Dpad), static_cast<uint64_t>(n_rows), KCHUNK,
                             static_cast<uint32_t>(p.shift_mode ? p.xrows : p.NT))) ||
      !make_map_2d(&tmW, Wt, static_cast<uint64_t>(k) * Dpad, static_cast<uint64_t>(Fpad), KCHUNK,
                   static_cast<uint32_t>(p.shift_mode ? (cluster ? MT / 2 : MT) : p.wbox))) {
    clsm_set_error("conv_pool_fwd(tcgen05): cuTensorMapEncodeTiled failed");
    return CLSM_ECUDA;
  }
  static int prof_on = -1;
  if (prof_on < 0) { const char* e = getenv("CLSM_CONV_TC_PROF"); prof_on = (e != nullptr && e[0] == '1') ? 1 : 0; }
  if (prof_on) {
    void* sym = nullptr;
    CLSM_CUDA(cudaGetSymbolAddress(&sym, g_conv_tc_prof));
    p.prof = static_cast<long long*>(sym);
    CLSM_CUDA(cudaMemsetAsync(sym, 0, sizeof(long long) * 8, st));
  }
  int grid = clsm_sm_count();
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  if (cluster) {
    const int n_pairs = (p.n_tiles + 1) / 2;
    grid = (grid / 2) * 2;
    if (grid > 2 * n_pairs) grid = 2 * n_pairs;
    cfg.attrs = attr; cfg.numAttrs = 1;
  } else if (grid > p.n_tiles) {
    grid = p.n_tiles;
  }
  cfg.gridDim = dim3(grid);
#define TC_LAUNCH(CL, N)                                                                                                 \
  do {                                                                                                                   \
    CLSM_CUDA(cudaFuncSetAttribute(conv_pool_tc_kernel<CL, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)); \
    CLSM_CUDA(cudaLaunchKernelEx(&cfg, conv_pool_tc_kernel<CL, N>, tmX, tmW, p));                                        \
  } while (0)
  const int nch = p.NT / 32;
#define TC_LAUNCH_G(N)                                                                                                          \
  do {                                                                                                                          \
    CLSM_CUDA(cudaFuncSetAttribute(conv_pool_tc_kernel<false, N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT)); \
    CLSM_CUDA(cudaLaunchKernelEx(&cfg, conv_pool_tc_kernel<false, N, true>, tmX, tmW, p));                                      \
  } while (0)
  if (gather) {
    cfg.blockDim = dim3(NUM_THREADS + 32);
    switch (nch) {
      case 2: TC_LAUNCH_G(2); break;
      case 3: TC_LAUNCH_G(3); break;
      case 4: TC_LAUNCH_G(4); break;
      case 5: TC_LAUNCH_G(5); break;
      case 6: TC_LAUNCH_G(6); break;
      case 7: TC_LAUNCH_G(7); break;
      default: TC_LAUNCH_G(8); break;
    }
  } else if (cluster) {
    switch (nch) {
      case 2: TC_LAUNCH(true, 2); break;
      case 3: TC_LAUNCH(true, 3); break;
      case 4: TC_LAUNCH(true, 4); break;
      case 5: TC_LAUNCH(true, 5); break;
      case 6: TC_LAUNCH(true, 6); break;
      case 7: TC_LAUNCH(true, 7); break;
      default: TC_LAUNCH(true, 8); break;
    }
  }
  else {
    switch (nch) {
      case 2: TC_LAUNCH(false, 2); break;
      case 3: TC_LAUNCH(false, 3); break;
      case 4: TC_LAUNCH(false, 4); break;
      case 5: TC_LAUNCH(false, 5); break;
      case 6: TC_LAUNCH(false, 6); break;
      case 7: TC_LAUNCH(false, 7); break;
      default: TC_LAUNCH(false, 8); break;
    }
  }
#undef TC_LAUNCH
#undef TC_LAUNCH_G
  CLSM_LAUNCH_CHECK();
  if (prof_on) {           // diagnostics only: synchronises and prints average cycles per CTA
    long long h[8];
    CLSM_CUDA(cudaStreamSynchronize(st));
    CLSM_CUDA(cudaMemcpy(h, p.prof, sizeof(h), cudaMemcpyDeviceToHost));
    const double g = static_cast<double>(grid);
    fprintf(stderr, "[conv_tc prof] mode %d NT %d tiles %d grid %d | MMA thread: total %.0f, wait TMEM %.0f, wait feed %.0f | "
            "epilogue warp: wait accumulator %.0f, work %.0f (cycles per CTA)\n", p.shift_mode, p.NT, p.n_tiles, grid,
            h[0] / g, h[1] / g, h[2] / g, h[3] / g, h[4] / g);
  }
  return CLSM_OK;
}
