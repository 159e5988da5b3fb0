// Cosine scoring of a bf16 page-vector corpus on the tensor cores, feeding the exact top-k selection of
// topk.cu.  (No retrieval code exists in the reference; the score is RTF's cosine, cnn_dssm_tf.py:63-77,
// on L2-normalised vectors.)
//
//   scores[q, row] = Qn[q, :] . Cn[row, :]          Qn (nQ, H), Cn (N, H) bf16, H a multiple of 64
//
// One CTA per SM, persistent over 256-row corpus tiles.  A operand = a block of 128 queries, resident in
// shared memory for the whole launch; B operand = the corpus tile, streamed by TMA (all H/64 K-chunks of a
// tile form one pipeline stage); D = 128 queries (TMEM lanes) x 256 rows (TMEM columns), double buffered so
// the epilogue of tile i overlaps the MMAs of tile i+1.  The epilogue turns every score into the 64-bit key
// (ordered(score) << 32) | ~row of topk.cu and writes it to keys[q, row - seg_begin]; with a threshold array
// it writes only keys >= thr[q] (the matrix is all-empty on entry and the selection zeroes what it reads), which
// is how later corpus segments are pruned by the k-th best key found so far.  The corpus is read exactly once per 128-query block: HBM-bound.
#include "common.cuh"
#include "tc_ptx.cuh"

using namespace tcptx;

namespace {

constexpr int QB = 128;                       // queries per block (UMMA M)
constexpr int NT = 256;                       // corpus rows per tile (UMMA N)
constexpr int KCHUNK = 64;
constexpr int NUM_EPI_WARPS = 8;              // lane quarter = warp % 4, half of the tile's columns = (warp - 2) / 4
constexpr int NUM_THREADS = 64 + NUM_EPI_WARPS * 32;
constexpr int SMEM_LIMIT = 227 * 1024;

typedef unsigned long long u64;

struct ScoreParams {
  int64_t row_begin, row_end;   // corpus rows of this segment
  int64_t q_begin;              // first query of this block
  int64_t nQ;
  int hchunks;                  // H / 64
  int n_tiles;
  u64* keys;                    // (QB, seg_cap): keys[(q - q_begin) * seg_cap + (row - row_begin)]
  int64_t seg_cap;
  const u64* thr;               // (128) or nullptr: only keys >= thr[q - q_begin] are kept
  int* cnt;                     // nullptr = DENSE mode: keys[q * seg_cap + (row - row_begin)].  Non-null = COMPACT mode:
                                // every (CTA, column half) owns a private sub-list per query, so appending a survivor
                                // is a register increment and a store -- no atomic on the scan's critical path.
                                // Sub-list s = 2 * blockIdx + half of query q starts at keys[q * seg_cap + s * subcap]
                                // and its length goes to cnt[q * nsub + s] (order irrelevant: keys are unique and the
                                // selection is exact).  subcap = rows one (CTA, half) can see in this segment.
  int64_t subcap; int nsub;
  int stages;
};

// One 16-row chunk in which SOME lane (query) has a survivor.  Out of line and SMALL on purpose: ncu showed the scan of
// a segment with survivors stalled on instruction fetch (stall_no_inst 36 % of samples) -- the slow path, unrolled over
// the 64 rows of a batch with the key construction inlined, was ~20 KB of code that eight epilogue warps walked at
// different places; that, not atomics, stores or votes (all measured), halved the scan rate.  Here: the lane's hits as
// a 16-bit mask from 16 independent compares, then one iteration per hit with the score picked by a select tree.
// Values travel in registers (by value: a reference would put them in local memory).
__device__ __noinline__ int chunk_hits(uint32_t v0, uint32_t v1, uint32_t v2, uint32_t v3, uint32_t v4, uint32_t v5, uint32_t v6, uint32_t v7,
                                       uint32_t v8, uint32_t v9, uint32_t v10, uint32_t v11, uint32_t v12, uint32_t v13, uint32_t v14, uint32_t v15,
                                       float thr_f, u64 thr, int64_t row0, int64_t row_end, u64* klist, int n) {
  const uint32_t v[16] = {v0, v1, v2, v3, v4, v5, v6, v7, v8, v9, v10, v11, v12, v13, v14, v15};
  unsigned hm = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) hm |= (__uint_as_float(v[i]) >= thr_f) ? (1u << i) : 0u;
  while (hm != 0u) {
    const int pos = __ffs(hm) - 1;
    hm &= hm - 1;
    uint32_t a[8], b[4], c[2];                                      // select tree on the bits of pos: static register indices only
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = (pos & 1) ? v[2 * i + 1] : v[2 * i];
#pragma unroll
    for (int i = 0; i < 4; ++i) b[i] = (pos & 2) ? a[2 * i + 1] : a[2 * i];
#pragma unroll
    for (int i = 0; i < 2; ++i) c[i] = (pos & 4) ? b[2 * i + 1] : b[2 * i];
    const float x = __uint_as_float((pos & 8) ? c[1] : c[0]);
    const int64_t row = row0 + pos;
    // a row that ties the threshold score but loses on the id is dropped by the exact 64-bit compare
    const u64 key = (static_cast<u64>(f32_to_ordered(x + 0.0f)) << 32) | static_cast<u64>(0xFFFFFFFFu - static_cast<uint32_t>(row));
    if (row < row_end && key >= thr) klist[n++] = key;
  }
  return n;
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
score_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmC, const ScoreParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_bytes = static_cast<uint32_t>(p.hchunks) * QB * 128u;          // resident query block
  const uint32_t stage_bytes = static_cast<uint32_t>(p.hchunks) * NT * 128u;      // one corpus tile, all K chunks
  const uint32_t stage0 = base + q_bytes;
  const uint32_t bars = stage0 + static_cast<uint32_t>(p.stages) * stage_bytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (p.stages + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * p.stages + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * p.stages + 2 + a); };
  const uint32_t q_bar = bars + 8u * (2 * p.stages + 4);
  const uint32_t tmem_slot = bars + 8u * (2 * p.stages + 5);

  const int warp = warp_id_uniform(), lane = threadIdx.x % 32;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), NUM_EPI_WARPS); }
    mbar_init(q_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmQ) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmC) : "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    if (lane == 0) {
      // the query block, once
      mbar_expect_tx(q_bar, q_bytes);
      for (int hc = 0; hc < p.hchunks; ++hc)
        tma_load_2d(base + static_cast<uint32_t>(hc) * QB * 128u, &tmQ, hc * KCHUNK, static_cast<int>(p.q_begin), q_bar);
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        const int64_t r0 = p.row_begin + static_cast<int64_t>(tile) * NT;
        mbar_wait(empty_bar(stage), phase ^ 1u);
        mbar_expect_tx(full_bar(stage), stage_bytes);
        for (int hc = 0; hc < p.hchunks; ++hc)
          tma_load_2d(stage0 + stage * stage_bytes + static_cast<uint32_t>(hc) * NT * 128u, &tmC, hc * KCHUNK, static_cast<int>(r0), full_bar(stage));
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // whole warp, warp-uniform issue loop (tc_ptx.cuh)
    {
      const uint32_t idesc = make_idesc(NT);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      mbar_wait(q_bar, 0);
      const uint32_t q_lo = sw128_desc_lo(base);
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t b_lo = sw128_desc_lo(stage0 + stage * stage_bytes);
        for (int hc = 0; hc < p.hchunks; ++hc) {
#pragma unroll
          for (int ks = 0; ks < KCHUNK / 16; ++ks)
            tc_mma_bf16_u(tmem_base + static_cast<uint32_t>(acc * NT), q_lo + static_cast<uint32_t>(hc) * (QB * 128u / 16u) + ks * 2,
                          b_lo + static_cast<uint32_t>(hc) * (NT * 128u / 16u) + ks * 2, idesc, (hc | ks) != 0 ? 1u : 0u);
        }
        tc_commit_u(empty_bar(stage));
        tc_commit_u(tfull_bar(acc));
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    const int lq = warp & 3;
    const int ql = lq * 32 + lane;                              // query inside the block = TMEM lane
    const int64_t q = p.q_begin + ql;
    const bool qok = q < p.nQ;
    const u64 thr = (p.thr != nullptr && qok) ? p.thr[ql] : 0ull;      // thresholds are indexed inside the query block
    // Compact mode rejects on the SCORE alone (one float compare per row; a tree max per 64 rows decides whether
    // anything in the batch needs a key at all).  A row that ties the threshold score but loses on the id passes
    // this test and is dropped by the exact 64-bit compare below; the selection is exact either way.
    const float thr_f = !qok ? INFINITY : (thr != 0ull ? ordered_to_f32(static_cast<uint32_t>(thr >> 32)) : -INFINITY);
    u64* krow = p.keys + static_cast<int64_t>(ql) * p.seg_cap;
    const int sub = 2 * static_cast<int>(blockIdx.x) + ((warp - 2) >> 2);
    // compact mode: this thread's private sub-list (query-major: measured against a sub-list-major layout, whose 983 k-row
    // segment ran 160 us instead of 102 us)
    u64* klist = krow + static_cast<int64_t>(sub) * p.subcap;
    int n_mine = 0;
    int acc = 0; uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t r0 = p.row_begin + static_cast<int64_t>(tile) * NT;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + static_cast<uint32_t>(acc * NT);
      const int cb_begin = ((warp - 2) >> 2) * (NT / 2);               // this warp's half of the tile's columns
#pragma unroll 1
      for (int cb = cb_begin; cb < cb_begin + NT / 2; cb += 64) {
        uint32_t v[4][16];
#pragma unroll
        for (int c = 0; c < 4; ++c) tc_ld16_async(taddr + cb + c * 16, v[c]);
        tc_wait_ld();
        if (cb + 64 >= cb_begin + NT / 2) { tc_fence_before(); if (lane == 0) mbar_arrive(tempty_bar(acc)); }
        if (p.cnt != nullptr) {
          // Lanes are different queries.  A tree max per 16 rows and two levels of warp-uniform votes (all 32 lanes are
          // converged here; lanes of queries past nQ carry thr_f = +inf and never vote yes) decide whether any lane has a
          // survivor in a chunk; only then the small out-of-line routine above runs.
          float m[4];
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            float x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = __uint_as_float(v[c][i]);
            m[c] = fmaxf(fmaxf(fmaxf(fmaxf(x[0], x[1]), fmaxf(x[2], x[3])), fmaxf(fmaxf(x[4], x[5]), fmaxf(x[6], x[7]))),
                         fmaxf(fmaxf(fmaxf(x[8], x[9]), fmaxf(x[10], x[11])), fmaxf(fmaxf(x[12], x[13]), fmaxf(x[14], x[15]))));
          }
          if (__any_sync(0xffffffffu, fmaxf(fmaxf(m[0], m[1]), fmaxf(m[2], m[3])) >= thr_f)) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              if (__any_sync(0xffffffffu, m[c] >= thr_f))
                n_mine = chunk_hits(v[c][0], v[c][1], v[c][2], v[c][3], v[c][4], v[c][5], v[c][6], v[c][7], v[c][8], v[c][9], v[c][10],
                                    v[c][11], v[c][12], v[c][13], v[c][14], v[c][15], thr_f, thr, r0 + cb + c * 16, p.row_end, klist, n_mine);
            }
          }
        } else if (qok) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
#pragma unroll
            for (int i = 0; i < 16; i += 2) {
              const int64_t row = r0 + cb + c * 16 + i;
              u64 k0 = 0ull, k1 = 0ull;
              if (row < p.row_end) {
                k0 = (static_cast<u64>(f32_to_ordered(__uint_as_float(v[c][i]) + 0.0f)) << 32) | static_cast<u64>(0xFFFFFFFFu - static_cast<uint32_t>(row));
                if (k0 < thr) k0 = 0ull;
              }
              if (row + 1 < p.row_end) {
                k1 = (static_cast<u64>(f32_to_ordered(__uint_as_float(v[c][i + 1]) + 0.0f)) << 32) | static_cast<u64>(0xFFFFFFFFu - static_cast<uint32_t>(row + 1));
                if (k1 < thr) k1 = 0ull;
              }
              if ((k0 | k1) != 0ull)       // dense, survivors only: the matrix is all-empty on entry (topk.cu clears what it reads)
                *reinterpret_cast<ulonglong2*>(krow + (row - p.row_begin)) = make_ulonglong2(k0, k1);
            }
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (p.cnt != nullptr && qok) p.cnt[static_cast<int64_t>(ql) * p.nsub + sub] = n_mine;
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
bool make_map(CUtensorMap* map, const void* ptr, uint64_t cols, uint64_t rows, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {KCHUNK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

// Scores queries [q_begin, q_begin+128) against corpus rows [row_begin, row_end) into keys (128, seg_cap).
// Compact-mode geometry of a segment of `rows` rows: CTAs launched, sub-lists per query, rows per sub-list.
void clsm_score_tc_geometry(int64_t rows, int* grid, int* nsub, int64_t* subcap) {
  const int64_t n_tiles = ceil_div64(rows, NT);
  int g = clsm_sm_count();
  if (g > n_tiles) g = static_cast<int>(n_tiles);
  if (g < 1) g = 1;
  *grid = g; *nsub = 2 * g; *subcap = ceil_div64(n_tiles, g) * (NT / 2);
}

int clsm_score_keys_tc(const void* Qn, int64_t nQ, int64_t q_begin, const void* Cn, int64_t N, int H,
                       int64_t row_begin, int64_t row_end, unsigned long long* keys, int64_t seg_cap,
                       const unsigned long long* thr, int* cnt, cudaStream_t st) {
  CLSM_REQUIRE(H == 64 || H == 128, "cos_topk(bf16): the tensor-core scan takes H = 64 or 128 (use clsm_cos_topk otherwise)");
  CLSM_REQUIRE(N < (1ll << 31) && seg_cap % 2 == 0, "cos_topk(bf16): corpus shard too large or odd segment");
  ScoreParams p{};
  p.row_begin = row_begin; p.row_end = row_end; p.q_begin = q_begin; p.nQ = nQ;
  p.hchunks = H / 64;
  p.n_tiles = static_cast<int>(ceil_div64(row_end - row_begin, NT));
  p.keys = keys; p.seg_cap = seg_cap; p.thr = thr; p.cnt = cnt;
  int geo_grid = 0;
  clsm_score_tc_geometry(row_end - row_begin, &geo_grid, &p.nsub, &p.subcap);
  CLSM_REQUIRE(cnt == nullptr || static_cast<int64_t>(p.nsub) * p.subcap <= seg_cap, "cos_topk(bf16): survivor lists too small for this segment");
  const int q_bytes = p.hchunks * QB * 128;
  const int stage_bytes = p.hchunks * NT * 128;
  int stages = (SMEM_LIMIT - 2048 - q_bytes) / stage_bytes;
  if (stages > 4) stages = 4;
  CLSM_REQUIRE(stages >= 2, "cos_topk(bf16): H too large for a 2-stage pipeline");
  p.stages = stages;
  const int smem_bytes = q_bytes + stages * stage_bytes + 1024 + 8 * (2 * stages + 6);
  CUtensorMap tmQ, tmC;
  if (!make_map(&tmQ, Qn, static_cast<uint64_t>(H), static_cast<uint64_t>(nQ), QB) ||
      !make_map(&tmC, Cn, static_cast<uint64_t>(H), static_cast<uint64_t>(N), NT)) {
    clsm_set_error("cos_topk(bf16): cuTensorMapEncodeTiled failed");
    return CLSM_ECUDA;
  }
  CLSM_CUDA(cudaFuncSetAttribute(score_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  int grid = clsm_sm_count();
  if (grid > p.n_tiles) grid = p.n_tiles;
  if (grid < 1) return CLSM_OK;
  score_tc_kernel<<<grid, NUM_THREADS, smem_bytes, st>>>(tmQ, tmC, p);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}
