// Data-parallel table optimizer over NVLink peer memory (one process per GPU, buffers mapped with CUDA IPC by the
// host side; no counterpart in the reference, which has no data parallelism -- SURVEY 8e).
//
// Rank r owns table rows [r*R, (r+1)*R).  Per step and per tower:
//   1. dp_push: every rank cuts its row-sparse gradient (ids ascending, rows compact) at the ownership
//      boundaries and STORES each piece straight into the owner's receive buffer over NVLink -- rows into
//      rows_in[src][0..], and the slice of its id -> compact-row map into map_in[src][0..R) (rebased to the
//      piece).  No counts travel and nothing is synchronised on the host.
//   2. (any collective on the same stream -- the allreduce of the small gradients -- orders the pushes)
//   3. adam_table_dp: the owner walks ITS rows once: g = sum over sources s = 0..world-1 (fixed order) of the
//      pushed row if map_in[s][row] >= 0; dense Keras Adam on p, m, v; the refreshed bf16 compute row is stored
//      into EVERY rank's compute table (local + peers) -- if its bf16 image changed.  The HBM-bound optimizer pass and the NVLink-bound
//      "allgather" are one kernel, so the transfer overlaps the math row by row.
#include "common.cuh"
#include <math.h>

namespace {

constexpr int MAX_WORLD = 16;

struct PushP {
  const int32_t* uniq; const int32_t* n_uniq; const float* dE; int ldg; int ncol4;
  const int32_t* row_of_id; int64_t n_rows; int64_t R; int world; int64_t cap;
  float* peer_rows[MAX_WORLD]; int32_t* peer_map[MAX_WORLD];
  int32_t* seg;            // world + 1 ints: first compact row of each owner's piece
  int row_blocks;          // blocks [0, row_blocks) copy rows, the rest write the maps
};

__global__ void dp_seg_kernel(const int32_t* __restrict__ uniq, const int32_t* __restrict__ n_uniq, int64_t R, int world,
                              int32_t* __restrict__ seg) {
  const int o = threadIdx.x;
  if (o > world) return;
  const int n = n_uniq[0];
  if (o == world) { seg[o] = n; return; }
  const int64_t bound = static_cast<int64_t>(o) * R;
  int lo = 0, hi = n;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (uniq[mid] < bound) lo = mid + 1; else hi = mid; }
  seg[o] = lo;
}

__global__ void __launch_bounds__(256) dp_push_kernel(PushP p) {
  const int lane = threadIdx.x & 31;
  if (static_cast<int>(blockIdx.x) < p.row_blocks) {
    // one warp per compact row, grid-stride
    const int n = p.n_uniq[0];
    const int warps = p.row_blocks * 8;
    for (int i = blockIdx.x * 8 + (threadIdx.x >> 5); i < n; i += warps) {
      const int id = p.uniq[i];
      const int o = static_cast<int>(id / p.R);
      const int64_t dst_row = i - p.seg[o];
      if (dst_row >= p.cap) continue;                        // cannot happen when cap >= min(R, tokens); never write past the buffer
      const float4* src = reinterpret_cast<const float4*>(p.dE + static_cast<int64_t>(i) * p.ldg);
      float4* dst = reinterpret_cast<float4*>(p.peer_rows[o] + dst_row * p.ldg);
      for (int c = lane; c < p.ncol4; c += 32) dst[c] = __ldg(src + c);
    }
  } else {
    const int64_t total = p.R * p.world;
    const int64_t stride = static_cast<int64_t>(gridDim.x - p.row_blocks) * blockDim.x;
    for (int64_t idx = static_cast<int64_t>(blockIdx.x - p.row_blocks) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
      const int o = static_cast<int>(idx / p.R);
      const int64_t r = idx - o * p.R;
      const int32_t v = idx < p.n_rows ? p.row_of_id[idx] : -1;
      p.peer_map[o][r] = v < 0 ? -1 : v - p.seg[o];
    }
  }
}

struct AdamDpP {
  float* p; float* m; float* v; int64_t n_rows; int D; int world;
  const int32_t* maps; int64_t R; const float* rows; int64_t cap; int ldg;
  float b1, b2, omb1, omb2, lr_t, eps;
  __nv_bfloat16* copies[MAX_WORLD]; int ld_copy;
  int self;                // index of this rank's own table in copies[]
};

__device__ __forceinline__ void adam1(float& p, float& m, float& v, float g, const AdamDpP& c) {
  adam_step1(p, m, v, g, c.b1, c.b2, c.omb1, c.omb2, c.lr_t, c.eps);      // common.cuh: the arithmetic of adam.cu
}

// D % 4 == 0.  One WARP per owned row (same arithmetic as adam_table_kernel<true> in adam.cu).  Lane s looks up
// source s's map entry for the row -- ONE look-up per row and source, shared by a vote, where a thread per float4
// did `world` dependent look-ups each (at 8 ranks that, not memory, was what the pass waited for: 1.35 ms against a
// 0.3 ms bandwidth floor) -- and rows nobody touched skip the gradient gather altogether.
__global__ void __launch_bounds__(256) adam_table_dp_kernel(AdamDpP c) {
  const int per_row = c.D / 4;
  const int lane = threadIdx.x & 31;
  const int64_t r = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  if (r >= c.n_rows) return;
  const int32_t my_g = lane < c.world ? __ldg(c.maps + static_cast<int64_t>(lane) * c.R + r) : -1;
  const unsigned have = __ballot_sync(0xffffffffu, my_g >= 0);          // sources that hold a piece of this row's gradient
  int32_t gs[MAX_WORLD];                                                    // every lane gets every source's entry (all 32 lanes
#pragma unroll                                                              // are still converged here; the q loop below is not)
  for (int s = 0; s < MAX_WORLD; ++s) gs[s] = __shfl_sync(0xffffffffu, my_g, s);
  for (int q = lane; q < per_row; q += 32) {
    float4 gi = make_float4(0.f, 0.f, 0.f, 0.f);
    if (have != 0u) {
#pragma unroll
      for (int s = 0; s < MAX_WORLD; ++s) {                                 // ascending source order: fixed summation order
        if ((have >> s) & 1u) {
          const float4 x = __ldg(reinterpret_cast<const float4*>(c.rows + (static_cast<int64_t>(s) * c.cap + gs[s]) * c.ldg + q * 4));
          gi.x += x.x; gi.y += x.y; gi.z += x.z; gi.w += x.w;
        }
      }
    }
    const int64_t off = r * c.D + q * 4;
    float4 pi = *reinterpret_cast<float4*>(c.p + off);
    float4 mi = *reinterpret_cast<float4*>(c.m + off);
    float4 vi = *reinterpret_cast<float4*>(c.v + off);
    adam1(pi.x, mi.x, vi.x, gi.x, c); adam1(pi.y, mi.y, vi.y, gi.y, c);
    adam1(pi.z, mi.z, vi.z, gi.z, c); adam1(pi.w, mi.w, vi.w, gi.w, c);
    *reinterpret_cast<float4*>(c.p + off) = pi;
    *reinterpret_cast<float4*>(c.m + off) = mi;
    *reinterpret_cast<float4*>(c.v + off) = vi;
    __nv_bfloat162 lo = __floats2bfloat162_rn(pi.x, pi.y), hi = __floats2bfloat162_rn(pi.z, pi.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    // Replicas hold identical compute tables, so the old local value is also what every peer holds: only rows
    // whose bf16 image actually changed cross NVLink.  A row without a recent gradient moves by lr * m / sqrt(v)
    // with m decaying 10 % per step -- far below bf16 resolution after a dozen steps -- so most of the table
    // stays put and the broadcast shrinks to the rows touched recently.
    const int64_t coff = r * c.ld_copy + q * 4;
    uint2* mine = reinterpret_cast<uint2*>(c.copies[c.self] + coff);
    const uint2 old = *mine;
    if (old.x != pk.x || old.y != pk.y) {
      *mine = pk;
      for (int s = 0; s < c.world; ++s)
        if (s != c.self) *reinterpret_cast<uint2*>(c.copies[s] + coff) = pk;
    }
  }
}

}  // namespace

extern "C" {

int clsm_dp_push_rows(const int32_t* uniq_ids, const int32_t* n_uniq, int64_t max_uniq, const float* dE_rows, int ldg, int D,
                      const int32_t* row_of_id, int64_t n_rows, int64_t rows_per_owner, int world,
                      void* const* peer_rows, void* const* peer_map, int64_t cap, int32_t* seg_scratch, void* stream) {
  CLSM_REQUIRE(world >= 1 && world <= MAX_WORLD, "dp_push_rows: world must be 1..16");
  CLSM_REQUIRE(max_uniq > 0 && D > 0 && ldg >= D && ldg % 4 == 0 && rows_per_owner > 0 && cap > 0 && n_rows > 0,
               "dp_push_rows: bad sizes (ldg must be a multiple of 4)");
  CLSM_REQUIRE(rows_per_owner * world >= n_rows, "dp_push_rows: rows_per_owner * world must cover the table");
  CLSM_REQUIRE(peer_rows != nullptr && peer_map != nullptr && seg_scratch != nullptr, "dp_push_rows: null pointer array");
  cudaStream_t st = as_stream(stream);
  PushP p{};
  p.uniq = uniq_ids; p.n_uniq = n_uniq; p.dE = dE_rows; p.ldg = ldg; p.ncol4 = (D + 3) / 4;
  p.row_of_id = row_of_id; p.n_rows = n_rows; p.R = rows_per_owner; p.world = world; p.cap = cap;
  for (int o = 0; o < world; ++o) {
    CLSM_REQUIRE(peer_rows[o] != nullptr && peer_map[o] != nullptr, "dp_push_rows: null peer pointer");
    p.peer_rows[o] = static_cast<float*>(peer_rows[o]);
    p.peer_map[o] = static_cast<int32_t*>(peer_map[o]);
  }
  p.seg = seg_scratch;
  dp_seg_kernel<<<1, 32, 0, st>>>(uniq_ids, n_uniq, rows_per_owner, world, seg_scratch);
  CLSM_LAUNCH_CHECK();
  const int sms = clsm_sm_count();
  int64_t rb = ceil_div64(max_uniq, 8);
  if (rb > 4 * sms) rb = 4 * sms;
  p.row_blocks = static_cast<int>(rb);
  int64_t mb = ceil_div64(rows_per_owner * world, 256);
  if (mb > 2 * sms) mb = 2 * sms;
  dp_push_kernel<<<static_cast<unsigned>(rb + mb), 256, 0, st>>>(p);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_adam_table_dp(float* p, float* m, float* v, int64_t n_rows, int D, int world,
                       const int32_t* maps_in, int64_t rows_per_owner, const float* rows_in, int64_t cap, int ldg,
                       int step, float lr, float b1, float b2, float eps,
                       void* const* bf16_copies, int self_rank, int ld_copy, void* stream) {
  CLSM_REQUIRE(world >= 1 && world <= MAX_WORLD, "adam_table_dp: world must be 1..16");
  CLSM_REQUIRE(self_rank >= 0 && self_rank < world, "adam_table_dp: self_rank out of range");
  CLSM_REQUIRE(n_rows > 0 && n_rows <= rows_per_owner && D > 0 && D % 4 == 0 && step >= 1 && ldg >= D && ldg % 4 == 0 && cap > 0,
               "adam_table_dp: bad sizes (D and ldg must be multiples of 4)");
  CLSM_REQUIRE(bf16_copies != nullptr && ld_copy >= D && ld_copy % 4 == 0, "adam_table_dp: bad compute-copy arguments");
  AdamDpP c{};
  c.p = p; c.m = m; c.v = v; c.n_rows = n_rows; c.D = D; c.world = world;
  c.maps = maps_in; c.R = rows_per_owner; c.rows = rows_in; c.cap = cap; c.ldg = ldg;
  c.b1 = b1; c.b2 = b2; c.omb1 = 1.0f - b1; c.omb2 = 1.0f - b2; c.eps = eps;
  const double t = static_cast<double>(step);
  c.lr_t = static_cast<float>(static_cast<double>(lr) * sqrt(1.0 - pow(static_cast<double>(b2), t)) / (1.0 - pow(static_cast<double>(b1), t)));
  for (int s = 0; s < world; ++s) {
    CLSM_REQUIRE(bf16_copies[s] != nullptr, "adam_table_dp: null compute-copy pointer");
    c.copies[s] = static_cast<__nv_bfloat16*>(bf16_copies[s]);
  }
  c.ld_copy = ld_copy; c.self = self_rank;
  adam_table_dp_kernel<<<static_cast<unsigned>(ceil_div64(n_rows * 32, 256)), 256, 0, as_stream(stream)>>>(c);     // a warp per row
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

}  // extern "C"
