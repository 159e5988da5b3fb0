// Embedding gradient: dE[id] = sum over token positions with that id of dX[pos], deterministic.
//   replaces the gradient of Embedding (Theano AdvancedIncSubtensor1 / TF UnsortedSegmentSum) that
//   Keras builds for dssm_cnn_v2/cnn_dssm_tf.py:111-116.
// 1. stable radix sort of (id, position)            (cub::DeviceRadixSort, utility primitive)
// 2. segment heads + inclusive scan -> segment index per sorted entry (cub::DeviceScan)
// 3. segmented reduce by RECURSIVE CHUNKING (hand-written): level 0 sums 64 consecutive sorted
//    entries per CTA; a run that lies wholly inside its chunk is final, a run cut by a chunk edge
//    leaves a partial row in one of the chunk's two slots; level l+1 repeats over the slots of 32
//    chunks.  The shape of the tree depends only on n, and every run is added left to right, so
//    the result is bit-identical from run to run; a 40 %-PAD batch costs no more than a uniform one.
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace {

constexpr int CH = 64;          // entries (level 0) or slots (level >= 1) per chunk
constexpr int FAN = CH / 2;     // chunks of the previous level covered by one chunk

__global__ void iota_kernel(int32_t* p, int64_t n) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) p[i] = static_cast<int32_t>(i);
}

__global__ void heads_kernel(const int32_t* __restrict__ st, int64_t n, int32_t* __restrict__ head) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) head[i] = (i == 0 || st[i] != st[i - 1]) ? 1 : 0;
}

// seg[i] = inclusive_scan(head)[i] - 1 is computed in place by the caller (scan output minus one
// is applied here).  Writes uniq ids, segment starts (+ sentinel), row_of_id, n_uniq.
__global__ void segments_kernel(const int32_t* __restrict__ st, const int32_t* __restrict__ head,
                                int32_t* __restrict__ seg, int64_t n, int32_t* __restrict__ uniq,
                                int32_t* __restrict__ seg_start, int32_t* __restrict__ row_of_id,
                                int32_t* __restrict__ n_uniq) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t g = seg[i] - 1;
  seg[i] = g;
  if (head[i]) { uniq[g] = st[i]; seg_start[g] = static_cast<int32_t>(i); row_of_id[st[i]] = g; }
  if (i == n - 1) { seg_start[g + 1] = static_cast<int32_t>(n); n_uniq[0] = g + 1; }
}

// level 0: one CTA per chunk of CH sorted entries; thread = one float4 column of the row.
// MASK: 0 = no dropout, 1 = regenerate the Philox stream of the forward, 2 = read the 1-bit-per-element mask the
// forward gather stored (mask_bits (n_tok, ncol4 / 4) uint16, bit c of word (row, blk) = column 16 blk + c kept).
// DXBF16: dX rows are bf16 (the single-window tensor-core backward writes them so); sums stay fp32.
template <int MASK, bool DXBF16>
__global__ void __launch_bounds__(128)
seg_reduce_l0_kernel(const int32_t* __restrict__ spos, const int32_t* __restrict__ seg,
                     const int32_t* __restrict__ seg_start, int64_t n,
                     const void* __restrict__ dXv, int ldx4, int ncol4,
                     float inv_keep, uint32_t thr, uint64_t seed, const uint16_t* __restrict__ mask_bits,
                     float4* __restrict__ dE, int lde4,
                     int32_t* __restrict__ slot_key, float4* __restrict__ slot_rows) {
  __shared__ int32_t s_pos[CH], s_seg[CH];
  const int64_t c = blockIdx.x;
  const int64_t i0 = c * CH;
  const int cnt = static_cast<int>((i0 + CH <= n) ? CH : n - i0);
  if (threadIdx.x < CH) {
    const bool ok = threadIdx.x < cnt;
    s_pos[threadIdx.x] = ok ? spos[i0 + threadIdx.x] : 0;
    s_seg[threadIdx.x] = ok ? seg[i0 + threadIdx.x] : -1;
  }
  if (threadIdx.x == 0) { slot_key[2 * c] = -1; slot_key[2 * c + 1] = -1; }
  __syncthreads();
  const int col = threadIdx.x;
  const bool active = col < ncol4;
  const int nblk = ncol4 >> 2;
  int e = 0;
  while (e < cnt) {
    const int g = s_seg[e];
    int e1 = e + 1;
    while (e1 < cnt && s_seg[e1] == g) ++e1;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    {
      const int lane = threadIdx.x & 31;
      const int sub = lane & 3;                       // which float4 of the 16-column block this lane owns
      auto load_row = [&](int q) {
        if (!active) return make_float4(0.f, 0.f, 0.f, 0.f);
        if (DXBF16) {
          const uint2 u = __ldcs(reinterpret_cast<const uint2*>(dXv) + static_cast<int64_t>(s_pos[q]) * ldx4 + col);
          return make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u),
                             __uint_as_float(u.y << 16), __uint_as_float(u.y & 0xffff0000u));
        }
        return __ldcs(reinterpret_cast<const float4*>(dXv) + static_cast<int64_t>(s_pos[q]) * ldx4 + col);
      };
      auto apply = [&](float4 v, uint32_t word) {                 // Philox decision bytes
        v.x = ((word & 0xffu) < thr) ? v.x * inv_keep : 0.f;
        v.y = (((word >> 8) & 0xffu) < thr) ? v.y * inv_keep : 0.f;
        v.z = (((word >> 16) & 0xffu) < thr) ? v.z * inv_keep : 0.f;
        v.w = ((word >> 24) < thr) ? v.w * inv_keep : 0.f;
        return v;
      };
      auto apply_bits = [&](float4 v, int q) {                    // stored mask bits
        const uint32_t w = active ? static_cast<uint32_t>(__ldg(mask_bits + static_cast<int64_t>(s_pos[q]) * nblk + (col >> 2))) >> (4 * sub) : 0u;
        v.x = (w & 1u) ? v.x * inv_keep : 0.f;
        v.y = (w & 2u) ? v.y * inv_keep : 0.f;
        v.z = (w & 4u) ? v.z * inv_keep : 0.f;
        v.w = (w & 8u) ? v.w * inv_keep : 0.f;
        return v;
      };
      int q = e;
      for (; q + 4 <= e1; q += 4) {            // 4 independent row loads in flight, added in order
        float4 v0 = load_row(q), v1 = load_row(q + 1), v2 = load_row(q + 2), v3 = load_row(q + 3);
        if (MASK == 1) {
          // One Philox call yields the 16 decision bytes of a (row, 16-column block); the 4 lanes that own that
          // block's 4 float4 columns would each recompute it.  Instead lane (4b + j) computes the block for row
          // q + j of the 4 rows in flight and the words are exchanged with shuffles: 1 call per lane per 4 rows.
          uint32_t bytes[4];
          dropout_bytes16(seed, static_cast<uint64_t>(s_pos[q + sub]), static_cast<uint32_t>(col >> 2), bytes);
          const int base = lane & ~3;
          uint32_t w[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t b0 = __shfl_sync(0xffffffffu, bytes[0], base + j), b1 = __shfl_sync(0xffffffffu, bytes[1], base + j);
            const uint32_t b2 = __shfl_sync(0xffffffffu, bytes[2], base + j), b3 = __shfl_sync(0xffffffffu, bytes[3], base + j);
            w[j] = sub == 0 ? b0 : sub == 1 ? b1 : sub == 2 ? b2 : b3;
          }
          v0 = apply(v0, w[0]); v1 = apply(v1, w[1]); v2 = apply(v2, w[2]); v3 = apply(v3, w[3]);
        } else if (MASK == 2) {
          v0 = apply_bits(v0, q); v1 = apply_bits(v1, q + 1); v2 = apply_bits(v2, q + 2); v3 = apply_bits(v3, q + 3);
        }
        acc.x += v0.x; acc.y += v0.y; acc.z += v0.z; acc.w += v0.w;
        acc.x += v1.x; acc.y += v1.y; acc.z += v1.z; acc.w += v1.w;
        acc.x += v2.x; acc.y += v2.y; acc.z += v2.z; acc.w += v2.w;
        acc.x += v3.x; acc.y += v3.y; acc.z += v3.z; acc.w += v3.w;
      }
      for (; q < e1; ++q) {
        float4 v = load_row(q);
        if (MASK == 1) {
          uint32_t bytes[4];
          dropout_bytes16(seed, static_cast<uint64_t>(s_pos[q]), static_cast<uint32_t>(col >> 2), bytes);
          v = apply(v, bytes[sub]);
        } else if (MASK == 2) {
          v = apply_bits(v, q);
        }
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
    }
    const bool final_run = (seg_start[g] == i0 + e) && (seg_start[g + 1] == i0 + e1);
    if (final_run) {
      if (active) dE[static_cast<int64_t>(g) * lde4 + col] = acc;
    } else {
      const int64_t slot = 2 * c + (e == 0 ? 0 : 1);
      if (active) slot_rows[slot * ncol4 + col] = acc;
      if (threadIdx.x == 0) slot_key[slot] = g;
    }
    e = e1;
  }
}

// level >= 1: chunk c covers slots [c*CH, c*CH+CH) = level-0 entries [c*R, (c+1)*R)
__global__ void __launch_bounds__(128)
seg_reduce_ln_kernel(const int32_t* __restrict__ in_key, const float4* __restrict__ in_rows, int64_t n_slots,
                     const int32_t* __restrict__ seg_start, int64_t n, int64_t R, int ncol4,
                     float4* __restrict__ dE, int lde4,
                     int32_t* __restrict__ out_key, float4* __restrict__ out_rows) {
  __shared__ int32_t s_key[CH];
  const int64_t c = blockIdx.x;
  const int64_t b0 = c * CH;
  if (threadIdx.x < CH) s_key[threadIdx.x] = (b0 + threadIdx.x < n_slots) ? in_key[b0 + threadIdx.x] : -1;
  if (threadIdx.x == 0) { out_key[2 * c] = -1; out_key[2 * c + 1] = -1; }
  __syncthreads();
  const int64_t lo = c * R;
  const int64_t hi = ((c + 1) * R < n) ? (c + 1) * R : n;
  const int col = threadIdx.x;
  const bool active = col < ncol4;
  int e = 0;
  bool first = true;
  while (e < CH) {
    const int g = s_key[e];
    if (g < 0) { ++e; continue; }
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int e1 = e;
    // up to four partial rows of the run in flight at a time, added left to right (same order as one by one)
    while (e1 < CH && (s_key[e1] == g || s_key[e1] < 0)) {
      int idx[4], n = 0;
      while (n < 4 && e1 < CH && (s_key[e1] == g || s_key[e1] < 0)) {
        if (s_key[e1] == g) idx[n++] = e1;
        ++e1;
      }
      float4 v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
        v[i] = (i < n && active) ? in_rows[(b0 + idx[i]) * ncol4 + col] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (i < n) { acc.x += v[i].x; acc.y += v[i].y; acc.z += v[i].z; acc.w += v[i].w; }
    }
    const bool final_run = (seg_start[g] >= lo) && (seg_start[g + 1] <= hi);
    if (final_run) {
      if (active) dE[static_cast<int64_t>(g) * lde4 + col] = acc;
    } else {
      const int64_t slot = 2 * c + (first ? 0 : 1);
      if (active) out_rows[slot * ncol4 + col] = acc;
      if (threadIdx.x == 0) out_key[slot] = g;
    }
    first = false;
    e = e1;
  }
}

struct ScatterPlan {
  size_t off_keys_out, off_vals_in, off_vals_out, off_head, off_seg, off_seg_start;
  size_t off_cub, cub_bytes;
  size_t off_slot_key[2], off_slot_rows[2];
  size_t total;
};

ScatterPlan make_plan(int64_t n, int ncol4) {
  ScatterPlan p{};
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o += align_up(bytes, 256); return r; };
  p.off_keys_out = take(n * 4);
  p.off_vals_in = take(n * 4);
  p.off_vals_out = take(n * 4);
  p.off_head = take(n * 4);
  p.off_seg = take(n * 4);
  p.off_seg_start = take((n + 1) * 4);
  size_t sort_bytes = 0, scan_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const int32_t*)nullptr, (int32_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, static_cast<int>(n), 0, 32);
  cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (const int32_t*)nullptr, (int32_t*)nullptr, static_cast<int>(n));
  p.cub_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
  p.off_cub = take(p.cub_bytes);
  const int64_t chunks0 = ceil_div64(n, CH);
  const int64_t slots1 = 2 * chunks0;
  const int64_t slots2 = 2 * ceil_div64(slots1, CH);
  p.off_slot_key[0] = take(slots1 * 4);
  p.off_slot_rows[0] = take(static_cast<size_t>(slots1) * ncol4 * 16);
  p.off_slot_key[1] = take(slots2 * 4);
  p.off_slot_rows[1] = take(static_cast<size_t>(slots2) * ncol4 * 16);
  p.total = o;
  return p;
}

}  // namespace

extern "C" {

size_t clsm_emb_scatter_workspace_bytes(int64_t n_tok, int D) {
  if (n_tok <= 0) return 256;
  const int Dpad = (D + 63) / 64 * 64;
  return make_plan(n_tok, Dpad / 4).total;
}

// plan: everything that depends on the token ids alone (sort, segments, unique ids, row_of_id).  It does not need
// dX, so a host can run it at the START of the step on another stream, underneath the forward pass.
int clsm_emb_scatter_plan(const int32_t* tok, int64_t n_tok, int Dpad, int D, int64_t n_rows,
                          int32_t* uniq_ids, int32_t* n_uniq_out, int32_t* row_of_id,
                          void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(n_tok > 0 && n_tok < (1ll << 31), "emb_scatter: n_tok out of range");
  CLSM_REQUIRE(Dpad == (D + 63) / 64 * 64 && Dpad <= 512, "emb_scatter: Dpad must be D rounded up to 64 (<= 512)");
  CLSM_REQUIRE(n_rows > 0, "emb_scatter: bad n_rows");
  const int ncol4 = Dpad / 4;
  const ScatterPlan p = make_plan(n_tok, ncol4);
  if (ws_bytes < p.total) {
    clsm_set_error("emb_scatter: workspace too small (%zu < %zu)", ws_bytes, p.total);
    return CLSM_EWORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  char* w = static_cast<char*>(ws);
  int32_t* keys_out = reinterpret_cast<int32_t*>(w + p.off_keys_out);
  int32_t* vals_in = reinterpret_cast<int32_t*>(w + p.off_vals_in);
  int32_t* vals_out = reinterpret_cast<int32_t*>(w + p.off_vals_out);
  int32_t* head = reinterpret_cast<int32_t*>(w + p.off_head);
  int32_t* seg = reinterpret_cast<int32_t*>(w + p.off_seg);
  int32_t* seg_start = reinterpret_cast<int32_t*>(w + p.off_seg_start);
  void* cub_ws = w + p.off_cub;
  const int n = static_cast<int>(n_tok);
  const unsigned g256 = static_cast<unsigned>(ceil_div64(n_tok, 256));

  int end_bit = 1;
  while (end_bit < 32 && (static_cast<int64_t>(1) << end_bit) < n_rows) ++end_bit;

  CLSM_CUDA(cudaMemsetAsync(row_of_id, 0xFF, static_cast<size_t>(n_rows) * 4, st));
  iota_kernel<<<g256, 256, 0, st>>>(vals_in, n_tok);
  CLSM_LAUNCH_CHECK();
  size_t cb = p.cub_bytes;
  CLSM_CUDA(cub::DeviceRadixSort::SortPairs(cub_ws, cb, tok, keys_out, vals_in, vals_out, n, 0, end_bit, st));
  heads_kernel<<<g256, 256, 0, st>>>(keys_out, n_tok, head);
  CLSM_LAUNCH_CHECK();
  cb = p.cub_bytes;
  CLSM_CUDA(cub::DeviceScan::InclusiveSum(cub_ws, cb, head, seg, n, st));
  segments_kernel<<<g256, 256, 0, st>>>(keys_out, head, seg, n_tok, uniq_ids, seg_start, row_of_id, n_uniq_out);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

// reduce: the segmented sum over dX, on the plan left in `ws` by clsm_emb_scatter_plan for the same tokens.
int clsm_emb_scatter_reduce(int64_t n_tok, const void* dX, int dx_dtype, int Dpad, int D,
                            float keep, uint64_t seed, const uint16_t* mask_bits, float* dE_rows,
                            void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(n_tok > 0 && n_tok < (1ll << 31), "emb_scatter: n_tok out of range");
  CLSM_REQUIRE(Dpad == (D + 63) / 64 * 64 && Dpad <= 512, "emb_scatter: Dpad must be D rounded up to 64 (<= 512)");
  CLSM_REQUIRE(dx_dtype == CLSM_F32 || dx_dtype == CLSM_BF16, "emb_scatter: dX must be fp32 or bf16");
  CLSM_REQUIRE(keep > 0.0f, "emb_scatter: bad keep");
  const int ncol4 = Dpad / 4;
  const ScatterPlan p = make_plan(n_tok, ncol4);
  if (ws_bytes < p.total) {
    clsm_set_error("emb_scatter: workspace too small (%zu < %zu)", ws_bytes, p.total);
    return CLSM_EWORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  char* w = static_cast<char*>(ws);
  const int32_t* vals_out = reinterpret_cast<const int32_t*>(w + p.off_vals_out);
  const int32_t* seg = reinterpret_cast<const int32_t*>(w + p.off_seg);
  const int32_t* seg_start = reinterpret_cast<const int32_t*>(w + p.off_seg_start);

  const int threads = (ncol4 + 31) / 32 * 32 < 64 ? 64 : (ncol4 + 31) / 32 * 32;
  int32_t* skey[2] = {reinterpret_cast<int32_t*>(w + p.off_slot_key[0]), reinterpret_cast<int32_t*>(w + p.off_slot_key[1])};
  float4* srow[2] = {reinterpret_cast<float4*>(w + p.off_slot_rows[0]), reinterpret_cast<float4*>(w + p.off_slot_rows[1])};
  const int64_t chunks0 = ceil_div64(n_tok, CH);
  const bool drop = keep < 1.0f;
  const int mask = !drop ? 0 : (mask_bits != nullptr ? 2 : 1);
  const float inv_keep = drop ? dropout_scale(keep) : 1.0f;
  const uint32_t thr = drop ? keep_threshold(keep) : 256u;
  float4* dE4 = reinterpret_cast<float4*>(dE_rows);
  const unsigned g0 = static_cast<unsigned>(chunks0);
#define L0(M, B) seg_reduce_l0_kernel<M, B><<<g0, threads, 0, st>>>(vals_out, seg, seg_start, n_tok, dX, ncol4, ncol4, \
        inv_keep, thr, seed, mask_bits, dE4, ncol4, skey[0], srow[0])
  if (dx_dtype == CLSM_BF16) { if (mask == 0) L0(0, true); else if (mask == 1) L0(1, true); else L0(2, true); }
  else { if (mask == 0) L0(0, false); else if (mask == 1) L0(1, false); else L0(2, false); }
#undef L0
  CLSM_LAUNCH_CHECK();
  // higher levels ping-pong between the two slot buffers
  int64_t n_chunks_prev = chunks0;
  int64_t R = static_cast<int64_t>(CH) * FAN;     // level-0 entries covered by one level-1 chunk
  int src = 0;
  while (n_chunks_prev > 1) {
    const int64_t n_slots = 2 * n_chunks_prev;
    const int64_t n_chunks = ceil_div64(n_slots, CH);
    const int dst = src ^ 1;
    seg_reduce_ln_kernel<<<static_cast<unsigned>(n_chunks), threads, 0, st>>>(skey[src], srow[src], n_slots, seg_start, n_tok, R,
        ncol4, dE4, ncol4, skey[dst], srow[dst]);
    CLSM_LAUNCH_CHECK();
    n_chunks_prev = n_chunks;
    R *= FAN;
    src = dst;
  }
  return CLSM_OK;
}

int clsm_emb_scatter(const int32_t* tok, int64_t n_tok, const float* dX, int Dpad, int D,
                     float keep, uint64_t seed, int64_t n_rows,
                     int32_t* uniq_ids, int32_t* n_uniq_out, float* dE_rows, int32_t* row_of_id,
                     void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(keep > 0.0f && n_rows > 0, "emb_scatter: bad keep / n_rows");
  const int rc = clsm_emb_scatter_plan(tok, n_tok, Dpad, D, n_rows, uniq_ids, n_uniq_out, row_of_id, ws, ws_bytes, stream);
  if (rc != CLSM_OK) return rc;
  return clsm_emb_scatter_reduce(n_tok, dX, CLSM_F32, Dpad, D, keep, seed, nullptr, dE_rows, ws, ws_bytes, stream);
}

}  // extern "C"
