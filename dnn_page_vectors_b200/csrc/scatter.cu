// Embedding gradient: dE[id] = sum over token positions with that id of dX[pos], deterministic.
//   replaces the gradient of Embedding (Theano AdvancedIncSubtensor1 / TF UnsortedSegmentSum) that
//   Keras builds for dssm_cnn_v2/cnn_dssm_tf.py:111-116.
// plan (clsm_emb_scatter_plan, needs only the token ids, so the host runs it under the forward pass):
//   1. stable radix sort of (id, position)            (cub::DeviceRadixSort, utility primitive)
//   2. segment heads + inclusive scan -> segment index per sorted entry (cub::DeviceScan)
// reduce (clsm_emb_scatter_reduce, ONE launch):
//   3. segmented reduce by RECURSIVE CHUNKING (hand-written): level 0 sums 64 consecutive sorted entries per thread
//      group; a run that lies wholly inside its chunk is final, a run cut by a chunk edge leaves a partial row in
//      one of the chunk's two slots; level l+1 repeats over the slots of 32 chunks and is executed by whichever group
//      finishes last among its children.  The shape of the tree depends only on n, and every run is added left to
//      right, so the result is bit-identical from run to run; a 40 %-PAD batch costs no more than a uniform one.
//      dX rows (fp32, or bf16 from the single-window tensor-core backward) and the forward's keep bits are staged
//      through shared memory with cp.async.  (No shared-memory carve-out preference is set on purpose: with one, the
//      kernel could not share an SM with the other stream's kernels and the two-stream step lost 0.6 ms.)
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

// ---- the whole reduction tree in ONE launch -------------------------------------------------------------------
// Level 0 sums CH consecutive sorted entries per chunk, level l + 1 the two slots of FAN chunks of level l.  A thread
// GROUP (64 or 128 threads of a 128-thread CTA) owns one level-0 chunk; when it is done it counts itself in at the
// chunk's parent, and the group that arrives LAST at a parent goes on to reduce that parent (and so on up), so the
// upper levels need no launches of their own and overlap the rest of level 0.  The tree's shape and every addition
// order depend on n only -- which group happens to execute a chunk does not change a bit of the result.
constexpr int MAXLV = 8;
struct TreeParams {
  const int32_t* spos; const int32_t* seg; const int32_t* seg_start; int64_t n;
  const void* dX; int ncol4;
  float inv_keep; uint32_t thr; uint64_t seed; const uint16_t* mask_bits;
  float4* dE;
  int n_levels;                       // level l has n_chunks[l] chunks; n_chunks[n_levels - 1] == 1
  int64_t n_chunks[MAXLV];
  int32_t* slot_key[MAXLV];           // outputs of level l: 2 slots per chunk
  float4* slot_rows[MAXLV];
  int* counters[MAXLV];               // counters[l][c]: children of chunk c of level l that have finished (self-resetting)
};

__device__ __forceinline__ void group_sync(int grp, int gthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(gthreads) : "memory");
}
__device__ __forceinline__ float4 f4add(float4 a, const float4& b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; return a; }

// MASK: 0 = no dropout, 1 = regenerate the Philox stream of the forward, 2 = read the 1-bit-per-element mask the
// forward gather stored (mask_bits (n_tok, ncol4 / 4) uint16, bit c of word (row, blk) = column 16 blk + c kept).
// DXBF16: dX rows are bf16 (the single-window tensor-core backward writes them so), thread = 8 columns (one 16-byte
// load per row), 64 threads per group; otherwise fp32 rows, thread = one float4 column, 128 threads per group.
template <int MASK, bool DXBF16>
__global__ void __launch_bounds__(128)
seg_reduce_tree_kernel(const TreeParams p) {
  constexpr int GT = DXBF16 ? 64 : 128;                 // threads per group
  constexpr int NG = 128 / GT;                          // groups per CTA
  constexpr int NF4 = DXBF16 ? 2 : 1;                   // float4 accumulators per thread
  constexpr int DEPTH = 16;                             // dX rows in flight per thread (shared-memory ring)
  extern __shared__ __align__(128) uint8_t s_rows_all[];   // NG x DEPTH staged dX rows
  __shared__ int32_t s_pos[NG][CH], s_seg[NG][CH];
  __shared__ uint16_t s_bits[NG][CH * 32];              // keep bits of the staged rows (MASK == 2), nblk <= 32 words per row
  __shared__ int s_edge[NG][2];                         // [0]: the chunk's first run starts its segment; [1]: its last run ends it
  __shared__ int s_last[NG];
  const int grp = threadIdx.x / GT, gt = threadIdx.x % GT;
  int64_t c = static_cast<int64_t>(blockIdx.x) * NG + grp;
  if (c >= p.n_chunks[0]) return;
  const int ncol = DXBF16 ? p.ncol4 / 2 : p.ncol4;      // 16-byte thread columns per row
  const bool active = gt < ncol;
  const int ncol4 = p.ncol4;
  const int nblk = ncol4 >> 2;
  const uint32_t row_bytes = static_cast<uint32_t>(ncol) * 16u;

  // =============================== level 0: chunk c of the sorted (id, position) list
  // The chunk's CH rows of dX sit at CH unrelated addresses.  They are staged through shared memory with cp.async (16
  // bytes per thread and row): a register-staged loop with 8 rows in flight spent two thirds of its time waiting for
  // them (ncu: stall_long_scoreboard) and more rows in flight cost registers, i.e. occupancy.
  {
    const int64_t i0 = c * CH;
    const int cnt = static_cast<int>((i0 + CH <= p.n) ? CH : p.n - i0);
    uint8_t* rows = s_rows_all + static_cast<size_t>(grp) * DEPTH * row_bytes;
    if (gt < CH) {
      const bool ok = gt < cnt;
      s_pos[grp][gt] = ok ? p.spos[i0 + gt] : 0;
      s_seg[grp][gt] = ok ? p.seg[i0 + gt] : -1;
    }
    if (gt == 0) { p.slot_key[0][2 * c] = -1; p.slot_key[0][2 * c + 1] = -1; }
    group_sync(grp, GT);
    const int32_t* pos = s_pos[grp];
    const int32_t* sg = s_seg[grp];
    // Each thread owns the same 16 bytes of every row: it copies them and it alone reads them, so the staging needs
    // no barrier -- a private ring of DEPTH slots, refilled one row at a time as rows are consumed, keeps DEPTH rows
    // in flight for the whole chunk (a register-staged loop drains to zero after every batch).
    const uint32_t ring = static_cast<uint32_t>(__cvta_generic_to_shared(rows)) + gt * 16;
    const uint8_t* src0 = static_cast<const uint8_t*>(p.dX) + gt * 16;
    auto issue = [&](int q) {                           // one commit group per row, empty past the end (uniform counting)
      if (active && q < cnt)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring + (q % DEPTH) * row_bytes), "l"(src0 + static_cast<int64_t>(pos[q]) * row_bytes) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if constexpr (MASK == 2) {
      // the rows' keep bits (nblk 16-bit words per row, a multiple of 8 bytes) go first, as the OLDEST commit group:
      // 8-byte cp.async pieces, all in flight at once (plain loads through registers were 4 deep and a third of the kernel)
      const int ppr = nblk >> 2;                        // 8-byte pieces per row
      const uint32_t sb = static_cast<uint32_t>(__cvta_generic_to_shared(&s_bits[grp][0]));
      for (int i = gt; i < cnt * ppr; i += GT) {
        const int q = i / ppr, pc = i - q * ppr;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sb + q * 64 + pc * 8),
                     "l"(reinterpret_cast<const uint8_t*>(p.mask_bits) + (static_cast<int64_t>(pos[q]) * nblk) * 2 + pc * 8) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
#pragma unroll
    for (int q = 0; q < DEPTH; ++q) issue(q);
    if (gt == 0) {
      s_edge[grp][0] = p.seg_start[sg[0]] == i0 ? 1 : 0;
      s_edge[grp][1] = p.seg_start[sg[cnt - 1] + 1] == i0 + cnt ? 1 : 0;
    }
    asm volatile("cp.async.wait_group %0;" ::"n"(DEPTH) : "memory");      // this thread's pieces of the keep bits have landed
    group_sync(grp, GT);
    const int sub = gt & 3;
    struct Row { float4 v[NF4]; };
    auto load_row = [&](int q) {                        // from the staged copy
      Row r;
      asm volatile("cp.async.wait_group %0;" ::"n"(DEPTH - 1) : "memory");      // the oldest row in flight has landed
      if constexpr (DXBF16) {
        const uint4 u = active ? *reinterpret_cast<const uint4*>(rows + static_cast<size_t>(q % DEPTH) * row_bytes + gt * 16) : make_uint4(0u, 0u, 0u, 0u);
        r.v[0] = make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u), __uint_as_float(u.y << 16), __uint_as_float(u.y & 0xffff0000u));
        r.v[1] = make_float4(__uint_as_float(u.z << 16), __uint_as_float(u.z & 0xffff0000u), __uint_as_float(u.w << 16), __uint_as_float(u.w & 0xffff0000u));
      } else {
        r.v[0] = active ? *reinterpret_cast<const float4*>(rows + static_cast<size_t>(q % DEPTH) * row_bytes + gt * 16) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      issue(q + DEPTH);                                 // refill the slot just read
      return r;
    };
    auto keep4 = [&](float4 v, uint32_t w) {                    // low 4 bits of w: keep decisions of the 4 columns
      v.x = (w & 1u) ? v.x * p.inv_keep : 0.f;
      v.y = (w & 2u) ? v.y * p.inv_keep : 0.f;
      v.z = (w & 4u) ? v.z * p.inv_keep : 0.f;
      v.w = (w & 8u) ? v.w * p.inv_keep : 0.f;
      return v;
    };
    auto bytes_to_nibble = [&](uint32_t word) -> uint32_t {     // 4 Philox decision bytes -> 4 keep bits
      return ((word & 0xffu) < p.thr ? 1u : 0u) | (((word >> 8) & 0xffu) < p.thr ? 2u : 0u) |
             (((word >> 16) & 0xffu) < p.thr ? 4u : 0u) | ((word >> 24) < p.thr ? 8u : 0u);
    };
    // keep bits of this thread's columns of row q: NF4 nibbles, lowest first
    auto row_bits = [&](int q) -> uint32_t {
      if (!active) return 0u;
      if constexpr (MASK == 2) {
        if constexpr (DXBF16) return (static_cast<uint32_t>(s_bits[grp][q * 32 + (gt >> 1)]) >> (8 * (gt & 1))) & 0xFFu;
        else return (static_cast<uint32_t>(s_bits[grp][q * 32 + (gt >> 2)]) >> (4 * sub)) & 0xFu;
      } else {
        uint32_t b[4];
        if constexpr (DXBF16) {
          dropout_bytes16(p.seed, static_cast<uint64_t>(pos[q]), static_cast<uint32_t>(gt >> 1), b);
          return bytes_to_nibble(b[2 * (gt & 1)]) | (bytes_to_nibble(b[2 * (gt & 1) + 1]) << 4);
        } else {
          dropout_bytes16(p.seed, static_cast<uint64_t>(pos[q]), static_cast<uint32_t>(gt >> 2), b);
          return bytes_to_nibble(sub == 0 ? b[0] : sub == 1 ? b[1] : sub == 2 ? b[2] : b[3]);
        }
      }
    };
    // A run of equal ids that lies wholly inside the chunk is final; a run cut by a chunk edge leaves a partial row
    // in one of the chunk's two slots.  Only the first and the last run can be cut (s_edge).
    auto flush = [&](int g, int rs, int re, const Row& acc) {
      const bool final_run = (rs > 0 || s_edge[grp][0] != 0) && (re < cnt || s_edge[grp][1] != 0);
      float4* dst = final_run ? p.dE + static_cast<int64_t>(g) * ncol4 : p.slot_rows[0] + (2 * c + (rs == 0 ? 0 : 1)) * ncol4;
      if (active) {
#pragma unroll
        for (int i = 0; i < NF4; ++i) dst[gt * NF4 + i] = acc.v[i];
      }
      if (!final_run && gt == 0) p.slot_key[0][2 * c + (rs == 0 ? 0 : 1)] = g;
    };
    // rows are added strictly left to right within a run: the same bits whatever the staging
    int g = sg[0], rs = 0;
    Row acc;
#pragma unroll
    for (int i = 0; i < NF4; ++i) acc.v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int idx = 0; idx < cnt; ++idx) {
      Row v = load_row(idx);
      if constexpr (MASK != 0) {
        const uint32_t w = row_bits(idx);
#pragma unroll
        for (int j = 0; j < NF4; ++j) v.v[j] = keep4(v.v[j], w >> (4 * j));
      }
      const int gi = sg[idx];
      if (gi != g) {
        flush(g, rs, idx, acc);
        g = gi; rs = idx;
#pragma unroll
        for (int j = 0; j < NF4; ++j) acc.v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int j = 0; j < NF4; ++j) acc.v[j] = f4add(acc.v[j], v.v[j]);
    }
    flush(g, rs, cnt, acc);
  }

  // =============================== levels >= 1: whoever arrives last at a parent reduces it
  int64_t R = static_cast<int64_t>(CH) * FAN;           // level-0 entries covered by one chunk of the current level
  for (int lv = 1; lv < p.n_levels; ++lv, R *= FAN) {
    const int64_t parent = c / FAN;
    const int64_t first_child = parent * FAN;
    const int n_children = static_cast<int>((first_child + FAN <= p.n_chunks[lv - 1]) ? FAN : p.n_chunks[lv - 1] - first_child);
    __threadfence();                                    // this group's slot rows / keys / dE rows are visible ...
    group_sync(grp, GT);
    if (gt == 0) {                                      // ... before it counts itself in
      const int old = atomicAdd(p.counters[lv] + parent, 1);
      const int last = old == n_children - 1 ? 1 : 0;
      if (last) p.counters[lv][parent] = 0;             // ready for the next call on this plan
      s_last[grp] = last;
    }
    group_sync(grp, GT);
    if (!s_last[grp]) return;
    __threadfence();
    c = parent;
    // slots [c * CH, c * CH + 2 * n_children) of level lv - 1
    int32_t* s_key = s_seg[grp];
    const int32_t* in_key = p.slot_key[lv - 1];
    const float4* in_rows = p.slot_rows[lv - 1];
    const int64_t b0 = c * CH;
    if (gt < CH) s_key[gt] = gt < 2 * n_children ? __ldcg(in_key + b0 + gt) : -1;
    if (gt == 0) { p.slot_key[lv][2 * c] = -1; p.slot_key[lv][2 * c + 1] = -1; }
    group_sync(grp, GT);
    const int64_t lo = c * R;
    const int64_t hi = ((c + 1) * R < p.n) ? (c + 1) * R : p.n;
    int e = 0;
    bool first = true;
    while (e < CH) {
      const int g = s_key[e];
      if (g < 0) { ++e; continue; }
      int e1 = e;
      while (e1 < CH && (s_key[e1] == g || s_key[e1] < 0)) ++e1;
      const bool final_run = (p.seg_start[g] >= lo) && (p.seg_start[g + 1] <= hi);
      float4* dst = final_run ? p.dE + static_cast<int64_t>(g) * ncol4 : p.slot_rows[lv] + (2 * c + (first ? 0 : 1)) * ncol4;
      for (int col = gt; col < ncol4; col += GT) {      // partial rows of the run, four in flight, added left to right
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int q = e;
        while (q < e1) {
          int idx[4], m = 0;
          while (m < 4 && q < e1) { if (s_key[q] == g) idx[m++] = q; ++q; }
          float4 v[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) v[i] = i < m ? __ldcg(in_rows + (b0 + idx[i]) * ncol4 + col) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int i = 0; i < 4; ++i) if (i < m) acc = f4add(acc, v[i]);
        }
        dst[col] = acc;
      }
      if (!final_run && gt == 0) p.slot_key[lv][2 * c + (first ? 0 : 1)] = g;
      first = false;
      e = e1;
    }
  }
}

struct ScatterPlan {
  size_t off_keys_out, off_vals_in, off_vals_out, off_head, off_seg, off_seg_start;
  size_t off_cub, cub_bytes;
  int n_levels;
  int64_t n_chunks[MAXLV];
  size_t off_slot_key[MAXLV], off_slot_rows[MAXLV], off_counters[MAXLV];
  size_t off_counters_all, counters_bytes;
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
  // the reduction tree: level 0 = chunks of CH entries, level l + 1 = FAN chunks of level l, until one chunk is left
  int lv = 0;
  int64_t nc = ceil_div64(n, CH);
  for (;; ++lv) {
    p.n_chunks[lv] = nc;
    p.off_slot_key[lv] = take(static_cast<size_t>(2 * nc) * 4);
    p.off_slot_rows[lv] = take(static_cast<size_t>(2 * nc) * ncol4 * 16);
    if (nc == 1) break;
    nc = ceil_div64(nc, FAN);
  }
  p.n_levels = lv + 1;
  p.off_counters_all = o;
  for (int l = 0; l < p.n_levels; ++l) p.off_counters[l] = take(static_cast<size_t>(p.n_chunks[l]) * 4);
  p.counters_bytes = o - p.off_counters_all;
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
  CLSM_CUDA(cudaMemsetAsync(w + p.off_counters_all, 0, p.counters_bytes, st));     // arrival counters of the reduction tree
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

  const bool drop = keep < 1.0f && keep_threshold(keep) < 256u;      // a keep that rounds to 256/256 drops nothing
  const int mask = !drop ? 0 : (mask_bits != nullptr ? 2 : 1);
  TreeParams t{};
  t.spos = vals_out; t.seg = seg; t.seg_start = seg_start; t.n = n_tok;
  t.dX = dX; t.ncol4 = ncol4;
  t.inv_keep = drop ? dropout_scale(keep) : 1.0f;
  t.thr = drop ? keep_threshold(keep) : 256u;
  t.seed = seed; t.mask_bits = mask_bits;
  t.dE = reinterpret_cast<float4*>(dE_rows);
  t.n_levels = p.n_levels;
  for (int l = 0; l < p.n_levels; ++l) {
    t.n_chunks[l] = p.n_chunks[l];
    t.slot_key[l] = reinterpret_cast<int32_t*>(w + p.off_slot_key[l]);
    t.slot_rows[l] = reinterpret_cast<float4*>(w + p.off_slot_rows[l]);
    t.counters[l] = reinterpret_cast<int*>(w + p.off_counters[l]);
  }
  const int64_t chunks0 = p.n_chunks[0];
  // staged rows: 16 rows per group; bf16: 2 groups x 16 x Dpad*2 B, fp32: 1 group x 16 x Dpad*4 B -- the same bytes
  const size_t smem = static_cast<size_t>(16) * Dpad * 4;
#define TREE(M, B, G)                                                                                                   \
  do {                                                                                                                  \
    CLSM_CUDA(cudaFuncSetAttribute(seg_reduce_tree_kernel<M, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem))); \
    seg_reduce_tree_kernel<M, B><<<G, 128, smem, st>>>(t);                                                              \
  } while (0)
  if (dx_dtype == CLSM_BF16) {
    const unsigned g0 = static_cast<unsigned>(ceil_div64(chunks0, 2));       // two 64-thread groups per CTA
    if (mask == 0) TREE(0, true, g0); else if (mask == 1) TREE(1, true, g0); else TREE(2, true, g0);
  } else {
    const unsigned g0 = static_cast<unsigned>(chunks0);
    if (mask == 0) TREE(0, false, g0); else if (mask == 1) TREE(1, false, g0); else TREE(2, false, g0);
  }
#undef TREE
  CLSM_LAUNCH_CHECK();
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
