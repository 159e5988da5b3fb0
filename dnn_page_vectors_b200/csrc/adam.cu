// Keras-1 Adam (optimizer="adam", dssm_cnn_v2/cnn_dssm_tf.py:182) as fused HBM-bound passes.
//   t = step (1-based); lr_t = lr*sqrt(1-b2^t)/(1-b1^t)
//   m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; p -= lr_t m / (sqrt(v) + eps)
// adam_table walks the WHOLE embedding table every step (Keras updates every row; rows without
// gradient still move by momentum) but takes the gradient row-sparsely, so no (V, D) gradient
// is ever materialised: 24 B/param of p,m,v traffic (+2 B for the bf16 compute copy).
#include "common.cuh"
#include <math.h>

namespace {

struct AdamC { float b1, b2, omb1, omb2, lr_t, eps; };

__device__ __forceinline__ void adam1(float& p, float& m, float& v, float g, const AdamC& c) {
  adam_step1(p, m, v, g, c.b1, c.b2, c.omb1, c.omb2, c.lr_t, c.eps);      // common.cuh
}

__global__ void __launch_bounds__(256)
adam_dense_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                  const float* __restrict__ g, int64_t n, AdamC c) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float pi = p[i], mi = m[i], vi = v[i];
  adam1(pi, mi, vi, g[i], c);
  p[i] = pi; m[i] = mi; v[i] = vi;
}

// D % 4 == 0: one thread = one float4 of one row.  rows_sel == nullptr: all rows (dense Keras
// semantics, gradient looked up through row_of_id); else only rows uniq[0:n_uniq] (lazy Adam).
template <bool VEC4>
__global__ void __launch_bounds__(256)
adam_table_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                  int64_t n_rows, int D, const int32_t* __restrict__ row_of_id,
                  const int32_t* __restrict__ uniq, const int32_t* __restrict__ n_uniq,
                  const float* __restrict__ dE, int ldg, AdamC c,
                  __nv_bfloat16* __restrict__ copy, int ld_copy) {
  constexpr int W = VEC4 ? 4 : 1;
  const int per_row = D / W;
  const int64_t item = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  int64_t r;
  if (item < 0xffffffffll) r = static_cast<uint32_t>(item) / static_cast<uint32_t>(per_row);   // 32-bit divide: ~4x cheaper
  else r = item / per_row;
  const int q = static_cast<int>(item - r * per_row);
  int64_t grow;
  if (uniq != nullptr) {
    if (r >= n_uniq[0]) return;
    grow = r;
    r = uniq[r];
  } else {
    if (r >= n_rows) return;
    grow = row_of_id[r];
  }
  const int64_t off = r * D + q * W;
  if (VEC4) {
    // (streaming ld.cs / st.cs hints on p/m/v were measured: 1.23 -> 1.34 ms per table, so plain accesses stay)
    float4 pi = *reinterpret_cast<float4*>(p + off);
    float4 mi = *reinterpret_cast<float4*>(m + off);
    float4 vi = *reinterpret_cast<float4*>(v + off);
    float4 gi = make_float4(0.f, 0.f, 0.f, 0.f);
    if (grow >= 0) gi = __ldg(reinterpret_cast<const float4*>(dE + grow * ldg + q * 4));
    adam1(pi.x, mi.x, vi.x, gi.x, c); adam1(pi.y, mi.y, vi.y, gi.y, c);
    adam1(pi.z, mi.z, vi.z, gi.z, c); adam1(pi.w, mi.w, vi.w, gi.w, c);
    *reinterpret_cast<float4*>(p + off) = pi;
    *reinterpret_cast<float4*>(m + off) = mi;
    *reinterpret_cast<float4*>(v + off) = vi;
    if (copy) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pi.x, pi.y), hi = __floats2bfloat162_rn(pi.z, pi.w);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(copy + r * ld_copy + q * 4) = pk;
    }
  } else {
    float pi = p[off], mi = m[off], vi = v[off];
    const float gi = grow >= 0 ? dE[grow * ldg + q] : 0.0f;
    adam1(pi, mi, vi, gi, c);
    p[off] = pi; m[off] = mi; v[off] = vi;
    if (copy) copy[r * ld_copy + q] = __float2bfloat16_rn(pi);
  }
}

// ---- deferred rows ------------------------------------------------------------------------------------------
// Keras moves EVERY table row every step: a row without gradient still takes the step with g = 0 (momentum).  That
// step reads nothing but the row's own p, m, v and the step number, so it can be taken LATER, in registers, when the
// row is next needed -- by running the same adam1() with g = 0 once per owed step, in order, with that step's lr_t.
// `last[row]` is the last step the stored p, m, v include; -1 marks a row whose m and v are still exactly zero
// (a zero-gradient step leaves such a row unchanged, so it owes nothing).  The arithmetic and its order are those of
// the dense pass, so the table is bit-identical to it once every row is settled -- while a training step touches
// only the rows of its batch (cfg2: ~190 k of 2 M rows) instead of streaming 7.2 GB of p, m, v per table.
struct DeferC { float b1, b2, omb1, omb2, eps, zero; };     // `zero` is 0.0f passed at run time: g = 0 goes through the SAME
                                                           // instruction sequence as a gradient (no constant folding)
constexpr int EPL = 12;                                    // elements per lane and pass: a pass covers 384 columns

__device__ __forceinline__ AdamC with_lr(const DeferC& d, float lr_t) {
  AdamC c; c.b1 = d.b1; c.b2 = d.b2; c.omb1 = d.omb1; c.omb2 = d.omb2; c.eps = d.eps; c.lr_t = lr_t; return c;
}

// One warp brings one row from step `from` to step `to` (zero-gradient steps from+1 .. to), then, if g_row != nullptr,
// takes step to+1 with that gradient.  lr_tab[s] = lr_t of step s.
__device__ __forceinline__ void settle_row(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, int D,
                                           int from, int to, const float* __restrict__ g_row,
                                           const float* __restrict__ lr_tab, const DeferC& d,
                                           __nv_bfloat16* __restrict__ copy_row, int lane) {
  for (int c0 = 0; c0 < D; c0 += 32 * EPL) {
    float pe[EPL], me[EPL], ve[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) {
      const int e = c0 + lane + 32 * i;
      pe[i] = e < D ? p[e] : 0.0f; me[i] = e < D ? m[e] : 0.0f; ve[i] = e < D ? v[e] : 0.0f;
    }
    for (int s = from + 1; s <= to; ++s) {
      const AdamC c = with_lr(d, __ldg(lr_tab + s));
#pragma unroll
      for (int i = 0; i < EPL; ++i) adam1(pe[i], me[i], ve[i], d.zero, c);
    }
    if (g_row != nullptr) {
      const AdamC c = with_lr(d, __ldg(lr_tab + to + 1));
#pragma unroll
      for (int i = 0; i < EPL; ++i) {
        const int e = c0 + lane + 32 * i;
        adam1(pe[i], me[i], ve[i], e < D ? __ldg(g_row + e) : 0.0f, c);
      }
    }
#pragma unroll
    for (int i = 0; i < EPL; ++i) {
      const int e = c0 + lane + 32 * i;
      if (e < D) {
        p[e] = pe[i]; m[e] = me[i]; v[e] = ve[i];
        if (copy_row != nullptr) copy_row[e] = __float2bfloat16_rn(pe[i]);
      }
    }
  }
}

// pass 1 of "bring the rows of this batch up to step `to`": one thread per token claims its row (one claimer per row,
// however often the id occurs) and appends (row, last) to the work list.  ctl[0] = list length, ctl[1] = ticket.
__global__ void __launch_bounds__(256)
defer_claim_kernel(const int32_t* __restrict__ tok, int64_t n_tok, int64_t n_rows, int32_t* __restrict__ last, int to,
                   int32_t* __restrict__ work, int32_t* __restrict__ ctl) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  int row = -1, old = 0;
  if (i < n_tok) {
    const int r = tok[i];
    if (r >= 0 && r < n_rows) {
      old = last[r];                                   // plain read first: hot rows are current and cost no atomic
      if (old >= 0 && old < to && atomicCAS(last + r, old, to) == old) row = r;
    }
  }
  const unsigned won = __ballot_sync(0xffffffffu, row >= 0);
  if (won == 0u) return;
  const int lane = threadIdx.x & 31;
  int base = 0;
  if (lane == 0) base = atomicAdd(ctl, __popc(won));
  base = __shfl_sync(0xffffffffu, base, 0);
  if (row >= 0) {
    const int at = base + __popc(won & ((1u << lane) - 1u));
    work[2 * at] = row; work[2 * at + 1] = old;
  }
}

// pass 2: warps take list entries by ticket (rows owe different numbers of steps) and settle them
__global__ void __launch_bounds__(256)
defer_settle_list_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, int D, int to,
                         const int32_t* __restrict__ work, int32_t* __restrict__ ctl, const float* __restrict__ lr_tab,
                         DeferC d, __nv_bfloat16* __restrict__ copy, int ld_copy) {
  const int lane = threadIdx.x & 31;
  const int n = ctl[0];
  for (;;) {
    int i = 0;
    if (lane == 0) i = atomicAdd(ctl + 1, 1);
    i = __shfl_sync(0xffffffffu, i, 0);
    if (i >= n) return;
    const int64_t row = work[2 * i];
    settle_row(p + row * D, m + row * D, v + row * D, D, work[2 * i + 1], to, nullptr, lr_tab, d,
               copy != nullptr ? copy + row * ld_copy : nullptr, lane);
  }
}

// the optimizer step of the rows that HAVE a gradient: settle to step-1 if the row owes anything, take step `step`
__global__ void __launch_bounds__(256)
defer_adam_rows_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, int D, int32_t* __restrict__ last,
                       const int32_t* __restrict__ uniq, const int32_t* __restrict__ n_uniq, const float* __restrict__ dE, int ldg,
                       int step, const float* __restrict__ lr_tab, DeferC d, __nv_bfloat16* __restrict__ copy, int ld_copy) {
  const int lane = threadIdx.x & 31;
  const int n = n_uniq[0];
  const int warps = gridDim.x * (blockDim.x >> 5);
  for (int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n; i += warps) {
    const int64_t row = uniq[i];
    const int old = last[row];
    const int from = old < 0 ? step - 1 : old;         // a row with m = v = 0 owes nothing
    settle_row(p + row * D, m + row * D, v + row * D, D, from, step - 1, dE + static_cast<int64_t>(i) * ldg, lr_tab, d,
               copy != nullptr ? copy + row * ld_copy : nullptr, lane);
    if (lane == 0) last[row] = step;
  }
}

// every row that owes steps, up to `to` (before the table is read as a whole: export, inference, checkpoint)
__global__ void __launch_bounds__(256, 4)      // 64 registers: the replay is MUFU-bound, more warps per SM hide it better (6.9 -> 6.5 ms)
defer_settle_all_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, int64_t n_rows, int D,
                        int32_t* __restrict__ last, int to, const float* __restrict__ lr_tab, DeferC d,
                        __nv_bfloat16* __restrict__ copy, int ld_copy) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5); row < n_rows; row += warps) {
    const int old = last[row];
    if (old < 0 || old >= to) continue;
    settle_row(p + row * D, m + row * D, v + row * D, D, old, to, nullptr, lr_tab, d,
               copy != nullptr ? copy + row * ld_copy : nullptr, lane);
    if (lane == 0) last[row] = to;
  }
}

DeferC make_d(float b1, float b2, float eps) {
  DeferC d; d.b1 = b1; d.b2 = b2; d.omb1 = 1.0f - b1; d.omb2 = 1.0f - b2; d.eps = eps; d.zero = 0.0f; return d;
}

AdamC make_c(int step, float lr, float b1, float b2, float eps) {
  AdamC c;
  c.b1 = b1; c.b2 = b2; c.omb1 = 1.0f - b1; c.omb2 = 1.0f - b2; c.eps = eps;
  const double t = static_cast<double>(step);
  c.lr_t = static_cast<float>(static_cast<double>(lr) * sqrt(1.0 - pow(static_cast<double>(b2), t)) /
                              (1.0 - pow(static_cast<double>(b1), t)));
  return c;
}

}  // namespace

extern "C" {

int clsm_adam_dense(float* p, float* m, float* v, const float* g, int64_t n, int step,
                    float lr, float b1, float b2, float eps, void* stream) {
  CLSM_REQUIRE(n >= 0 && step >= 1, "adam_dense: bad n/step");
  if (n == 0) return CLSM_OK;
  adam_dense_kernel<<<static_cast<unsigned>(ceil_div64(n, 256)), 256, 0, as_stream(stream)>>>(p, m, v, g, n, make_c(step, lr, b1, b2, eps));
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_adam_table(float* p, float* m, float* v, int64_t n_rows, int D,
                    const int32_t* row_of_id, const float* dE_rows, int ldg, int step,
                    float lr, float b1, float b2, float eps,
                    void* bf16_copy, int ld_copy, void* stream) {
  CLSM_REQUIRE(n_rows > 0 && D > 0 && step >= 1 && ldg >= D, "adam_table: bad sizes");
  CLSM_REQUIRE(bf16_copy == nullptr || ld_copy >= D, "adam_table: ld_copy < D");
  const AdamC c = make_c(step, lr, b1, b2, eps);
  cudaStream_t st = as_stream(stream);
  __nv_bfloat16* cp = static_cast<__nv_bfloat16*>(bf16_copy);
  if (D % 4 == 0 && ldg % 4 == 0 && (cp == nullptr || ld_copy % 4 == 0)) {
    const int64_t items = n_rows * (D / 4);
    adam_table_kernel<true><<<static_cast<unsigned>(ceil_div64(items, 256)), 256, 0, st>>>(p, m, v, n_rows, D, row_of_id, nullptr, nullptr, dE_rows, ldg, c, cp, ld_copy);
  } else {
    const int64_t items = n_rows * D;
    adam_table_kernel<false><<<static_cast<unsigned>(ceil_div64(items, 256)), 256, 0, st>>>(p, m, v, n_rows, D, row_of_id, nullptr, nullptr, dE_rows, ldg, c, cp, ld_copy);
  }
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_adam_rows(float* p, float* m, float* v, int D, const int32_t* uniq_ids,
                   const int32_t* n_uniq, int64_t max_uniq, const float* dE_rows, int ldg, int step,
                   float lr, float b1, float b2, float eps, void* bf16_copy, int ld_copy, void* stream) {
  CLSM_REQUIRE(D > 0 && step >= 1 && max_uniq > 0 && ldg >= D, "adam_rows: bad sizes");
  const AdamC c = make_c(step, lr, b1, b2, eps);
  cudaStream_t st = as_stream(stream);
  __nv_bfloat16* cp = static_cast<__nv_bfloat16*>(bf16_copy);
  if (D % 4 == 0 && ldg % 4 == 0 && (cp == nullptr || ld_copy % 4 == 0)) {
    const int64_t items = max_uniq * (D / 4);
    adam_table_kernel<true><<<static_cast<unsigned>(ceil_div64(items, 256)), 256, 0, st>>>(p, m, v, 0, D, nullptr, uniq_ids, n_uniq, dE_rows, ldg, c, cp, ld_copy);
  } else {
    const int64_t items = max_uniq * D;
    adam_table_kernel<false><<<static_cast<unsigned>(ceil_div64(items, 256)), 256, 0, st>>>(p, m, v, 0, D, nullptr, uniq_ids, n_uniq, dE_rows, ldg, c, cp, ld_copy);
  }
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

/* ---- deferred rows (see the kernels above) ---- */
int clsm_adam_lr_table(float lr, float b1, float b2, int first_step, int n, float* lr_t_host) {
  CLSM_REQUIRE(first_step >= 0 && n >= 0 && lr_t_host != nullptr, "adam_lr_table: bad arguments");
  for (int i = 0; i < n; ++i) {
    const int s = first_step + i;
    lr_t_host[i] = s >= 1 ? make_c(s, lr, b1, b2, 0.0f).lr_t : 0.0f;     // the value clsm_adam_table uses at step s
  }
  return CLSM_OK;
}

int clsm_adam_defer_settle_tokens(float* p, float* m, float* v, int64_t n_rows, int D, int32_t* last,
                                  const int32_t* tok, int64_t n_tok, int to_step, const float* lr_tab,
                                  float b1, float b2, float eps, void* bf16_copy, int ld_copy,
                                  void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(n_rows > 0 && D > 0 && n_tok >= 0 && to_step >= 0, "adam_defer_settle_tokens: bad sizes");
  CLSM_REQUIRE(ws_bytes >= clsm_adam_defer_workspace_bytes(n_tok, n_rows), "adam_defer_settle_tokens: workspace too small");
  if (n_tok == 0 || to_step == 0) return CLSM_OK;
  cudaStream_t st = as_stream(stream);
  int32_t* ctl = static_cast<int32_t*>(ws);
  int32_t* work = ctl + 64;
  CLSM_CUDA(cudaMemsetAsync(ctl, 0, 8, st));
  defer_claim_kernel<<<static_cast<unsigned>(ceil_div64(n_tok, 256)), 256, 0, st>>>(tok, n_tok, n_rows, last, to_step, work, ctl);
  CLSM_LAUNCH_CHECK();
  defer_settle_list_kernel<<<clsm_sm_count() * 8, 256, 0, st>>>(p, m, v, D, to_step, work, ctl, lr_tab, make_d(b1, b2, eps),
                                                                static_cast<__nv_bfloat16*>(bf16_copy), ld_copy);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

size_t clsm_adam_defer_workspace_bytes(int64_t n_tok, int64_t n_rows) {
  const int64_t cap = n_tok < n_rows ? n_tok : n_rows;
  return 256 + static_cast<size_t>(cap > 0 ? cap : 0) * 8;
}

int clsm_adam_defer_rows(float* p, float* m, float* v, int D, int32_t* last, const int32_t* uniq_ids,
                         const int32_t* n_uniq, int64_t max_uniq, const float* dE_rows, int ldg, int step,
                         const float* lr_tab, float b1, float b2, float eps, void* bf16_copy, int ld_copy, void* stream) {
  CLSM_REQUIRE(D > 0 && step >= 1 && max_uniq > 0 && ldg >= D, "adam_defer_rows: bad sizes");
  int64_t blocks = ceil_div64(max_uniq, 8);
  const int64_t cap = static_cast<int64_t>(clsm_sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  defer_adam_rows_kernel<<<static_cast<unsigned>(blocks), 256, 0, as_stream(stream)>>>(p, m, v, D, last, uniq_ids, n_uniq, dE_rows, ldg, step,
                                                                                       lr_tab, make_d(b1, b2, eps),
                                                                                       static_cast<__nv_bfloat16*>(bf16_copy), ld_copy);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_adam_defer_settle_all(float* p, float* m, float* v, int64_t n_rows, int D, int32_t* last, int to_step,
                               const float* lr_tab, float b1, float b2, float eps, void* bf16_copy, int ld_copy, void* stream) {
  CLSM_REQUIRE(n_rows > 0 && D > 0 && to_step >= 0, "adam_defer_settle_all: bad sizes");
  if (to_step == 0) return CLSM_OK;
  defer_settle_all_kernel<<<clsm_sm_count() * 8, 256, 0, as_stream(stream)>>>(p, m, v, n_rows, D, last, to_step, lr_tab, make_d(b1, b2, eps),
                                                                              static_cast<__nv_bfloat16*>(bf16_copy), ld_copy);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

}  // extern "C"
