// Cosine top-k retrieval over a page-vector corpus.  No script in the reference does this; the
// score is RTF's cosine (dssm_cnn_v2/cnn_dssm_tf.py:63-77) on already-normalised vectors.
// Order is total: (score desc, id asc) -- every candidate is a 64-bit key
//   (ordered(score) << 32) | (0xFFFFFFFF - id32), all keys distinct, larger key = better,
// so the result does not depend on chunking, on the number of shards or on thread timing.
// Selection = exact radix select (8 passes of 8 bits) of the k-th largest key among the 2048
// candidates a CTA holds in registers, then a bitonic sort of the k survivors; levels repeat on
// the survivors until one list per query is left.
#include "common.cuh"
#include <stdlib.h>
#include <mutex>

namespace {

constexpr int TPB = 256;
constexpr int KPT = 8;                 // keys per thread
constexpr int R = TPB * KPT;           // candidates per CTA
constexpr int QG = 8;                  // queries per CTA at level 0
constexpr int KMAX = 128;

typedef unsigned long long u64;

__device__ __forceinline__ u64 make_key(float score, uint32_t id) {
  return (static_cast<u64>(f32_to_ordered(score)) << 32) | static_cast<u64>(0xFFFFFFFFu - id);
}

// Select the k largest of the CTA's keys (KPT per thread, 0 = empty) and write them sorted
// descending into s_out[0:k] (zero padded).  Uses s_hist[256], s_ctl[4]; all threads call.
__device__ void select_topk(const u64 (&keys)[KPT], int k, unsigned* s_hist, unsigned* s_ctl, u64* s_out) {
  const int tid = threadIdx.x;
  u64 prefix = 0, mask = 0;
  unsigned need = static_cast<unsigned>(k);
  for (int pass = 7; pass >= 0; --pass) {
    const int shift = pass * 8;
    s_hist[tid] = 0;                     // TPB == 256 bins
    __syncthreads();
#pragma unroll
    for (int i = 0; i < KPT; ++i)
      if ((keys[i] & mask) == prefix) atomicAdd(&s_hist[static_cast<unsigned>(keys[i] >> shift) & 255u], 1u);
    __syncthreads();
    if (tid < 32) {
      // lane l owns bins 255-8l .. 248-8l (descending digit order)
      unsigned c[8], tot = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) { c[i] = s_hist[255 - 8 * tid - i]; tot += c[i]; }
      unsigned incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const unsigned n = __shfl_up_sync(0xffffffffu, incl, o); if (tid >= o) incl += n; }
      const unsigned excl = incl - tot;   // count of keys in strictly higher bins
      const bool mine = excl < need && need <= incl;
      const unsigned any = __ballot_sync(0xffffffffu, mine);
      if (any == 0) { if (tid == 0) { s_ctl[0] = 0; s_ctl[1] = need; } }   // fewer than `need` candidates
      else if (mine) {
        unsigned run = excl;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (run < need && need <= run + c[i]) { s_ctl[0] = 255 - 8 * tid - i; s_ctl[1] = need - run; }
          run += c[i];
        }
      }
    }
    __syncthreads();
    prefix |= static_cast<u64>(s_ctl[0]) << shift;
    mask |= static_cast<u64>(0xFF) << shift;
    need = s_ctl[1];
  }
  // prefix is now the k-th largest key (or 0 when there are fewer than k candidates)
  if (tid < KMAX) s_out[tid] = 0;
  if (tid == 0) s_ctl[2] = 0;
  __syncthreads();
#pragma unroll
  for (int i = 0; i < KPT; ++i) {
    if (keys[i] != 0 && keys[i] >= prefix) {
      const unsigned slot = atomicAdd(&s_ctl[2], 1u);
      if (slot < KMAX) s_out[slot] = keys[i];
    }
  }
  __syncthreads();
  // bitonic sort of KMAX keys, descending
  for (int size = 2; size <= KMAX; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      if (tid < KMAX / 2) {
        const int lo = (tid / stride) * stride * 2 + (tid % stride);
        const int hi = lo + stride;
        const bool desc = ((lo & size) == 0);
        const u64 a = s_out[lo], b = s_out[hi];
        if ((a < b) == desc) { s_out[lo] = b; s_out[hi] = a; }
      }
      __syncthreads();
    }
  }
}

// level 0: grid (ceil(nQ/QG), n_chunks).  Each thread scores KPT corpus rows against QG queries.
__global__ void __launch_bounds__(TPB)
topk_score_kernel(const float* __restrict__ Qn, int64_t nQ, const float* __restrict__ Cn, int64_t N, int H, int k,
                  u64* __restrict__ out, int64_t out_stride_q /* = n_chunks*k */) {
  extern __shared__ __align__(16) float s_q[];              // QG * H
  __shared__ unsigned s_hist[256];
  __shared__ unsigned s_ctl[4];
  __shared__ u64 s_out[KMAX];
  const int tid = threadIdx.x;
  const int64_t q0 = static_cast<int64_t>(blockIdx.x) * QG;
  const int64_t chunk = blockIdx.y;
  for (int i = tid; i < QG * H; i += TPB) {
    const int64_t q = q0 + i / H;
    s_q[i] = q < nQ ? Qn[q * H + (i % H)] : 0.0f;
  }
  __syncthreads();
  float sc[KPT][QG];
#pragma unroll
  for (int i = 0; i < KPT; ++i)
#pragma unroll
    for (int q = 0; q < QG; ++q) sc[i][q] = 0.0f;
  const int64_t row0 = chunk * R;
#pragma unroll
  for (int i = 0; i < KPT; ++i) {
    const int64_t row = row0 + tid + i * TPB;
    if (row < N) {
      if ((H & 3) == 0) {
        const float4* c4 = reinterpret_cast<const float4*>(Cn + row * H);
        for (int h = 0; h < H / 4; ++h) {
          const float4 c = __ldg(c4 + h);
#pragma unroll
          for (int q = 0; q < QG; ++q) {
            const float4 w = *reinterpret_cast<const float4*>(s_q + q * H + h * 4);
            sc[i][q] = fmaf(c.x, w.x, fmaf(c.y, w.y, fmaf(c.z, w.z, fmaf(c.w, w.w, sc[i][q]))));
          }
        }
      } else {
        for (int h = 0; h < H; ++h) {
          const float c = __ldg(Cn + row * H + h);
#pragma unroll
          for (int q = 0; q < QG; ++q) sc[i][q] = fmaf(c, s_q[q * H + h], sc[i][q]);
        }
      }
    }
  }
  for (int q = 0; q < QG; ++q) {
    if (q0 + q >= nQ) break;
    u64 keys[KPT];
#pragma unroll
    for (int i = 0; i < KPT; ++i) {
      const int64_t row = row0 + tid + i * TPB;
      float s = 0.0f;
#pragma unroll
      for (int qq = 0; qq < QG; ++qq) if (qq == q) s = sc[i][qq];
      keys[i] = row < N ? make_key(s, static_cast<uint32_t>(row)) : 0ull;
    }
    select_topk(keys, k, s_hist, s_ctl, s_out);
    if (tid < k) out[(q0 + q) * out_stride_q + chunk * k + tid] = s_out[tid];
    __syncthreads();
  }
}

// level >= 1: grid (nQ, n_chunks); candidates in[q*in_stride + i], i < n_in
// CLEAR: the (thresholded) key matrix of the tensor-core scan is mostly empty; non-empty entries are zeroed
// after reading so that the matrix is all-empty again for the next segment (the scan only WRITES survivors).
template <bool CLEAR>
__global__ void __launch_bounds__(TPB)
topk_merge_kernel(const u64* __restrict__ in, int64_t in_stride, int64_t n_in, int k,
                  u64* __restrict__ out, int64_t out_stride) {
  __shared__ unsigned s_hist[256];
  __shared__ unsigned s_ctl[4];
  __shared__ u64 s_out[KMAX];
  const int tid = threadIdx.x;
  const int64_t q = blockIdx.x, chunk = blockIdx.y;
  u64 keys[KPT];
  int nz = 0;
#pragma unroll
  for (int i = 0; i < KPT; ++i) {
    const int64_t idx = chunk * R + tid + i * TPB;
    keys[i] = idx < n_in ? in[q * in_stride + idx] : 0ull;
    if (keys[i] != 0ull) {
      ++nz;
      if (CLEAR) const_cast<u64*>(in)[q * in_stride + idx] = 0ull;
    }
  }
  // few survivors (the common case behind a threshold): no radix passes, just gather and sort them
  const int total = __syncthreads_count(nz > 0) == 0 ? 0 : -1;
  if (total == 0) {
    if (tid < k) out[q * out_stride + chunk * k + tid] = 0ull;
    return;
  }
  if (tid == 0) s_ctl[3] = 0;
  __syncthreads();
  if (nz) atomicAdd(&s_ctl[3], static_cast<unsigned>(nz));
  __syncthreads();
  if (s_ctl[3] <= static_cast<unsigned>(k)) {
    if (tid < KMAX) s_out[tid] = 0;
    if (tid == 0) s_ctl[2] = 0;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < KPT; ++i)
      if (keys[i] != 0ull) s_out[atomicAdd(&s_ctl[2], 1u)] = keys[i];
    __syncthreads();
    for (int size = 2; size <= KMAX; size <<= 1) {
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        if (tid < KMAX / 2) {
          const int lo = (tid / stride) * stride * 2 + (tid % stride);
          const int hi = lo + stride;
          const bool desc = ((lo & size) == 0);
          const u64 a = s_out[lo], b = s_out[hi];
          if ((a < b) == desc) { s_out[lo] = b; s_out[hi] = a; }
        }
        __syncthreads();
      }
    }
  } else {
    select_topk(keys, k, s_hist, s_ctl, s_out);
  }
  if (tid < k) out[q * out_stride + chunk * k + tid] = s_out[tid];
}

__global__ void topk_decode_kernel(const u64* __restrict__ keys, int64_t n, int64_t id_base,
                                   float* __restrict__ score, int64_t* __restrict__ id) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const u64 key = keys[i];
  if (key == 0) { score[i] = -INFINITY; id[i] = -1; return; }
  score[i] = ordered_to_f32(static_cast<uint32_t>(key >> 32));
  id[i] = id_base + static_cast<int64_t>(0xFFFFFFFFu - static_cast<uint32_t>(key & 0xFFFFFFFFull));
}

// Merge keys carry 32 bits of id, the lists carry int64 ids (shard row + id_base, and id_base is unrestricted):
// keys hold id - lo, lo = the smallest valid id of the call (found by the pass below), and the decode adds lo back.
// ctl[0] = lo (starts at ~0), ctl[1] = 1 if some id lies 2^32 - 1 or more above lo (the call then reports failure
// in its outputs, see clsm_b200.h).
__global__ void __launch_bounds__(256)
topk_min_id_kernel(const int64_t* __restrict__ id, int64_t n, u64* __restrict__ ctl) {
  __shared__ u64 s_min[8];
  u64 m = ~0ull;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t v = id[i];
    if (v >= 0 && static_cast<u64>(v) < m) m = static_cast<u64>(v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { const u64 n2 = __shfl_xor_sync(0xffffffffu, m, o); m = n2 < m ? n2 : m; }
  if ((threadIdx.x & 31) == 0) s_min[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) m = s_min[w] < m ? s_min[w] : m;
    if (m != ~0ull) atomicMin(ctl, m);
  }
}

__global__ void topk_encode_kernel(const float* __restrict__ score, const int64_t* __restrict__ id,
                                   int n_parts, int64_t nQ, int k, u64* __restrict__ keys, u64* __restrict__ ctl) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t n = static_cast<int64_t>(n_parts) * nQ * k;
  if (i >= n) return;
  const int64_t part = i / (nQ * k);
  const int64_t rem = i - part * nQ * k;
  const int64_t q = rem / k;
  const int j = static_cast<int>(rem - q * k);
  const int64_t idv = id[i];
  u64 key = 0ull;
  if (idv >= 0) {
    const u64 rel = static_cast<u64>(idv) - ctl[0];
    if (rel >= 0xFFFFFFFFull) ctl[1] = 1ull;                    // benign race: every writer stores the same value
    else key = make_key(score[i], static_cast<uint32_t>(rel));
  }
  keys[q * (static_cast<int64_t>(n_parts) * k) + part * k + j] = key;
}

__global__ void topk_decode_merge_kernel(const u64* __restrict__ keys, int64_t n, const u64* __restrict__ ctl,
                                         float* __restrict__ score, int64_t* __restrict__ id) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (ctl[1] != 0ull) { score[i] = __uint_as_float(0x7fc00000u); id[i] = -2; return; }      // id span too wide: refuse loudly
  const u64 key = keys[i];
  if (key == 0) { score[i] = -INFINITY; id[i] = -1; return; }
  score[i] = ordered_to_f32(static_cast<uint32_t>(key >> 32));
  id[i] = static_cast<int64_t>(ctl[0]) + static_cast<int64_t>(0xFFFFFFFFu - static_cast<uint32_t>(key & 0xFFFFFFFFull));
}

// runs merge levels on keys (nQ, n_in) in buf[src] until one list of k per query is left;
// returns the buffer index that holds (nQ, k).
int run_merge_levels(u64* buf[2], int src, int64_t nQ, int64_t n_in, int k, cudaStream_t st) {
  while (true) {
    const int64_t n_chunks = ceil_div64(n_in, R);
    const int dst = src ^ 1;
    dim3 grid(static_cast<unsigned>(nQ), static_cast<unsigned>(n_chunks));
    topk_merge_kernel<false><<<grid, TPB, 0, st>>>(buf[src], n_in, n_in, k, buf[dst], n_chunks * k);
    ++g_clsm_launches;
    src = dst;
    n_in = n_chunks * k;
    if (n_chunks == 1) break;
  }
  return src;
}

size_t level_bytes(int64_t nQ, int64_t n_lists, int k) { return align_up(static_cast<size_t>(nQ) * n_lists * k * 8, 256); }

}  // namespace

extern "C" {

size_t clsm_cos_topk_workspace_bytes(int64_t nQ, int64_t N, int k) {
  const int64_t c0 = ceil_div64(N, R);
  const int64_t c1 = ceil_div64(c0 * k, R);
  return level_bytes(nQ, c0, k) + level_bytes(nQ, c1 > 1 ? c1 : 1, k);
}

int clsm_cos_topk(const float* Qn, int64_t nQ, const float* Cn, int64_t N, int H, int k,
                  int64_t id_base, float* out_score, int64_t* out_id,
                  void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(nQ > 0 && N > 0 && N < (1ll << 32) && H > 0, "cos_topk: bad sizes");
  CLSM_REQUIRE(k >= 1 && k <= KMAX, "cos_topk: k must be in [1, %d]", KMAX);
  CLSM_REQUIRE(static_cast<size_t>(QG) * H * 4 <= 40 * 1024, "cos_topk: H too large");
  if (ws_bytes < clsm_cos_topk_workspace_bytes(nQ, N, k)) {
    clsm_set_error("cos_topk: workspace too small (%zu < %zu)", ws_bytes, clsm_cos_topk_workspace_bytes(nQ, N, k));
    return CLSM_EWORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  const int64_t c0 = ceil_div64(N, R);
  u64* buf[2];
  buf[0] = static_cast<u64*>(ws);
  buf[1] = reinterpret_cast<u64*>(static_cast<char*>(ws) + level_bytes(nQ, c0, k));
  dim3 grid(static_cast<unsigned>(ceil_div64(nQ, QG)), static_cast<unsigned>(c0));
  topk_score_kernel<<<grid, TPB, static_cast<size_t>(QG) * H * 4, st>>>(Qn, nQ, Cn, N, H, k, buf[0], c0 * k);
  CLSM_LAUNCH_CHECK();
  int src = 0;
  if (c0 > 1) { src = run_merge_levels(buf, 0, nQ, c0 * k, k, st); CLSM_CUDA(cudaGetLastError()); }
  const int64_t n = nQ * k;
  topk_decode_kernel<<<static_cast<unsigned>(ceil_div64(n, 256)), 256, 0, st>>>(buf[src], n, id_base, out_score, out_id);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

size_t clsm_topk_merge_workspace_bytes(int n_parts, int64_t nQ, int k) {
  const int64_t c1 = ceil_div64(static_cast<int64_t>(n_parts) * k, R);
  return level_bytes(nQ, n_parts, k) + level_bytes(nQ, c1 > 1 ? c1 : 1, k) + 256 /* id base + range flag */;
}

int clsm_topk_merge(const float* scores, const int64_t* ids, int n_parts, int64_t nQ, int k,
                    float* out_score, int64_t* out_id, void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(n_parts >= 1 && nQ > 0 && k >= 1 && k <= KMAX, "topk_merge: bad sizes");
  if (ws_bytes < clsm_topk_merge_workspace_bytes(n_parts, nQ, k)) {
    clsm_set_error("topk_merge: workspace too small");
    return CLSM_EWORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  const int64_t c1 = ceil_div64(static_cast<int64_t>(n_parts) * k, R);
  u64* buf[2];
  buf[0] = static_cast<u64*>(ws);
  buf[1] = reinterpret_cast<u64*>(static_cast<char*>(ws) + level_bytes(nQ, n_parts, k));
  u64* ctl = reinterpret_cast<u64*>(static_cast<char*>(ws) + level_bytes(nQ, n_parts, k) + level_bytes(nQ, c1 > 1 ? c1 : 1, k));
  const int64_t n_in = static_cast<int64_t>(n_parts) * nQ * k;
  CLSM_CUDA(cudaMemsetAsync(ctl, 0xFF, 8, st));
  CLSM_CUDA(cudaMemsetAsync(ctl + 1, 0, 8, st));
  const int64_t min_blocks = ceil_div64(n_in, 256);
  topk_min_id_kernel<<<static_cast<unsigned>(min_blocks < 1024 ? min_blocks : 1024), 256, 0, st>>>(ids, n_in, ctl);
  CLSM_LAUNCH_CHECK();
  topk_encode_kernel<<<static_cast<unsigned>(ceil_div64(n_in, 256)), 256, 0, st>>>(scores, ids, n_parts, nQ, k, buf[0], ctl);
  CLSM_LAUNCH_CHECK();
  const int src = run_merge_levels(buf, 0, nQ, static_cast<int64_t>(n_parts) * k, k, st);
  CLSM_LAUNCH_CHECK();
  const int64_t n = nQ * k;
  topk_decode_merge_kernel<<<static_cast<unsigned>(ceil_div64(n, 256)), 256, 0, st>>>(buf[src], n, ctl, out_score, out_id);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

}  // extern "C"

void clsm_score_tc_geometry(int64_t rows, int* grid, int* nsub, int64_t* subcap);       // topk_tc.cu

namespace {

// ---- bf16 corpus: tensor-core scoring (topk_tc.cu) + the same exact selection --------------------------
// Per 128-query block the corpus is scanned once, in segments.  Every segment is scanned behind the k-th best key
// found so far (none for the first): a row that beats it is appended to a compact per-query list -- one private
// sub-list per (CTA, column half), so an append is a register increment and a store -- and one CTA per query folds the
// lists into the running best-k, whose k-th key is the next segment's threshold.  A segment boundary costs a scan
// launch and a fold (~20 us), so segments grow 16-fold: 4 096 rows (scanned DENSE: no threshold exists yet), then to
// 65 536, to 1 Mi, and 1 Mi at a time from there on.  (Round 1 scanned the first 65 536 rows into a dense
// 67 MB key matrix and appended survivors with a global atomic each: 47 % of a 1 M-row call went to the dense rows
// and a segment whose threshold let 1 row in 600 through ran 3x slower than one without survivors.)
constexpr int64_t SEG0 = 4096;                  // first segment: no threshold yet
constexpr int64_t SEG_MAX = 1 << 20;            // 1 Mi rows
constexpr int64_t GROW = 16;
inline int64_t next_seg_len(int64_t done) {     // length of the segment that starts at row `done`
  const int64_t target = done * GROW;           // scan up to 16 x what has been scanned
  const int64_t len = target - done;
  return len < SEG_MAX ? len : SEG_MAX;
}

__global__ void kth_key_kernel(const u64* __restrict__ best, int64_t nq, int k, u64* __restrict__ thr) {
  const int64_t q = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (q < nq) thr[q] = best[q * k + (k - 1)];          // 0 while fewer than k candidates exist
}
__global__ void normalize_rows_bf16_kernel(const float* __restrict__ sem, int64_t N, int H, __nv_bfloat16* __restrict__ vec) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * (blockDim.x / 32) + warp;
  if (r >= N) return;
  float s = 0.0f;
  for (int h = lane; h < H; h += 32) { const float x = sem[r * H + h]; s = fmaf(x, x, s); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float n = sqrtf(fmaxf(s, FLT_MIN));
  for (int h = lane; h < H; h += 32) vec[r * H + h] = __float2bfloat16_rn(sem[r * H + h] / n);
}

// best (k, sorted descending, 0 = empty) <- k largest of best U (the query's nsub sub-lists); thr <- its k-th key
constexpr int NSUB_MAX = 1024;
__global__ void __launch_bounds__(TPB)
select_compact_kernel(const u64* __restrict__ lists, int64_t cap, const int* __restrict__ cnt, int nsub, int64_t subcap, int k,
                      u64* __restrict__ best, u64* __restrict__ thr) {
  __shared__ unsigned s_hist[256];
  __shared__ unsigned s_ctl[4];
  __shared__ u64 s_out[KMAX];
  __shared__ u64 s_best[KMAX];
  __shared__ int s_off[NSUB_MAX + 1];           // exclusive prefix sums of the sub-list lengths
  const int tid = threadIdx.x;
  const int64_t q = blockIdx.x;
  const u64* list = lists + q * cap;
  for (int i = tid; i < nsub; i += TPB) {
    int c = cnt[q * nsub + i];
    s_off[i + 1] = c < subcap ? c : static_cast<int>(subcap);
  }
  if (tid == 0) s_off[0] = 0;
  if (tid < KMAX) s_best[tid] = tid < k ? best[q * k + tid] : 0ull;
  __syncthreads();
  if (tid < 32) {                               // inclusive scan of s_off[1..nsub] by one warp, 32 entries at a time
    int carry = 0;
    for (int b0 = 1; b0 <= nsub; b0 += 32) {
      const int i = b0 + tid;
      int v = i <= nsub ? s_off[i] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int n2 = __shfl_up_sync(0xffffffffu, v, o); if (tid >= o) v += n2; }
      if (i <= nsub) s_off[i] = v + carry;
      carry += __shfl_sync(0xffffffffu, v, 31);
    }
  }
  __syncthreads();
  const int64_t n = s_off[nsub];
  for (int64_t off = 0; off < n; off += R - KMAX) {
    u64 keys[KPT];
#pragma unroll
    for (int i = 0; i < KPT; ++i) {
      const int idx = tid + i * TPB;
      if (idx < KMAX) keys[i] = s_best[idx];
      else {
        const int64_t j = off + idx - KMAX;
        keys[i] = 0ull;
        if (j < n) {                            // flat index -> (sub-list, position): the last s with s_off[s] <= j
          int lo = 0, hi = nsub;
          while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (s_off[mid] <= j) lo = mid; else hi = mid; }
          keys[i] = list[static_cast<int64_t>(lo) * subcap + (j - s_off[lo])];
        }
      }
    }
    select_topk(keys, k, s_hist, s_ctl, s_out);
    __syncthreads();
    if (tid < KMAX) s_best[tid] = tid < k ? s_out[tid] : 0ull;
    __syncthreads();
  }
  if (tid < k) best[q * k + tid] = s_best[tid];
  if (tid == 0) thr[q] = s_best[k - 1];
}

struct TcPlan { size_t off_keys, off_l1, off_best, off_thr, off_cnt, total; int64_t cap; int max_nsub; };
TcPlan tc_plan(int k, int64_t N) {
  TcPlan p{};
  size_t o = 0;
  auto take = [&](size_t b) { size_t r = o; o += align_up(b, 256); return r; };
  // list capacity per query = the largest sub-list footprint any segment of this corpus needs
  p.cap = 0; p.max_nsub = 0;
  for (int64_t done = SEG0; done < N;) {
    int64_t len = next_seg_len(done);
    if (len > N - done) len = N - done;
    int grid, nsub; int64_t subcap;
    clsm_score_tc_geometry(len, &grid, &nsub, &subcap);
    if (nsub * subcap > p.cap) p.cap = nsub * subcap;
    if (nsub > p.max_nsub) p.max_nsub = nsub;
    done += len;
  }
  if (p.cap < SEG0) p.cap = SEG0;                                 // the dense first segment uses the same region
  p.cap = (p.cap + 1) / 2 * 2;
  p.off_keys = take(static_cast<size_t>(128) * p.cap * 8);
  p.off_l1 = take(static_cast<size_t>(128) * (SEG0 / R) * k * 8);
  p.off_cnt = take(static_cast<size_t>(128) * p.max_nsub * 4);
  p.off_best = take(static_cast<size_t>(128) * k * 8);
  p.off_thr = take(128 * 8);
  p.total = o;
  (void)k;
  return p;
}

}  // namespace

int clsm_score_keys_tc(const void* Qn, int64_t nQ, int64_t q_begin, const void* Cn, int64_t N, int H,
                       int64_t row_begin, int64_t row_end, unsigned long long* keys, int64_t seg_cap,
                       const unsigned long long* thr, int* cnt, cudaStream_t st);     // topk_tc.cu

extern "C" {

int clsm_normalize_rows_bf16(const float* sem, int64_t N, int H, void* vec, void* stream) {
  CLSM_REQUIRE(N >= 0 && H > 0, "normalize_rows_bf16: bad sizes");
  if (N == 0) return CLSM_OK;
  normalize_rows_bf16_kernel<<<static_cast<unsigned>(ceil_div64(N, 8)), 256, 0, as_stream(stream)>>>(sem, N, H, static_cast<__nv_bfloat16*>(vec));
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

size_t clsm_cos_topk_bf16_workspace_bytes(int64_t nQ, int64_t N, int k) { (void)nQ; return tc_plan(k, N > 0 ? N : 1).total; }

int clsm_cos_topk_bf16(const void* Qn, int64_t nQ, const void* Cn, int64_t N, int H, int k,
                       int64_t id_base, float* out_score, int64_t* out_id,
                       void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(nQ > 0 && N > 0 && N < (1ll << 31), "cos_topk_bf16: bad sizes");
  CLSM_REQUIRE(k >= 1 && k <= KMAX, "cos_topk_bf16: k must be in [1, %d]", KMAX);
  const TcPlan p = tc_plan(k, N);
  if (ws_bytes < p.total) { clsm_set_error("cos_topk_bf16: workspace too small (%zu < %zu)", ws_bytes, p.total); return CLSM_EWORKSPACE; }
  cudaStream_t st = as_stream(stream);
  char* w = static_cast<char*>(ws);
  u64* keys = reinterpret_cast<u64*>(w + p.off_keys);
  u64* l1 = reinterpret_cast<u64*>(w + p.off_l1);
  u64* best = reinterpret_cast<u64*>(w + p.off_best);
  u64* thr = reinterpret_cast<u64*>(w + p.off_thr);
  int* cnt = reinterpret_cast<int*>(w + p.off_cnt);
  CLSM_REQUIRE(p.max_nsub <= NSUB_MAX, "cos_topk_bf16: too many SMs for the fold kernel's table");
  // CLSM_TOPK_PROF=1 (diagnostics, synchronises): device time of every launch of the first query block, from events
  static int prof_on = -1;
  if (prof_on < 0) { const char* e = getenv("CLSM_TOPK_PROF"); prof_on = (e != nullptr && e[0] == '1') ? 1 : 0; }
  cudaEvent_t pev[64]; int64_t plen[64]; int npev = 0;
  auto mark = [&](int64_t len) { if (prof_on && npev < 64) { cudaEventCreate(&pev[npev]); cudaEventRecord(pev[npev], st); plen[npev++] = len; } };
  for (int64_t qb = 0; qb < nQ; qb += 128) {
    const int64_t nq = (nQ - qb < 128) ? nQ - qb : 128;
    CLSM_CUDA(cudaMemsetAsync(best, 0, static_cast<size_t>(nq) * k * 8, st));
    CLSM_CUDA(cudaMemsetAsync(thr, 0, 128 * 8, st));          // no threshold yet: the first segment keeps every row
    // first segment, DENSE (no threshold exists yet, so every row would survive: compact lists are the wrong tool): the
    // scan writes each key at its row position, two selection levels reduce the 2048-key chunks to the block's best-k.
    // The level-1 select zeroes what it reads; the region must start all-empty (compact lists were in it).
    {
      const int64_t r1 = SEG0 < N ? SEG0 : N;
      CLSM_CUDA(cudaMemsetAsync(keys, 0, static_cast<size_t>(128) * SEG0 * 8, st));
      if (qb == 0) mark(r1);
      int rc = clsm_score_keys_tc(Qn, nQ, qb, Cn, N, H, 0, r1, keys, SEG0, nullptr, nullptr, st);
      if (rc != CLSM_OK) return rc;
      if (qb == 0) mark(-1);
      const int64_t n_chunks = ceil_div64(r1, R);
      dim3 g1(static_cast<unsigned>(nq), static_cast<unsigned>(n_chunks));
      topk_merge_kernel<true><<<g1, TPB, 0, st>>>(keys, SEG0, ((r1 + 1) / 2) * 2, k, l1, n_chunks * k);
      CLSM_LAUNCH_CHECK();
      dim3 g2(static_cast<unsigned>(nq), 1);             // SEG0 / R lists of k keys: one more level
      topk_merge_kernel<false><<<g2, TPB, 0, st>>>(l1, n_chunks * k, n_chunks * k, k, best, k);
      CLSM_LAUNCH_CHECK();
      kth_key_kernel<<<1, 128, 0, st>>>(best, nq, k, thr);
      CLSM_LAUNCH_CHECK();
    }
    for (int64_t done = SEG0; done < N;) {
      int64_t len = next_seg_len(done);
      if (len > N - done) len = N - done;
      int grid, nsub; int64_t subcap;
      clsm_score_tc_geometry(len, &grid, &nsub, &subcap);
      if (qb == 0) mark(len);
      int rc = clsm_score_keys_tc(Qn, nQ, qb, Cn, N, H, done, done + len, keys, p.cap, thr, cnt, st);
      if (rc != CLSM_OK) return rc;
      if (qb == 0) mark(-1);
      select_compact_kernel<<<static_cast<unsigned>(nq), TPB, 0, st>>>(keys, p.cap, cnt, nsub, subcap, k, best, thr);
      CLSM_LAUNCH_CHECK();
      done += len;
    }
    if (qb == 0 && prof_on) {
      mark(0);
      cudaStreamSynchronize(st);
      for (int i = 0; i + 1 < npev; ++i) {
        float ms = 0.f; cudaEventElapsedTime(&ms, pev[i], pev[i + 1]);
        if (plen[i] > 0) fprintf(stderr, "[topk prof] scan %8lld rows %.1f us", static_cast<long long>(plen[i]), ms * 1e3f);
        else fprintf(stderr, " | fold %.1f us\n", ms * 1e3f);
      }
      for (int i = 0; i < npev; ++i) cudaEventDestroy(pev[i]);
    }
    const int64_t n = nq * k;
    topk_decode_kernel<<<static_cast<unsigned>(ceil_div64(n, 256)), 256, 0, st>>>(best, n, id_base, out_score + qb * k, out_id + qb * k);
    CLSM_LAUNCH_CHECK();
  }
  return CLSM_OK;
}

// ---- clsm_topk_merge_static: the same merge, for callers that must not allocate ---------------------
// Built-in scratch.  It lives in every GPU context that loads this library, from load until the process
// exits, so it is kept small (CLSM_TOPK_MERGE_STATIC_MAX_BYTES = 2 MiB: e.g. 8 shards x 128 queries x
// k=100 needs 0.9 MiB).  In host code the NAME of a __device__ array is only a host-side stand-in and
// must never be handed to a kernel; its real device address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_topk_merge_scratch[CLSM_TOPK_MERGE_STATIC_MAX_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxDevices];          // device address of g_topk_merge_scratch, per device ordinal

// One-time set-up for the CURRENT device: resolve the array's device address and remember it.
int topk_merge_static_setup(void** out) {
  int dev = -1;
  CLSM_CUDA(cudaGetDevice(&dev));
  CLSM_REQUIRE(dev >= 0 && dev < kMaxDevices, "topk_merge_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    CLSM_CUDA(cudaGetSymbolAddress(&p, g_topk_merge_scratch));      // device `dev` is current here
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return CLSM_OK;
}
}  // namespace

int clsm_topk_merge_static(const float* scores, const int64_t* ids, int n_parts, int64_t nQ, int k,
                           float* out_score, int64_t* out_id, void* stream) {
  CLSM_REQUIRE(n_parts >= 1 && nQ > 0 && k >= 1 && k <= KMAX, "topk_merge_static: bad sizes");
  const size_t need = clsm_topk_merge_workspace_bytes(n_parts, nQ, k);
  if (need > static_cast<size_t>(CLSM_TOPK_MERGE_STATIC_MAX_BYTES)) {
    clsm_set_error("topk_merge_static: %d lists x %lld queries x k=%d needs %zu bytes of scratch but the built-in "
                   "array holds %d; call clsm_topk_merge with a caller-provided workspace instead",
                   n_parts, static_cast<long long>(nQ), k, need, CLSM_TOPK_MERGE_STATIC_MAX_BYTES);
    return CLSM_EWORKSPACE;
  }
  void* scratch = nullptr;
  const int rc = topk_merge_static_setup(&scratch);
  if (rc != CLSM_OK) return rc;
  // Every byte the merge kernels read from the scratch is written earlier in the same call (the two control
  // words are re-initialised first, the key encoding pass fills level 0 completely, each level fills the next),
  // so the previous call's contents never matter.
  return clsm_topk_merge(scores, ids, n_parts, nQ, k, out_score, out_id, scratch,
                         static_cast<size_t>(CLSM_TOPK_MERGE_STATIC_MAX_BYTES), stream);
}

}  // extern "C"
