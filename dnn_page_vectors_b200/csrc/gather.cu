// Embedding gather (+ inverted dropout) and the table format conversions.
//   replaces Embedding + Dropout of dssm_cnn_v2/cnn_dssm_tf.py:111-122.
// HBM-bound: one thread moves one 16-column block of one token row with 128-bit loads/stores;
// consecutive threads take consecutive blocks of the same row, so a warp reads and writes
// contiguous 512 B (bf16) / 1 KB (fp32) runs.
#include "common.cuh"

namespace {

// bf16: 16 columns = 2 x uint4.  fp32: 16 columns = 4 x uint4.
template <bool BF16, bool DROPOUT>
__global__ void __launch_bounds__(256)
emb_gather_kernel(const int32_t* __restrict__ tok, int64_t n_tok, const uint4* __restrict__ table,
                  int in_blocks,          // readable 16-col blocks per table row (ceil(ld_table/16))
                  int ld_table_v,         // table row pitch in uint4
                  int in_vec_valid,       // valid uint4 per table row (ld_table*esize/16)
                  uint4* __restrict__ X, int out_blocks, int ldx_v,
                  float inv_keep, uint32_t thr, uint64_t seed,
                  uint16_t* __restrict__ mask_bits) {    // optional (n_tok, out_blocks): bit c = column 16*blk + c kept
  constexpr int VPB = BF16 ? 2 : 4;       // uint4 per 16-column block
  const int64_t item = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t total = n_tok * out_blocks;
  if (item >= total) return;
  const int64_t r = item / out_blocks;
  const int blk = static_cast<int>(item - r * out_blocks);
  const int64_t id = __ldg(tok + r);
  uint4 v[VPB];
#pragma unroll
  for (int i = 0; i < VPB; ++i) {
    const int vi = blk * VPB + i;
    v[i] = (blk < in_blocks && vi < in_vec_valid) ? __ldg(table + id * ld_table_v + vi)
                                                  : make_uint4(0u, 0u, 0u, 0u);
  }
  if (DROPOUT) {
    uint32_t bytes[4];
    dropout_bytes16(seed, static_cast<uint64_t>(r), static_cast<uint32_t>(blk), bytes);
    if (mask_bits != nullptr) {              // the backward reads these bits instead of regenerating the stream
      uint32_t w = 0;
#pragma unroll
      for (int c = 0; c < 16; ++c) w |= dropout_keep(bytes, c, thr) ? (1u << c) : 0u;
      mask_bits[item] = static_cast<uint16_t>(w);
    }
    if (BF16) {
      __nv_bfloat16* e = reinterpret_cast<__nv_bfloat16*>(v);
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        float f = __bfloat162float(e[c]);
        f = dropout_keep(bytes, c, thr) ? f * inv_keep : 0.0f;
        e[c] = __float2bfloat16_rn(f);
      }
    } else {
      float* e = reinterpret_cast<float*>(v);
#pragma unroll
      for (int c = 0; c < 16; ++c) e[c] = dropout_keep(bytes, c, thr) ? e[c] * inv_keep : 0.0f;
    }
  }
  uint4* dst = X + r * ldx_v + blk * VPB;
#pragma unroll
  for (int i = 0; i < VPB; ++i) dst[i] = v[i];
}

// bf16 table + dropout, the training forward: the generic kernel above spends ~20 instructions per element on the
// mask (byte extraction, compare, select, two conversions); here the 16 keep decisions of a block come from four SIMD
// byte compares, are applied to PACKED bf16 pairs with one permute + one AND per pair, and the stored mask bits fall
// out of the same compare results.  Same values bit for bit (fp32 multiply by 256/t, round to nearest even, dropped -> +0).
__global__ void __launch_bounds__(256)
emb_gather_bf16_drop_kernel(const int32_t* __restrict__ tok, int64_t n_tok, const uint4* __restrict__ table,
                            int in_blocks, int ld_table_v, int in_vec_valid,
                            uint4* __restrict__ X, int out_blocks, int ldx_v,
                            float inv_keep, uint32_t thr, uint64_t seed, uint16_t* __restrict__ mask_bits) {
  const int64_t item = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (item >= n_tok * out_blocks) return;
  const int64_t r = item / out_blocks;
  const int blk = static_cast<int>(item - r * out_blocks);
  const int64_t id = __ldg(tok + r);
  uint4 v[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int vi = blk * 2 + i;
    v[i] = (blk < in_blocks && vi < in_vec_valid) ? __ldg(table + id * ld_table_v + vi) : make_uint4(0u, 0u, 0u, 0u);
  }
  uint32_t bytes[4];
  dropout_bytes16(seed, static_cast<uint64_t>(r), static_cast<uint32_t>(blk), bytes);
  const uint32_t thr4 = thr * 0x01010101u;                // thr <= 255 here
  uint32_t* x = reinterpret_cast<uint32_t*>(v);           // 8 words of two bf16 each; word 2w, 2w+1 <-> decision bytes of bytes[w]
  uint32_t bits = 0;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const uint32_t vm = __vcmpltu4(bytes[w], thr4);       // 0xFF per kept column
    bits |= ((((vm & 0x01010101u) * 0x01020408u) >> 24) & 0xFu) << (4 * w);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const uint32_t p = x[2 * w + h];
      const float lo = __uint_as_float(p << 16) * inv_keep, hi = __uint_as_float(p & 0xffff0000u) * inv_keep;
      const __nv_bfloat162 q = __floats2bfloat162_rn(lo, hi);
      x[2 * w + h] = *reinterpret_cast<const uint32_t*>(&q) & __byte_perm(vm, 0, h == 0 ? 0x1100 : 0x3322);
    }
  }
  if (mask_bits != nullptr) mask_bits[item] = static_cast<uint16_t>(bits);
  uint4* dst = X + r * ldx_v + blk * 2;
  dst[0] = v[0];
  dst[1] = v[1];
}

// out[r, c] = x[r, c] * mask(r, c) * inv_keep: the same counter-based mask as the embedding dropout, on any (n, H) matrix
__global__ void dropout_rows_kernel(const float* __restrict__ x, int64_t n, int H, float inv_keep, uint32_t thr, uint64_t seed,
                                    float* __restrict__ out) {
  const int nblk = (H + 15) / 16;
  const int64_t item = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (item >= n * nblk) return;
  const int64_t r = item / nblk;
  const int blk = static_cast<int>(item - r * nblk);
  uint32_t bytes[4];
  dropout_bytes16(seed, static_cast<uint64_t>(r), static_cast<uint32_t>(blk), bytes);
  for (int c = 0; c < 16; ++c) {
    const int col = blk * 16 + c;
    if (col < H) out[r * H + col] = dropout_keep(bytes, c, thr) ? x[r * H + col] * inv_keep : 0.0f;
  }
}

// fp32 MASTER table -> bf16 X (training, single GPU): the optimizer then need not maintain a bf16 copy of the table at
// all (0.64 GB less written per table and step).  X = bf16(E * keep_bit * 256/t): ONE rounding, where the bf16-table
// path rounds the table first.  Same Philox stream, same stored bits.
template <bool DROPOUT>
__global__ void __launch_bounds__(256)
emb_gather_master_kernel(const int32_t* __restrict__ tok, int64_t n_tok, const float4* __restrict__ table, int D4 /* float4 per row */,
                         uint4* __restrict__ X, int out_blocks, int ldx_v,
                         float inv_keep, uint32_t thr, uint64_t seed, uint16_t* __restrict__ mask_bits) {
  const int64_t item = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (item >= n_tok * out_blocks) return;
  const int64_t r = item / out_blocks;
  const int blk = static_cast<int>(item - r * out_blocks);
  const int64_t id = __ldg(tok + r);
  float4 f[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c4 = blk * 4 + i;
    f[i] = c4 < D4 ? __ldg(table + id * D4 + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  uint32_t x[8];
  uint32_t bits = 0xFFFFu;
  if (DROPOUT) {
    uint32_t bytes[4];
    dropout_bytes16(seed, static_cast<uint64_t>(r), static_cast<uint32_t>(blk), bytes);
    const uint32_t thr4 = thr * 0x01010101u;
    bits = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const uint32_t vm = __vcmpltu4(bytes[w], thr4);
      bits |= ((((vm & 0x01010101u) * 0x01020408u) >> 24) & 0xFu) << (4 * w);
      const __nv_bfloat162 q0 = __floats2bfloat162_rn(f[w].x * inv_keep, f[w].y * inv_keep);
      const __nv_bfloat162 q1 = __floats2bfloat162_rn(f[w].z * inv_keep, f[w].w * inv_keep);
      x[2 * w] = *reinterpret_cast<const uint32_t*>(&q0) & __byte_perm(vm, 0, 0x1100);
      x[2 * w + 1] = *reinterpret_cast<const uint32_t*>(&q1) & __byte_perm(vm, 0, 0x3322);
    }
    if (mask_bits != nullptr) mask_bits[item] = static_cast<uint16_t>(bits);
  } else {
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const __nv_bfloat162 q0 = __floats2bfloat162_rn(f[w].x, f[w].y), q1 = __floats2bfloat162_rn(f[w].z, f[w].w);
      x[2 * w] = *reinterpret_cast<const uint32_t*>(&q0);
      x[2 * w + 1] = *reinterpret_cast<const uint32_t*>(&q1);
    }
  }
  uint4* dst = X + r * ldx_v + blk * 2;
  dst[0] = make_uint4(x[0], x[1], x[2], x[3]);
  dst[1] = make_uint4(x[4], x[5], x[6], x[7]);
}

__global__ void check_ids_kernel(const int32_t* __restrict__ tok, int64_t n, int64_t n_rows, int32_t* __restrict__ n_bad) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool bad = i < n && (tok[i] < 0 || tok[i] >= n_rows);
  const unsigned m = __ballot_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && m != 0) atomicAdd(n_bad, __popc(m));
}

__global__ void dropout_mask_kernel(int64_t n_tok, int D, uint32_t thr, uint64_t seed, uint8_t* mask) {
  const int nblk = (D + 15) / 16;
  const int64_t item = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (item >= n_tok * nblk) return;
  const int64_t r = item / nblk;
  const int blk = static_cast<int>(item - r * nblk);
  uint32_t bytes[4];
  dropout_bytes16(seed, static_cast<uint64_t>(r), static_cast<uint32_t>(blk), bytes);
  for (int c = 0; c < 16; ++c) {
    const int col = blk * 16 + c;
    if (col < D) mask[r * D + col] = dropout_keep(bytes, c, thr) ? 1 : 0;
  }
}

__global__ void table_to_bf16_kernel(const float* __restrict__ p, int64_t n_rows, int D,
                                     __nv_bfloat16* __restrict__ out, int ld) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_rows * ld) return;
  const int64_t r = i / ld;
  const int c = static_cast<int>(i - r * ld);
  out[i] = __float2bfloat16_rn(c < D ? p[r * D + c] : 0.0f);
}

// Wt[f, j*Dpad + c] = W_ref[j, c, f]
template <typename T>
__global__ void pack_conv_w_kernel(const float* __restrict__ W, int k, int D, int F, int Dpad, int Fpad,
                                   T* __restrict__ Wt) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t K = static_cast<int64_t>(k) * Dpad;
  if (i >= K * Fpad) return;
  const int f = static_cast<int>(i / K);
  const int kk = static_cast<int>(i - f * K);
  const int j = kk / Dpad, c = kk - j * Dpad;
  float v = (f < F && c < D) ? W[(static_cast<int64_t>(j) * D + c) * F + f] : 0.0f;
  if constexpr (sizeof(T) == 2) Wt[i] = __float2bfloat16_rn(v); else Wt[i] = v;
}

}  // namespace

extern "C" {

static int emb_gather_impl(const int32_t* tok, int64_t n_tok, const void* table, int64_t n_rows,
                           int D, int ld_table, int dtype, void* X, int ldx,
                           float keep, uint64_t seed, uint16_t* mask_bits, void* stream) {
  CLSM_REQUIRE(dtype == CLSM_BF16 || dtype == CLSM_F32, "emb_gather: bad dtype %d", dtype);
  CLSM_REQUIRE(n_tok >= 0 && n_rows > 0 && D > 0 && ld_table >= D && ldx >= D, "emb_gather: bad sizes");
  const int es = dtype == CLSM_BF16 ? 2 : 4;
  CLSM_REQUIRE((ld_table * es) % 16 == 0 && ldx % 16 == 0,
               "emb_gather: rows must be 16-byte aligned and ldx a multiple of 16 (ld_table=%d ldx=%d)", ld_table, ldx);
  CLSM_REQUIRE(keep > 0.0f, "emb_gather: keep must be > 0");
  if (n_tok == 0) return CLSM_OK;
  const int vec_per_row_in = ld_table * es / 16;
  const int in_blocks = (ld_table + 15) / 16;
  const int out_blocks = ldx / 16;
  const int64_t total = n_tok * out_blocks;
  const unsigned grid = static_cast<unsigned>(ceil_div64(total, 256));
  const uint32_t thr = keep_threshold(keep);
  const bool drop = keep < 1.0f && thr < 256u;            // a keep that rounds to 256/256 drops nothing
  CLSM_REQUIRE(!drop || thr > 0, "emb_gather: keep %g rounds to a keep probability of 0/256", keep);
  const float inv_keep = drop ? dropout_scale(keep) : 1.0f;          // 1 / the REALISED keep probability thr/256
  const uint4* t = static_cast<const uint4*>(table);
  uint4* x = static_cast<uint4*>(X);
  cudaStream_t st = as_stream(stream);
  if (dtype == CLSM_BF16) {
    if (drop) emb_gather_bf16_drop_kernel<<<grid, 256, 0, st>>>(tok, n_tok, t, in_blocks, vec_per_row_in, vec_per_row_in, x, out_blocks, ldx * es / 16, inv_keep, thr, seed, mask_bits);
    else      emb_gather_kernel<true, false><<<grid, 256, 0, st>>>(tok, n_tok, t, in_blocks, vec_per_row_in, vec_per_row_in, x, out_blocks, ldx * es / 16, inv_keep, thr, seed, nullptr);
  } else {
    if (drop) emb_gather_kernel<false, true><<<grid, 256, 0, st>>>(tok, n_tok, t, in_blocks, vec_per_row_in, vec_per_row_in, x, out_blocks, ldx * es / 16, inv_keep, thr, seed, mask_bits);
    else      emb_gather_kernel<false, false><<<grid, 256, 0, st>>>(tok, n_tok, t, in_blocks, vec_per_row_in, vec_per_row_in, x, out_blocks, ldx * es / 16, inv_keep, thr, seed, nullptr);
  }
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_emb_gather(const int32_t* tok, int64_t n_tok, const void* table, int64_t n_rows,
                    int D, int ld_table, int dtype, void* X, int ldx,
                    float keep, uint64_t seed, void* stream) {
  return emb_gather_impl(tok, n_tok, table, n_rows, D, ld_table, dtype, X, ldx, keep, seed, nullptr, stream);
}

int clsm_emb_gather_mask(const int32_t* tok, int64_t n_tok, const void* table, int64_t n_rows,
                         int D, int ld_table, int dtype, void* X, int ldx,
                         float keep, uint64_t seed, uint16_t* mask_bits, void* stream) {
  return emb_gather_impl(tok, n_tok, table, n_rows, D, ld_table, dtype, X, ldx, keep, seed, mask_bits, stream);
}

int clsm_emb_gather_master(const int32_t* tok, int64_t n_tok, const float* table, int64_t n_rows, int D,
                           void* X, int ldx, float keep, uint64_t seed, uint16_t* mask_bits, void* stream) {
  CLSM_REQUIRE(n_tok >= 0 && n_rows > 0 && D > 0 && D % 4 == 0 && ldx >= D && ldx % 16 == 0,
               "emb_gather_master: D must be a multiple of 4 and ldx a multiple of 16 (D=%d ldx=%d)", D, ldx);
  CLSM_REQUIRE(keep > 0.0f, "emb_gather_master: keep must be > 0");
  if (n_tok == 0) return CLSM_OK;
  const int out_blocks = ldx / 16;
  const unsigned grid = static_cast<unsigned>(ceil_div64(n_tok * out_blocks, 256));
  const uint32_t thr = keep_threshold(keep);
  const bool drop = keep < 1.0f && thr < 256u;
  CLSM_REQUIRE(!drop || thr > 0, "emb_gather_master: keep %g rounds to a keep probability of 0/256", keep);
  cudaStream_t st = as_stream(stream);
  if (drop) emb_gather_master_kernel<true><<<grid, 256, 0, st>>>(tok, n_tok, reinterpret_cast<const float4*>(table), D / 4,
      static_cast<uint4*>(X), out_blocks, ldx * 2 / 16, dropout_scale(keep), thr, seed, mask_bits);
  else emb_gather_master_kernel<false><<<grid, 256, 0, st>>>(tok, n_tok, reinterpret_cast<const float4*>(table), D / 4,
      static_cast<uint4*>(X), out_blocks, ldx * 2 / 16, 1.0f, 256u, seed, nullptr);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_check_token_ids(const int32_t* tok, int64_t n_tok, int64_t n_rows, int32_t* n_bad, void* stream) {
  CLSM_REQUIRE(n_tok >= 0 && n_rows > 0 && n_bad != nullptr, "check_token_ids: bad arguments");
  cudaStream_t st = as_stream(stream);
  CLSM_CUDA(cudaMemsetAsync(n_bad, 0, sizeof(int32_t), st));
  if (n_tok == 0) return CLSM_OK;
  check_ids_kernel<<<static_cast<unsigned>(ceil_div64(n_tok, 256)), 256, 0, st>>>(tok, n_tok, n_rows, n_bad);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_dropout_rows(const float* x, int64_t n, int H, float keep, uint64_t seed, float* out, void* stream) {
  CLSM_REQUIRE(n >= 0 && H > 0 && keep > 0.0f, "dropout_rows: bad arguments");
  if (n == 0) return CLSM_OK;
  const uint32_t thr = keep_threshold(keep);
  const bool drop = keep < 1.0f && thr < 256u;
  CLSM_REQUIRE(!drop || thr > 0, "dropout_rows: keep %g rounds to a keep probability of 0/256", keep);
  const int64_t total = n * ((H + 15) / 16);
  dropout_rows_kernel<<<static_cast<unsigned>(ceil_div64(total, 256)), 256, 0, as_stream(stream)>>>(
      x, n, H, drop ? dropout_scale(keep) : 1.0f, drop ? thr : 256u, seed, out);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_copy_rows_strided(const void* src, size_t src_pitch, void* dst, size_t dst_pitch, size_t width_bytes,
                           int64_t rows, void* stream) {
  CLSM_REQUIRE(src != nullptr && dst != nullptr && rows >= 0 && src_pitch >= width_bytes && dst_pitch >= width_bytes,
               "copy_rows_strided: bad arguments");
  if (rows == 0 || width_bytes == 0) return CLSM_OK;
  CLSM_CUDA(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width_bytes, static_cast<size_t>(rows), cudaMemcpyDefault,
                              as_stream(stream)));
  return CLSM_OK;
}

float clsm_dropout_scale(float keep) { return keep < 1.0f ? dropout_scale(keep) : 1.0f; }

int clsm_emb_gather_h(const int32_t* tok_h, int32_t* tok_d, int64_t n_tok, const void* table,
                      int64_t n_rows, int D, int ld_table, int dtype, void* X, int ldx,
                      float keep, uint64_t seed, void* stream) {
  CLSM_REQUIRE(tok_h != nullptr && tok_d != nullptr, "emb_gather_h: null token buffer");
  CLSM_CUDA(cudaMemcpyAsync(tok_d, tok_h, static_cast<size_t>(n_tok) * sizeof(int32_t),
                            cudaMemcpyHostToDevice, as_stream(stream)));
  return clsm_emb_gather(tok_d, n_tok, table, n_rows, D, ld_table, dtype, X, ldx, keep, seed, stream);
}

int clsm_dropout_mask(int64_t n_tok, int D, float keep, uint64_t seed, uint8_t* mask, void* stream) {
  CLSM_REQUIRE(n_tok >= 0 && D > 0 && keep > 0.0f, "dropout_mask: bad arguments");
  if (n_tok == 0) return CLSM_OK;
  const int64_t total = n_tok * ((D + 15) / 16);
  dropout_mask_kernel<<<static_cast<unsigned>(ceil_div64(total, 256)), 256, 0, as_stream(stream)>>>(
      n_tok, D, keep_threshold(keep), seed, mask);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_table_to_bf16(const float* p, int64_t n_rows, int D, void* bf16_copy, int ld_copy, void* stream) {
  CLSM_REQUIRE(n_rows > 0 && D > 0 && ld_copy >= D, "table_to_bf16: bad sizes");
  const int64_t total = n_rows * ld_copy;
  table_to_bf16_kernel<<<static_cast<unsigned>(ceil_div64(total, 256)), 256, 0, as_stream(stream)>>>(
      p, n_rows, D, static_cast<__nv_bfloat16*>(bf16_copy), ld_copy);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_pack_conv_w(const float* W_ref, int k, int D, int F, int Dpad, int Fpad,
                     int dtype, void* Wt, void* stream) {
  CLSM_REQUIRE(k > 0 && D > 0 && F > 0 && Dpad >= D && Fpad >= F, "pack_conv_w: bad sizes");
  const int64_t total = static_cast<int64_t>(k) * Dpad * Fpad;
  const unsigned grid = static_cast<unsigned>(ceil_div64(total, 256));
  if (dtype == CLSM_BF16)
    pack_conv_w_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>(W_ref, k, D, F, Dpad, Fpad, static_cast<__nv_bfloat16*>(Wt));
  else if (dtype == CLSM_F32)
    pack_conv_w_kernel<float><<<grid, 256, 0, as_stream(stream)>>>(W_ref, k, D, F, Dpad, Fpad, static_cast<float*>(Wt));
  else CLSM_REQUIRE(false, "pack_conv_w: bad dtype %d", dtype);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

}  // extern "C"
