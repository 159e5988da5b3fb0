// Dense semantic projection, forward and backward (CUDA-core fp32 GEMM, 64x64x16 tiles).
//   replaces Dense(hidden_dims) + Activation('relu'), dssm_cnn_v2/cnn_dssm_tf.py:124-126,
//   and its gradient.  These GEMMs are <0.2 % of the step's FLOPs.
#include "common.cuh"
#include <stdlib.h>

namespace {

constexpr int TM = 64, TN = 64, TK = 16;
enum { EPI_NONE = 0, EPI_BIAS_ACT = 1, EPI_MUL_ACTGRAD = 2 };

struct GemmP {
  const float* A; int64_t sam, sak;     // A(m,k) = A[m*sam + k*sak]
  const float* B; int64_t sbk, sbn;     // B(k,n) = B[k*sbk + n*sbn]
  float* C; int64_t ldc;                // C(m,n) = C[m*ldc + n]  (split z writes at C + z*M*ldc)
  int64_t M; int N; int64_t K;
  int64_t k_per_split;
  int epi; int act;
  const float* bias;                    // EPI_BIAS_ACT
  const float* aux; int64_t ldaux;      // EPI_MUL_ACTGRAD: multiply by act'(aux[m,n])
};

__global__ void __launch_bounds__(256) gemm_f32_kernel(GemmP p) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int tid = threadIdx.x;
  const int64_t m0 = static_cast<int64_t>(blockIdx.x) * TM;
  const int n0 = blockIdx.y * TN;
  const int64_t kbeg = static_cast<int64_t>(blockIdx.z) * p.k_per_split;
  const int64_t kend = (kbeg + p.k_per_split < p.K) ? kbeg + p.k_per_split : p.K;
  const int ty = tid / 16, tx = tid % 16;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  // loader mappings: pick the one whose fastest thread index walks the contiguous stride
  const bool a_kfast = (p.sak == 1);
  const bool b_nfast = (p.sbn == 1);
  for (int64_t kk = kbeg; kk < kend; kk += TK) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int m, q;
      if (a_kfast) { q = tid % 16; m = tid / 16 + e * 16; }
      else         { m = tid % 64; q = tid / 64 + e * 4; }
      const int64_t gm = m0 + m, gk = kk + q;
      As[q][m] = (gm < p.M && gk < kend) ? p.A[gm * p.sam + gk * p.sak] : 0.0f;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int n, q;
      if (b_nfast) { n = tid % 64; q = tid / 64 + e * 4; }
      else         { q = tid % 16; n = tid / 16 + e * 16; }
      const int gn = n0 + n; const int64_t gk = kk + q;
      Bs[q][n] = (gn < p.N && gk < kend) ? p.B[gk * p.sbk + gn * p.sbn] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < TK; ++q) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[q][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[q][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* C = p.C + static_cast<int64_t>(blockIdx.z) * p.M * p.ldc;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= p.N) continue;
      float v = acc[i][j];
      if (p.epi == EPI_BIAS_ACT) v = act_apply(v + p.bias[n], p.act);
      else if (p.epi == EPI_MUL_ACTGRAD) v *= act_grad_out(p.aux[m * p.ldaux + n], p.act);
      C[m * p.ldc + n] = v;
    }
  }
}

// ---- the same GEMM on the tensor cores with TF32 operands (act | CLSM_MATH_TF32; the bf16-precision mode of the
// host model passes it).  128 x 64 x 32 tiles, 8 warps of 32 x 32, mma.sync m16n8k8, fp32 accumulation.  Operands
// are rounded to TF32 (10-bit mantissa, cvt.rna) when they are staged; both tiles sit in shared memory as
// [row][k] with a row pitch of 36 floats, which makes every fragment load conflict-free.  This projection is
// 0.2 % of the step's flops: the legacy warp-level MMA is ample, no TMEM pipeline needed.
constexpr int GM = 128, GN = 64, GK = 32, GP = GK + 4;

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

__global__ void __launch_bounds__(256) gemm_tf32_kernel(GemmP p) {
  __shared__ uint32_t As[GM][GP];
  __shared__ uint32_t Bs[GN][GP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;
  const int g = lane >> 2, t = lane & 3;
  const int64_t m0 = static_cast<int64_t>(blockIdx.x) * GM;
  const int n0 = blockIdx.y * GN;
  const int64_t kbeg = static_cast<int64_t>(blockIdx.z) * p.k_per_split;
  const int64_t kend = (kbeg + p.k_per_split < p.K) ? kbeg + p.k_per_split : p.K;
  float acc[2][4][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[i][j][c] = 0.0f;
  const bool a_kfast = (p.sak == 1);
  const bool b_kfast = (p.sbk == 1);
  float ra[16], rb[8];
  auto load_tiles = [&](int64_t kk) {
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      int m, q;
      if (a_kfast) { q = tid & 31; m = (tid >> 5) + e * 8; }
      else         { m = tid & 127; q = (tid >> 7) + e * 2; }
      const int64_t gm = m0 + m, gk = kk + q;
      ra[e] = (gm < p.M && gk < kend) ? __ldg(p.A + gm * p.sam + gk * p.sak) : 0.0f;
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      int n, q;
      if (b_kfast) { q = tid & 31; n = (tid >> 5) + e * 8; }
      else         { n = tid & 63; q = (tid >> 6) + e * 4; }
      const int gn = n0 + n; const int64_t gk = kk + q;
      rb[e] = (gn < p.N && gk < kend) ? __ldg(p.B + gk * p.sbk + gn * p.sbn) : 0.0f;
    }
  };
  auto store_tiles = [&]() {
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      int m, q;
      if (a_kfast) { q = tid & 31; m = (tid >> 5) + e * 8; }
      else         { m = tid & 127; q = (tid >> 7) + e * 2; }
      As[m][q] = to_tf32(ra[e]);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      int n, q;
      if (b_kfast) { q = tid & 31; n = (tid >> 5) + e * 8; }
      else         { n = tid & 63; q = (tid >> 6) + e * 4; }
      Bs[n][q] = to_tf32(rb[e]);
    }
  };
  if (kbeg < kend) load_tiles(kbeg);
  for (int64_t kk = kbeg; kk < kend; kk += GK) {
    store_tiles();
    __syncthreads();
    if (kk + GK < kend) load_tiles(kk + GK);          // next tile's global loads fly during the MMAs
#pragma unroll
    for (int ks = 0; ks < GK; ks += 8) {
      uint32_t a[2][4], b[4][2];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        a[i][0] = As[wm + i * 16 + g][ks + t];
        a[i][1] = As[wm + i * 16 + g + 8][ks + t];
        a[i][2] = As[wm + i * 16 + g][ks + t + 4];
        a[i][3] = As[wm + i * 16 + g + 8][ks + t + 4];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        b[j][0] = Bs[wn + j * 8 + g][ks + t];
        b[j][1] = Bs[wn + j * 8 + g][ks + t + 4];
      }
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                       : "+f"(acc[i][j][0]), "+f"(acc[i][j][1]), "+f"(acc[i][j][2]), "+f"(acc[i][j][3])
                       : "r"(a[i][0]), "r"(a[i][1]), "r"(a[i][2]), "r"(a[i][3]), "r"(b[j][0]), "r"(b[j][1]));
    }
    __syncthreads();
  }
  float* C = p.C + static_cast<int64_t>(blockIdx.z) * p.M * p.ldc;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int64_t m = m0 + wm + i * 16 + g + (c >> 1) * 8;
        const int n = n0 + wn + j * 8 + 2 * t + (c & 1);
        if (m >= p.M || n >= p.N) continue;
        float v = acc[i][j][c];
        if (p.epi == EPI_BIAS_ACT) v = act_apply(v + p.bias[n], p.act);
        else if (p.epi == EPI_MUL_ACTGRAD) v *= act_grad_out(p.aux[m * p.ldaux + n], p.act);
        C[m * p.ldc + n] = v;
      }
}

// ---- TF32 GEMM, cp.async-pipelined (the default of the bf16 mode whenever every row is 16-byte aligned) ----------
// gemm_tf32_kernel above stages its tiles through registers with scalar loads, one tile ahead: the document tower's
// projection (24 576 x 300 x 128) took 54 us against a 7 us memory floor.  Here 64 x 128 x 32 tiles move with 16-byte
// cp.async into a 3-stage ring (nothing held in registers, two tiles in flight behind the one being multiplied), 8
// warps of 32 x 32 run mma.sync m16n8k8, operands are rounded to TF32 (cvt.rna) as fragments are read -- the same
// arithmetic as above.  Layouts: A is (m, k) with k contiguous, or (k, m) with m contiguous (AM: pooled^T for dWd);
// B is (k, n) with n contiguous, or (n, k) with k contiguous (BK: Wd^T for the pooled gradient).  Every tile row is
// padded so that all fragment loads are conflict-free.
constexpr int QM = 64, QN = 128, QK = 32, QST = 3;
template <bool AM, bool BK>
struct AsyncTiles {
  static constexpr int A_ROWS = AM ? QK : QM, A_PITCH = AM ? QM + 8 : QK + 4;     // [k][m] or [m][k]
  static constexpr int B_ROWS = BK ? QN : QK, B_PITCH = BK ? QK + 4 : QN + 8;     // [n][k] or [k][n]
  static constexpr int A_FLOATS = A_ROWS * A_PITCH, B_FLOATS = B_ROWS * B_PITCH;
  static constexpr int STAGE_FLOATS = A_FLOATS + B_FLOATS;
};

__device__ __forceinline__ void cp16(uint32_t dst, const void* src, int src_bytes) {      // src_bytes < 16: the rest is zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}

template <bool AM, bool BK>
__global__ void __launch_bounds__(256) gemm_tf32_async_kernel(GemmP p) {
  using T = AsyncTiles<AM, BK>;
  extern __shared__ __align__(16) float smem_dyn[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = (warp >> 2) * 32, wn = (warp & 3) * 32;
  const int g = lane >> 2, t = lane & 3;
  const int64_t m0 = static_cast<int64_t>(blockIdx.x) * QM;
  const int n0 = blockIdx.y * QN;
  const int64_t kbeg = static_cast<int64_t>(blockIdx.z) * p.k_per_split;
  const int64_t kend = (kbeg + p.k_per_split < p.K) ? kbeg + p.k_per_split : p.K;
  const int n_kt = static_cast<int>((kend - kbeg + QK - 1) / QK);
  const uint32_t sbase = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dyn));
  float acc[2][4][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[i][j][c] = 0.0f;

  auto load_stage = [&](int kt, int stage) {
    const int64_t kk = kbeg + static_cast<int64_t>(kt) * QK;
    const uint32_t a_s = sbase + static_cast<uint32_t>(stage * T::STAGE_FLOATS) * 4u;
    const uint32_t b_s = a_s + T::A_FLOATS * 4u;
    // A tile: 16-byte pieces along the contiguous dimension
    if (AM) {                                      // rows = k (32), 16 pieces of 4 m per row
      for (int i = tid; i < QK * (QM / 4); i += 256) {
        const int r = i / (QM / 4), c4 = (i - r * (QM / 4)) * 4;
        const int64_t gk = kk + r, gm = m0 + c4;
        int bytes = (gk < kend && gm < p.M) ? static_cast<int>((p.M - gm < 4 ? p.M - gm : 4) * 4) : 0;
        cp16(a_s + static_cast<uint32_t>(r * T::A_PITCH + c4) * 4u, p.A + (bytes ? gk * p.sak + gm : 0), bytes);
      }
    } else {                                       // rows = m (64), 8 pieces of 4 k per row
      for (int i = tid; i < QM * (QK / 4); i += 256) {
        const int r = i / (QK / 4), c4 = (i - r * (QK / 4)) * 4;
        const int64_t gm = m0 + r, gk = kk + c4;
        int bytes = (gm < p.M && gk < kend) ? static_cast<int>((kend - gk < 4 ? kend - gk : 4) * 4) : 0;
        cp16(a_s + static_cast<uint32_t>(r * T::A_PITCH + c4) * 4u, p.A + (bytes ? gm * p.sam + gk : 0), bytes);
      }
    }
    if (BK) {                                      // rows = n (128), 8 pieces of 4 k per row
      for (int i = tid; i < QN * (QK / 4); i += 256) {
        const int r = i / (QK / 4), c4 = (i - r * (QK / 4)) * 4;
        const int gn = n0 + r; const int64_t gk = kk + c4;
        int bytes = (gn < p.N && gk < kend) ? static_cast<int>((kend - gk < 4 ? kend - gk : 4) * 4) : 0;
        cp16(b_s + static_cast<uint32_t>(r * T::B_PITCH + c4) * 4u, p.B + (bytes ? static_cast<int64_t>(gn) * p.sbn + gk : 0), bytes);
      }
    } else {                                       // rows = k (32), 32 pieces of 4 n per row
      for (int i = tid; i < QK * (QN / 4); i += 256) {
        const int r = i / (QN / 4), c4 = (i - r * (QN / 4)) * 4;
        const int64_t gk = kk + r; const int gn = n0 + c4;
        int bytes = (gk < kend && gn < p.N) ? ((p.N - gn < 4 ? p.N - gn : 4) * 4) : 0;
        cp16(b_s + static_cast<uint32_t>(r * T::B_PITCH + c4) * 4u, p.B + (bytes ? gk * p.sbk + gn : 0), bytes);
      }
    }
  };

#pragma unroll
  for (int s = 0; s < QST - 1; ++s) {
    if (s < n_kt) load_stage(s, s);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (int kt = 0; kt < n_kt; ++kt) {
    asm volatile("cp.async.wait_group %0;" ::"n"(QST - 2) : "memory");
    __syncthreads();                               // tile kt has landed for everyone; the slot of tile kt-1 is free
    if (kt + QST - 1 < n_kt) load_stage(kt + QST - 1, (kt + QST - 1) % QST);
    asm volatile("cp.async.commit_group;" ::: "memory");
    const float* As = smem_dyn + (kt % QST) * T::STAGE_FLOATS;
    const float* Bs = As + T::A_FLOATS;
#pragma unroll
    for (int ks = 0; ks < QK; ks += 8) {
      uint32_t a[2][4], b[4][2];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int r = wm + i * 16 + g;
        if (AM) {
          a[i][0] = to_tf32(As[(ks + t) * T::A_PITCH + r]);     a[i][1] = to_tf32(As[(ks + t) * T::A_PITCH + r + 8]);
          a[i][2] = to_tf32(As[(ks + t + 4) * T::A_PITCH + r]); a[i][3] = to_tf32(As[(ks + t + 4) * T::A_PITCH + r + 8]);
        } else {
          a[i][0] = to_tf32(As[r * T::A_PITCH + ks + t]);       a[i][1] = to_tf32(As[(r + 8) * T::A_PITCH + ks + t]);
          a[i][2] = to_tf32(As[r * T::A_PITCH + ks + t + 4]);   a[i][3] = to_tf32(As[(r + 8) * T::A_PITCH + ks + t + 4]);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = wn + j * 8 + g;
        if (BK) { b[j][0] = to_tf32(Bs[c * T::B_PITCH + ks + t]); b[j][1] = to_tf32(Bs[c * T::B_PITCH + ks + t + 4]); }
        else    { b[j][0] = to_tf32(Bs[(ks + t) * T::B_PITCH + c]); b[j][1] = to_tf32(Bs[(ks + t + 4) * T::B_PITCH + c]); }
      }
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                       : "+f"(acc[i][j][0]), "+f"(acc[i][j][1]), "+f"(acc[i][j][2]), "+f"(acc[i][j][3])
                       : "r"(a[i][0]), "r"(a[i][1]), "r"(a[i][2]), "r"(a[i][3]), "r"(b[j][0]), "r"(b[j][1]));
    }
  }
  float* C = p.C + static_cast<int64_t>(blockIdx.z) * p.M * p.ldc;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int64_t m = m0 + wm + i * 16 + g + (c >> 1) * 8;
        const int n = n0 + wn + j * 8 + 2 * t + (c & 1);
        if (m >= p.M || n >= p.N) continue;
        float v = acc[i][j][c];
        if (p.epi == EPI_BIAS_ACT) v = act_apply(v + p.bias[n], p.act);
        else if (p.epi == EPI_MUL_ACTGRAD) v *= act_grad_out(p.aux[m * p.ldaux + n], p.act);
        C[m * p.ldc + n] = v;
      }
}

// the async kernel takes a GEMM whose contiguous operand dimensions are unit-stride and whose rows are 16-byte aligned
bool async_ok(const GemmP& p, bool* am, bool* bk) {
  const bool a_k = p.sak == 1, a_m = p.sam == 1 && !a_k;
  const bool b_n = p.sbn == 1, b_k = p.sbk == 1 && !b_n;
  if (!(a_k || a_m) || !(b_n || b_k)) return false;
  const int64_t a_pitch = a_k ? p.sam : p.sak, b_pitch = b_n ? p.sbk : p.sbn;
  if (a_pitch % 4 != 0 || b_pitch % 4 != 0) return false;
  if ((reinterpret_cast<uintptr_t>(p.A) | reinterpret_cast<uintptr_t>(p.B)) & 15) return false;
  if (p.k_per_split % 4 != 0) return false;       // split starts stay 16-byte aligned along k
  *am = a_m; *bk = b_k;
  return true;
}

template <bool AM, bool BK>
void launch_async(const GemmP& p, int splits, cudaStream_t st) {
  using T = AsyncTiles<AM, BK>;
  const int smem = QST * T::STAGE_FLOATS * 4;
  static bool configured = false;
  if (!configured) { cudaFuncSetAttribute(gemm_tf32_async_kernel<AM, BK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); configured = true; }
  dim3 grid(static_cast<unsigned>(ceil_div64(p.M, QM)), (p.N + QN - 1) / QN, splits);
  gemm_tf32_async_kernel<AM, BK><<<grid, 256, smem, st>>>(p);
}

void launch_gemm(const GemmP& p, int splits, bool tf32, cudaStream_t st) {
  bool am = false, bk = false;
  static int legacy = -1;                          // CLSM_DENSE_LEGACY=1: the register-staged kernel (A/B measurements)
  if (legacy < 0) { const char* e = getenv("CLSM_DENSE_LEGACY"); legacy = (e != nullptr && e[0] == '1') ? 1 : 0; }
  if (tf32 && !legacy && async_ok(p, &am, &bk)) {
    if (am && bk) launch_async<true, true>(p, splits, st);
    else if (am) launch_async<true, false>(p, splits, st);
    else if (bk) launch_async<false, true>(p, splits, st);
    else launch_async<false, false>(p, splits, st);
  } else if (tf32) {
    dim3 grid(static_cast<unsigned>(ceil_div64(p.M, GM)), (p.N + GN - 1) / GN, splits);
    gemm_tf32_kernel<<<grid, 256, 0, st>>>(p);
  } else {
    dim3 grid(static_cast<unsigned>(ceil_div64(p.M, TM)), (p.N + TN - 1) / TN, splits);
    gemm_f32_kernel<<<grid, 256, 0, st>>>(p);
  }
}

// out[i] = sum_z part[z*stride + i], z ascending (deterministic)
__global__ void reduce_splits_kernel(const float* __restrict__ part, int n_splits, int64_t stride,
                                     int64_t n, float* __restrict__ out) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float s = 0.0f;
  for (int z = 0; z < n_splits; ++z) s += part[z * stride + i];
  out[i] = s;
}

// dpre = dsem * act'(sem)
__global__ void dpre_kernel(const float* __restrict__ dsem, const float* __restrict__ sem, int64_t n, int act,
                            float* __restrict__ dpre) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) dpre[i] = dsem[i] * act_grad_out(sem[i], act);
}

}  // namespace

// column sums of a (rows, ld) matrix over columns [col0, col0+n): two fixed-shape stages.
// part must hold n_chunks*n floats.  Shared with conv_bwd.cu.
__global__ void colsum_stage1_kernel(const float* __restrict__ A, int64_t rows, int ld, int col0, int n,
                                     int64_t rows_per_chunk, float* __restrict__ part) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  const int64_t rb = static_cast<int64_t>(blockIdx.y) * rows_per_chunk;
  const int64_t re = (rb + rows_per_chunk < rows) ? rb + rows_per_chunk : rows;
  float s = 0.0f;
  for (int64_t r = rb; r < re; ++r) s += A[r * ld + col0 + c];
  part[static_cast<int64_t>(blockIdx.y) * n + c] = s;
}

int clsm_colsum(const float* A, int64_t rows, int ld, int col0, int n, float* out, float* part,
                int n_chunks, cudaStream_t st) {
  const int64_t rpc = ceil_div64(rows, n_chunks);
  dim3 g1((n + 127) / 128, n_chunks);
  colsum_stage1_kernel<<<g1, 128, 0, st>>>(A, rows, ld, col0, n, rpc, part);
  reduce_splits_kernel<<<(n + 255) / 256, 256, 0, st>>>(part, n_chunks, n, n, out);
  g_clsm_launches += 2;
  return cudaGetLastError() == cudaSuccess ? CLSM_OK : CLSM_ECUDA;
}

static int dense_bwd_splits(int64_t S) {
  int64_t s = S / 512;
  if (s < 1) s = 1;
  if (s > 64) s = 64;
  return static_cast<int>(s);
}
static const int kColsumChunks = 64;

extern "C" {

int clsm_dense_fwd(const float* pooled, int64_t S, int P, int ldp, const float* Wd, const float* bd,
                   int H, int act, float* sem, void* stream) {
  CLSM_REQUIRE(S >= 0 && P > 0 && H > 0 && ldp >= P, "dense_fwd: bad sizes");
  if (S == 0) return CLSM_OK;
  const bool tf32 = (act & CLSM_MATH_TF32) != 0;
  act &= ~CLSM_MATH_TF32;
  GemmP p{};
  p.A = pooled; p.sam = ldp; p.sak = 1;
  p.B = Wd; p.sbk = H; p.sbn = 1;
  p.C = sem; p.ldc = H; p.M = S; p.N = H; p.K = P; p.k_per_split = P;
  p.epi = EPI_BIAS_ACT; p.act = act; p.bias = bd;
  launch_gemm(p, 1, tf32, as_stream(stream));
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

size_t clsm_dense_bwd_workspace_bytes(int64_t S, int P, int H) {
  const size_t dpre = align_up(static_cast<size_t>(S) * H * 4, 256);
  const size_t parts = align_up(static_cast<size_t>(dense_bwd_splits(S)) * P * H * 4, 256);
  const size_t cs = align_up(static_cast<size_t>(kColsumChunks) * H * 4, 256);
  return dpre + parts + cs;
}

int clsm_dense_bwd(const float* pooled, const float* sem, const float* dsem, int64_t S, int P, int ldp,
                   const float* Wd, int H, int act, float* dWd, float* dbd, float* G,
                   void* ws, size_t ws_bytes, void* stream) {
  CLSM_REQUIRE(S > 0 && P > 0 && H > 0 && ldp >= P, "dense_bwd: bad sizes");
  const bool tf32 = (act & CLSM_MATH_TF32) != 0;
  act &= ~CLSM_MATH_TF32;
  if (ws_bytes < clsm_dense_bwd_workspace_bytes(S, P, H)) {
    clsm_set_error("dense_bwd: workspace too small (%zu < %zu)", ws_bytes, clsm_dense_bwd_workspace_bytes(S, P, H));
    return CLSM_EWORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  char* w = static_cast<char*>(ws);
  float* dpre = reinterpret_cast<float*>(w);
  w += align_up(static_cast<size_t>(S) * H * 4, 256);
  float* parts = reinterpret_cast<float*>(w);
  const int splits = dense_bwd_splits(S);
  w += align_up(static_cast<size_t>(splits) * P * H * 4, 256);
  float* cs = reinterpret_cast<float*>(w);

  const int64_t n = S * H;
  dpre_kernel<<<static_cast<unsigned>(ceil_div64(n, 256)), 256, 0, st>>>(dsem, sem, n, act, dpre);
  CLSM_LAUNCH_CHECK();
  // dbd = colsum(dpre)
  if (clsm_colsum(dpre, S, H, 0, H, dbd, cs, kColsumChunks, st) != CLSM_OK) { clsm_set_error("dense_bwd: colsum launch failed"); return CLSM_ECUDA; }
  // dWd (P,H) = pooled^T (P,S) . dpre (S,H), split over S
  {
    GemmP p{};
    p.A = pooled; p.sam = 1; p.sak = ldp;
    p.B = dpre; p.sbk = H; p.sbn = 1;
    p.C = (splits > 1) ? parts : dWd; p.ldc = H; p.M = P; p.N = H; p.K = S;
    p.k_per_split = ceil_div64(ceil_div64(S, splits), GK) * GK;      // a multiple of both kernels' K tiles
    p.epi = EPI_NONE;
    launch_gemm(p, splits, tf32, st);
    CLSM_LAUNCH_CHECK();
    if (splits > 1) {
      const int64_t ne = static_cast<int64_t>(P) * H;
      reduce_splits_kernel<<<static_cast<unsigned>(ceil_div64(ne, 256)), 256, 0, st>>>(parts, splits, ne, ne, dWd);
      CLSM_LAUNCH_CHECK();
    }
  }
  // G (S, ldp) = (dpre (S,H) . Wd^T (H,P)) * act'(pooled)
  {
    GemmP p{};
    p.A = dpre; p.sam = H; p.sak = 1;
    p.B = Wd; p.sbk = 1; p.sbn = H;
    p.C = G; p.ldc = ldp; p.M = S; p.N = P; p.K = H; p.k_per_split = H;
    p.epi = EPI_MUL_ACTGRAD; p.act = act; p.aux = pooled; p.ldaux = ldp;
    launch_gemm(p, 1, tf32, st);
    CLSM_LAUNCH_CHECK();
  }
  return CLSM_OK;
}

}  // extern "C"
