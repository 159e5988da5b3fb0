// Clipped cosine, gamma-softmax, loss and its gradient in one warp-per-query kernel.
//   replaces RTF (dssm_cnn_v2/cnn_dssm_tf.py:63-77), x*GAMMA / exp / p0 (:164-176) and
//   binary_crossentropy + accuracy (:182), forward and backward.
// All arithmetic fp32 with IEEE sqrt/div (no approximations: a zero semantic vector must give
// cosine exactly 0, SURVEY hard part 7).
#include "common.cuh"

namespace {

constexpr int MAXJ1 = 128;   // 1 + J
constexpr float BCE_EPS = 1e-7f;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(128)
sim_loss_kernel(const float* __restrict__ semQ, const float* __restrict__ semD, int64_t B, int J1, int H,
                float gamma, float inv_bt, float* __restrict__ cosv, float* __restrict__ prob,
                float* __restrict__ dQ, float* __restrict__ dD) {
  __shared__ float s_raw[4][MAXJ1];
  __shared__ float s_nd[4][MAXJ1];
  __shared__ float s_sd[4][MAXJ1];
  __shared__ float s_k1[4][MAXJ1];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t b = static_cast<int64_t>(blockIdx.x) * 4 + warp;
  if (b >= B) return;
  const float tiny = FLT_MIN;
  const float* q = semQ + b * H;
  float sq = 0.0f;
  for (int h = lane; h < H; h += 32) { const float x = q[h]; sq = fmaf(x, x, sq); }
  sq = warp_sum(sq);
  const float nq = sqrtf(fmaxf(sq, tiny));
  float esum = 0.0f;
  for (int j = 0; j < J1; ++j) {
    const float* d = semD + (b * J1 + j) * H;
    float dot = 0.0f, sd = 0.0f;
    for (int h = lane; h < H; h += 32) { const float y = d[h]; dot = fmaf(q[h], y, dot); sd = fmaf(y, y, sd); }
    dot = warp_sum(dot); sd = warp_sum(sd);
    const float nd = sqrtf(fmaxf(sd, tiny));
    const float raw = dot / (nq * nd);
    const float c = fminf(fmaxf(raw, 0.0f), 1.0f);
    esum += expf(gamma * c);
    if (lane == 0) { s_raw[warp][j] = raw; s_nd[warp][j] = nd; s_sd[warp][j] = sd; cosv[b * J1 + j] = c; }
  }
  __syncwarp();
  float p0 = 0.0f;
  for (int j = lane; j < J1; j += 32) {
    const float raw = s_raw[warp][j];
    const float c = fminf(fmaxf(raw, 0.0f), 1.0f);
    const float pj = expf(gamma * c) / esum;
    prob[b * J1 + j] = pj;
    if (j == 0) p0 = pj;
  }
  p0 = __shfl_sync(0xffffffffu, p0, 0);
  if (dQ == nullptr) return;
  const float live = (p0 >= BCE_EPS && p0 <= 1.0f - BCE_EPS) ? 1.0f : 0.0f;
  const float fq = sq > tiny ? 1.0f : 0.0f;
  // dq accumulators for this lane's h values are rebuilt per h-chunk to keep registers bounded
  float coef_q = 0.0f;   // sum_j dc_j * raw_j
  for (int j = 0; j < J1; ++j) {
    const float raw = s_raw[warp][j];
    const float c = fminf(fmaxf(raw, 0.0f), 1.0f);
    const float pj = expf(gamma * c) / esum;
    float dc = gamma * (pj - (j == 0 ? 1.0f : 0.0f)) * live * inv_bt;
    if (!(raw >= 0.0f && raw <= 1.0f)) dc = 0.0f;
    coef_q = fmaf(dc, raw, coef_q);
    const float nd = s_nd[warp][j];
    const float inv = 1.0f / (nq * nd);
    const float fd = s_sd[warp][j] > tiny ? 1.0f : 0.0f;
    const float k1 = dc * inv;
    const float k2 = dc * raw * fd / (nd * nd);
    if (lane == 0) s_k1[warp][j] = k1;
    const float* d = semD + (b * J1 + j) * H;
    float* gd = dD + (b * J1 + j) * H;
    for (int h = lane; h < H; h += 32) gd[h] = k1 * q[h] - k2 * d[h];
  }
  const float kq = coef_q * fq / (nq * nq);
  __syncwarp();
  for (int h = lane; h < H; h += 32) {
    float g = 0.0f;
    for (int j = 0; j < J1; ++j) g = fmaf(s_k1[warp][j], semD[(b * J1 + j) * H + h], g);
    dQ[b * H + h] = g - kq * q[h];
  }
}

// mean loss / accuracy over B rows in a fixed order (single block, fixed tree): deterministic.
__global__ void __launch_bounds__(256)
loss_reduce_kernel(const float* __restrict__ prob, int64_t B, int J1, float* __restrict__ loss_acc) {
  __shared__ float s_l[256], s_a[256];
  float l = 0.0f, a = 0.0f;
  for (int64_t b = threadIdx.x; b < B; b += 256) {
    const float p0 = prob[b * J1];
    const float ph = fminf(fmaxf(p0, BCE_EPS), 1.0f - BCE_EPS);
    l += -logf(ph);
    a += p0 > 0.5f ? 1.0f : 0.0f;
  }
  s_l[threadIdx.x] = l; s_a[threadIdx.x] = a;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { s_l[threadIdx.x] += s_l[threadIdx.x + o]; s_a[threadIdx.x] += s_a[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { loss_acc[0] = s_l[0] / static_cast<float>(B); loss_acc[1] = s_a[0] / static_cast<float>(B); }
}

__global__ void normalize_rows_kernel(const float* __restrict__ sem, int64_t N, int H, float* __restrict__ vec) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * (blockDim.x / 32) + warp;
  if (r >= N) return;
  float s = 0.0f;
  for (int h = lane; h < H; h += 32) { const float x = sem[r * H + h]; s = fmaf(x, x, s); }
  s = warp_sum(s);
  const float n = sqrtf(fmaxf(s, FLT_MIN));
  for (int h = lane; h < H; h += 32) vec[r * H + h] = sem[r * H + h] / n;
}

}  // namespace

extern "C" {

int clsm_sim_loss(const float* semQ, const float* semD, int64_t B, int J, int H, float gamma,
                  int64_t batch_total, float* cosv, float* prob, float* loss_acc,
                  float* dSemQ, float* dSemD, void* stream) {
  CLSM_REQUIRE(B > 0 && J >= 0 && H > 0, "sim_loss: bad sizes");
  CLSM_REQUIRE(J + 1 <= MAXJ1, "sim_loss: at most %d negatives", MAXJ1 - 1);
  CLSM_REQUIRE((dSemQ == nullptr) == (dSemD == nullptr), "sim_loss: pass both gradient buffers or neither");
  if (batch_total <= 0) batch_total = B;
  cudaStream_t st = as_stream(stream);
  sim_loss_kernel<<<static_cast<unsigned>(ceil_div64(B, 4)), 128, 0, st>>>(
      semQ, semD, B, J + 1, H, gamma, 1.0f / static_cast<float>(batch_total), cosv, prob, dSemQ, dSemD);
  CLSM_LAUNCH_CHECK();
  loss_reduce_kernel<<<1, 256, 0, st>>>(prob, B, J + 1, loss_acc);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

int clsm_normalize_rows(const float* sem, int64_t N, int H, float* vec, void* stream) {
  CLSM_REQUIRE(N >= 0 && H > 0, "normalize_rows: bad sizes");
  if (N == 0) return CLSM_OK;
  normalize_rows_kernel<<<static_cast<unsigned>(ceil_div64(N, 8)), 256, 0, as_stream(stream)>>>(sem, N, H, vec);
  CLSM_LAUNCH_CHECK();
  return CLSM_OK;
}

}  // extern "C"
