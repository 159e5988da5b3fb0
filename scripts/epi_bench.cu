// micro-benchmark: per-warp column max + first-arg-max over 32 lanes for 32 register columns
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ float credux_max(float v) { float m; asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(m) : "f"(v)); return m; }
__device__ __forceinline__ uint32_t ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }

__global__ void k_credux(const float* in, float* out, unsigned* outb, long long* cyc, int iters) {
  const int lane = threadIdx.x % 32;
  float v[32];
  for (int i = 0; i < 32; ++i) v[i] = in[(threadIdx.x * 32 + i) % 4096];
  float accm = 0; unsigned accb = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    float mym = 0; unsigned myb = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float x = v[i] + (float)it;
      const float m = credux_max(x);
      const unsigned b = __ballot_sync(0xffffffffu, x == m);
      if (lane == i) { mym = m; myb = b; }
    }
    accm += mym; accb ^= myb;
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = accm; outb[blockIdx.x * blockDim.x + threadIdx.x] = accb;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

// butterfly transposed reduction on 64-bit keys (ordered value << 32 | ~lane)
__global__ void k_bfly(const float* in, float* out, unsigned* outb, long long* cyc, int iters) {
  const int lane = threadIdx.x % 32;
  float v[32];
  for (int i = 0; i < 32; ++i) v[i] = in[(threadIdx.x * 32 + i) % 4096];
  float accm = 0; unsigned accb = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    unsigned long long key[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) key[i] = ((unsigned long long)ord(v[i] + (float)it) << 32) | (unsigned)(31 - lane);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      const bool up = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < off; ++i) {
        const unsigned long long mine = up ? key[i + off] : key[i];
        const unsigned long long send = up ? key[i] : key[i + off];
        const unsigned long long recv = __shfl_xor_sync(0xffffffffu, send, off);
        key[i] = mine > recv ? mine : recv;
      }
    }
    accm += __uint_as_float((unsigned)(key[0] >> 32)); accb ^= (unsigned)key[0];
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = accm; outb[blockIdx.x * blockDim.x + threadIdx.x] = accb;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

// value-only butterfly, then arg-max by equality ballot against the broadcast column max
__global__ void k_bfly_val(const float* in, float* out, unsigned* outb, long long* cyc, int iters) {
  const int lane = threadIdx.x % 32;
  float v[32];
  for (int i = 0; i < 32; ++i) v[i] = in[(threadIdx.x * 32 + i) % 4096];
  float accm = 0; unsigned accb = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    float x[32], r[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) { x[i] = v[i] + (float)it; r[i] = x[i]; }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      const bool up = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < off; ++i) {
        const float mine = up ? r[i + off] : r[i];
        const float send = up ? r[i] : r[i + off];
        r[i] = fmaxf(mine, __shfl_xor_sync(0xffffffffu, send, off));
      }
    }
    // r[0] = max of column `lane`.  first arg-max lane: for each column broadcast its max, ballot equality
    unsigned myb = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float m = __shfl_sync(0xffffffffu, r[0], i);
      const unsigned b = __ballot_sync(0xffffffffu, x[i] == m);
      if (lane == i) myb = b;
    }
    accm += r[0]; accb ^= myb;
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = accm; outb[blockIdx.x * blockDim.x + threadIdx.x] = accb;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

int main() {
  float *in, *out; unsigned* outb; long long* cyc;
  cudaMalloc(&in, 4096 * 4); cudaMalloc(&out, 148 * 128 * 4); cudaMalloc(&outb, 148 * 128 * 4); cudaMalloc(&cyc, 8);
  float h[4096]; for (int i = 0; i < 4096; ++i) h[i] = (float)((i * 2654435761u) % 1000) / 7.0f;
  cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
  const int iters = 200;
  for (int warps = 1; warps <= 4; warps *= 4) {
    long long c;
    k_credux<<<148, 32 * warps>>>(in, out, outb, cyc, iters); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("credux+ballot   warps/SM=%d : %.1f cycles per 32x32 chunk\n", warps, (double)c / iters);
    k_bfly<<<148, 32 * warps>>>(in, out, outb, cyc, iters); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("butterfly u64   warps/SM=%d : %.1f cycles per 32x32 chunk\n", warps, (double)c / iters);
    k_bfly_val<<<148, 32 * warps>>>(in, out, outb, cyc, iters); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("butterfly+ballot warps/SM=%d : %.1f cycles per 32x32 chunk\n", warps, (double)c / iters);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
