// micro-benchmark: sustained tcgen05.mma issue rate for the operand shapes the conv kernels use.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I dnn_page_vectors_b200/csrc scripts/mma_bench.cu -o /tmp/mma_bench
// Every SM runs one CTA (or one CTA of a pair) that issues ITERS x 4 K-steps x nA MMAs on resident
// shared-memory tiles (no TMA traffic, no epilogue), so the cycles per MMA show what the tensor pipe plus its
// shared-memory operand reads sustain for (cta_group, M, N, number of A tiles sharing one B tile).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "tc_ptx.cuh"
using namespace tcptx;

__device__ __forceinline__ void mma_cg2(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit_cg2(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"(mask) : "memory");
}

template <int CG>
__global__ void __launch_bounds__(128, 1) k_mma(int N, int nA, int iters, long long* cyc, int shift_rows) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* bp = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_bytes = 128 * 128;                      // 128 rows x 64 bf16, SW128
  const uint32_t b_rows = CG == 2 ? N / 2 : N;
  const uint32_t b_off = nA * a_bytes;
  const uint32_t total = b_off + ((b_rows * 128 + 1023) & ~1023u) + 1024;
  const uint32_t bar = base + total;
  const uint32_t slot = bar + 16;
  // pseudo-random small bf16 values (so the datapath toggles like real data)
  for (uint32_t i = threadIdx.x; i < total / 2; i += blockDim.x) {
    uint32_t h = (i + blockIdx.x * 7919u) * 2654435761u;
    float f = static_cast<float>(static_cast<int>(h >> 24) - 128) * (1.0f / 256.0f);
    reinterpret_cast<__nv_bfloat16*>(bp)[i] = __float2bfloat16_rn(f);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (threadIdx.x == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 1) {
    if (CG == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "n"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "n"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(slot));
  const bool leader = CG == 1 || cluster_ctarank() == 0;
  long long t0 = 0, t1 = 0;
  if (warp == 1 && lane == 0 && leader) {
    uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) |
                     (static_cast<uint32_t>((CG == 2 ? 256 : 128) >> 4) << 24);
    const int dstride = (nA * N <= 512) ? N : 0;
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint64_t bdesc = make_sw128_desc(base + b_off + ks * 32 + ((it % 3) * shift_rows) * 128);
        for (int m = 0; m < nA; ++m) {
          const uint64_t adesc = make_sw128_desc(base + m * a_bytes + ks * 32);
          if (CG == 2) mma_cg2(tmem_base + m * dstride, adesc, bdesc, idesc, (it | ks) ? 1u : 0u);
          else tc_mma_bf16(tmem_base + m * dstride, adesc, bdesc, idesc, (it | ks) ? 1u : 0u);
        }
      }
    }
    if (CG == 2) commit_cg2(bar, 0x3); else tc_commit(bar);
  }
  if (warp == 1 && lane == 0) {
    mbar_wait(bar, 0);
    t1 = clock64();
    if (leader) cyc[blockIdx.x] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    if (CG == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
  }
}

template <int CG>
void run(int N, int nA, int iters, int sms, long long* d_cyc, int shift_rows = 0) {
  const int smem = nA * 16384 + N * 128 + 8192;
  cudaFuncSetAttribute(k_mma<CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(sms); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaMemset(d_cyc, 0, sizeof(long long) * sms);
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    cudaError_t rc = cudaLaunchKernelEx(&cfg, k_mma<CG>, N, nA, iters, d_cyc, shift_rows);
    cudaEventRecord(e1);
    if (rc != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) { printf("cg%d N=%d nA=%d: FAILED %s\n", CG, N, nA, cudaGetErrorString(cudaGetLastError())); return; }
  }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[256]; cudaMemcpy(h, d_cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
  long long mx = 0; for (int i = 0; i < sms; ++i) if (h[i] > mx) mx = h[i];
  const double n_mma = static_cast<double>(iters) * 4 * nA;
  const double flops = n_mma * 2.0 * 128 * N * 16 * sms;          // per CTA: its own 128 rows x N
  printf("shift %d cg%d M=%d N=%3d nA=%d : %.1f cyc/MMA (floor %.0f)  kernel %.3f ms  %.0f TFLOP/s\n", shift_rows, CG, CG * 128, N, nA, mx / n_mma,
         CG == 2 ? N / 2.0 : N / 2.0, ms, flops / (ms * 1e-3) / 1e12);
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}

// Variant U: the WHOLE MMA warp runs the issue loop (warp-uniform control flow and addresses); only the
// tcgen05.mma itself sits under elect.sync.  Descriptors are assembled from a 32-bit running address plus a
// constant high word, so the per-MMA work is two 32-bit adds.
__global__ void __launch_bounds__(128, 1) k_mma_u(int N, int nA, int iters, long long* cyc) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* bp = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_bytes = 128 * 128;
  const uint32_t b_off = nA * a_bytes;
  const uint32_t total = b_off + ((N * 128 + 1023) & ~1023u) + 1024;
  const uint32_t bar = base + total;
  const uint32_t slot = bar + 16;
  for (uint32_t i = threadIdx.x; i < total / 2; i += blockDim.x) {
    uint32_t h = (i + blockIdx.x * 7919u) * 2654435761u;
    float f = static_cast<float>(static_cast<int>(h >> 24) - 128) * (1.0f / 256.0f);
    reinterpret_cast<__nv_bfloat16*>(bp)[i] = __float2bfloat16_rn(f);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x / 32, 0);
  if (threadIdx.x == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(slot));
  if (warp == 1) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
    const int dstride = (nA * N <= 512) ? N : 0;
    const uint32_t hi = static_cast<uint32_t>(make_sw128_desc(0) >> 32);
    const uint32_t lo_const = 1u << 16;
    const uint32_t a_lo = ((base & 0x3FFFFu) >> 4) | lo_const;
    const uint32_t b_lo = (((base + b_off) & 0x3FFFFu) >> 4) | lo_const;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint32_t acc = (it | ks) ? 1u : 0u;
        for (int m = 0; m < nA; ++m) {
          const uint32_t alo = a_lo + (m * a_bytes + ks * 32) / 16;
          const uint32_t blo = b_lo + (ks * 32) / 16;
          const uint32_t d = tmem_base + m * dstride;
          if (elect_one()) {
            asm volatile(
                "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %6, 0;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
                ::"r"(d), "r"(alo), "r"(hi), "r"(blo), "r"(hi), "r"(idesc), "r"(acc) : "memory");
          }
        }
      }
    }
    if (elect_one()) tc_commit(bar);
    mbar_wait(bar, 0);
    const long long t1 = clock64();
    if (threadIdx.x == 32) cyc[blockIdx.x] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
  }
}

void run_u(int N, int nA, int iters, int sms, long long* d_cyc) {
  const int smem = nA * 16384 + N * 128 + 8192;
  cudaFuncSetAttribute(k_mma_u, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaMemset(d_cyc, 0, sizeof(long long) * sms);
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    k_mma_u<<<sms, 128, smem>>>(N, nA, iters, d_cyc);
    cudaEventRecord(e1);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("U N=%d nA=%d: FAILED %s\n", N, nA, cudaGetErrorString(cudaGetLastError())); return; }
  }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[256]; cudaMemcpy(h, d_cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
  long long mx = 0; for (int i = 0; i < sms; ++i) if (h[i] > mx) mx = h[i];
  const double n_mma = static_cast<double>(iters) * 4 * nA;
  const double flops = n_mma * 2.0 * 128 * N * 16 * sms;
  printf("uniform-issue M=128 N=%3d nA=%d : %.1f cyc/MMA (floor %.0f)  kernel %.3f ms  %.0f TFLOP/s\n", N, nA, mx / n_mma, N / 2.0, ms,
         flops / (ms * 1e-3) / 1e12);
}

int main() {
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  const int sms = prop.multiProcessorCount / 2 * 2;
  long long* d_cyc; cudaMalloc(&d_cyc, sizeof(long long) * 256);
  const int iters = 4000;
  for (int nA : {1, 3}) for (int N : {64, 128, 160, 224, 256}) run_u(N, nA, iters, sms, d_cyc);
  for (int nA : {1, 3}) for (int N : {64, 128, 160, 256}) if (nA * 16384 + N * 128 < 200000) run<1>(N, nA, iters, sms, d_cyc);
  for (int nA : {1, 3}) for (int N : {128, 160, 256}) run<2>(N, nA, iters, sms, d_cyc);
  for (int sh : {0, 1}) { run<1>(224, 1, iters, sms, d_cyc, sh); run<1>(160, 3, iters, sms, d_cyc, sh); run<1>(192, 1, iters, sms, d_cyc, sh); }
  return 0;
}
