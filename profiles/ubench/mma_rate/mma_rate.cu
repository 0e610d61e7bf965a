// Microbenchmark: how fast can ONE thread feed the tensor core with M=128 (or M=256 cta-pair off) N=256 K=16 bf16 MMAs out
// of shared memory, alone and next to busy CUDA-core warps on the same schedulers?
//   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I flow-q-learning_b200/csrc -o mma_rate mma_rate.cu
//   variants: 0 bare MMAs, 1 + commit to an mbarrier every 2 MMAs
//   busy warps: 0 none, 1 = 16 warps of dependent FMAs, 2 = 16 warps of st.shared, 3 = 16 warps of MUFU+FMA
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "ptx_sm100.cuh"
using namespace fqlpop;

constexpr int kThreads = 640;
constexpr int kSmem = 200 * 1024;

__global__ void __launch_bounds__(kThreads, 1) mma_rate_kernel(int variant, int busy, int n_mma, int n_cols, long long* out, float* sink) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bars[8];
  __shared__ uint32_t tmem_slot;
  __shared__ volatile int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1);
    stop = 0;
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += kThreads) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_proxy_async_smem();
  if (warp == 1) { tmem_alloc(&tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (warp == 1) {
    if (lane == 0) {
      const uint64_t adesc0 = make_smem_desc_sw128(smem_u32(smem), 16, 1024);
      const uint64_t bdesc0 = make_smem_desc_sw128(smem_u32(smem + 128 * 1024), 4096, 1024);
      const uint32_t idesc = make_idesc_bf16(128, n_cols, 0, 1);
      const long long t0 = clock64();
      for (int i = 0; i < n_mma; i += 2) {
        const int kb = (i >> 2) & 7, h = (i >> 1) & 1, st = (i >> 1) & 1;
        const uint64_t a = adesc0 + (uint64_t)((kb * 16384 + h * 64) >> 4);
        const uint64_t b = bdesc0 + (uint64_t)((st * 16384) >> 4);
        umma_bf16(tmem_base, a, b, idesc, i ? 1u : 0u);
        umma_bf16(tmem_base, a + 2, b + 128, idesc, 1u);
        if (variant >= 1) umma_commit(&bars[st]);
      }
      const long long t1 = clock64();
      umma_commit(&bars[7]);
      mbar_wait(&bars[7], 0);
      const long long t2 = clock64();
      out[0] = t1 - t0;
      out[1] = t2 - t0;
      stop = 1;
    }
  } else if (warp >= 4 && busy) {
    float x = threadIdx.x * 1e-3f, y = 1.0001f;
    uint32_t a = smem_u32(smem + 160 * 1024) + threadIdx.x * 16;
    long long it = 0;
    while (!stop) {
      if (busy == 1) {
#pragma unroll
        for (int i = 0; i < 64; ++i) x = fmaf(x, y, 0.25f);
      } else if (busy == 2) {
#pragma unroll
        for (int i = 0; i < 16; ++i) asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(a), "r"(i) : "memory");
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) { x = fmaf(x, y, 0.25f); x = tanh_fast(x); x = fmaf(x, x, y); x = x * y; }
      }
      ++it;
    }
    if (x == 123.456f) sink[0] = x + it;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

// cluster of two CTAs, tcgen05.mma.cta_group::2 (M = 256): each CTA holds its 128 rows of A and half of B's columns
__global__ void __launch_bounds__(kThreads, 1) mma_rate_pair_kernel(int n_mma, int n_cols, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bars[8];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += kThreads) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_proxy_async_smem();
  if (warp == 1) { tmem_alloc_2sm(&tmem_slot, 512); tmem_relinquish_2sm(); }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (warp == 1 && lane == 0 && rank == 0) {
    const uint64_t adesc0 = make_smem_desc_sw128(smem_u32(smem), 16, 1024);
    const uint64_t bdesc0 = make_smem_desc_sw128(smem_u32(smem + 128 * 1024), 4096, 1024);
    const uint32_t idesc = make_idesc_bf16(256, n_cols, 0, 1);
    const long long t0 = clock64();
    for (int i = 0; i < n_mma; i += 2) {
      const int kb = (i >> 2) & 7, h = (i >> 1) & 1, st = (i >> 1) & 1;
      const uint64_t a = adesc0 + (uint64_t)((kb * 16384 + h * 64) >> 4);
      const uint64_t b = bdesc0 + (uint64_t)((st * 8192) >> 4);
      umma_bf16_2sm(tmem_base, a, b, idesc, i ? 1u : 0u);
      umma_bf16_2sm(tmem_base, a + 2, b + 128, idesc, 1u);
      umma_commit_2sm(&bars[st], 3);
    }
    const long long t1 = clock64();
    umma_commit_2sm(&bars[7], 3);
    mbar_wait(&bars[7], 0);
    const long long t2 = clock64();
    out[0] = t1 - t0;
    out[1] = t2 - t0;
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) { tc_fence_after(); tmem_dealloc_2sm(tmem_base, 512); }
}

// queue depth: clock after each of the first 16 back-to-back MMA issues (N = 256) on an idle tensor core
__global__ void __launch_bounds__(kThreads, 1) mma_queue_kernel(long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bars[2];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&bars[0], 1); fence_barrier_init(); }
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += kThreads) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_proxy_async_smem();
  if (warp == 1) { tmem_alloc(&tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (warp == 1 && lane == 0) {
    const uint64_t adesc0 = make_smem_desc_sw128(smem_u32(smem), 16, 1024);
    const uint64_t bdesc0 = make_smem_desc_sw128(smem_u32(smem + 128 * 1024), 4096, 1024);
    const uint32_t idesc = make_idesc_bf16(128, 256, 0, 1);
    long long t[17];
    t[0] = clock64();
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      umma_bf16(tmem_base, adesc0 + 2 * (i & 1), bdesc0 + 128 * (i & 1), idesc, i ? 1u : 0u);
      t[i + 1] = clock64();
    }
    umma_commit(&bars[0]);
    mbar_wait(&bars[0], 0);
    const long long t_end = clock64();
    for (int i = 0; i < 17; ++i) out[i] = t[i] - t[0];
    out[17] = t_end - t[0];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

int main(int argc, char** argv) {
  const int n_mma = 4096;
  long long* out; float* sink;
  cudaMalloc(&out, 64); cudaMalloc(&sink, 64);
  cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  for (int ncols : {256, 128}) for (int variant = 0; variant < 2; ++variant) for (int busy = 0; busy < 4; ++busy) {
    long long h[2];
    for (int rep = 0; rep < 2; ++rep) {
      mma_rate_kernel<<<1, kThreads, kSmem>>>(variant, busy, n_mma, ncols, out, sink);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    }
    printf("N=%d variant %d busy %d: issue %.1f cycles/MMA, complete %.1f cycles/MMA\n", ncols, variant, busy, (double)h[0] / n_mma, (double)h[1] / n_mma);
  }
  {
    cudaFuncSetAttribute(mma_queue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
    long long* q; cudaMalloc(&q, 18 * 8);
    long long h[18];
    for (int rep = 0; rep < 2; ++rep) {
      mma_queue_kernel<<<1, kThreads, kSmem>>>(q);
      cudaDeviceSynchronize();
      cudaMemcpy(h, q, 18 * 8, cudaMemcpyDeviceToHost);
    }
    printf("issue times of 16 back-to-back MMAs (cycles):");
    for (int i = 1; i < 17; ++i) printf(" %lld", h[i]);
    printf("; all complete at %lld\n", h[17]);
  }
  cudaFuncSetAttribute(mma_rate_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  for (int ncols : {256, 128}) {
    long long h[2];
    for (int rep = 0; rep < 2; ++rep) {
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(2); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = kSmem;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      cudaLaunchKernelEx(&cfg, mma_rate_pair_kernel, n_mma, ncols, out);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("pair error %s\n", cudaGetErrorString(e)); return 1; }
      cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    }
    printf("PAIR M=256 N=%d: issue %.1f cycles/MMA, complete %.1f cycles/MMA\n", ncols, (double)h[0] / n_mma, (double)h[1] / n_mma);
  }
  return 0;
}
