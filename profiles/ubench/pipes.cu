// Micro-benchmark of the per-SM pipe rates the chain-kernel epilogue depends on (B200, sm_100a):
// MUFU.TANH, MUFU.EX2, FFMA (scalar), FFMA2 (packed), F2FP bf16 pack.  One CTA per SM, W warps, clock64 timing.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu && ./pipes
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ float tanh_fast(float x) { float y; asm volatile("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float ex2_fast(float x) { float y; asm volatile("ex2.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
template <int KIND>
__global__ void bench(float* out, long long* cyc, int iters) {
  float x[16];
  u64 y[16];
  for (int i = 0; i < 16; ++i) { x[i] = threadIdx.x * 0.001f + i; y[i] = (u64)__float_as_uint(x[i]) | ((u64)__float_as_uint(x[i] + 0.5f) << 32); }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (KIND == 0) x[i] = tanh_fast(x[i]);
      if (KIND == 1) x[i] = ex2_fast(x[i]);
      if (KIND == 2) x[i] = ffma(x[i], 1.0001f, 0.5f);
      if (KIND == 3) y[i] = fma2(y[i], y[(i + 1) & 15], y[(i + 2) & 15]);
      if (KIND == 4) x[i] = ffma(x[i], x[(i + 1) & 15], x[(i + 2) & 15]);
      if (KIND == 5) { __nv_bfloat162 h = __floats2bfloat162_rn(x[i], x[(i + 1) & 15]); x[i] = __uint_as_float(*reinterpret_cast<unsigned*>(&h)); }
    }
  }
  long long t1 = clock64();
  __syncthreads();
  float s = 0;
  for (int i = 0; i < 16; ++i) s += x[i] + __uint_as_float((unsigned)y[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  const char* names[] = {"MUFU.TANH", "MUFU.EX2", "FFMA imm", "FFMA2 3-reg", "FFMA 3-reg", "F2FP.BF16 pack"};
  const int iters = 2000;
  for (int kind = 0; kind < 6; ++kind)
    for (int warps = 4; warps <= 16; warps *= 2) {
      for (int rep = 0; rep < 2; ++rep) {
        switch (kind) {
          case 0: bench<0><<<148, warps * 32>>>(out, cyc, iters); break;
          case 1: bench<1><<<148, warps * 32>>>(out, cyc, iters); break;
          case 2: bench<2><<<148, warps * 32>>>(out, cyc, iters); break;
          case 3: bench<3><<<148, warps * 32>>>(out, cyc, iters); break;
          case 4: bench<4><<<148, warps * 32>>>(out, cyc, iters); break;
          case 5: bench<5><<<148, warps * 32>>>(out, cyc, iters); break;
        }
        cudaDeviceSynchronize();
      }
      long long h[148];
      cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      double ops = (double)iters * 16 * warps * 32;  // thread-level ops per SM
      printf("%-16s warps/SM=%2d  cycles=%8lld  lane-ops/clk/SM=%7.1f\n", names[kind], warps, h[0], ops / h[0]);
    }
  return 0;
}
