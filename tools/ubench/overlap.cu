// microbenchmark: do packed-FP32 math and 64-bit shared-memory traffic overlap on one SM?
// per iteration and thread: NL LDS.64 + NL STS.64 (conflict free) and NM FADD2/FFMA2; modes: mem only,
// math only, both (independent), both with the math depending on the loads (FFT-like item).
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2000
typedef unsigned long long u64;
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(float* out, int stride) {
  extern __shared__ u64 sm[];
  const int t = threadIdx.x;
  for (int i = t; i < 512 * 17; i += blockDim.x) sm[i] = (u64)i * 0x3f8000013f800001ull;
  __syncthreads();
  u64 z[16], acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) { z[i] = 0; acc[i] = (u64)(t + i) << 20; }
  u64* base = sm + (t >> 5) * 32 * 17 + (t & 31);
  for (int it = 0; it < ITERS; it++) {
    if (MODE != 1) {
#pragma unroll
      for (int i = 0; i < 16; i++) z[i] = base[i * 32 * stride + 0];
    }
    if (MODE == 3) {
#pragma unroll
      for (int r = 0; r < 5; r++)
#pragma unroll
        for (int i = 0; i < 16; i++) z[i] = add2(z[i], z[(i + 1 + r) & 15]);
    } else if (MODE != 0) {
#pragma unroll
      for (int r = 0; r < 5; r++)
#pragma unroll
        for (int i = 0; i < 16; i++) acc[i] = add2(acc[i], acc[(i + 1 + r) & 15]);
    }
    if (MODE != 1) {
#pragma unroll
      for (int i = 0; i < 16; i++) base[i * 32 * stride] = z[i] + (MODE == 3 ? 0 : it);
    }
  }
  u64 s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += z[i] + acc[i];
  out[blockIdx.x * blockDim.x + t] = (float)s;
}
template <int MODE>
void run(const char* name, int threads) {
  float* d;
  cudaMalloc(&d, 148 * 1024 * 4);
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 512 * 17 * 8);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148, threads, 512 * 17 * 8>>>(d, 1);
  cudaEventRecord(e0);
  k<MODE><<<148, threads, 512 * 17 * 8>>>(d, 1);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double cyc = ms * 1e-3 * 1.9e9 / ITERS;
  printf("%-28s threads %4d: %.3f ms, %.0f clk/iter/SM  (smem wavefronts/iter %d, FADD2 pipe clk/iter/SMSP %d) err %d\n", name,
         threads, ms, cyc, threads / 32 * 32 * 2, threads / 32 / 4 * 80 * 2, (int)cudaGetLastError());
  cudaFree(d);
}
int main() {
  for (int t : {256, 512}) {
    run<0>("smem only", t); run<1>("math only", t); run<2>("both, independent", t); run<3>("both, math on loaded data", t);
  }
}
