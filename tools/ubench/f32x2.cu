// microbenchmark: scalar FADD/FFMA vs packed add.f32x2 / fma.rn.f32x2 issue throughput on sm_100a
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
template <int MODE>
__global__ void k(float* out, float seed) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; i++) a[i] = seed + i + threadIdx.x;
  const float c = seed * 0.5f;
  unsigned long long cc;
  asm volatile("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
  for (int it = 0; it < ITERS; it++) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 16; i++) a[i] = a[i] + c;
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 16; i++) a[i] = fmaf(a[i], c, c);
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        unsigned long long v;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[i]), "f"(a[i + 1]));
        asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(v) : "l"(cc));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(a[i + 1]) : "l"(v));
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        unsigned long long v;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[i]), "f"(a[i + 1]));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(v) : "l"(cc));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(a[i + 1]) : "l"(v));
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name, int threads) {
  float* d;
  cudaMalloc(&d, 148 * 1024 * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148, threads>>>(d, 1.0f);
  cudaEventRecord(e0);
  k<MODE><<<148, threads>>>(d, 1.0f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double lane_ops = 148.0 * threads * 16.0 * ITERS;
  printf("%-10s threads/SM %4d: %.3f ms, %.1f lane-ops/clk/SM (at 1.9 GHz), err %d\n", name, threads, ms,
         lane_ops / (ms * 1e-3) / 148 / 1.9e9, (int)cudaGetLastError());
  cudaFree(d);
}
int main() {
  for (int t : {128, 256, 512, 1024}) {
    run<0>("FADD", t); run<1>("FFMA", t); run<2>("FADD2", t); run<3>("FFMA2", t);
  }
}
