// microbenchmark: the real row/column stage functions of pc_pfa.cuh in a bare loop (no barriers, no
// global traffic), 15 main warps on the N = 240 section array -> cycles per stage call per SM,
// next to the shared-memory and FMA-pipe floors.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../mrs_optic_flow_b200/csrc/pc_pfa.cuh"
using namespace mof;
using namespace mof::pfa;
#define ITERS 200
typedef PG<16> G;
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(float* out) {
  extern __shared__ __align__(16) unsigned char raw[];
  c32* A = reinterpret_cast<c32*>(raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < G::A_ELEMS; i += blockDim.x) A[i] = mk((float)(i % 251), (float)(i % 127));
  __syncthreads();
  if (warp < G::TASKS) {
    const int task = warp;
    const int nu = task == G::TASKS - 1 ? G::UPT - 1 : G::UPT;
    MaxTrack mt; mt.vmax = -1e30f; mt.which = 0; mt.multi = false;
    for (int it = 0; it < ITERS; it++) {
      if (MODE == 0) stage_x16<16, 1, false>(A, task, lane, nullptr);
      if (MODE == 1) stage_ix15<16>(A, task, lane);
      if (MODE == 2) stage_y16<16, false>(A, task * G::UPT, nu, lane);
      if (MODE == 3) stage_y16<16, true>(A, task * G::UPT, nu, lane);
      if (MODE == 4) stage_x16<16, -1, true>(A, task, lane, &mt);
      __syncwarp();
    }
    if (mt.vmax == 12345.f) out[tid] = mt.vmax;
  }
  __syncthreads();
  if (tid == 0) out[blockIdx.x] = A[blockIdx.x].x;
}
template <int MODE>
void run(const char* name, double wavefronts, double packed_instr_per_warp) {
  float* d;
  cudaMalloc(&d, 1 << 16);
  const size_t smem = sizeof(c32) * G::A_ELEMS;
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148, 512, smem>>>(d);
  cudaEventRecord(e0);
  k<MODE><<<148, 512, smem>>>(d);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("%-22s %.3f ms -> %.0f clk/call/SM   floors: smem %.0f clk, FMA pipe %.0f clk  err %d\n", name, ms,
         ms * 1e-3 * 1.9e9 / ITERS, wavefronts, packed_instr_per_warp * 2 * 4, (int)cudaGetLastError());
  cudaFree(d);
}
int main() {
  // one section pass = 28800 c32 read + written = 2 * 1800 wavefronts
  run<0>("FX16 (dft16 fwd)", 3600, 4 * 80);    // 4 rounds of dft16 (80 packed) per warp; 4 warps / SMSP
  run<1>("IX15 (dft15 inv)", 3600, 4 * 78);
  run<2>("FY16 (unpack+dft16)", 3600, 4 * 96);
  run<3>("IY16 (dft16+repack)", 3600, 4 * 96);
  run<4>("IX16 tracked", 3600, 4 * 96);
}
