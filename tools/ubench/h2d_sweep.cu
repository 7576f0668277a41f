// h2d_sweep.cu -- ceiling of concurrent pinned host -> device copies on one box: plain cudaMemcpyAsync, one
// copy per buffer, G GPUs at once (G = 1, 2, 4, 8 ...).  The e2e leg of bench.py (mof_fft_pairs with host
// buffers) is bounded by this number divided by 115 216 B per frame pair.
// build: nvcc -O2 -o tools/ubench/h2d_sweep tools/ubench/h2d_sweep.cu ; run: tools/ubench/h2d_sweep [MB per buffer]
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

int main(int argc, char** argv) {
  const size_t mb = argc > 1 ? (size_t)std::atoi(argv[1]) : 256;
  const size_t bytes = mb << 20;
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  std::vector<void*> h(ndev), d(ndev);
  std::vector<cudaStream_t> st(ndev);
  for (int g = 0; g < ndev; g++) {
    CK(cudaSetDevice(g));
    CK(cudaHostAlloc(&h[g], bytes, cudaHostAllocPortable));
    std::memset(h[g], g + 1, bytes);
    CK(cudaMalloc(&d[g], bytes));
    CK(cudaStreamCreateWithFlags(&st[g], cudaStreamNonBlocking));
  }
  std::printf("{\"buffer_mb\": %zu, \"gpus_on_box\": %d, \"h2d\": [", mb, ndev);
  bool first = true;
  for (int G = 1; G <= ndev; G *= 2) {
    const int reps = 6;
    double best = 0;
    for (int trial = 0; trial < 3; trial++) {
      for (int g = 0; g < G; g++) { CK(cudaSetDevice(g)); CK(cudaStreamSynchronize(st[g])); }
      auto t0 = std::chrono::steady_clock::now();
      for (int r = 0; r < reps; r++)
        for (int g = 0; g < G; g++) {
          CK(cudaSetDevice(g));
          CK(cudaMemcpyAsync(d[g], h[g], bytes, cudaMemcpyHostToDevice, st[g]));
        }
      for (int g = 0; g < G; g++) { CK(cudaSetDevice(g)); CK(cudaStreamSynchronize(st[g])); }
      const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      const double gbs = (double)bytes * reps * G / dt / 1e9;
      if (gbs > best) best = gbs;
    }
    std::printf("%s{\"gpus\": %d, \"aggregate_gbs\": %.1f, \"per_gpu_gbs\": %.1f, \"pairs_per_s_ceiling\": %.0f}", first ? "" : ", ", G,
                best, best / G, best * 1e9 / 115216.0);
    first = false;
  }
  std::printf("]}\n");
  return 0;
}
