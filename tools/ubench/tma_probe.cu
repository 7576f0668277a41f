// tma_probe.cu -- isolates the TMA box load used by bm_kernels.cu: one CTA, one 3-D box, compare with the image.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void probe2(const __grid_constant__ CUtensorMap map, int x, int y, int bytes, unsigned char* out) {
  extern __shared__ __align__(128) unsigned char sm[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + bytes);
  const uint32_t ba = (uint32_t)__cvta_generic_to_shared(bar);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ba) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ba), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"((uint32_t)__cvta_generic_to_shared(sm)), "l"(&map), "r"(x), "r"(y), "r"(ba) : "memory");
  }
  unsigned done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(ba), "r"(0) : "memory");
  } while (!done);
  for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = sm[i];
}
__global__ void probe(const __grid_constant__ CUtensorMap map, int x, int y, int z, int bw, int bh, unsigned char* out) {
  extern __shared__ __align__(128) unsigned char sm[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + bw * bh);
  const uint32_t ba = (uint32_t)__cvta_generic_to_shared(bar);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ba) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ba), "r"(bw * bh) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"((uint32_t)__cvta_generic_to_shared(sm)), "l"(&map), "r"(x), "r"(y), "r"(z), "r"(ba) : "memory");
  }
  unsigned done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(ba), "r"(0) : "memory");
  } while (!done);
  for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = sm[i];
}
int main(int argc, char** argv) {
  const int rank = argc > 1 ? atoi(argv[1]) : 3, eb = argc > 2 ? atoi(argv[2]) : 1;
  const int W = 240, H = 240, NP = 2, bw = 64, bh = 60;
  if (rank == 2) {
    std::vector<unsigned char> h2((size_t)W * H * NP, 3);
    unsigned char *d2, *o2;
    cudaMalloc(&d2, h2.size()); cudaMalloc(&o2, bw * bh);
    cudaMemcpy(d2, h2.data(), h2.size(), cudaMemcpyHostToDevice);
    void* fn2 = nullptr; cudaDriverEntryPointQueryResult q2;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn2, cudaEnableDefault, &q2);
    CUtensorMap m2;
    const cuuint64_t dims2[2] = {(cuuint64_t)(W / eb), (cuuint64_t)H * NP}; const cuuint64_t str2[1] = {W};
    const cuuint32_t box2[2] = {(cuuint32_t)(bw / eb), (cuuint32_t)bh}; const cuuint32_t es2[2] = {1, 1};
    CUresult r2 = ((EncodeTiledFn)fn2)(&m2, eb == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d2, dims2, str2,
                                       box2, es2, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    std::printf("2-D eb=%d encode: %d\n", eb, (int)r2);
    cudaFuncSetAttribute(probe2, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    probe2<<<1, 128, bw * bh + 64>>>(m2, argc > 3 ? atoi(argv[3]) : 0, argc > 4 ? atoi(argv[4]) : 0, bw * bh, o2);
    std::printf("2-D launch: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
  }
  std::vector<unsigned char> h((size_t)W * H * NP);
  for (size_t i = 0; i < h.size(); i++) h[i] = (unsigned char)(i * 7 + (i >> 8));
  unsigned char *d, *o;
  cudaMalloc(&d, h.size()); cudaMalloc(&o, bw * bh);
  cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  std::printf("entry point: %s q=%d fn=%p\n", cudaGetErrorString(e), (int)q, fn);
  CUtensorMap m;
  const cuuint64_t dims[3] = {W, H, NP}; const cuuint64_t strides[2] = {W, (cuuint64_t)W * H};
  const cuuint32_t box[3] = {bw, bh, 1}; const cuuint32_t es[3] = {1, 1, 1};
  CUresult r = ((EncodeTiledFn)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  std::printf("encode: %d\n", (int)r);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int t = 0; t < 3; t++) {
    const int x = argc > 3 ? atoi(argv[3]) : (t == 0 ? 8 : (t == 1 ? 21 : 200)), y = t == 2 ? 200 : 8, z = t & 1;
    probe<<<1, 128, bw * bh + 64>>>(m, x, y, z, bw, bh, o);
    e = cudaDeviceSynchronize();
    std::printf("launch %d (x=%d y=%d z=%d): %s\n", t, x, y, z, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<unsigned char> g(bw * bh);
    cudaMemcpy(g.data(), o, g.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int j = 0; j < bh; j++) for (int i = 0; i < bw; i++) {
      const int xx = x + i, yy = y + j;
      const unsigned char want = (xx < W && yy < H) ? h[(size_t)z * W * H + (size_t)yy * W + xx] : 0;
      bad += g[j * bw + i] != want;
    }
    std::printf("  mismatches %d\n", bad);
  }
  return 0;
}
