// lp_kernels.cu -- log-polar resample of uint8 frames for sm_100a (scale / rotation path).
//
// Restates what scaleRotationEstimator::processImage asks OpenCV to do
// (src/scaleRotationEstimator.cpp:45 first frame INTER_CUBIC, :112 later frames INTER_LANCZOS4):
// cv::logPolar -> cv::warpPolar(WARP_POLAR_LOG) -> cv::remap on CV_8UC1 with float maps, i.e.
//   * source coordinates quantised to 1/32 px (INTER_BITS = 5),
//   * separable 4x4 / 8x8 weights from a 32x32 table of 15-bit fixed-point coefficients,
//   * int32 accumulation, (sum + 2^14) >> 15, saturate to uint8,
//   * BORDER_TRANSPARENT: destination untouched when the centre tap is outside the source, taps
//     outside the source mirrored (BORDER_REFLECT_101).
// The per-pixel fixed-point coordinates depend only on (res, M); the backend computes them once on
// the host in double precision exactly as warpPolar does and uploads them (LpMap), so the device
// work is pure integer arithmetic and bit-exact.
#include <cuda_runtime.h>

#include "mof_internal.h"

namespace mof {

__device__ __forceinline__ int reflect101(int p, int len) {
  // one reflection suffices for |overshoot| < len (taps reach at most 4 px outside)
  if (p < 0) p = -p;
  if (p >= len) p = 2 * (len - 1) - p;
  if (p < 0) p = 0;
  return p;
}

template <int KS>
__global__ void __launch_bounds__(256) lp_kernel(const LpMap m, const LpArgs a) {
  const int res = m.res;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= res * res) return;
  const int img = blockIdx.y;
  const int ix = m.ix[p], iy = m.iy[p];
  if ((unsigned)ix >= (unsigned)res || (unsigned)iy >= (unsigned)res) return;  // transparent
  const int16_t* w = (KS == 4 ? m.tab_cubic : m.tab_lanczos) + (size_t)m.fxy[p] * KS * KS;
  const uint8_t* src = a.src + (size_t)img * a.src_image_stride;
  const int x0 = ix - (KS / 2 - 1), y0 = iy - (KS / 2 - 1);
  int xs[KS];
#pragma unroll
  for (int k = 0; k < KS; k++) xs[k] = reflect101(x0 + k, res);
  int sum = 0;
#pragma unroll
  for (int k1 = 0; k1 < KS; k1++) {
    const uint8_t* row = src + (size_t)reflect101(y0 + k1, res) * a.src_row_stride;
#pragma unroll
    for (int k2 = 0; k2 < KS; k2++) sum += (int)row[xs[k2]] * (int)w[k1 * KS + k2];
  }
  int v = (sum + (1 << 14)) >> 15;
  v = v < 0 ? 0 : (v > 255 ? 255 : v);
  const int oy = p / res, ox = p - oy * res;
  a.dst[(size_t)img * a.dst_image_stride + (size_t)oy * a.dst_row_stride + ox] = (uint8_t)v;
}

int lp_launch(const LpMap& map, const LpArgs& args, cudaStream_t stream) {
  if (args.n_images <= 0) return 0;
  const int npx = map.res * map.res;
  for (int i0 = 0; i0 < args.n_images; i0 += 65535) {
    LpArgs b = args;
    const int n = (args.n_images - i0 < 65535) ? args.n_images - i0 : 65535;
    b.src = args.src + (size_t)i0 * args.src_image_stride;
    b.dst = args.dst + (size_t)i0 * args.dst_image_stride;
    b.n_images = n;
    dim3 grid((npx + 255) / 256, n);
    if (args.interp == 2) lp_kernel<4><<<grid, 256, 0, stream>>>(map, b);
    else if (args.interp == 4) lp_kernel<8><<<grid, 256, 0, stream>>>(map, b);
    else return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}

}  // namespace mof
