// fe_kernels.cu -- camera front end on the device for sm_100a: integer down-scale, crop, grayscale.
//
// Restates what the nodelet does to every camera frame before the per-section search
// (src/optic_flow.cpp:1603-1622):
//   cv::resize(image, Size(w / scale_factor, h / scale_factor))         (:1603-1604, INTER_LINEAR)
//   crop F x F at (cx - F/2, h_scaled/2 - F/2)                            (:1610-1617)
//   cv::cvtColor(crop, imCurr, CV_RGB2GRAY)                               (:1622)
// and FftMethod::processImageLongRange's cv::resize(im, 1/4) (src/FftMethod.cpp:1931-1932), which is
// the same kernel on one channel with k = 4.
//
// Arithmetic (pinned bit-exact against cv2 4.13, oracle/gen_golden.py):
//   * INTER_LINEAR with an exact even integer ratio k samples at k*dx + (k-1)/2, i.e. half way
//     between source pixels k*dx + k/2 - 1 and k*dx + k/2 on both axes: (a + b + c + d + 2) >> 2;
//     k = 1 is a copy.  Other ratios are rejected by the C ABI.
//   * RGB2GRAY on 8-bit: (9798*c0 + 19235*c1 + 3735*c2 + 16384) >> 15 (c0 = first channel in memory;
//     the nodelet applies CV_RGB2GRAY to BGR data and so do we).  One channel passes through.
// Pure byte/integer work, HBM bound: every source byte inside the crop is read once, every gray
// byte written once; a thread makes four adjacent output pixels and stores them as one word.
#include <cuda_runtime.h>

#include "mof_internal.h"

namespace mof {

template <int CH, int K>
__device__ __forceinline__ uint32_t fe_pixel(const uint8_t* __restrict__ img, int row_stride, int xs, int ys) {
  // (xs, ys): pixel of the scaled image
  uint32_t c[CH];
  if (K == 1) {
    const uint8_t* p = img + (size_t)ys * row_stride + (size_t)xs * CH;
#pragma unroll
    for (int i = 0; i < CH; i++) c[i] = p[i];
  } else {
    const uint8_t* p = img + (size_t)(ys * K + K / 2 - 1) * row_stride + (size_t)(xs * K + K / 2 - 1) * CH;
#pragma unroll
    for (int i = 0; i < CH; i++)
      c[i] = ((uint32_t)p[i] + p[CH + i] + p[row_stride + i] + p[row_stride + CH + i] + 2u) >> 2;
  }
  if (CH == 1) return c[0];
  return (9798u * c[0] + 19235u * c[CH > 1 ? 1 : 0] + 3735u * c[CH > 2 ? 2 : 0] + 16384u) >> 15;
}

template <int CH, int K>
__global__ void __launch_bounds__(256) fe_kernel(const FeArgs a) {
  const int wq = (a.out_w + 3) >> 2;   // 4-pixel groups per output row
  const long long total = (long long)a.n_images * a.out_h * wq;
  const bool vec = ((((size_t)a.dst) | a.dst_image_stride | (size_t)a.dst_row_stride) & 3) == 0;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(t % wq);
    const long long r = t / wq;
    const int y = (int)(r % a.out_h);
    const long long img = r / a.out_h;
    const uint8_t* src = a.src + (size_t)img * a.src_image_stride;
    uint8_t* drow = a.dst + (size_t)img * a.dst_image_stride + (size_t)y * a.dst_row_stride;
    const int x = 4 * q;
    if (vec && x + 3 < a.out_w) {
      uint32_t w = 0;
#pragma unroll
      for (int j = 0; j < 4; j++) w |= fe_pixel<CH, K>(src, a.src_row_stride, a.x0 + x + j, a.y0 + y) << (8 * j);
      *reinterpret_cast<uint32_t*>(drow + x) = w;
    } else {
      for (int j = 0; j < 4 && x + j < a.out_w; j++)
        drow[x + j] = (uint8_t)fe_pixel<CH, K>(src, a.src_row_stride, a.x0 + x + j, a.y0 + y);
    }
  }
}

template <int CH>
static int fe_launch_ch(const FeArgs& a, int grid, cudaStream_t st) {
  switch (a.k) {
    case 1: fe_kernel<CH, 1><<<grid, 256, 0, st>>>(a); break;
    case 2: fe_kernel<CH, 2><<<grid, 256, 0, st>>>(a); break;
    case 4: fe_kernel<CH, 4><<<grid, 256, 0, st>>>(a); break;
    default: return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}

int fe_launch(const FeArgs& a, int sms, cudaStream_t st) {
  if (a.n_images <= 0 || a.out_w <= 0 || a.out_h <= 0) return 0;
  const long long total = (long long)a.n_images * a.out_h * ((a.out_w + 3) >> 2);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)(sms > 0 ? sms : 148) * 8;   // 8 resident CTAs of 256 threads per SM
  const int grid = (int)(blocks < cap ? blocks : cap);
  if (a.channels == 1) return fe_launch_ch<1>(a, grid, st);
  if (a.channels == 3) return fe_launch_ch<3>(a, grid, st);
  return (int)cudaErrorInvalidValue;
}

}  // namespace mof
