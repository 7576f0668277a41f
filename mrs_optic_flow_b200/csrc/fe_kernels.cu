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
#include <stdlib.h>

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

// gray value of pixel i (0..3) of a 12-byte BGR run held in three words
__device__ __forceinline__ uint32_t fe_gray3(uint32_t c0, uint32_t c1, uint32_t c2) {
  return (9798u * c0 + 19235u * c1 + 3735u * c2 + 16384u) >> 15;
}
__device__ __forceinline__ uint32_t fe_gray4(uint32_t b0, uint32_t b1, uint32_t b2) {
  const uint32_t g0 = fe_gray3(b0 & 255u, (b0 >> 8) & 255u, (b0 >> 16) & 255u);
  const uint32_t g1 = fe_gray3(b0 >> 24, b1 & 255u, (b1 >> 8) & 255u);
  const uint32_t g2 = fe_gray3((b1 >> 16) & 255u, b1 >> 24, b2 & 255u);
  const uint32_t g3 = fe_gray3((b2 >> 8) & 255u, (b2 >> 16) & 255u, b2 >> 24);
  return g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
}

// BGR8 at scale 1, 16 output pixels per thread: the 48 source bytes come as four 16-byte loads from the
// 16-byte boundary below them (the crop of cfg5 starts 8 bytes past one), are re-aligned with funnel
// shifts, and the 16 gray bytes leave as one 16-byte store.  WSH = whole words to skip (0..3, uniform over
// the launch), sh = remaining byte shift * 8.  Four times the bytes in flight per thread of the 4-pixel
// path: 3.3 -> see DESIGN 4.4.
template <int WSH>
__device__ __forceinline__ uint4 fe_bgr16(const uint4* __restrict__ p, int sh) {
  const uint4 v0 = p[0], v1 = p[1], v2 = p[2], v3 = p[3];
  const uint32_t w[16] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w, v3.x, v3.y, v3.z, v3.w};
  uint32_t b[12];
#pragma unroll
  for (int i = 0; i < 12; i++) b[i] = __funnelshift_r(w[i + WSH], i + WSH + 1 < 16 ? w[i + WSH + 1] : 0u, sh);
  return make_uint4(fe_gray4(b[0], b[1], b[2]), fe_gray4(b[3], b[4], b[5]), fe_gray4(b[6], b[7], b[8]),
                    fe_gray4(b[9], b[10], b[11]));
}

__global__ void __launch_bounds__(256) fe_bgr16_kernel(const FeArgs a, int off16) {
  const int wq = a.out_w >> 4;   // 16-pixel groups per output row
  const unsigned per_image = (unsigned)a.out_h * (unsigned)wq;
  const size_t img = blockIdx.y;
  const uint8_t* src = a.src + img * a.src_image_stride;
  uint8_t* dst = a.dst + img * a.dst_image_stride;
  const int wsh = off16 >> 2, sh = (off16 & 3) * 8;
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < per_image; t += gridDim.x * blockDim.x) {
    const int q = (int)(t % (unsigned)wq), y = (int)(t / (unsigned)wq);
    const size_t o = (size_t)(a.y0 + y) * a.src_row_stride + (size_t)(a.x0 + 16 * q) * 3;
    const uint4* p = reinterpret_cast<const uint4*>(src + (o - (size_t)off16));
    uint4 g;
    switch (wsh) {
      case 0: g = fe_bgr16<0>(p, sh); break;
      case 1: g = fe_bgr16<1>(p, sh); break;
      case 2: g = fe_bgr16<2>(p, sh); break;
      default: g = fe_bgr16<3>(p, sh); break;
    }
    *reinterpret_cast<uint4*>(dst + (size_t)y * a.dst_row_stride + 16 * q) = g;
  }
}

template <int CH, int K>
__global__ void __launch_bounds__(256) fe_kernel(const FeArgs a) {
  const int wq = (a.out_w + 3) >> 2;   // 4-pixel groups per output row
  const unsigned per_image = (unsigned)a.out_h * (unsigned)wq;   // 32-bit index math: images ride on blockIdx.y
  const bool vec = ((((size_t)a.dst) | a.dst_image_stride | (size_t)a.dst_row_stride) & 3) == 0;
#if defined(MOF_FE_NO_WORDS)
  const bool src_words = false;
#else
  const bool src_words = ((((size_t)a.src) | a.src_image_stride | (size_t)a.src_row_stride) & 3) == 0;
#endif
  const size_t img = blockIdx.y;
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < per_image; t += gridDim.x * blockDim.x) {
    const int q = (int)(t % (unsigned)wq);
    const int y = (int)(t / (unsigned)wq);
    const uint8_t* src = a.src + (size_t)img * a.src_image_stride;
    uint8_t* drow = a.dst + (size_t)img * a.dst_image_stride + (size_t)y * a.dst_row_stride;
    const int x = 4 * q;
    // BGR8 at scale 1 (cfg5): the 12 source bytes of the four pixels come from aligned words and a
    // funnel shift -- a byte load per channel makes 12 load instructions of 3 cache lines each per
    // warp.  The words may run up to 3 bytes past the last pixel: the last group of an image's
    // last row takes the byte path.
    if (CH == 3 && K == 1 && vec && src_words && x + 3 < a.out_w &&
        !(y == a.out_h - 1 && x + 4 >= a.out_w)) {
      const size_t o = (size_t)(a.y0 + y) * a.src_row_stride + (size_t)(a.x0 + x) * 3;
      const uint32_t* w = reinterpret_cast<const uint32_t*>(src + (o & ~(size_t)3));
      const int sh = (int)(o & 3) * 8;
      const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = sh ? w[3] : 0u;
      const uint32_t b0 = __funnelshift_r(w0, w1, sh), b1 = __funnelshift_r(w1, w2, sh),
                     b2 = __funnelshift_r(w2, w3, sh);
      const uint32_t g0 = (9798u * (b0 & 255u) + 19235u * ((b0 >> 8) & 255u) + 3735u * ((b0 >> 16) & 255u) + 16384u) >> 15;
      const uint32_t g1 = (9798u * (b0 >> 24) + 19235u * (b1 & 255u) + 3735u * ((b1 >> 8) & 255u) + 16384u) >> 15;
      const uint32_t g2 = (9798u * ((b1 >> 16) & 255u) + 19235u * (b1 >> 24) + 3735u * (b2 & 255u) + 16384u) >> 15;
      const uint32_t g3 = (9798u * ((b2 >> 8) & 255u) + 19235u * ((b2 >> 16) & 255u) + 3735u * (b2 >> 24) + 16384u) >> 15;
      *reinterpret_cast<uint32_t*>(drow + x) = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
    } else if (vec && x + 3 < a.out_w) {
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
static int fe_launch_ch(const FeArgs& a, dim3 grid, cudaStream_t st) {
  switch (a.k) {
    case 1: fe_kernel<CH, 1><<<grid, 256, 0, st>>>(a); break;
    case 2: fe_kernel<CH, 2><<<grid, 256, 0, st>>>(a); break;
    case 4: fe_kernel<CH, 4><<<grid, 256, 0, st>>>(a); break;
    default: return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}

int fe_launch(const FeArgs& a, int sms, cudaStream_t st) {
  (void)sms;
  if (a.n_images <= 0 || a.out_w <= 0 || a.out_h <= 0) return 0;
  if (a.channels != 1 && a.channels != 3) return (int)cudaErrorInvalidValue;
  const long long per_image = (long long)a.out_h * ((a.out_w + 3) >> 2);
  if (per_image > 0x7fffffffLL) return (int)cudaErrorInvalidValue;
  const int bx = (int)((per_image + 255) / 256);
  // 16 pixels per thread when every address of the launch is 16-byte regular: the offset of a group's
  // first byte from the 16-byte boundary below it is then the same for all groups, rows and images.  The
  // four 16-byte loads may run up to 16 + 15 bytes past a group's last pixel: the source allocation must
  // cover them (row pitch >= crop end + 31 bytes, the crop not ending in the frame's last row, or slack
  // after the last image that the caller vouches for)
  const size_t first = (size_t)a.y0 * a.src_row_stride + (size_t)a.x0 * 3;
  const int off16 = (int)(((size_t)a.src + first) & 15);
  const bool tail_ok = (size_t)(a.x0 + a.out_w) * 3 + 32 <= (size_t)a.src_row_stride ||
                       (size_t)(a.y0 + a.out_h) * a.src_row_stride + 32 <= a.src_image_stride || a.src_slack >= 32;
  if (a.channels == 3 && a.k == 1 && (a.out_w & 15) == 0 && (a.src_row_stride & 15) == 0 &&
      (a.src_image_stride & 15) == 0 && ((((size_t)a.dst) | a.dst_image_stride | (size_t)a.dst_row_stride) & 15) == 0 &&
      tail_ok && !getenv("MOF_FE_NO_BGR16")) {
    const int bx16 = (int)(((long long)a.out_h * (a.out_w >> 4) + 255) / 256);
    for (int i0 = 0; i0 < a.n_images; i0 += 65535) {
      FeArgs b = a;
      b.n_images = a.n_images - i0 < 65535 ? a.n_images - i0 : 65535;
      b.src = a.src + (size_t)i0 * a.src_image_stride;
      b.dst = a.dst + (size_t)i0 * a.dst_image_stride;
      fe_bgr16_kernel<<<dim3(bx16, b.n_images), 256, 0, st>>>(b, off16);
    }
    return (int)cudaGetLastError();
  }
  for (int i0 = 0; i0 < a.n_images; i0 += 65535) {   // gridDim.y limit
    FeArgs b = a;
    b.n_images = a.n_images - i0 < 65535 ? a.n_images - i0 : 65535;
    b.src = a.src + (size_t)i0 * a.src_image_stride;
    b.dst = a.dst + (size_t)i0 * a.dst_image_stride;
    const int rc = a.channels == 1 ? fe_launch_ch<1>(b, dim3(bx, b.n_images), st)
                                   : fe_launch_ch<3>(b, dim3(bx, b.n_images), st);
    if (rc) return rc;
  }
  return 0;
}

}  // namespace mof

// -------------------------------------------------------------------------------------------------
// Point-pair assembly for OpticFlow::getRT / get2DT (src/optic_flow.cpp:527-550, :402-416): per frame,
// the sections with finite shifts, in section order (j outer, i inner), as
//   initial = (i*N + N/2, j*N + N/2),  shifted = initial + shift,
// plus their count (the caller compares it with shifted_pts_thr, :552).  One warp per frame,
// order-preserving compaction with a ballot prefix.  Unused tail entries are set to NaN.
// -------------------------------------------------------------------------------------------------
namespace mof {

__global__ void __launch_bounds__(128) pp_kernel(const double* __restrict__ shifts, int n_frames, int sq, int N,
                                                 double* __restrict__ initial, double* __restrict__ shifted,
                                                 int* __restrict__ counts) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n_frames) return;
  const int S = sq * sq;
  const double* sh = shifts + (size_t)warp * S * 2;
  double* oi = initial + (size_t)warp * S * 2;
  double* os = shifted + (size_t)warp * S * 2;
  int base = 0;
  for (int s0 = 0; s0 < S; s0 += 32) {
    const int s = s0 + lane;
    double x = 0.0, y = 0.0;
    bool ok = false;
    if (s < S) {
      x = sh[2 * s];
      y = sh[2 * s + 1];
      ok = isfinite(x) && isfinite(y);
    }
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const int k = base + __popc(m & ((1u << lane) - 1u));
      const int j = s / sq, i = s - j * sq;
      const double xi = (double)(i * N + N / 2), yi = (double)(j * N + N / 2);
      oi[2 * k] = xi; oi[2 * k + 1] = yi;
      os[2 * k] = xi + x; os[2 * k + 1] = yi + y;
    }
    base += __popc(m);
  }
  for (int k = base + lane; k < S; k += 32) {
    oi[2 * k] = oi[2 * k + 1] = os[2 * k] = os[2 * k + 1] = __longlong_as_double(0x7ff8000000000000LL);
  }
  if (lane == 0) counts[warp] = base;
}

int pp_launch(const double* shifts, int n_frames, int sq, int N, double* initial, double* shifted, int* counts,
              cudaStream_t st) {
  if (n_frames <= 0 || sq <= 0) return 0;
  const int blocks = (n_frames + 3) / 4;
  pp_kernel<<<blocks, 128, 0, st>>>(shifts, n_frames, sq, N, initial, shifted, counts);
  return (int)cudaGetLastError();
}

}  // namespace mof
