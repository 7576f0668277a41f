// rf_kernels.cu -- BlockMethod::Refine on the device for sm_100a (src/BlockMethod.cpp:96-147).
//
// What the reference does, restated (quirks included -- they define the result):
//   * both working images are made from imCurr: `cv::resize(imCurr2x, imPrev2x, 2x)` overwrites the
//     previous image with the up-scaled CURRENT one (:110), then imCurr2x is up-scaled in place (:111);
//   * the target size is always twice the ORIGINAL size, so pass 2 "resizes" 2F x 2F to 2F x 2F (a
//     copy): both passes search the same 2x image, only the offset and the scale double;
//   * per pass: offset *= 2; nine SADs between the cut-out of the image at (1, 1) and the cut-out at
//     startpoint + (n, m), n, m in {-1, 0, 1}; startpoint = (-off + 1 if off < 0 else 1) per axis;
//     cut-out = 2F - (|off| + 2) per axis (:113-137); cv::minMaxLoc keeps the first minimum in
//     row-major (m outer, n inner) order (:142); offset += argmin - (1, 1) (:144);
//   * result = offset / (float)2^passes (:146).
// cv::resize (INTER_LINEAR, 8-bit, exact factor 2) is pinned bit-exact against cv2 4.13
// (oracle/gen_golden.py): 11-bit coefficients {512, 1536} (2048/0 at the left/right border, where
// the fraction is zeroed; rows are clamped instead, keeping the fraction) and
//   D = (((b0 * (H0 >> 4)) >> 16) + ((b1 * (H1 >> 4)) >> 16) + 2) >> 2,  H = S[sx]*a0 + S[sx+1]*a1.
// Pure byte/integer work.  rf_up2_kernel is HBM bound (reads F^2, writes 4 F^2 per frame);
// rf_sad_kernel runs one CTA per frame on the L2-resident 2x image (the two passes are dependent).
#include <cuda_runtime.h>
#include <math.h>

#include "mof_internal.h"

namespace mof {

__device__ __forceinline__ void rf_coef(int d, int n, int& s0, int& s1, int& c0, int& c1, bool zero_frac_at_border) {
  // output index d of a 2x up-scale of n samples: sample pair (s0, s1) and 11-bit weights (c0, c1)
  int s = (d & 1) ? (d - 1) >> 1 : (d >> 1) - 1;
  c1 = (d & 1) ? 512 : 1536;
  if (zero_frac_at_border) {
    if (s < 0) { s = 0; c1 = 0; }
    if (s >= n - 1) { s = n - 1; c1 = 0; }
    s0 = s;
    s1 = s + 1 < n ? s + 1 : n - 1;
  } else {
    s0 = s < 0 ? 0 : (s > n - 1 ? n - 1 : s);
    s1 = s + 1 < 0 ? 0 : (s + 1 > n - 1 ? n - 1 : s + 1);
  }
  c0 = 2048 - c1;
}

__global__ void __launch_bounds__(256) rf_up2_kernel(const RfArgs a) {
  const int W2 = 2 * a.w, H2 = 2 * a.h;
  const int wq = (W2 + 3) >> 2;
  const long long total = (long long)a.n_frames * H2 * wq;
  const bool vec = (W2 & 3) == 0;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(t % wq);
    const long long r = t / wq;
    const int dy = (int)(r % H2);
    const long long f = r / H2;
    const uint8_t* src = a.curr + (size_t)f * a.frame_stride;
    uint8_t* drow = a.up + ((size_t)f * H2 + dy) * W2;
    int y0, y1, b0, b1;
    rf_coef(dy, a.h, y0, y1, b0, b1, false);
    const uint8_t* r0 = src + (size_t)y0 * a.row_stride;
    const uint8_t* r1 = src + (size_t)y1 * a.row_stride;
    uint32_t word = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int dx = 4 * q + j;
      if (dx < W2) {
        int x0, x1, a0, a1;
        rf_coef(dx, a.w, x0, x1, a0, a1, true);
        const int h0 = r0[x0] * a0 + r0[x1] * a1;
        const int h1 = r1[x0] * a0 + r1[x1] * a1;
        const uint32_t v = (uint32_t)((((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2);
        if (vec) word |= v << (8 * j);
        else drow[dx] = (uint8_t)v;
      }
    }
    if (vec) *reinterpret_cast<uint32_t*>(drow + 4 * q) = word;
  }
}

constexpr int RF_THREADS = 1024;

__global__ void __launch_bounds__(RF_THREADS) rf_sad_kernel(const RfArgs a) {
  __shared__ unsigned int s_part[RF_THREADS / 32][9];
  __shared__ int s_off[2];
  __shared__ int s_bad;
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int W2 = 2 * a.w, H2 = 2 * a.h;
  const uint8_t* up = a.up + (size_t)f * H2 * W2;
  if (tid == 0) {
    s_off[0] = a.fullpix[2 * f];
    s_off[1] = a.fullpix[2 * f + 1];
    s_bad = 0;
  }
  __syncthreads();
  int scale = 1;
  for (int pass = 0; pass < a.passes; pass++) {
    scale *= 2;
    const int ox = 2 * s_off[0], oy = 2 * s_off[1];
    const int sx = ox < 0 ? -ox + 1 : 1, sy = oy < 0 ? -oy + 1 : 1;
    const int cw = W2 - (abs(ox) + 2), ch = H2 - (abs(oy) + 2);
    unsigned int acc[9];
#pragma unroll
    for (int k = 0; k < 9; k++) acc[k] = 0u;
    if (cw > 0 && ch > 0) {
      for (int y = warp; y < ch; y += RF_THREADS / 32) {
        const uint8_t* pa = up + (size_t)(1 + y) * W2 + 1;
        const uint8_t* pb = up + (size_t)(sy + y) * W2 + sx;
        for (int x = lane; x < cw; x += 32) {
          const int va = pa[x];
#pragma unroll
          for (int m = -1; m <= 1; m++)
#pragma unroll
            for (int n = -1; n <= 1; n++)
              acc[(m + 1) * 3 + (n + 1)] += (unsigned int)abs(va - (int)pb[(ptrdiff_t)m * W2 + x + n]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 9; k++) {
      const unsigned int s = __reduce_add_sync(0xffffffffu, acc[k]);
      if (lane == 0) s_part[warp][k] = s;
    }
    __syncthreads();
    if (tid == 0) {
      if (cw <= 0 || ch <= 0) {
        s_bad = 1;   // the reference's cv::Rect would be empty / throw: no refinement possible
      } else {
        unsigned int best = 0xffffffffu;
        int bk = 0;
        for (int k = 0; k < 9; k++) {
          unsigned int s = 0;
          for (int w = 0; w < RF_THREADS / 32; w++) s += s_part[w][k];
          if (s < best) { best = s; bk = k; }   // first minimum, row-major (cv::minMaxLoc)
        }
        s_off[0] = ox + (bk % 3) - 1;
        s_off[1] = oy + (bk / 3) - 1;
      }
    }
    __syncthreads();
    if (s_bad) break;
  }
  if (tid == 0) {
    if (s_bad) {
      a.out[2 * f] = a.out[2 * f + 1] = (double)NAN;
    } else {
      a.out[2 * f] = (double)((float)s_off[0] / (float)scale);
      a.out[2 * f + 1] = (double)((float)s_off[1] / (float)scale);
    }
  }
}

int rf_launch(const RfArgs& a, int sms, cudaStream_t st, int* launches) {
  if (a.n_frames <= 0) return 0;
  if (a.w < 2 || a.h < 2 || a.passes < 1 || a.passes > 8) return (int)cudaErrorInvalidValue;
  const long long total = (long long)a.n_frames * 2 * a.h * ((2 * a.w + 3) >> 2);
  const long long blocks = (total + 255) / 256;
  const long long cap = (long long)(sms > 0 ? sms : 148) * 8;
  rf_up2_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, st>>>(a);
  if (launches) (*launches)++;
  rf_sad_kernel<<<a.n_frames, RF_THREADS, 0, st>>>(a);
  if (launches) (*launches)++;
  return (int)cudaGetLastError();
}

}  // namespace mof
