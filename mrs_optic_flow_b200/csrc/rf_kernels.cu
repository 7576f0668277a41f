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

// Up-scale.  The horizontal pass is exact integer arithmetic, so zeroing the fraction at the left/right
// border equals clamping the index; weights depend on the parity of the output index only.  One
// thread makes the 4 x 2 output block (columns 4q..4q+3, rows 2r+1 and 2r+2) that shares source
// columns 2q-1..2q+2 of source rows r and r+1 (clamped): 8 bytes in, 8 bytes out.
__device__ __forceinline__ uint32_t rf_vmix(int b0, int h0, int b1, int h1) {
  return (uint32_t)((((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2);
}

__global__ void __launch_bounds__(256) rf_up2_kernel(const RfArgs a) {
  const int W2 = 2 * a.w, H2 = 2 * a.h;
  const int wq = (W2 + 3) >> 2;
  const int hr = a.h + 1;   // r = -1 .. h-1
  const unsigned per_frame = (unsigned)hr * (unsigned)wq;   // 32-bit index math: frames ride on blockIdx.y
  const bool vec = (W2 & 3) == 0;
  const size_t f = blockIdx.y;
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < per_frame; t += gridDim.x * blockDim.x) {
    const int q = (int)(t % (unsigned)wq);
    const int r = (int)(t / (unsigned)wq) - 1;
    const uint8_t* src = a.curr + (size_t)f * a.frame_stride;
    const uint8_t* r0 = src + (size_t)(r < 0 ? 0 : r) * a.row_stride;
    const uint8_t* r1 = src + (size_t)(r + 1 > a.h - 1 ? a.h - 1 : r + 1) * a.row_stride;
    int c[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int x = 2 * q - 1 + k;
      c[k] = x < 0 ? 0 : (x > a.w - 1 ? a.w - 1 : x);
    }
    int h0[4], h1[4];   // horizontal pass of output columns 4q .. 4q+3 on the two source rows
    {
      const int p0 = r0[c[0]], p1 = r0[c[1]], p2 = r0[c[2]], p3 = r0[c[3]];
      h0[0] = p0 * 512 + p1 * 1536; h0[1] = p1 * 1536 + p2 * 512;
      h0[2] = p1 * 512 + p2 * 1536; h0[3] = p2 * 1536 + p3 * 512;
      const int s0 = r1[c[0]], s1 = r1[c[1]], s2 = r1[c[2]], s3 = r1[c[3]];
      h1[0] = s0 * 512 + s1 * 1536; h1[1] = s1 * 1536 + s2 * 512;
      h1[2] = s1 * 512 + s2 * 1536; h1[3] = s2 * 1536 + s3 * 512;
    }
    uint32_t wo = 0, we = 0;   // odd row 2r+1: weights (1536, 512); even row 2r+2: (512, 1536)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      wo |= rf_vmix(1536, h0[j], 512, h1[j]) << (8 * j);
      we |= rf_vmix(512, h0[j], 1536, h1[j]) << (8 * j);
    }
    uint8_t* base = a.up + (size_t)f * H2 * W2 + 4 * q;
    const int dyo = 2 * r + 1, dye = 2 * r + 2;
    if (vec) {
      if (dyo >= 0) *reinterpret_cast<uint32_t*>(base + (size_t)dyo * W2) = wo;
      if (dye < H2) *reinterpret_cast<uint32_t*>(base + (size_t)dye * W2) = we;
    } else {
      for (int j = 0; j < 4 && 4 * q + j < W2; j++) {
        if (dyo >= 0) base[(size_t)dyo * W2 + j] = (uint8_t)(wo >> (8 * j));
        if (dye < H2) base[(size_t)dye * W2 + j] = (uint8_t)(we >> (8 * j));
      }
    }
  }
}

constexpr int RF_THREADS = 1024;

// sum of absolute byte differences added to acc in one instruction (VABSDIFF4.U8.ACC)
__device__ __forceinline__ unsigned rf_sad4(unsigned a, unsigned b, unsigned acc) {
  unsigned r;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(acc));
  return r;
}
// 4 bytes at byte offset o (0..5) of the 12 bytes w0 w1 w2
__device__ __forceinline__ unsigned rf_bytes(unsigned w0, unsigned w1, unsigned w2, int o) {
  const unsigned lo = o < 4 ? w0 : w1, hi = o < 4 ? w1 : w2;
  return __funnelshift_r(lo, hi, (o & 3) * 8);
}

// One CTA per frame; both passes.  A warp walks rows of the cut-out, a lane takes 4 pixels: the base
// word and, per candidate row, the 6 bytes around it come from aligned words (the byte offsets are
// the same for every group of a row), nine VABSDIFF4.ACC per group.  Word reads may run up to 8
// bytes past the last row: the scratch image is allocated with padding.
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
      const int groups = (cw + 3) >> 2;
      for (int y = warp; y < ch; y += RF_THREADS / 32) {
        const uint8_t* pa = up + (size_t)(1 + y) * W2 + 1;
        const int oa = (int)((size_t)pa & 3);
        const unsigned* wa = reinterpret_cast<const unsigned*>(pa - oa);
        const unsigned* wb[3];
        int ob[3];
#pragma unroll
        for (int m = 0; m < 3; m++) {
          const uint8_t* pb = up + (size_t)(sy + m - 1 + y) * W2 + sx - 1;   // byte x-1 of candidate row m
          ob[m] = (int)((size_t)pb & 3);
          wb[m] = reinterpret_cast<const unsigned*>(pb - ob[m]);
        }
        for (int g = lane; g < groups; g += 32) {
          const int rem = cw - 4 * g;
          const unsigned mask = rem >= 4 ? 0xffffffffu : ((1u << (8 * rem)) - 1u);
          const unsigned va = __funnelshift_r(wa[g], wa[g + 1], oa * 8) & mask;
#pragma unroll
          for (int m = 0; m < 3; m++) {
            const unsigned w0 = wb[m][g], w1 = wb[m][g + 1], w2 = wb[m][g + 2];
#pragma unroll
            for (int n = 0; n < 3; n++)
              acc[m * 3 + n] = rf_sad4(va, rf_bytes(w0, w1, w2, ob[m] + n) & mask, acc[m * 3 + n]);
          }
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
  (void)sms;
  const long long per_frame = (long long)(a.h + 1) * ((2 * a.w + 3) >> 2);
  if (per_frame > 0x7fffffffLL) return (int)cudaErrorInvalidValue;
  const int bx = (int)((per_frame + 255) / 256);
  for (int f0 = 0; f0 < a.n_frames; f0 += 65535) {   // gridDim.y limit
    RfArgs b = a;
    b.n_frames = a.n_frames - f0 < 65535 ? a.n_frames - f0 : 65535;
    b.curr = a.curr + (size_t)f0 * a.frame_stride;
    b.fullpix = a.fullpix + 2 * (size_t)f0;
    b.out = a.out + 2 * (size_t)f0;
    // the 2x scratch of a chunk is reused by the next one (same stream: ordered)
    rf_up2_kernel<<<dim3(bx, b.n_frames), 256, 0, st>>>(b);
    if (launches) (*launches)++;
    rf_sad_kernel<<<b.n_frames, RF_THREADS, 0, st>>>(b);
    if (launches) (*launches)++;
  }
  return (int)cudaGetLastError();
}

}  // namespace mof
