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
#include <stdlib.h>

#include "mof_internal.h"

namespace mof {

// Up-scale.  The horizontal pass is exact integer arithmetic, so zeroing the fraction at the left/right
// border equals clamping the index; weights depend on the parity of the output index only.  One
// thread makes the 4 x 2 output block (columns 4q..4q+3, rows 2r+1 and 2r+2) that shares source
// columns 2q-1..2q+2 of source rows r and r+1 (clamped): 8 bytes in, 8 bytes out.
__device__ __forceinline__ uint32_t rf_vmix(int b0, int h0, int b1, int h1) {
  return (uint32_t)((((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2);
}

// one item of the up-scale: output columns 4q .. 4q + 3 of rows 2r + 1 and 2r + 2 of frame `src` into `dst` (2h x 2w bytes)
__device__ __forceinline__ void rf_up2_item(const RfArgs& a, const uint8_t* src, uint8_t* dst, int q, int r, bool vec) {
  const int W2 = 2 * a.w, H2 = 2 * a.h;
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
  uint8_t* base = dst + 4 * q;
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
    rf_up2_item(a, a.curr + (size_t)f * a.frame_stride, a.up + (size_t)f * H2 * W2, q, r, vec);
  }
}

constexpr int RF_THREADS = 1024;
#if !defined(RF_STRIPS)
#define RF_STRIPS 1   // fused variant: up-scale by column strips (0: one 4 x 2 block per item as in rf_up2_kernel)
#endif
#if !defined(RF_ALIGNED_ROWS)
#define RF_ALIGNED_ROWS 1   // 0: the per-group path for every width (A/B)
#endif

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

// ---- word-aligned rows (W2 % 4 == 0: every row of the 2x image starts on a word) ----------------------------------
// The nine candidates of a pass are three rows (m) x three byte offsets (n) of the SAME image.  Walking the candidate
// rows by their aligned words needs no shuffling on that side at all; the cut-out row is seen through three views, one
// per n, that are shared by the three candidate rows: per word 3 + 3 loads, 3 funnel shifts, 9 VABSDIFF4.ACC (the first
// version built nine shifted candidate views per group: 11 loads and 10 shifts).  R = (-1 - bx0) mod 4 fixes the three
// shifts at compile time.  Bytes outside the cut-out are masked on the first / last words only.
template <int R>
__device__ __forceinline__ unsigned rf_view(unsigned a0, unsigned a1, unsigned a2, int n) {   // bytes R + 2 - n ..
  const int o = R + 2 - n;
  return o < 4 ? (o == 0 ? a0 : __funnelshift_r(a0, a1, o * 8)) : (o == 4 ? a1 : __funnelshift_r(a1, a2, (o - 4) * 8));
}
template <int R, int KW>
__device__ __forceinline__ void rf_rows_aligned(const unsigned* upw, int WW, int y0, int ystep, int ch, int sy, int bx0, int cw,
                                                int lane, unsigned (&acc)[9]) {
  const int dw = (-1 - bx0 - R) >> 2;                  // word of the lowest cut-out byte a word needs, relative to w
  const int w_first = bx0 >> 2, w_last = (bx0 + cw + 1) >> 2;
  const int wf_lo = (bx0 + 2 + 3) >> 2, wf_hi = ((bx0 + cw) >> 2) - 1;   // words whose four bytes count for every n
  const int rstep = ystep * WW;
  // a lane owns KW words (wb, wb + 32, ...) of every row of its warp: which of them exist and which are edge words
  // is the same for all rows; the four words of a row are loaded before any of its SADs
  for (int wb = w_first + lane; wb <= w_last; wb += 32 * KW) {
    bool ok[KW], edge[KW];
#pragma unroll
    for (int k = 0; k < KW; k++) {
      const int w = wb + 32 * k;
      ok[k] = w <= w_last;
      edge[k] = w < wf_lo || w > wf_hi;
    }
    const unsigned* ap = upw + (1 + y0) * WW + dw + wb;
    const unsigned* bp0 = upw + (sy - 1 + y0) * WW + wb;
    const unsigned* bp1 = bp0 + WW;
    const unsigned* bp2 = bp1 + WW;
    for (int y = y0; y < ch; y += ystep, ap += rstep, bp0 += rstep, bp1 += rstep, bp2 += rstep) {
      unsigned a0[KW], a1[KW], a2[KW], b[KW][3];
#pragma unroll
      for (int k = 0; k < KW; k++) {
        a0[k] = ok[k] ? ap[32 * k] : 0u;
        a1[k] = ok[k] ? ap[32 * k + 1] : 0u;
        a2[k] = (R >= 3 && ok[k]) ? ap[32 * k + 2] : 0u;
        b[k][0] = ok[k] ? bp0[32 * k] : 0u;
        b[k][1] = ok[k] ? bp1[32 * k] : 0u;
        b[k][2] = ok[k] ? bp2[32 * k] : 0u;
      }
#pragma unroll
      for (int k = 0; k < KW; k++) {
        if (!ok[k]) continue;
        unsigned v[3];
#pragma unroll
        for (int n = 0; n < 3; n++) v[n] = rf_view<R>(a0[k], a1[k], a2[k], n);
        if (!edge[k]) {
#pragma unroll
          for (int m = 0; m < 3; m++)
#pragma unroll
            for (int n = 0; n < 3; n++) acc[m * 3 + n] = rf_sad4(v[n], b[k][m], acc[m * 3 + n]);
        } else {
          const int w = wb + 32 * k;
#pragma unroll
          for (int n = 0; n < 3; n++) {
            // candidate n covers bytes [bx0 + n, bx0 + n + cw) of its rows
            int jlo = bx0 + n - 4 * w, jhi = bx0 + n + cw - 4 * w;
            jlo = jlo < 0 ? 0 : jlo;
            jhi = jhi > 4 ? 4 : jhi;
            if (jlo < jhi) {
              const unsigned mask = (jhi == 4 ? 0xffffffffu : ((1u << (8 * jhi)) - 1u)) & ~((1u << (8 * jlo)) - 1u);
              const unsigned vn = v[n] & mask;
#pragma unroll
              for (int m = 0; m < 3; m++) acc[m * 3 + n] = rf_sad4(vn, b[k][m] & mask, acc[m * 3 + n]);
            }
          }
        }
      }
    }
  }
}

// One CTA per frame; both passes.  A warp walks rows of the cut-out, a lane takes 4 pixels: the base
// word and, per candidate row, the 6 bytes around it come from aligned words (the byte offsets are
// the same for every group of a row), nine VABSDIFF4.ACC per group.  Word reads may run up to 8
// bytes past the last row: the scratch image is allocated with padding.
// FUSED (frames whose 2x image fits beside the kernel's 1.2 KB of scratch in the 227 KB of an SM, i.e. up to 240 x 240,
// and word-aligned rows): the CTA up-scales its frame straight into shared memory and runs both SAD passes there --
// the 2x image (4 F^2 bytes per frame) never travels to HBM and back, the passes read at shared-memory latency.
template <bool FUSED>
__global__ void __launch_bounds__(RF_THREADS) rf_sad_kernel(const RfArgs a) {
  extern __shared__ __align__(16) unsigned char rf_smem[];
  __shared__ unsigned int s_part[RF_THREADS / 32][9];
  __shared__ int s_off[2];
  __shared__ int s_bad;
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int W2 = 2 * a.w, H2 = 2 * a.h;
  const uint8_t* up = FUSED ? rf_smem : a.up + (size_t)f * H2 * W2;
  if (FUSED) {
    const int wq = W2 >> 2;
    const int per_frame = (a.h + 1) * wq;
    const uint8_t* src = a.curr + (size_t)f * a.frame_stride;
    if (RF_STRIPS && wq <= RF_THREADS) {
      // column strips: a thread owns output columns 4q .. 4q + 3 over a slab of source rows and walks down, so every
      // source row's horizontal pass is made once; with the weights 512 and 1536 = 3 * 512 the vertical mix
      // ((b * (H >> 4)) >> 16 summed, + 2, >> 2) is (3 g >> 7) + (g' >> 7) with g = H >> 4: adds and shifts only
      const int slabs = RF_THREADS / wq, rows = (a.h + 1 + slabs - 1) / slabs;
      const int slab = tid / wq, q = tid - slab * wq;
      if (slab < slabs) {
        int c[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int x = 2 * q - 1 + k;
          c[k] = x < 0 ? 0 : (x > a.w - 1 ? a.w - 1 : x);
        }
        auto horiz = [&](int row, int (&t3)[4], int (&t1)[4]) {   // (3 g) >> 7 and g >> 7 of the row's four columns
          const uint8_t* rp = src + (size_t)(row < 0 ? 0 : (row > a.h - 1 ? a.h - 1 : row)) * a.row_stride;
          const int p0 = rp[c[0]], p1 = rp[c[1]], p2 = rp[c[2]], p3 = rp[c[3]];
          int g[4];
          g[0] = (p0 * 512 + p1 * 1536) >> 4; g[1] = (p1 * 1536 + p2 * 512) >> 4;
          g[2] = (p1 * 512 + p2 * 1536) >> 4; g[3] = (p2 * 1536 + p3 * 512) >> 4;
#pragma unroll
          for (int j = 0; j < 4; j++) { t3[j] = (3 * g[j]) >> 7; t1[j] = g[j] >> 7; }
        };
        const int r_lo = slab * rows - 1, r_hi = min(a.h - 1, r_lo + rows - 1);   // items r = r_lo .. r_hi
        int u3[4], u1[4], d3[4], d1[4];
        horiz(r_lo, u3, u1);
        unsigned char* base = rf_smem + 4 * q;
        for (int r = r_lo; r <= r_hi; r++) {
          horiz(r + 1, d3, d1);
          uint32_t wo = 0, we = 0;   // row 2r + 1: weights (1536, 512); row 2r + 2: (512, 1536)
#pragma unroll
          for (int j = 0; j < 4; j++) {
            wo |= (uint32_t)((u3[j] + d1[j] + 2) >> 2) << (8 * j);
            we |= (uint32_t)((u1[j] + d3[j] + 2) >> 2) << (8 * j);
            u3[j] = d3[j];
            u1[j] = d1[j];
          }
          if (r >= 0) *reinterpret_cast<uint32_t*>(base + (size_t)(2 * r + 1) * W2) = wo;
          if (2 * r + 2 < H2) *reinterpret_cast<uint32_t*>(base + (size_t)(2 * r + 2) * W2) = we;
        }
      }
    } else
    for (int t = tid; t < per_frame; t += RF_THREADS) {
      const int r = t / wq - 1, q = t - (r + 1) * wq;
      rf_up2_item(a, src, rf_smem, q, r, true);
    }
    if (tid < 4) reinterpret_cast<unsigned*>(rf_smem + (size_t)H2 * W2)[tid] = 0u;   // word reads may run past the last row
  }
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
    if (cw > 0 && ch > 0 && (W2 & 3) == 0 && RF_ALIGNED_ROWS) {
      const unsigned* upw = reinterpret_cast<const unsigned*>(up);
      const int bx0 = sx - 1;
      switch ((-1 - bx0) & 3) {
        case 0: rf_rows_aligned<0, FUSED ? 2 : 3>(upw, W2 >> 2, warp, RF_THREADS / 32, ch, sy, bx0, cw, lane, acc); break;
        case 1: rf_rows_aligned<1, FUSED ? 2 : 3>(upw, W2 >> 2, warp, RF_THREADS / 32, ch, sy, bx0, cw, lane, acc); break;
        case 2: rf_rows_aligned<2, FUSED ? 2 : 3>(upw, W2 >> 2, warp, RF_THREADS / 32, ch, sy, bx0, cw, lane, acc); break;
        default: rf_rows_aligned<3, FUSED ? 2 : 3>(upw, W2 >> 2, warp, RF_THREADS / 32, ch, sy, bx0, cw, lane, acc); break;
      }
    } else if (cw > 0 && ch > 0) {
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
    if (warp == 0) {   // lane k < 9 adds candidate k's partial sums of the 32 warps; first minimum, row-major (cv::minMaxLoc)
      unsigned int sum = 0xffffffffu;
      if (lane < 9) {
        sum = 0;
#pragma unroll 8
        for (int w = 0; w < RF_THREADS / 32; w++) sum += s_part[w][lane];
      }
      const unsigned int best = __reduce_min_sync(0xffffffffu, sum);
      const int bk = __ffs(__ballot_sync(0xffffffffu, sum == best && lane < 9)) - 1;
      if (lane == 0) {
        if (cw <= 0 || ch <= 0) {
          s_bad = 1;   // the reference's cv::Rect would be empty / throw: no refinement possible
        } else {
          s_off[0] = ox + (bk % 3) - 1;
          s_off[1] = oy + (bk / 3) - 1;
        }
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

// dynamic shared memory of the fused variant: everything an SM has minus the kernel's static scratch (1.2 KB)
constexpr size_t RF_FUSED_MAX_BYTES = (size_t)227 * 1024 - 1536;

int rf_setup(int device) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  return (int)cudaFuncSetAttribute(rf_sad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RF_FUSED_MAX_BYTES);
}

// true when rf_launch takes the fused variant for this frame size (no 2x scratch in global memory needed)
bool rf_fused(int w, int h) {
  return RF_ALIGNED_ROWS && ((2 * w) & 3) == 0 && (size_t)4 * w * h + 16 <= RF_FUSED_MAX_BYTES && !getenv("MOF_RF_NO_FUSED");
}

int rf_launch(const RfArgs& a, int sms, cudaStream_t st, int* launches) {
  if (a.n_frames <= 0) return 0;
  if (a.w < 2 || a.h < 2 || a.passes < 1 || a.passes > 8) return (int)cudaErrorInvalidValue;
  (void)sms;
  const long long per_frame = (long long)(a.h + 1) * ((2 * a.w + 3) >> 2);
  if (per_frame > 0x7fffffffLL) return (int)cudaErrorInvalidValue;
  const int bx = (int)((per_frame + 255) / 256);
  // fused variant: the whole 2x image of a frame in shared memory (RF_NO_FUSED=1: A/B, tests)
  const size_t fused_bytes = (size_t)4 * a.w * a.h + 16;
  const bool fused = rf_fused(a.w, a.h);
  for (int f0 = 0; f0 < a.n_frames; f0 += 65535) {   // gridDim.y limit
    RfArgs b = a;
    b.n_frames = a.n_frames - f0 < 65535 ? a.n_frames - f0 : 65535;
    b.curr = a.curr + (size_t)f0 * a.frame_stride;
    b.fullpix = a.fullpix + 2 * (size_t)f0;
    b.out = a.out + 2 * (size_t)f0;
    if (fused) {
      rf_sad_kernel<true><<<b.n_frames, RF_THREADS, fused_bytes, st>>>(b);
      if (launches) (*launches)++;
      continue;
    }
    // the 2x scratch of a chunk is reused by the next one (same stream: ordered)
    rf_up2_kernel<<<dim3(bx, b.n_frames), 256, 0, st>>>(b);
    if (launches) (*launches)++;
    rf_sad_kernel<false><<<b.n_frames, RF_THREADS, 0, st>>>(b);
    if (launches) (*launches)++;
  }
  return (int)cudaGetLastError();
}

}  // namespace mof
