// lp_kernels.cu -- log-polar resample of uint8 frames for sm_100a (scale / rotation path).
//
// Restates what scaleRotationEstimator::processImage asks OpenCV to do
// (src/scaleRotationEstimator.cpp:45 first frame INTER_CUBIC, :112 later frames INTER_LANCZOS4):
// cv::logPolar -> cv::warpPolar(WARP_POLAR_LOG) -> cv::remap on CV_8UC1 with float maps, i.e.
//   * source coordinates quantised to 1/32 px (INTER_BITS = 5),
//   * 4x4 / 8x8 weights from a 32x32 table of 15-bit fixed-point coefficients,
//   * int32 accumulation, (sum + 2^14) >> 15, saturate to uint8,
//   * BORDER_TRANSPARENT: destination untouched when the centre tap is outside the source, taps
//     outside the source mirrored (BORDER_REFLECT_101).
// The per-pixel fixed-point coordinates depend only on (res, M); the backend computes them once on
// the host in double precision exactly as warpPolar does and uploads them (LpMap), so the device
// work is pure integer arithmetic and bit-exact.
//
// Kernel design (everything on chip):
//   * a persistent CTA keeps the WHOLE weight table in shared memory (Lanczos4: 1024 entries x 64
//     taps = 128 KB; cubic 32 KB), each int16 weight split into a signed high byte and an unsigned
//     low byte so that four taps cost two DP4A instead of four IMAD;
//   * the source image of the current work item is staged into shared memory once, with a 4-pixel
//     BORDER_REFLECT_101 apron, so the tap loop has no border logic: a tap row is 2-3 aligned
//     32-bit shared loads funnel-shifted to the tap origin;
//   * lanes walk the tap rows in rotated order ((k + lane) mod KS), which spreads the 16-byte
//     weight reads of a warp over all banks (the accumulation is an integer sum: order-free);
//   * work item = (image, slab of output rows); large batches use one slab per image, single
//     images (stream push) are split over many CTAs.
#include <cuda_runtime.h>

#include "mof_internal.h"

namespace mof {

#if !defined(LP_THREADS_K8)
#define LP_THREADS_K8 1024
#endif
#if !defined(LP_CTAS_K4)
#define LP_CTAS_K4 2
#endif
// threads per CTA / resident CTAs per SM by kernel size.  The Lanczos4 kernel (193 KB of shared memory,
// one CTA per SM) is bound by shared-memory wavefronts and wants all 32 warps (64 registers, no spills);
// the cubic kernel needs 93 KB, so two CTAs fit (measured: 1024 threads / 2 CTAs = +13 % on scale/rot)
template <int KS> struct LpCfg { static constexpr int THREADS = 512, CTAS = LP_CTAS_K4; };
template <> struct LpCfg<8> { static constexpr int THREADS = LP_THREADS_K8, CTAS = 1; };
constexpr int LP_PAD = 4;   // apron (pixels) on every side of the staged source image

template <int KS>
__host__ __device__ constexpr size_t lp_tab_bytes() { return (size_t)1024 * KS * KS * 2; }
static size_t lp_img_bytes(int res) { return (size_t)(res + 2 * LP_PAD) * (res + 2 * LP_PAD) + 16; }

__device__ __forceinline__ int reflect101(int p, int len) {
  // one reflection suffices for |overshoot| < len (taps reach at most 4 px outside)
  if (p < 0) p = -p;
  if (p >= len) p = 2 * (len - 1) - p;
  if (p < 0) p = 0;
  return p;
}

__device__ __forceinline__ int dp4a_us(uint32_t pix, uint32_t w, int acc) {
  int r;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(pix), "r"(w), "r"(acc));
  return r;
}
__device__ __forceinline__ int dp4a_uu(uint32_t pix, uint32_t w, int acc) {
  int r;
  asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(pix), "r"(w), "r"(acc));
  return r;
}

// weight table -> shared memory, per entry and tap row: [KS high bytes][KS low bytes]; one thread
// moves one tap row (KS int16 in, 2*KS bytes out) with vector accesses
template <int KS>
__device__ void lp_load_table(unsigned char* s_tab, const int16_t* tab, int tid) {
  for (int i = tid; i < 1024 * KS; i += LpCfg<KS>::THREADS) {
    uint32_t v[KS / 2];
    if (KS == 8) {
      const uint4 t = reinterpret_cast<const uint4*>(tab)[i];
      v[0] = t.x; v[1] = t.y; v[KS / 2 - 2] = t.z; v[KS / 2 - 1] = t.w;
    } else {
      const uint2 t = reinterpret_cast<const uint2*>(tab)[i];
      v[0] = t.x; v[1] = t.y;
    }
    // v[j] = weights 2j (low half) and 2j+1 (high half); PRMT gathers the high / low bytes
    uint32_t hi[KS / 4], lo[KS / 4];
#pragma unroll
    for (int j = 0; j < KS / 4; j++) {
      hi[j] = __byte_perm(v[2 * j], v[2 * j + 1], 0x7531);
      lo[j] = __byte_perm(v[2 * j], v[2 * j + 1], 0x6420);
    }
    if (KS == 8) reinterpret_cast<uint4*>(s_tab)[i] = make_uint4(hi[0], hi[KS / 4 - 1], lo[0], lo[KS / 4 - 1]);
    else reinterpret_cast<uint2*>(s_tab)[i] = make_uint2(hi[0], lo[0]);
  }
}

// source image -> shared memory with a reflected apron; pitch = res + 2*LP_PAD bytes
__device__ void lp_load_image(unsigned char* s_img, const uint8_t* src, int row_stride, int res, int tid,
                              int nthreads) {
  const int pitch = res + 2 * LP_PAD;
  const bool al = ((((size_t)src) | (size_t)row_stride) & 3) == 0 && (res & 3) == 0;
  if (al) {
    const int wpr = pitch >> 2;   // words per padded row
    for (int i = tid; i < pitch * wpr; i += nthreads) {
      const int py = i / wpr, pw = i - py * wpr;
      const int y = reflect101(py - LP_PAD, res);
      const uint8_t* row = src + (size_t)y * row_stride;
      uint32_t v;
      if (pw >= 1 && pw <= (res >> 2)) {
        v = *reinterpret_cast<const uint32_t*>(row + 4 * (pw - 1));
      } else {
        const int x0 = 4 * pw - LP_PAD;
        v = (uint32_t)row[reflect101(x0, res)] | ((uint32_t)row[reflect101(x0 + 1, res)] << 8) |
            ((uint32_t)row[reflect101(x0 + 2, res)] << 16) | ((uint32_t)row[reflect101(x0 + 3, res)] << 24);
      }
      reinterpret_cast<uint32_t*>(s_img)[i] = v;
    }
  } else {
    for (int i = tid; i < pitch * pitch; i += nthreads) {
      const int py = i / pitch, px = i - py * pitch;
      s_img[i] = src[(size_t)reflect101(py - LP_PAD, res) * row_stride + reflect101(px - LP_PAD, res)];
    }
  }
  if (tid < 16) s_img[pitch * pitch + tid] = 0;   // the funnel shift reads one word past the last tap
}

template <int KS>
__global__ void __launch_bounds__(LpCfg<KS>::THREADS, LpCfg<KS>::CTAS) lp_kernel(const LpMap m, const LpArgs a, int slabs) {
  constexpr int NT = LpCfg<KS>::THREADS;
  extern __shared__ __align__(16) unsigned char lp_smem[];
  unsigned char* s_tab = lp_smem;
  unsigned char* s_img = lp_smem + lp_tab_bytes<KS>();
  const int tid = threadIdx.x, lane = tid & 31;
  const int res = m.res, pitch = res + 2 * LP_PAD;
  lp_load_table<KS>(s_tab, KS == 4 ? m.tab_cubic : m.tab_lanczos, tid);
  const int items = a.n_images * slabs;
  const int rows_per_slab = (res + slabs - 1) / slabs;
  int staged = -1;
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int img = item / slabs, slab = item - img * slabs;
    if (img != staged) {
      __syncthreads();   // everyone is done with the previous image (and, first time, nothing)
      lp_load_image(s_img, a.src + (size_t)img * a.src_image_stride, a.src_row_stride, res, tid, NT);
      staged = img;
    }
    __syncthreads();
    const int p_lo = slab * rows_per_slab * res;
    const int p_hi = min(res, (slab + 1) * rows_per_slab) * res;
    uint8_t* dst = a.dst + (size_t)img * a.dst_image_stride;
    const uint8_t* bg = a.bg ? a.bg + (size_t)img * a.bg_image_stride : nullptr;
    // packed map entry of pixel p: (byte offset of the centre tap in the padded image) << 10 | fxy,
    // 0xffffffff = centre outside the source; the next entry is fetched while this one is computed
    uint32_t pk_next = (p_lo + tid < p_hi) ? m.pk[p_lo + tid] : 0u;
    for (int p = p_lo + tid; p < p_hi; p += NT) {
      const uint32_t pk = pk_next;
      if (p + NT < p_hi) pk_next = m.pk[p + NT];
      const int oy = p / res, ox = p - oy * res;
      if (pk == 0xffffffffu) {   // BORDER_TRANSPARENT
        if (bg) dst[(size_t)oy * a.dst_row_stride + ox] = bg[(size_t)oy * a.bg_row_stride + ox];
        continue;
      }
      const unsigned char* wrow = s_tab + (size_t)(pk & 1023u) * (KS * KS * 2);
      const int off0 = (int)(pk >> 10) - (KS / 2 - 1) * (pitch + 1);   // top-left tap
      const unsigned char* prow = s_img + (off0 & ~3);
      const int sh = (off0 & 3) * 8;   // pitch is a multiple of 4
      int hi = 0, lo = 0;
#pragma unroll
      for (int k = 0; k < KS; k++) {
        const int r = (k + lane) & (KS - 1);
        const uint32_t* px = reinterpret_cast<const uint32_t*>(prow + r * pitch);
        if (KS == 8) {
          const uint4 w = *reinterpret_cast<const uint4*>(wrow + r * 16);
          const uint32_t c0 = px[0], c1 = px[1], c2 = px[2];
          const uint32_t p0 = __funnelshift_r(c0, c1, sh), p1 = __funnelshift_r(c1, c2, sh);
          hi = dp4a_us(p0, w.x, hi);
          hi = dp4a_us(p1, w.y, hi);
          lo = dp4a_uu(p0, w.z, lo);
          lo = dp4a_uu(p1, w.w, lo);
        } else {
          const uint2 w = *reinterpret_cast<const uint2*>(wrow + r * 8);
          const uint32_t p0 = __funnelshift_r(px[0], px[1], sh);
          hi = dp4a_us(p0, w.x, hi);
          lo = dp4a_uu(p0, w.y, lo);
        }
      }
      int v = (hi * 256 + lo + (1 << 14)) >> 15;
      v = v < 0 ? 0 : (v > 255 ? 255 : v);
      dst[(size_t)oy * a.dst_row_stride + ox] = (uint8_t)v;
    }
  }
}

static int g_lp_sms = 0;

int lp_setup(int device) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  if ((e = cudaDeviceGetAttribute(&g_lp_sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) return (int)e;
  if ((e = cudaFuncSetAttribute(lp_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(lp_tab_bytes<4>() + lp_img_bytes(240)))) != cudaSuccess) return (int)e;
  if ((e = cudaFuncSetAttribute(lp_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(lp_tab_bytes<8>() + lp_img_bytes(240)))) != cudaSuccess) return (int)e;
  return 0;
}

int lp_launch(const LpMap& map, const LpArgs& args, cudaStream_t stream) {
  if (args.n_images <= 0) return 0;
  if (map.res > 240 || map.res < 8 || (map.res & 3) || g_lp_sms <= 0) return (int)cudaErrorInvalidValue;
  // enough work items to fill the SMs: split single images into slabs of output rows
  int slabs = 1;
  if (args.n_images < g_lp_sms) {
    slabs = (g_lp_sms + args.n_images - 1) / args.n_images;
    if (slabs > map.res / 4) slabs = map.res / 4;
    if (slabs < 1) slabs = 1;
  }
  const long long items = (long long)args.n_images * slabs;
  const size_t img = lp_img_bytes(map.res);
  if (args.interp == 2) {
    const long long cap = (long long)g_lp_sms * LpCfg<4>::CTAS;
    const int grid = (int)(items < cap ? items : cap);
    lp_kernel<4><<<grid, LpCfg<4>::THREADS, lp_tab_bytes<4>() + img, stream>>>(map, args, slabs);
  } else if (args.interp == 4) {
    const int grid = (int)(items < g_lp_sms ? items : g_lp_sms);
    lp_kernel<8><<<grid, LpCfg<8>::THREADS, lp_tab_bytes<8>() + img, stream>>>(map, args, slabs);
  }
  else return (int)cudaErrorInvalidValue;
  return (int)cudaGetLastError();
}

}  // namespace mof
