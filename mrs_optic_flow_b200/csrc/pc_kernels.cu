// pc_kernels.cu -- batched phase-correlation kernel for sm_100a.
//
// One persistent CTA per resident slot; each CTA walks sections (frame pair x section) with a grid
// stride.  A section is transformed entirely inside shared memory (see pc_pipeline.cuh for the
// layout and the work items); the only global traffic is the uint8 pixels in, the previous image's
// half spectrum through an L2-resident scratch (N*N/2 complex floats: the two spectra of an
// N = 240 section do not fit one SM's 227 KB together), and 16 bytes of result out.
//
// Replaces: the per-section loop of FftMethod::processImage (src/FftMethod.cpp:1829-1856) and the
// OpenCL kernel phaseCorrelateField (cl/FftMethod.cl:1381-1485) it would otherwise dispatch to.
#include <cuda_runtime.h>
#include <limits.h>

#include "mof_internal.h"
#include "pc_pipeline.cuh"

namespace mof {

template <int P>
struct KCfg;
// THREADS is a multiple of Geo<P>::LD so that a thread keeps the same lane index l (= column k or
// row-pair q) in every phase and the butterfly index b stays warp-uniform (twiddles come from
// constant memory with a uniform address).
template <> struct KCfg<16> { static constexpr int THREADS = 512, MINB = 1; };
template <> struct KCfg<8>  { static constexpr int THREADS = 256, MINB = 2; };
template <> struct KCfg<4>  { static constexpr int THREADS = 128, MINB = 4; };

template <int P>
constexpr size_t pc_smem_bytes() {
  return sizeof(c32) * Geo<P>::A_ELEMS + sizeof(float) * (KCfg<P>::THREADS / 32) + 16;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}

template <int P>
__global__ void __launch_bounds__(KCfg<P>::THREADS, KCfg<P>::MINB) pc_kernel(const PcArgs a) {
  typedef Geo<P> G;
  constexpr int T = KCfg<P>::THREADS;
  constexpr int NB = P / 4;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  c32* A = reinterpret_cast<c32*>(smem_raw);
  float* s_wmax = reinterpret_cast<float*>(A + G::A_ELEMS);   // per-warp maxima
  int* s_peak = reinterpret_cast<int*>(s_wmax + T / 32);      // [0] first location of the maximum
  const int tid = threadIdx.x;
  const size_t spec_elems = (size_t)G::N * G::M;

  for (int sec = blockIdx.x; sec < a.n_sections; sec += gridDim.x) {
    const int f = sec / a.S, s = sec - f * a.S;
    const int j = s / a.sq, i = s - j * a.sq;
    const size_t off = (size_t)f * a.frame_stride + (size_t)(j * G::N) * a.row_stride + i * G::N;

    SecCtx<P> c;
    c.A = A;
    c.row_stride = a.row_stride;
    c.win = a.win;
    c.vmax = -INFINITY;
    c.spec_prev = nullptr;

    c32* spec;  // spectrum of the previous image of this section
    bool prev_pass;
    const uint8_t* prev_img;
    if (a.mode == PC_PAIRS) {
      spec = reinterpret_cast<c32*>(a.scratch) + (size_t)blockIdx.x * spec_elems;
      prev_pass = true;
      prev_img = a.prev + off;
    } else {
      spec = reinterpret_cast<c32*>(a.spec_cache) + ((size_t)a.stream_ids[f] * a.S + s) * spec_elems;
      // first call of a stream: imPrev := imCurr (src/FftMethod.cpp:1791-1793)
      prev_pass = (a.mode == PC_SPECTRUM) || (a.first[f] != 0);
      prev_img = a.curr + off;
    }
    c.spec_out = spec;

#pragma unroll 1
    for (int pass = prev_pass ? 0 : 1; pass < 2; pass++) {
      c.img = pass ? a.curr + off : prev_img;
      c.vec = (((size_t)c.img | (size_t)a.row_stride) & (NB - 1)) == 0;
      run_phase<P, PH_1A>(c, tid, T);
      __syncthreads();
      run_phase<P, PH_1B>(c, tid, T);
      __syncthreads();
      run_phase<P, PH_2>(c, tid, T);
      __syncthreads();
      if (pass == 0) {
        run_phase<P, PH_3S>(c, tid, T);
        __syncthreads();  // spectrum visible to the whole CTA; A free again
        if (a.mode == PC_SPECTRUM) break;
      }
    }
    if (a.mode == PC_SPECTRUM) continue;

    c.spec_prev = spec;
    c.spec_out = (a.mode == PC_STREAM) ? spec : nullptr;  // imPrev = imCurr.clone() (:1872)
    run_phase<P, PH_3C>(c, tid, T);
    __syncthreads();
    run_phase<P, PH_3Z>(c, tid, T);
    __syncthreads();
    run_phase<P, PH_3I>(c, tid, T);
    __syncthreads();
    run_phase<P, PH_4>(c, tid, T);
    __syncthreads();
    run_phase<P, PH_5A>(c, tid, T);
    __syncthreads();
    run_phase<P, PH_5B>(c, tid, T);

    // block maximum, then its first location in fft-shifted row-major order (cv::minMaxLoc)
    const float wm = warp_max(c.vmax);
    if ((tid & 31) == 0) s_wmax[tid >> 5] = wm;
    if (tid == 0) s_peak[0] = 0x7fffffff;
    __syncthreads();  // also: the surface in A is complete
    float bm = s_wmax[tid & (T / 32 - 1)];
    bm = warp_max(bm);
    int loc = run_search<P>(A, bm, tid, T);
    loc = __reduce_min_sync(0xffffffffu, loc);
    if ((tid & 31) == 0 && loc != 0x7fffffff) atomicMin(&s_peak[0], loc);
    __syncthreads();
    if (tid == 0) {
      const int pidx = s_peak[0];
      double xy[2];
      if (a.post == PC_POST_SHIFT) {
        centroid_shift<P>(A, pidx, a.max_px_speed_sq, xy);
      } else {
        // scaleRotationEstimator::processImage tail (src/scaleRotationEstimator.cpp:117-126):
        // pt = phaseCorrelate(cur, prev) = -(shift); reject |pt.x| > res/2 -> (1, 0)
        centroid_shift<P>(A, pidx, 1e300, xy);
        const double ptx = -xy[0], pty = -xy[1];
        if (fabs(ptx) > (double)(G::N / 2) || ptx != ptx || pty != pty) {
          xy[0] = a.sr_reject_nan ? (double)NAN : 1.0;
          xy[1] = a.sr_reject_nan ? (double)NAN : 0.0;
        } else {
          const double Ky = (double)G::N / 360.0;
          xy[0] = exp(ptx / a.sr_m);
          xy[1] = (pty / Ky) * (3.14159265358979323846 / 180.0);
        }
      }
      a.out[(size_t)sec * 2 + 0] = xy[0];
      a.out[(size_t)sec * 2 + 1] = xy[1];
    }
    __syncthreads();  // centroid done before the next section overwrites A
  }
}

// -------------------------------------------------------------------------------------------------
// host side
// -------------------------------------------------------------------------------------------------
template <int P>
static cudaError_t setup_one() {
  return cudaFuncSetAttribute(pc_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)pc_smem_bytes<P>());
}

int pc_setup(int device) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  fill_twiddles_host();
  e = cudaMemcpyToSymbol(c_tw240, h_tw240, sizeof(h_tw240));
  if (e != cudaSuccess) return (int)e;
  if ((e = setup_one<4>()) != cudaSuccess) return (int)e;
  if ((e = setup_one<8>()) != cudaSuccess) return (int)e;
  if ((e = setup_one<16>()) != cudaSuccess) return (int)e;
  return 0;
}

template <int P>
static int resident_one(int device) {
  int per_sm = 0, sms = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pc_kernel<P>, KCfg<P>::THREADS,
                                                    pc_smem_bytes<P>()) != cudaSuccess)
    return -1;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return -1;
  return per_sm * sms;
}

int pc_resident_ctas(int N, int device) {
  switch (N) {
    case 60: return resident_one<4>(device);
    case 120: return resident_one<8>(device);
    case 240: return resident_one<16>(device);
  }
  return -1;
}

size_t pc_spectrum_elems(int N) { return (size_t)N * (size_t)(N / 2); }

int pc_launch(int N, const PcArgs& args, int grid, cudaStream_t stream) {
  if (grid < 1) return (int)cudaErrorInvalidValue;
  switch (N) {
    case 60:
      pc_kernel<4><<<grid, KCfg<4>::THREADS, pc_smem_bytes<4>(), stream>>>(args);
      break;
    case 120:
      pc_kernel<8><<<grid, KCfg<8>::THREADS, pc_smem_bytes<8>(), stream>>>(args);
      break;
    case 240:
      pc_kernel<16><<<grid, KCfg<16>::THREADS, pc_smem_bytes<16>(), stream>>>(args);
      break;
    default:
      return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}

}  // namespace mof
