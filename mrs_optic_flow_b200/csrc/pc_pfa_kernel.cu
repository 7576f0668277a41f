// pc_pfa_kernel.cu -- batched phase-correlation kernel for sm_100a (twiddle-free prime-factor FFTs).
//
// One persistent CTA per resident slot walks sections (frame pair x section) with a grid stride.
// A CTA is one team of TASKS "main" warps plus one "aux" warp:
//   * row stages (FX15 -> FX16, IX15 -> IX16): main warp w owns RPT row pairs; the two stages of a
//     row transform chain with __syncwarp only;
//   * column stages (FY16 -> FY15 x cross-power x IY15 -> IY16): main warp w owns UPT column units,
//     again chained with __syncwarp only; the aux warp runs the packed unit (columns 0 and N/2),
//     whose cross-power needs bins v and -v together, concurrently with the main warps;
//   * block barriers remain only where ownership changes (rows <-> columns) and around the peak
//     search: 6 per frame pair.
// The section lives in shared memory for its whole life (231 KB at N = 240 -> one CTA per SM); the
// pixels of the NEXT image are prefetched into registers (30 words per thread = the whole section)
// one stage ahead, so DRAM latency is hidden behind the FFT stages.  The only other global traffic
// is the previous image's spectrum through an L2-resident per-CTA scratch (two spectra of a 240 px
// section do not fit one SM) and 16 bytes of result.
//
// Replaces: the per-section loop of FftMethod::processImage (src/FftMethod.cpp:1829-1856) and the
// OpenCL kernel phaseCorrelateField (cl/FftMethod.cl:1381-1485) it would otherwise dispatch to.
#include <cuda_runtime.h>
#include <limits.h>

#include "mof_internal.h"
#include "pc_pfa.cuh"

namespace mof {

using namespace pfa;

template <int P>
struct KCfg;
template <> struct KCfg<16> { static constexpr int MINB = 1; };
template <> struct KCfg<8>  { static constexpr int MINB = 3; };
template <> struct KCfg<4>  { static constexpr int MINB = 8; };

template <int P>
constexpr int pc_threads() { return (PG<P>::TASKS + 1) * 32; }
template <int P>
constexpr size_t pc_smem_bytes() {
  return sizeof(c32) * PG<P>::A_ELEMS + sizeof(float) * (PG<P>::TASKS + 1) + 16 + sizeof(uint16_t) * PG<P>::N;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

// what one section needs: its images and where its spectra live
struct SecDesc {
  const uint8_t* img0;   // first image to transform
  const uint8_t* img1;   // second image (correlate pass); == img0 for one-pass sections
  f4* spec0;             // pass 0 writes the spectrum here / pass 1 correlates against it
  f4* spec_out;          // pass 1 keeps the current spectrum here (streams) or nullptr
  bool two_pass;         // pass 0 (spectrum of img0) runs
  bool correlate;        // pass 1 runs
};

template <int P>
__device__ __forceinline__ SecDesc describe(const PcArgs& a, int sec) {
  typedef PG<P> G;
  const size_t spec_f4 = G::SPEC_F4;
  const int f = sec / a.S, s = sec - f * a.S;
  const int j = s / a.sq, i = s - j * a.sq;
  const size_t off = (size_t)f * a.frame_stride + (size_t)(j * G::N) * a.row_stride + i * G::N;
  SecDesc d;
  if (a.mode == PC_PAIRS) {
    d.img0 = a.prev + off;
    d.img1 = a.curr + off;
    d.spec0 = reinterpret_cast<f4*>(a.scratch) + (size_t)blockIdx.x * spec_f4;
    d.spec_out = nullptr;
    d.two_pass = true;
    d.correlate = true;
  } else {
    // stream cache: two halves per (stream, section); bit 1 of first[] = half holding the previous
    // frame's spectrum, bit 0 = first call of the stream (imPrev := imCurr, src/FftMethod.cpp:1791-1793)
    const unsigned fl = a.first[f];
    const unsigned par = (fl >> 1) & 1u;
    f4* base = reinterpret_cast<f4*>(a.spec_cache) + ((size_t)a.stream_ids[f] * a.S + s) * 2 * spec_f4;
    d.img0 = a.curr + off;
    d.img1 = d.img0;
    d.spec0 = base + par * spec_f4;
    d.spec_out = base + (par ^ 1u) * spec_f4;   // imPrev = imCurr.clone() (:1872)
    d.two_pass = (a.mode == PC_SPECTRUM) || (fl & 1u);
    d.correlate = (a.mode != PC_SPECTRUM);
  }
  return d;
}

template <int P, bool VEC>
__global__ void __launch_bounds__(pc_threads<P>(), KCfg<P>::MINB) pc_pfa_kernel(const PcArgs a) {
  typedef PG<P> G;
  constexpr int TASKS = G::TASKS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  c32* A = reinterpret_cast<c32*>(smem_raw);
  float* s_wmax = reinterpret_cast<float*>(A + G::A_ELEMS);   // per-main-warp maxima
  int* s_peak = reinterpret_cast<int*>(s_wmax + TASKS + 1);   // first location of the maximum
  uint16_t* s_rowy = reinterpret_cast<uint16_t*>(s_peak + 4);  // image row of (row pair, half)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool aux = warp == TASKS;
  const int task = warp;
  // column units of this warp: main warps take the regular units of their task, aux the packed one
  const int u0 = aux ? G::M - 1 : task * G::UPT;
  const int nu = aux ? 1 : ((task == TASKS - 1) ? G::UPT - 1 : G::UPT);

  for (int i = tid; i < G::N; i += (int)blockDim.x) fill_rowy<P>(s_rowy, i);
  if (tid == 0) *s_peak = 0x7fffffff;
  __syncthreads();

  // pixels of the next image to transform, parked in registers one stage ahead (VEC only)
  uint32_t pv[LoadCfg<P>::NREG];
#pragma unroll
  for (int i = 0; i < LoadCfg<P>::NREG; i++) pv[i] = 0;
  unsigned char* stage = reinterpret_cast<unsigned char*>(A + task * G::RPT * G::SQ);

  int sec = blockIdx.x;
  if (VEC && sec < a.n_sections && !aux)
    fx_prefetch<P>(describe<P>(a, sec).img0, a.row_stride, s_rowy, task, lane, pv);

  for (; sec < a.n_sections; sec += gridDim.x) {
    const SecDesc d = describe<P>(a, sec);
    const int nsec = sec + gridDim.x;

#pragma unroll 1
    for (int pass = d.two_pass ? 0 : 1; pass < 2; pass++) {
      // pass 0: spectrum of the first image -> spec0; pass 1: current image, correlate, invert columns
      if (!aux) {
        // rows: pixels -> staging (inside this warp's own rows of A) -> words -> FX15 -> FX16
        if (VEC) fx_stage<P>(stage, lane, pv);
        else fx_stage_bytes<P>(stage, pass ? d.img1 : d.img0, a.row_stride, s_rowy, task, lane);
        __syncwarp();
        uint32_t wa[15], wb[15];
        stage_fx15_read<P>(A, task, lane, wa, wb);
        __syncwarp();
        stage_fx15<P>(A, task, lane, wa, wb, a.win);
        __syncwarp();
        stage_x16<P, 1, false>(A, task, lane, nullptr);
      }
      __syncthreads();
      // prefetch the next image one stage ahead: the second image of this section, or (spectrum
      // only) the first image of this CTA's next section
      if (VEC && pass == 0 && !aux) {
        if (d.correlate) fx_prefetch<P>(d.img1, a.row_stride, s_rowy, task, lane, pv);
        else if (nsec < a.n_sections) fx_prefetch<P>(describe<P>(a, nsec).img0, a.row_stride, s_rowy, task, lane, pv);
      }
      // column stages on this warp's units (main: nu regular units; aux: the packed unit)
      stage_y16<P, false>(A, u0, nu, lane);
      __syncwarp();
      if (pass == 0) {
        stage_fy15_store<P>(A, u0, nu, lane, d.spec0);
      } else {
        if (!aux) {
          stage_fy15_corr<P>(A, u0, nu, lane, d.spec0, d.spec_out);
        } else {
          sp_dft15<P, 1>(A, lane);
          __syncwarp();
          sp_cross_power<P>(A, lane, d.spec0, d.spec_out);
          __syncwarp();
          sp_dft15<P, -1>(A, lane);
        }
        __syncwarp();
        stage_y16<P, true>(A, u0, nu, lane);
      }
      __syncthreads();   // pass 0: A free again, spectrum visible to the CTA; pass 1: columns done
      if (!d.correlate) break;
    }
    if (!d.correlate) continue;

    MaxTrack mt;
    mt.vmax = -INFINITY;
    mt.which = 0;
    mt.multi = false;
    if (!aux) {
      if (VEC && nsec < a.n_sections)
        fx_prefetch<P>(describe<P>(a, nsec).img0, a.row_stride, s_rowy, task, lane, pv);
      stage_ix15<P>(A, task, lane);
      __syncwarp();
      stage_x16<P, -1, true>(A, task, lane, &mt);
      const float wm = warp_max(mt.vmax);
      if (lane == 0) s_wmax[warp] = wm;
    }
    __syncthreads();   // B1: surface complete, warp maxima visible

    // block maximum, then its first location in fft-shifted row-major order (cv::minMaxLoc)
    float bm = s_wmax[lane < TASKS ? lane : 0];
    bm = warp_max(bm);
    if (!aux) {
      int loc = search_lane<P>(A, task, lane, mt, bm);
      loc = __reduce_min_sync(0xffffffffu, loc);
      if (lane == 0 && loc != 0x7fffffff) atomicMin(s_peak, loc);
    }
    __syncthreads();   // B2: peak location final

    double cx = 0.0, cy = 0.0, sum = 0.0;
    if (aux) {
      int pidx = *s_peak;
      if (pidx == 0x7fffffff) pidx = 0;   // all-NaN surface: minMaxLoc leaves the location at (0,0)
      if (lane < 25) centroid_tap<P>(A, pidx, lane, cx, cy, sum);
      __syncwarp();
      if (lane == 0) *s_peak = 0x7fffffff;
    }
    __syncthreads();   // B3: A may be overwritten by the next section

    if (aux) {
      cx = warp_sum(cx);
      cy = warp_sum(cy);
      sum = warp_sum(sum);
      if (lane == 0) {
        double xy[2];
        if (a.post == PC_POST_SHIFT) {
          centroid_finish<P>(cx, cy, sum, a.max_px_speed_sq, xy);
        } else {
          // scaleRotationEstimator::processImage tail (src/scaleRotationEstimator.cpp:117-126):
          // pt = phaseCorrelate(cur, prev) = -(shift); reject |pt.x| > res/2 -> (1, 0)
          centroid_finish<P>(cx, cy, sum, 1e300, xy);
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
    }
  }
}

// -------------------------------------------------------------------------------------------------
// host side
// -------------------------------------------------------------------------------------------------
template <int P>
static cudaError_t setup_one() {
  cudaError_t e = cudaFuncSetAttribute(pc_pfa_kernel<P, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)pc_smem_bytes<P>());
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(pc_pfa_kernel<P, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)pc_smem_bytes<P>());
}

int pc_setup(int device) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  if ((e = setup_one<4>()) != cudaSuccess) return (int)e;
  if ((e = setup_one<8>()) != cudaSuccess) return (int)e;
  if ((e = setup_one<16>()) != cudaSuccess) return (int)e;
  return 0;
}

template <int P>
static int resident_one(int device) {
  int per_sm = 0, sms = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pc_pfa_kernel<P, true>, pc_threads<P>(),
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

size_t pc_spectrum_bytes(int N) {
  switch (N) {
    case 60: return sizeof(f4) * PG<4>::SPEC_F4;
    case 120: return sizeof(f4) * PG<8>::SPEC_F4;
    case 240: return sizeof(f4) * PG<16>::SPEC_F4;
  }
  return 0;
}

template <int P>
static void launch_one(const PcArgs& args, int grid, cudaStream_t stream) {
  // P-byte vector pixel loads need P-byte aligned section origins (N is a multiple of P)
  const bool vec = (((size_t)args.curr | (size_t)args.prev | (size_t)args.row_stride |
                     (size_t)args.frame_stride) & (size_t)(P - 1)) == 0;
  if (vec) pc_pfa_kernel<P, true><<<grid, pc_threads<P>(), pc_smem_bytes<P>(), stream>>>(args);
  else pc_pfa_kernel<P, false><<<grid, pc_threads<P>(), pc_smem_bytes<P>(), stream>>>(args);
}

int pc_launch(int N, const PcArgs& args, int grid, cudaStream_t stream) {
  if (grid < 1) return (int)cudaErrorInvalidValue;
  switch (N) {
    case 60: launch_one<4>(args, grid, stream); break;
    case 120: launch_one<8>(args, grid, stream); break;
    case 240: launch_one<16>(args, grid, stream); break;
    default: return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}

}  // namespace mof
