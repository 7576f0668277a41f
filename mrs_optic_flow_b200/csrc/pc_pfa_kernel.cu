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
// pixels of the NEXT image are pulled into L2 one stage ahead with prefetch.global.L2 hints (the
// 25 KB of L1 left beside the section cannot hold an image in flight, and parking it in registers
// costs 30 registers per thread across the FFT stages), so the vector loads at the start of a row
// stage are L2 hits.  The only other global traffic
// is the previous image's spectrum through an L2-resident per-CTA scratch (two spectra of a 240 px
// section do not fit one SM) and 16 bytes of result.
//
// Replaces: the per-section loop of FftMethod::processImage (src/FftMethod.cpp:1829-1856) and the
// OpenCL kernel phaseCorrelateField (cl/FftMethod.cl:1381-1485) it would otherwise dispatch to.
#include <cuda_runtime.h>
#include <limits.h>

#include "mof_internal.h"
#include "pc_pfa.cuh"

// Optional per-stage cycle accounting (tools/stage_clocks.py builds a variant with
// -DMOF_STAGE_CLOCKS): lane 0 of every warp of CTA 0 adds the clock64() deltas between stage
// boundaries into the global counters mof_stage_clk[warp][stage].
#if defined(MOF_STAGE_CLOCKS)
__device__ unsigned long long mof_stage_clk[32][16];
#define CLK_INIT long long clk_t = clock64()
#define CLK(stage)                                                            \
  do {                                                                        \
    if (blockIdx.x == 0 && lane == 0) {                                       \
      const long long clk_n = clock64();                                      \
      atomicAdd(&mof_stage_clk[warp][stage], (unsigned long long)(clk_n - clk_t)); \
      clk_t = clk_n;                                                          \
    }                                                                         \
  } while (0)
#else
#define CLK_INIT
#define CLK(stage)
#endif

namespace mof {

using namespace pfa;

// Team geometry.  A team is the set of warps that shares one section in shared memory:
//   N = 240 / 120: TASKS main warps + 1 aux warp per CTA, one team per CTA, block barriers;
//   N = 60: the whole section is one warp task, so a team is ONE warp that also does the aux work
//   (SOLO); a CTA carries TEAMS independent warps and every barrier is a __syncwarp.
template <int P>
struct KCfg;
template <> struct KCfg<16> { static constexpr bool SOLO = false; static constexpr int TEAMS = 1, MINB = 1; };
template <> struct KCfg<8>  { static constexpr bool SOLO = false; static constexpr int TEAMS = 3, MINB = 1; };
template <> struct KCfg<4>  { static constexpr bool SOLO = true;  static constexpr int TEAMS = 5, MINB = 3; };

template <int P>
constexpr int pc_team_warps() { return KCfg<P>::SOLO ? 1 : PG<P>::TASKS + 1; }
template <int P>
constexpr int pc_threads() { return KCfg<P>::TEAMS * pc_team_warps<P>() * 32; }
// shared memory: TEAMS section arrays | warp maxima [16] | warp peak locations [16] | centroid taps
// [32] | peak index [4] | image row of (row pair, half) [N]
constexpr int SM_WMAX = 16, SM_WLOC = 16, SM_TAP = 32, SM_PIDX = 4;
template <int P>
constexpr size_t pc_smem_bytes() {
  return (sizeof(c32) * PG<P>::A_ELEMS + 4 * (SM_WMAX + SM_WLOC + SM_TAP + SM_PIDX)) * KCfg<P>::TEAMS +
         sizeof(uint16_t) * PG<P>::N;
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
__device__ __forceinline__ SecDesc describe(const PcArgs& a, int sec, int slot) {
  typedef PG<P> G;
  const size_t spec_f4 = G::SPEC_F4;
  const int f = sec / a.S, s = sec - f * a.S;
  const int j = s / a.sq, i = s - j * a.sq;
  const size_t off = (size_t)f * a.frame_stride + (size_t)(j * G::N) * a.row_stride + i * G::N;
  SecDesc d;
  if (a.mode == PC_PAIRS) {
    d.img0 = a.prev + off;
    d.img1 = a.curr + off;
    d.spec0 = reinterpret_cast<f4*>(a.scratch) + (size_t)slot * spec_f4;
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

// result of one section from the centroid sums (lane 0 of the finishing warp)
template <int P>
__device__ __forceinline__ void write_result(const PcArgs& a, int sec, double cx, double cy, double sum) {
  typedef PG<P> G;
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

template <int P, bool VEC>
__global__ void __launch_bounds__(pc_threads<P>(), KCfg<P>::MINB) pc_pfa_kernel(const PcArgs a) {
  typedef PG<P> G;
  typedef KCfg<P> K;
  constexpr int TASKS = G::TASKS;
  constexpr bool SOLO = K::SOLO;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  c32* A0 = reinterpret_cast<c32*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int WPT = SOLO ? 1 : TASKS + 1;   // warps per team
  const int team = warp / WPT;
  const int wl = warp - team * WPT;      // warp inside its team
  const int task = SOLO ? 0 : wl;
  const bool is_main = SOLO || wl < TASKS;
  const bool is_aux = SOLO || wl == TASKS;
  c32* A = A0 + (size_t)team * G::A_ELEMS;
  float* s_scal = reinterpret_cast<float*>(A0 + (size_t)G::A_ELEMS * K::TEAMS);
  float* s_wmax = s_scal + team * (SM_WMAX + SM_WLOC + SM_TAP + SM_PIDX);   // per-main-warp maxima
  int* s_wloc = reinterpret_cast<int*>(s_wmax + SM_WMAX);      // first location of each warp's maximum
  float* s_tap = reinterpret_cast<float*>(s_wloc + SM_WLOC);   // 5x5 window around the peak
  int* s_pidx = reinterpret_cast<int*>(s_tap + SM_TAP);        // peak location of the finished section
  uint16_t* s_rowy = reinterpret_cast<uint16_t*>(s_scal + K::TEAMS * (SM_WMAX + SM_WLOC + SM_TAP + SM_PIDX));
  // Teams of one CTA are independent, but they share the instruction cache.  In pairs mode every
  // team runs the same control flow, so the team barriers are made block barriers: all warps of the
  // SM then stay in the same stretch of this (large) kernel body.  A team without a section of its
  // own re-does the last section (identical result) to keep the barrier count uniform.
  const bool lockstep = K::TEAMS > 1 && a.mode == PC_PAIRS;
  auto team_bar = [&]() {
    if (K::TEAMS == 1 || lockstep) __syncthreads();
    else if (SOLO) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "n"(WPT * 32) : "memory");
  };
  // column units: main warps own the regular units of their task, the aux role owns the packed unit;
  // the P-point stages treat both kinds alike, so a SOLO warp runs them over all M units at once
  const int ur0 = task * G::UPT;
  const int nur = (task == TASKS - 1) ? G::UPT - 1 : G::UPT;
  const int uy0 = SOLO ? 0 : (is_main ? ur0 : G::M - 1);
  const int nuy = SOLO ? G::M : (is_main ? nur : 1);

  for (int i = tid; i < G::N; i += (int)blockDim.x) fill_rowy<P>(s_rowy, i);
  __syncthreads();

  // pixels of the image being staged (VEC only; live only between the loads and the staging stores)
  uint32_t pv[LoadCfg<P>::NREG];
#pragma unroll
  for (int i = 0; i < LoadCfg<P>::NREG; i++) pv[i] = 0;
  unsigned char* stage = reinterpret_cast<unsigned char*>(A + task * G::RPT * G::SQ);

  const int slot = blockIdx.x * K::TEAMS + team;
  const int nslots = gridDim.x * K::TEAMS;
  auto clampsec = [&](int s_) { return (lockstep && s_ >= a.n_sections) ? a.n_sections - 1 : s_; };
  int sec = clampsec(slot);

  CLK_INIT;
  bool pend = false;   // aux warp (team mode): the previous section's centroid is still to be finished
  int pend_sec = 0;
  // aux warp, after a team barrier: centroid of the previous section from the taps the main warps left
  auto finish_pending = [&]() {
    double cx = 0.0, cy = 0.0, sum = 0.0;
    int y, x;
    if (lane < 25 && tap_coords<P>(*s_pidx, lane, y, x)) centroid_acc(y, x, s_tap[lane], cx, cy, sum);
    cx = warp_sum(cx);
    cy = warp_sum(cy);
    sum = warp_sum(sum);
    if (lane == 0) write_result<P>(a, pend_sec, cx, cy, sum);
    pend = false;
  };

  const int iters = (a.n_sections + nslots - 1) / nslots;
  for (int it = 0; it < iters; it++, sec = clampsec(sec + nslots)) {
    if (K::TEAMS > 1) __syncthreads();   // (stream mode: at least one re-alignment per section)
    if (sec >= a.n_sections) {   // never taken in lockstep mode
      if (K::TEAMS == 1) break;
      continue;
    }
    const SecDesc d = describe<P>(a, sec, slot);
    const int nsec = clampsec(sec + nslots);

#pragma unroll 1
    for (int pass = d.two_pass ? 0 : 1; pass < 2; pass++) {
      CLK(0);
      // pass 0: spectrum of the first image -> spec0; pass 1: current image, correlate, invert columns
      if (is_main) {
        // rows: pixels -> staging (inside this warp's own rows of A) -> words -> FX15 -> FX16
        if (VEC) {
          fx_prefetch<P>(pass ? d.img1 : d.img0, a.row_stride, s_rowy, task, lane, pv);
          fx_stage<P>(stage, lane, pv);
        }
        else fx_stage_bytes<P>(stage, pass ? d.img1 : d.img0, a.row_stride, s_rowy, task, lane);
        __syncwarp();
        uint32_t wa[15], wb[15];
        stage_fx15_read<P>(A, task, lane, wa, wb);
        __syncwarp();
        stage_fx15<P>(A, task, lane, wa, wb, a.win, a.k23);
        __syncwarp();
        CLK(1);
        stage_x16<P, 1, false>(A, task, lane, nullptr);
        CLK(2);
      }
      team_bar();   // rows done
      CLK(3);
      if (!SOLO && is_aux && pend) finish_pending();
      // L2 hints for the next image one stage ahead: the second image of this section, or
      // (spectrum only) the first image of this team's next section
      if (VEC && pass == 0 && is_main) {
        const uint8_t* nx = d.correlate ? d.img1
                                        : (nsec < a.n_sections ? describe<P>(a, nsec, slot).img0 : d.img0);
        fx_prefetch_l2<P>(nx, a.row_stride, s_rowy, task, lane);
      }
      // column stages on this warp's units
      CLK(4);
      stage_y16<P, false>(A, uy0, nuy, lane);
      __syncwarp();
      CLK(5);
      if (pass == 0) {
        stage_fy15_store<P>(A, uy0, nuy, lane, d.spec0);
        CLK(6);
      } else {
        if (is_main) stage_fy15_corr<P>(A, ur0, nur, lane, d.spec0, d.spec_out);
        if (is_aux) {
          sp_dft15<P, 1>(A, lane);
          __syncwarp();
          sp_cross_power<P>(A, lane, d.spec0, d.spec_out);
          __syncwarp();
          sp_dft15<P, -1>(A, lane);
        }
        __syncwarp();
        CLK(7);
        stage_y16<P, true>(A, uy0, nuy, lane);
        CLK(8);
      }
      team_bar();   // pass 0: A free again, spectrum visible to the team; pass 1: columns done
      CLK(9);
      if (!d.correlate) break;
    }
    if (!d.correlate) continue;

    // inverse rows with a running maximum, then each warp's own maximum and its first location in
    // fft-shifted row-major order (cv::minMaxLoc)
    float wm = -INFINITY;
    int wloc = 0x7fffffff;
    if (is_main) {
      MaxTrack mt;
      mt.vmax = -INFINITY;
      mt.which = 0;
      mt.multi = false;
      if (VEC)
        fx_prefetch_l2<P>(nsec < a.n_sections ? describe<P>(a, nsec, slot).img0 : d.img0, a.row_stride, s_rowy,
                          task, lane);
      CLK(10);
      stage_ix15<P>(A, task, lane);
      __syncwarp();
      CLK(11);
      stage_x16<P, -1, true>(A, task, lane, &mt);
      CLK(12);
      wm = warp_max(mt.vmax);
      __syncwarp();
      wloc = __reduce_min_sync(0xffffffffu, search_lane<P>(A, task, lane, mt, wm));
    }
    if (SOLO) {
      const int pidx = wloc == 0x7fffffff ? 0 : wloc;   // all-NaN surface: minMaxLoc leaves (0,0)
      double cx = 0.0, cy = 0.0, sum = 0.0;
      if (lane < 25) centroid_tap<P>(A, pidx, lane, cx, cy, sum);
      cx = warp_sum(cx);
      cy = warp_sum(cy);
      sum = warp_sum(sum);
      if (lane == 0) write_result<P>(a, sec, cx, cy, sum);
      __syncwarp();   // taps read before the next section's pixels are staged into A
    } else {
      if (is_main && lane == 0) {
        s_wmax[wl] = wm;
        s_wloc[wl] = wloc;
      }
      CLK(13);
      team_bar();   // surface complete, warp maxima visible
      CLK(14);
      if (is_main) {
        // block maximum and its first location; every main warp picks the taps that live in its own
        // rows, so it may overwrite them (next section's pixels) without another block barrier
        const float v = lane < TASKS ? s_wmax[lane] : -INFINITY;
        const int l = lane < TASKS ? s_wloc[lane] : 0x7fffffff;
        const float bm = warp_max(v);
        int pidx = __reduce_min_sync(0xffffffffu, v == bm ? l : 0x7fffffff);
        if (pidx == 0x7fffffff) pidx = 0;
        int y, x;
        if (lane < 25 && tap_coords<P>(pidx, lane, y, x) && surf_rowpair<P>(y) / G::RPT == task)
          s_tap[lane] = surf_at<P>(A, y, x);
        if (wl == 0 && lane == 0) *s_pidx = pidx;
        __syncwarp();
      } else {
        pend = true;
        pend_sec = sec;
      }
    }
  }
  if (!SOLO) {
    team_bar();
    if (is_aux && pend) finish_pending();
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

#if defined(MOF_STAGE_CLOCKS)
}  // namespace mof
extern "C" int mof_debug_stage_clocks(unsigned long long* out, int reset) {
  cudaError_t e = cudaMemcpyFromSymbol(out, mof_stage_clk, sizeof(unsigned long long) * 32 * 16);
  if (e == cudaSuccess && reset) {
    static unsigned long long zeros[32 * 16];
    e = cudaMemcpyToSymbol(mof_stage_clk, zeros, sizeof(zeros));
  }
  return (int)e;
}
namespace mof {
#endif

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
  return per_sm * sms * KCfg<P>::TEAMS;
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
static void launch_one(const PcArgs& args_in, int slots, cudaStream_t stream) {
  PcArgs args = args_in;
  args.k23 = 0x4B000000u;
  const int grid = (slots + KCfg<P>::TEAMS - 1) / KCfg<P>::TEAMS;
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
