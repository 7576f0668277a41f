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
#include <stdlib.h>

#include "mof_internal.h"
#include "pc_pfa.cuh"

// Optional per-stage cycle accounting (tools/stage_clocks.py builds a variant with
// -DMOF_STAGE_CLOCKS): lane 0 of every warp of CTA 0 adds the clock64() deltas between stage
// boundaries into the global counters mof_stage_clk[warp][stage].
#if defined(MOF_STAGE_CLOCKS)
__device__ unsigned long long mof_stage_clk[32][24];
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
// TM: pairs mode keeps the first image's spectrum in tensor memory (needs one CTA per SM: the CTA
// takes all 512 columns, and <= 16 warps with four 32-column slots each).  Measured on B200:
// N = 240 4.16 -> 4.41 M pairs/s (+6 %), N = 120 (F = 480) 0.89 -> 0.97 M pairs/s (+8.5 %).
template <> struct KCfg<16> { static constexpr bool SOLO = false, TM = true; static constexpr int TEAMS = 1, MINB = 1; };
template <> struct KCfg<8>  { static constexpr bool SOLO = false, TM = true; static constexpr int TEAMS = 3, MINB = 1; };
// (N = 60 as one CTA of 15 one-warp teams with TM: 2 % faster than without TM, but the 15-warp CTA
// is 4.5 % slower than three CTAs of 5, which would need 768 columns -- stays on the scratch.)
template <> struct KCfg<4>  { static constexpr bool SOLO = true, TM = false; static constexpr int TEAMS = 5, MINB = 3; };

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

// -------------------------------------------------------------------------------------------------
// Tensor memory as the per-CTA spectrum scratch (pairs mode, N = 240 / 120).  The spectrum of the
// first image is written by a thread in pass 0 and read back by the SAME thread in pass 1 (the
// item maps of the two 15-point column stages are identical), so it needs no layout at all: every
// (round, half) of a thread is one 32-column slot of its own TMEM lane, moved with the 32x32b shape
// (thread i of warp w <-> lane 32*(w%4) + i).  15 main warps x 4 slots x 32 columns = the whole
// 512-column allocation.  This takes 2 x 230 KB per section off the L1/L2 path; TMEM is otherwise
// idle in this kernel.  The packed unit (aux warp) stays on the global scratch.
// tcgen05 instructions are warp-collective (.sync.aligned): the callers keep the warp converged.
// -------------------------------------------------------------------------------------------------
// 30 words (15 complex) of a thread <-> columns [taddr, taddr + 30) of its lane, as x16 + x8 + x4 + x2
__device__ __forceinline__ void tm_st30(uint32_t taddr, const uint32_t (&w)[30]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
               :: "r"(taddr + 0u), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]), "r"(w[8]), "r"(w[9]), "r"(w[10]), "r"(w[11]), "r"(w[12]), "r"(w[13]), "r"(w[14]), "r"(w[15]) : "memory");
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
               :: "r"(taddr + 16u), "r"(w[16]), "r"(w[17]), "r"(w[18]), "r"(w[19]), "r"(w[20]), "r"(w[21]), "r"(w[22]), "r"(w[23]) : "memory");
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
               :: "r"(taddr + 24u), "r"(w[24]), "r"(w[25]), "r"(w[26]), "r"(w[27]) : "memory");
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};"
               :: "r"(taddr + 28u), "r"(w[28]), "r"(w[29]) : "memory");
}
__device__ __forceinline__ void tm_ld30(uint32_t taddr, uint32_t (&w)[30]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
               : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]), "=r"(w[8]), "=r"(w[9]), "=r"(w[10]), "=r"(w[11]), "=r"(w[12]), "=r"(w[13]), "=r"(w[14]), "=r"(w[15]) : "r"(taddr + 0u) : "memory");
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(w[16]), "=r"(w[17]), "=r"(w[18]), "=r"(w[19]), "=r"(w[20]), "=r"(w[21]), "=r"(w[22]), "=r"(w[23]) : "r"(taddr + 16u) : "memory");
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(w[24]), "=r"(w[25]), "=r"(w[26]), "=r"(w[27]) : "r"(taddr + 24u) : "memory");
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];"
               : "=r"(w[28]), "=r"(w[29]) : "r"(taddr + 28u) : "memory");
}
// the loaded registers are tied to the wait so that no use of them can move above it
__device__ __forceinline__ void tm_wait_ld(uint32_t (&w)[30]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3]), "+r"(w[4]), "+r"(w[5]), "+r"(w[6]), "+r"(w[7]), "+r"(w[8]), "+r"(w[9]), "+r"(w[10]), "+r"(w[11]), "+r"(w[12]), "+r"(w[13]), "+r"(w[14]), "+r"(w[15]), "+r"(w[16]), "+r"(w[17]), "+r"(w[18]), "+r"(w[19]), "+r"(w[20]), "+r"(w[21]), "+r"(w[22]), "+r"(w[23]), "+r"(w[24]), "+r"(w[25]), "+r"(w[26]), "+r"(w[27]), "+r"(w[28]), "+r"(w[29]) :: "memory");
}
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t tm_alloc512(uint32_t* s_base) {   // one whole warp
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(s_base);
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(sa) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  return sa;
}
__device__ __forceinline__ void tm_dealloc512(uint32_t taddr) {       // the same warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(taddr) : "memory");
}
__device__ __forceinline__ void tm_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tm_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// slots of a warp (lane quarter of the warp in bits 31:16, column in bits 15:0), slot of (round, half)
__device__ __forceinline__ uint32_t tm_warp_slots(uint32_t tm_base, int warp) {
  return tm_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 128);
}
__device__ __forceinline__ uint32_t tm_slot(uint32_t tm_warp, int r, int half) {
  return tm_warp + (uint32_t)((r * 2 + half) * 32);
}

// stage FY15 of the first image with the spectrum kept in tensor memory (cf. stage_fy15_store).
// Lanes without an item (ul >= nu) run the butterfly of the task's last unit alongside its owner and
// park the result in their own slot: the body stays straight-line code for the whole warp.
template <int P>
__device__ __forceinline__ void stage_fy15_store_tm(const c32* A, int u0, int nu, int lane, uint32_t tm_warp) {
#pragma unroll 1
  for (int r = 0; r < ColMap<P>::rounds(nu); r++) {
    int ul, j;
    if (!ColMap<P>::y15(r, lane, nu, ul, j)) ul = nu - 1;
    int kcol, mcol;
    unit_cols<P>(u0 + ul, kcol, mcol);
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
      c32 z[15];
      col15_load<P>(A, va_base<P>(va_of<P>(j, half), kcol, mcol), z);
      dft15<1>(z);
      uint32_t w[30];
#pragma unroll
      for (int i = 0; i < 15; i++) {
        w[2 * i] = __float_as_uint(z[i].x);
        w[2 * i + 1] = __float_as_uint(z[i].y);
      }
      __syncwarp();
      tm_st30(tm_slot(tm_warp, r, half), w);
    }
  }
  tm_wait_st();
}

// exact-path redo (cf. fy15_corr_slow), entered by the whole warp; lanes with `need` redo their butterfly
template <int P>
__device__ __noinline__ void fy15_corr_slow_tm(c32* A, int base, bool need, uint32_t taddr) {
  uint32_t w[30];
  tm_ld30(taddr, w);
  tm_wait_ld(w);
  if (need) {
    c32 z[15];
    col15_load<P>(A, base, z);
    dft15<1>(z);
#pragma unroll
    for (int i = 0; i < 15; i++)
      z[i] = cross_power_exact(z[i], mk(__uint_as_float(w[2 * i]), __uint_as_float(w[2 * i + 1])),
                               16.0f * FLT_EPSILON);
    dft15<-1>(z);
    col15_store<P>(A, base, z);
  }
  __syncwarp();
}

// stage FY15 + cross-power + IY15 against the spectrum in tensor memory (cf. stage_fy15_corr).
// Lanes without an item compute on the last unit's column like in the store stage (whatever they
// read while its owner rewrites it is discarded): only the store back is conditional.
template <int P>
__device__ __forceinline__ void stage_fy15_corr_tm(c32* A, int u0, int nu, int lane, uint32_t tm_warp) {
#pragma unroll 1
  for (int r = 0; r < ColMap<P>::rounds(nu); r++) {
    int ul, j;
    const bool act = ColMap<P>::y15(r, lane, nu, ul, j);
    if (!act) ul = nu - 1;
    int kcol, mcol;
    unit_cols<P>(u0 + ul, kcol, mcol);
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
      const int base = va_base<P>(va_of<P>(j, half), kcol, mcol);
      const uint32_t taddr = tm_slot(tm_warp, r, half);
      uint32_t w[30];
      c32 z[15];
      __syncwarp();
      tm_ld30(taddr, w);   // in flight during the forward DFT
      col15_load<P>(A, base, z);
      dft15<1>(z);
      tm_wait_ld(w);
      float m2min = 3.0e38f;
#pragma unroll
      for (int i = 0; i < 15; i++)
        z[i] = cross_power_fast(z[i], mk(__uint_as_float(w[2 * i]), __uint_as_float(w[2 * i + 1])), m2min);
      dft15<-1>(z);
      const bool need = act && !(m2min >= MOF_CP_FAST_MIN);
      if (act && !need) col15_store<P>(A, base, z);
      if (__any_sync(0xffffffffu, need)) fy15_corr_slow_tm<P>(A, base, need, taddr);
    }
  }
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

// TM: pairs mode with the first image's spectrum of the regular units in tensor memory (N = 240 / 120)
template <int P, bool VEC, bool TM>
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
  __shared__ uint32_t s_tm_base;
  uint32_t tm_warp = 0;   // this warp's four slots
  if (TM) {
    if (warp == 0) tm_alloc512(&s_tm_base);
    tm_fence_before();
  }
  __syncthreads();
  if (TM) {
    tm_fence_after();
    tm_warp = tm_warp_slots(s_tm_base, warp);
  }

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
        CLK(15);
        uint32_t wa[15], wb[15];
        stage_fx15_read<P>(A, task, lane, wa, wb);
        __syncwarp();
        CLK(16);
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
      // team mode: the aux warp (idle through the row stages, short column stage) issues the hints
      // for every task (+1.2 % at N = 240, +1.8 % at N = 120 over each main warp hinting its own
      // rows); a solo warp is its own aux
      if (VEC && pass == 0 && is_aux) {
        const uint8_t* nx = d.correlate ? d.img1
                                        : (nsec < a.n_sections ? describe<P>(a, nsec, slot).img0 : d.img0);
        if (SOLO) fx_prefetch_l2<P>(nx, a.row_stride, s_rowy, task, lane);
        else
          for (int t = 0; t < TASKS; t++) fx_prefetch_l2<P>(nx, a.row_stride, s_rowy, t, lane);
      }
      // column stages on this warp's units
      CLK(4);
      stage_y16<P, false>(A, uy0, nuy, lane);
      __syncwarp();
      CLK(5);
      if (pass == 0) {
        if (TM) {   // regular units -> tensor memory, packed unit -> global scratch
          if (is_main) stage_fy15_store_tm<P>(A, ur0, nur, lane, tm_warp);
          if (is_aux) stage_fy15_store<P>(A, G::M - 1, 1, lane, d.spec0);
        } else {
          stage_fy15_store<P>(A, uy0, nuy, lane, d.spec0);
        }
        CLK(6);
      } else {
        if (TM) {
          if (is_main) stage_fy15_corr_tm<P>(A, ur0, nur, lane, tm_warp);
        } else if (is_main) {
          stage_fy15_corr<P>(A, ur0, nur, lane, d.spec0, d.spec_out);
        }
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
      if (VEC && SOLO)
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
    else if (VEC && !SOLO) {   // aux warp: L2 hints for the first image of this team's next section
      const uint8_t* nx = nsec < a.n_sections ? describe<P>(a, nsec, slot).img0 : d.img0;
      for (int t = 0; t < TASKS; t++) fx_prefetch_l2<P>(nx, a.row_stride, s_rowy, t, lane);
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
  if (TM) {   // every read of tensor memory precedes the barrier above (TM implies lockstep block barriers)
    __syncwarp();
    if (warp == 0) tm_dealloc512(s_tm_base);
  }
}

// -------------------------------------------------------------------------------------------------
// host side
// -------------------------------------------------------------------------------------------------
template <int P>
static cudaError_t setup_one() {
  cudaError_t e = cudaFuncSetAttribute(pc_pfa_kernel<P, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)pc_smem_bytes<P>());
  if (e != cudaSuccess) return e;
  if (KCfg<P>::TM) {
    e = cudaFuncSetAttribute(pc_pfa_kernel<P, true, KCfg<P>::TM>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)pc_smem_bytes<P>());
    if (e != cudaSuccess) return e;
  }
  return cudaFuncSetAttribute(pc_pfa_kernel<P, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)pc_smem_bytes<P>());
}

#if defined(MOF_STAGE_CLOCKS)
}  // namespace mof
extern "C" int mof_debug_stage_clocks(unsigned long long* out, int reset) {
  cudaError_t e = cudaMemcpyFromSymbol(out, mof_stage_clk, sizeof(unsigned long long) * 32 * 24);
  if (e == cudaSuccess && reset) {
    static unsigned long long zeros[32 * 24];
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
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pc_pfa_kernel<P, true, false>, pc_threads<P>(),
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
  const bool tm = KCfg<P>::TM && vec && args.mode == PC_PAIRS && !getenv("MOF_NO_TMEM");
  if (tm) pc_pfa_kernel<P, true, KCfg<P>::TM><<<grid, pc_threads<P>(), pc_smem_bytes<P>(), stream>>>(args);
  else if (vec) pc_pfa_kernel<P, true, false><<<grid, pc_threads<P>(), pc_smem_bytes<P>(), stream>>>(args);
  else pc_pfa_kernel<P, false, false><<<grid, pc_threads<P>(), pc_smem_bytes<P>(), stream>>>(args);
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
