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

// A/B switches of the N = 240 pairs-mode pipeline (nvcc -DMOF_SURF_TM=0 / -DMOF_PV_EARLY=0 build the older flow)
#if !defined(MOF_SPEC_PREFETCH)
#define MOF_SPEC_PREFETCH 0   // stream mode: L2 hints for the next step's cached spectrum (aux warp); measured -8 % at
                              // N = 240 (4.63 -> 4.25 M frames/s, profiles/r3h_spec_prefetch_ab.txt): off
#endif
#if !defined(MOF_SURF_TM)
#define MOF_SURF_TM 1   // correlation surface in tensor memory, next section's pixels staged behind IX16
#endif
#if !defined(MOF_LOCKSTEP)
#define MOF_LOCKSTEP 0     // N = 120 pairs mode: 1 = the three teams of a CTA share block barriers; 0 = team barriers, one
#endif                     // re-alignment per section; 2 = free-running teams
#if !defined(MOF_LOCKSTEP_SOLO)
#define MOF_LOCKSTEP_SOLO 1   // N = 60 pairs mode: the five one-warp teams of a CTA meet at every stage-group boundary
#endif
#if !defined(MOF_SOLO_GROUPS)
#define MOF_SOLO_GROUPS 3   // N = 60: lock-step groups of warps (0/1: one block-wide group; 2: 8 + 7; 3: 5 + 5 + 5; 4; 5), each a row
                            // stage behind the previous one: while one group waits for pixels the others compute
#endif
#if !defined(MOF_TEAM_SKEW)
#define MOF_TEAM_SKEW 1     // N = 120: the three teams free-run (no re-alignment per section), team t one row stage behind
                            // team t - 1 at the start (0: off, 1: pairs mode, 2: every mode)
#endif
#if !defined(MOF_SKEW_AT)
#define MOF_SKEW_AT 1       // where a team / group releases the one behind it: 1 = after its first row stage, 2 = after its first pass
#endif
#if !defined(MOF_SOLO_SYNC_ALL)
#define MOF_SOLO_SYNC_ALL 0   // N = 60 lock-step: block barrier between ALL stages (instruction-cache locality experiment)
#endif
#if !defined(MOF_COL_MIX)
#define MOF_COL_MIX 0      // (measured: -1 %, profiles/r2p_ab_column_interleave_rejected.txt) N = 240 pairs: half of the main warps interleave their column stages per unit half
#endif
#if !defined(MOF_TMA_FETCH)
#define MOF_TMA_FETCH 0    // N = 240 pairs: the aux warp fetches every image with bulk row copies (one mbarrier per main warp)
#endif
#if !defined(MOF_FETCH_ASYNC)
#define MOF_FETCH_ASYNC 1  // N = 240: pixels global -> staging with cp.async instead of LDG + STS
#endif
#if !defined(MOF_SOLO_PV_EARLY)
#define MOF_SOLO_PV_EARLY 0   // (measured: -8..-10 %, profiles/r2t_*: 400 B of spills) N = 60: bit 0 = second image requested before pass 0's column stages, bit 1 = next
#endif                        // section's first image requested before the inverse row stages (4-point: few registers)
#if !defined(MOF_OPAQUE_STAGES)
#define MOF_OPAQUE_STAGES 0
#endif
#if !defined(MOF_PARK2)
#define MOF_PARK2 0     // (measured: -2 %, profiles/r4b_*) N = 240 pairs: the second image's pixels are requested (cp.async) into the columns that pass 0's
#endif                  // 15-point column stage has already consumed, and picked up before the barrier that ends the pass
#if !defined(MOF_PV_EARLY)
#define MOF_PV_EARLY 0  // second image's pixels requested before the first image's last column stage
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
#if !defined(MOF_SOLO_TEAMS)
#define MOF_SOLO_TEAMS 15   // one CTA of 15 one-warp teams per SM, in lock-step (r2l: 750 -> 792 K pairs/s on cfg4 against 3 x 5)
#define MOF_SOLO_MINB 1
#endif
#if !defined(MOF_SOLO_TM)
#define MOF_SOLO_TM 1   // first spectrum in tensor memory at N = 60 too (r2m: 792 -> 817 K pairs/s on cfg4)
#endif
template <> struct KCfg<4>  { static constexpr bool SOLO = true, TM = (MOF_SOLO_TM != 0) && MOF_SOLO_MINB == 1; static constexpr int TEAMS = MOF_SOLO_TEAMS, MINB = MOF_SOLO_MINB; };

template <int P>
constexpr int pc_team_warps() { return KCfg<P>::SOLO ? 1 : PG<P>::TASKS + 1; }
template <int P>
constexpr int pc_threads() { return KCfg<P>::TEAMS * pc_team_warps<P>() * 32; }
// shared memory: TEAMS section arrays | warp maxima [16] | warp peak locations [16] | centroid taps
// [64] | peak index [4] || step descriptors [TEAMS][2] | image row of (row pair, half) [N] | (N = 120) pixel
// staging buffers [TEAMS][N*N]
constexpr int SM_WMAX = 16, SM_WLOC = 16, SM_TAP = 64, SM_PIDX = 4;   // taps: 5x5, or 7x7 in OCL-compatible mode
constexpr int SM_DESC = 2 * 64;   // bytes: two step descriptors per team (this step / the next one)
// N = 120 leaves 51 KB of the SM's shared memory unused (three 58 KB section arrays): each team gets a
// pixel staging buffer of its own (one image of the section, N*N bytes) OUTSIDE its section array, so that
// the next image can be requested while the current one is still being transformed.
// MEASURED AND SWITCHED OFF (profiles/r2r_ab_n120_staging_buffer_rejected.txt): 0.998 -> 0.90 M pairs/s at
// F = 480, whether the aux warp or every main warp itself requests the next image.  The 43 KB take the SM's
// L1 down from 52 KB to 9 KB, and this instantiation's register spills (150 B per thread) and the packed
// unit's spectrum traffic live there.
#if !defined(MOF_XSTAGE)
#define MOF_XSTAGE 0
#endif
template <int P>
__host__ __device__ constexpr size_t pc_xstage_bytes() { return (P == 8 && MOF_XSTAGE) ? (size_t)PG<P>::N * PG<P>::N : 0; }
template <int P>
constexpr size_t pc_smem_bytes() {
  return (sizeof(c32) * PG<P>::A_ELEMS + 4 * (SM_WMAX + SM_WLOC + SM_TAP + SM_PIDX) + SM_DESC + pc_xstage_bytes<P>()) *
             KCfg<P>::TEAMS +
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

// ---- second image parked in dead columns (N = 240 pairs, MOF_PARK2) ---------------------------------------------
// Pass 0 ends with the 15-point column stage, which moves the first spectrum to tensor memory: once round 0 / half 0
// of a warp task is through its loads, the k-columns of the task's first four units (4 x 120 complex slots = the
// 480 eight-byte pieces of the task's 16 image rows) are dead, and nobody but this warp touches them before the
// barrier that ends the pass.  The warp asks for its rows of the SECOND image there (cp.async, 8 bytes per lane:
// lane l moves piece l of a row, 30 pieces per row), the requests fly during the remaining three half rounds, and
// every lane picks its 2 x 15 pixel words from the columns before the barrier.  Pass 1 then starts with the forward
// butterflies: its pixel fetch (3 K of a section's 64 K cycles: every warp of every SM asked L2 for its rows at the
// same moment, after the barrier) is gone.  Piece c of row r (0..15) sits in row (r & 3) * 30 + c of column r >> 2.
__device__ __forceinline__ void park_issue(c32* A, int u0, const uint8_t* img, int row_stride, const uint16_t* rowy,
                                           int task, int lane) {
  typedef PG<16> G;
  asm volatile("" : "+r"(lane));   // (nothing derived from the lane is hoisted out of the section loop)
  if (lane >= 30) return;
  const uint8_t* src = img + 8 * lane;
  const uint16_t* ry = rowy + task * 2 * G::RPT;
#pragma unroll
  for (int c = 0; c < 4; c++) {
    int kcol, mcol;
    unit_cols<16>(u0 + c, kcol, mcol);
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(A + lane * G::SQ + kcol);
#pragma unroll
    for (int rr = 0; rr < 4; rr++) {
      const uint8_t* p = src + (unsigned long long)((unsigned)ry[4 * c + rr]) * (unsigned)row_stride;   // one IMAD.WIDE.U32
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d + rr * 30 * G::SQ * 8), "l"(p) : "memory");
    }
  }
}
// a definition without an instruction (ends the live range of whatever the registers held)
__device__ __forceinline__ void words_undef(uint32_t (&wa)[15], uint32_t (&wb)[15]) {
#pragma unroll
  for (int i = 0; i < 15; i++) asm volatile("" : "=r"(wa[i]), "=r"(wb[i]));
}
// the lane's words of the parked rows: (row pair, word group) = (lane / 4, lane % 4) as in fx15_lane / fx_read_words;
// word (g, t) of a row is word g & 1 of piece 2 t + (g >> 1)
__device__ __forceinline__ void park_read(const c32* A, int u0, int lane, uint32_t (&wa)[15], uint32_t (&wb)[15]) {
  typedef PG<16> G;
  asm volatile("" : "+r"(lane));
  const int qi = lane >> 2, g = lane & 3;
  int kcol, mcol;
  unit_cols<16>(u0 + (qi >> 1), kcol, mcol);
  // row a = 2 qi, row b = 2 qi + 1: same column, 30 rows apart
  const uint32_t* base = reinterpret_cast<const uint32_t*>(A + ((qi & 1) * 60 + (g >> 1)) * G::SQ + kcol) + (g & 1);
  int c = word_c<16>(g);
  asm volatile("" : "+r"(c));   // (keeps ptxas from hoisting the 15 lane-dependent addresses out of the section loop: they spill)
#pragma unroll
  for (int s = 0; s < 15; s++) {
    const int t = word_off<16>(s, c) >> 4;
    wa[s] = base[4 * t * G::SQ];
    wb[s] = base[(4 * t + 60) * G::SQ];
  }
}

// stage FY15 of the first image with the spectrum kept in tensor memory (cf. stage_fy15_store).
// Lanes without an item (ul >= nu) run the butterfly of the task's last unit alongside its owner and
// park the result in their own slot: the body stays straight-line code for the whole warp.
template <int P>
__device__ __forceinline__ void stage_fy15_store_tm(c32* A, int u0, int nu, int lane, uint32_t tm_warp, int r0 = 0,
                                                    const uint8_t* park_img = nullptr, int row_stride = 0,
                                                    const uint16_t* rowy = nullptr, int task = 0) {
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
      tm_st30(tm_slot(tm_warp, r0 + r, half), w);
      // every lane is through its loads of round 0 / half 0: the k-columns of units u0 .. u0 + 3 are dead
      if (P == 16 && park_img && r == 0 && half == 0) park_issue(A, u0, park_img, row_stride, rowy, task, lane);
    }
  }
  tm_wait_st();
}

// exact-path redo (cf. fy15_corr_slow), entered by the whole warp; lanes with `need` redo their butterfly
template <int P>
__device__ __noinline__ void fy15_corr_slow_tm(c32* A, int base, bool need, uint32_t taddr, int ocl) {
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
                               16.0f * FLT_EPSILON, ocl);
    dft15<-1>(z);
    col15_store<P>(A, base, z);
  }
  __syncwarp();
}

// stage FY15 + cross-power + IY15 against the spectrum in tensor memory (cf. stage_fy15_corr).
// Lanes without an item compute on the last unit's column like in the store stage (whatever they
// read while its owner rewrites it is discarded): only the store back is conditional.
template <int P>
__device__ __forceinline__ void stage_fy15_corr_tm(c32* A, int u0, int nu, int lane, uint32_t tm_warp, int ocl, int r0 = 0) {
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
      const uint32_t taddr = tm_slot(tm_warp, r0 + r, half);
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
      if (__any_sync(0xffffffffu, need)) fy15_corr_slow_tm<P>(A, base, need, taddr, ocl);
    }
  }
}

// 32 words (16 complex) of a thread <-> columns [taddr, taddr + 32) of its lane
// exact-path redo with the previous spectrum in registers (PC_SEQ: tensor memory already holds the
// current one); entered by the whole warp
template <int P>
__device__ __noinline__ void fy15_corr_slow_w(c32* A, int base, bool need, const uint32_t (&w)[30], int ocl) {
  if (need) {
    c32 z[15];
    col15_load<P>(A, base, z);
    dft15<1>(z);
#pragma unroll
    for (int i = 0; i < 15; i++)
      z[i] = cross_power_exact(z[i], mk(__uint_as_float(w[2 * i]), __uint_as_float(w[2 * i + 1])),
                               16.0f * FLT_EPSILON, ocl);
    dft15<-1>(z);
    col15_store<P>(A, base, z);
  }
  __syncwarp();
}

// PC_SEQ with the spectrum chain in tensor memory (mof_fft_stream_push_many, N = 240): the slot that walks a
// (stream, section) sequence correlates frame t against the spectrum its own threads left in their TMEM lanes
// one step earlier and leaves frame t's spectrum there in turn -- nothing of the regular units goes through
// L2 between frames.  prev_g: the previous spectrum comes from the stream cache instead (first step of a
// sequence whose stream already has a history); out_g: the stream's last frame also goes to the cache.
template <int P>
__device__ __forceinline__ void stage_fy15_corr_seqtm(c32* A, int u0, int nu, int lane, uint32_t tm_warp,
                                                      const f4* prev_g, f4* out_g) {
#pragma unroll 1
  for (int r = 0; r < ColMap<P>::rounds(nu); r++) {
    int ul, j;
    const bool act = ColMap<P>::y15(r, lane, nu, ul, j);
    if (!act) ul = nu - 1;
    int kcol, mcol;
    unit_cols<P>(u0 + ul, kcol, mcol);
    const int item = item_of<P>(u0 + ul, j);
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
      const int base = va_base<P>(va_of<P>(j, half), kcol, mcol);
      const uint32_t taddr = tm_slot(tm_warp, r, half);
      uint32_t w[30];
      c32 z[15];
      __syncwarp();
      if (!prev_g) tm_ld30(taddr, w);   // in flight during the forward DFT
      col15_load<P>(A, base, z);
      dft15<1>(z);
      if (prev_g) {
        c32 fp[15];
        spec_load15<P>(prev_g, item, half, fp);
#pragma unroll
        for (int i = 0; i < 15; i++) {
          w[2 * i] = __float_as_uint(fp[i].x);
          w[2 * i + 1] = __float_as_uint(fp[i].y);
        }
      } else {
        tm_wait_ld(w);
      }
      {   // the previous spectrum is in registers: this frame's takes its place
        uint32_t cw[30];
#pragma unroll
        for (int i = 0; i < 15; i++) {
          cw[2 * i] = __float_as_uint(z[i].x);
          cw[2 * i + 1] = __float_as_uint(z[i].y);
        }
        __syncwarp();
        tm_st30(taddr, cw);
      }
      if (out_g && act) spec_store15<P>(out_g, item, half, z);
      float m2min = 3.0e38f;
#pragma unroll
      for (int i = 0; i < 15; i++)
        z[i] = cross_power_fast(z[i], mk(__uint_as_float(w[2 * i]), __uint_as_float(w[2 * i + 1])), m2min);
      const bool need = act && !(m2min >= MOF_CP_FAST_MIN);
      if (__any_sync(0xffffffffu, need)) {   // rare: only here does the previous spectrum get an address (local memory)
        uint32_t wl[30];
#pragma unroll
        for (int i = 0; i < 30; i++) wl[i] = w[i];
        fy15_corr_slow_w<P>(A, base, need, wl, 0);
      }
      dft15<-1>(z);
      if (act && !need) col15_store<P>(A, base, z);
    }
  }
  tm_wait_st();
}

__device__ __forceinline__ void tm_st32(uint32_t taddr, const uint32_t (&w)[32]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
               "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
               :: "r"(taddr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]), "r"(w[8]),
                  "r"(w[9]), "r"(w[10]), "r"(w[11]), "r"(w[12]), "r"(w[13]), "r"(w[14]), "r"(w[15]), "r"(w[16]), "r"(w[17]),
                  "r"(w[18]), "r"(w[19]), "r"(w[20]), "r"(w[21]), "r"(w[22]), "r"(w[23]), "r"(w[24]), "r"(w[25]), "r"(w[26]),
                  "r"(w[27]), "r"(w[28]), "r"(w[29]), "r"(w[30]), "r"(w[31]) : "memory");
}
__device__ __forceinline__ void tm_ld32_wait(uint32_t taddr, uint32_t (&w)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
               "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
               "tcgen05.wait::ld.sync.aligned;"
               : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]), "=r"(w[8]),
                 "=r"(w[9]), "=r"(w[10]), "=r"(w[11]), "=r"(w[12]), "=r"(w[13]), "=r"(w[14]), "=r"(w[15]), "=r"(w[16]),
                 "=r"(w[17]), "=r"(w[18]), "=r"(w[19]), "=r"(w[20]), "=r"(w[21]), "=r"(w[22]), "=r"(w[23]), "=r"(w[24]),
                 "=r"(w[25]), "=r"(w[26]), "=r"(w[27]), "=r"(w[28]), "=r"(w[29]), "=r"(w[30]), "=r"(w[31])
               : "r"(taddr) : "memory");
}

// ---- bulk row copies + mbarriers: the aux warp as the pixel producer (N = 240 pairs mode) -------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(bar);
  unsigned done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(a), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void bulk_row(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src), "r"(bytes),
                 "r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
// One image for all main warps: main warp t's 2*RPT image rows (N bytes each, contiguous in its staging
// area) arrive on mbarrier t, in warp order, so that warp 0 starts its row stage while the rows of the
// later warps are still on their way.  Whole aux warp; lanes 0 .. 2*RPT-1 issue one row per main warp.
template <int P>
__device__ __forceinline__ void aux_fetch_image(c32* A, const uint8_t* img, int row_stride, const uint16_t* rowy,
                                                uint64_t* bars, int lane) {
  typedef PG<P> G;
  constexpr int ROWS = 2 * G::RPT;
#pragma unroll 1
  for (int t = 0; t < G::TASKS; t++) {
    if (lane == 0) mbar_expect_tx(&bars[t], (unsigned)(ROWS * G::N));
    __syncwarp();
    if (lane < ROWS) {
      unsigned char* dst = reinterpret_cast<unsigned char*>(A + t * G::RPT * G::SQ) + lane * G::N;
      bulk_row(dst, img + (size_t)rowy[t * ROWS + lane] * row_stride, (unsigned)G::N, &bars[t]);
    }
  }
}

// ---- N = 240, pairs mode: the correlation surface goes to tensor memory instead of back into A -------
// Stage IX16 is the last reader of A in a section.  With its output in the thread's own TMEM lane
// (round r -> columns [32r, 32r + 32) of the warp's block, free since the cross-power consumed the
// spectrum), a warp's first two row pairs are dead as soon as round 0 has read them -- exactly the
// staging area of the NEXT section's first image, which is then pulled in with cp.async while rounds
// 1..3 run (the pixel fetch, ~3.7 K of a section's 68 K cycles, used to sit exposed at the top of the
// next row stage).  The first maximum and its fft-shifted row-major location are tracked in registers.
template <int P>
__device__ __forceinline__ void stage_ix16_tm(c32* A, int task, int lane, uint32_t tm_warp, float& wm, int& wloc,
                                              const uint8_t* next_img, int row_stride, const uint16_t* rowy,
                                              unsigned char* stage) {
  typedef PG<P> G;
  static_assert(P == 16, "item map of N = 240: a half warp = one row pair");
#if MOF_OPAQUE_WORDC
  if (((MOF_OPAQUE_WORDC) & 16) != 0) asm volatile("" : "+r"(lane));   // (nothing derived from the lane is hoisted out of the section loop)
#endif
  float vmax = -INFINITY;
  int which = 0;
  bool multi = false;
  const int b = (lane & 15) < 15 ? (lane & 15) : 14;   // lanes 15 / 31 shadow their neighbour (collective stores)
  const bool act = (lane & 15) < 15;
  c32 zn[P];   // loads run one round ahead of the butterflies (cf. stage_x16)
  {
    const c32* a0 = A + (task * G::RPT + (lane >> 4)) * G::SQ + slot<P>(0, b);
#pragma unroll
    for (int i = 0; i < P; i++) zn[i] = a0[SlotStep<P>::A * i];
  }
#pragma unroll 1
  for (int r = 0; r < RowMap<P>::X16_ROUNDS; r++) {
    const int q = task * G::RPT + 2 * r + (lane >> 4);
    const c32* a = A + q * G::SQ + slot<P>(0, b);
    c32 z[P];
#pragma unroll
    for (int i = 0; i < P; i++) z[i] = zn[i];
    if (MOF_X16_PIPELINE && r + 1 < RowMap<P>::X16_ROUNDS) {
      const c32* an = a + 2 * G::SQ;
#pragma unroll
      for (int i = 0; i < P; i++) zn[i] = an[SlotStep<P>::A * i];
    }
    DftP<P, -1>::run(z);
    if (!MOF_X16_PIPELINE && r + 1 < RowMap<P>::X16_ROUNDS) {
      const c32* an = a + 2 * G::SQ;
#pragma unroll
      for (int i = 0; i < P; i++) zn[i] = an[SlotStep<P>::A * i];
    }
    if (r == 0) {   // rows 2r, 2r+1 of every lane's half warp have been consumed: the staging area is free
      __syncwarp();
      if (stage == nullptr) asm volatile("bar.arrive 3, %0;" ::"n"((PG<P>::TASKS + 1) * 32) : "memory");   // tell the aux warp (producer)
      else if (next_img) fx_stage_async<P>(stage, next_img, row_stride, rowy, task, lane);
    }
    float m = fmaxf(z[0].x, z[0].y);
#pragma unroll
    for (int i = 1; i < P; i++) m = fmaxf(m, fmaxf(z[i].x, z[i].y));
    if (act) {   // running maximum; its location is recovered afterwards from the slot of round `which`
      if (m > vmax) {
        vmax = m;
        which = r;
        multi = false;
      } else if (m == vmax) {
        multi = true;
      }
    }
    uint32_t w[32];
#pragma unroll
    for (int i = 0; i < P; i++) {
      w[2 * i] = __float_as_uint(z[i].x);
      w[2 * i + 1] = __float_as_uint(z[i].y);
    }
    __syncwarp();
    tm_st32(tm_warp + (uint32_t)(r * 32), w);
  }
  tm_wait_st();
  // the warp's maximum and its first location in fft-shifted row-major order (cv::minMaxLoc): only the
  // lanes that hold the maximum look again, at the round(s) that produced it
  wm = warp_max(vmax);
  int best = 0x7fffffff;
#pragma unroll 1
  for (int r = 0; r < RowMap<P>::X16_ROUNDS; r++) {
    const bool need = act && vmax == wm && (multi || which == r);
    if (__ballot_sync(0xffffffffu, need) == 0u) continue;
    uint32_t w[32];
    tm_ld32_wait(tm_warp + (uint32_t)(r * 32), w);
    if (need) {
      const int q = task * G::RPT + 2 * r + (lane >> 4);
      const int ya = rowy[2 * q];                          // image row of component x's partner: see rescan_item
      const int ysa = ya >= G::M ? ya - G::M : ya + G::M;
      int x = 16 * b;                                      // crt(ax, b) = (225 ax + 16 b) mod 240, walked incrementally
#pragma unroll
      for (int ax = 0; ax < P; ax++) {
        const int xs = x >= G::M ? x - G::M : x + G::M;
        if (__uint_as_float(w[2 * ax]) == wm) best = imin(best, ysa * G::N + xs);
        if (__uint_as_float(w[2 * ax + 1]) == wm) best = imin(best, ya * G::N + xs);
        x += 225;
        if (x >= G::N) x -= G::N;
      }
    }
  }
  wloc = __reduce_min_sync(0xffffffffu, best);
}

// centroid taps from the surface in tensor memory: lane L of round r holds (row pair q, bx = L & 15); it
// owns a tap when one of its two surface rows lies within +-2 of the peak row and one of the five peak
// columns has its bx.  Whole warp (collective loads); rounds without a tap are skipped.
template <int P>
__device__ __forceinline__ void taps_from_tm(int task, int lane, uint32_t tm_warp, int pidx, float* s_tap,
                                             const uint16_t* rowy) {
  typedef PG<P> G;
  const int py = pidx / G::N, px = pidx - py * G::N;
  const int b = lane & 15;
  int e_col = -1, dxs = 0;   // element (ax) whose shifted column is px + dxs
#pragma unroll
  for (int dx = -2; dx <= 2; dx++) {
    const int xs = px + dx;
    if (xs < 0 || xs >= G::N) continue;
    const int x = xs >= G::M ? xs - G::M : xs + G::M;
    if (x % 15 == b) { e_col = x % P; dxs = dx; }
  }
#pragma unroll 1
  for (int r = 0; r < RowMap<P>::X16_ROUNDS; r++) {
    const int q = task * G::RPT + 2 * r + (lane >> 4);
    const int ya = rowy[2 * q];
    const int ysa = ya >= G::M ? ya - G::M : ya + G::M;   // shifted row of component x; component y sits in row ya
    int comp = -1, dys = 0;
    if (ysa - py >= -2 && ysa - py <= 2) { comp = 0; dys = ysa - py; }
    else if (ya - py >= -2 && ya - py <= 2) { comp = 1; dys = ya - py; }
    const bool own = b < 15 && comp >= 0 && e_col >= 0;
    if (__ballot_sync(0xffffffffu, own) == 0u) continue;
    uint32_t w[32];
    tm_ld32_wait(tm_warp + (uint32_t)(r * 32), w);
    const int e = own ? 2 * e_col + comp : 0;
    uint32_t v = w[0];
#pragma unroll
    for (int i = 1; i < 32; i++) v = (e == i) ? w[i] : v;
    if (own) s_tap[(dys + 2) * 5 + dxs + 2] = __uint_as_float(v);
  }
}

// What one step of a resident slot needs: its images, where its spectra live, where the result goes.
// Descriptors live in shared memory, two per team: the team's aux warp (a solo warp itself) computes
// the descriptor of step it + 1 while the team works on step it, so that no warp carries the pointers
// in registers through the FFT stages and the address arithmetic (integer divisions, the stream
// flags from global memory) stays off the critical path.
struct StepDesc {
  const uint8_t* img0;      // first image to transform
  const uint8_t* img1;      // second image (correlate pass); == img0 for one-pass steps
  const uint8_t* next_img;  // first image of the slot's next step (L2 hints); img0 when there is none
  f4* spec0;                // pass 0 writes the spectrum here / pass 1 correlates against it
  f4* spec_out;             // pass 1 keeps the current spectrum here (streams) or nullptr
  int out_idx;              // index of the (x, y) result in PcArgs::out
  int s;                    // section inside its frame, i + j*sq
  int flags;                // SD_*
  int pad[3];
};
static_assert(sizeof(StepDesc) == 64, "two descriptors per team fit SM_DESC");
enum { SD_VALID = 1, SD_TWO_PASS = 2, SD_CORRELATE = 4,
       SD_PREV_GLOBAL = 8,   // PC_SEQ: the previous spectrum is the stream cache's (not this slot's own chain)
       SD_OUT_GLOBAL = 16 }; // PC_SEQ: the stream's last frame: its spectrum goes to the stream cache

// Step `it` of resident slot `slot` (flags = 0: the slot has nothing to do in this step).
//   pairs / stream / spectrum modes: section slot + it*nslots (lock-step teams re-do the last section);
//   PC_SEQ, tape cut (seq_C > 0): frame slot*C + it of the sequences laid end to end; segment cut: work item
//     slot + (it / L)*nslots = (stream k, section s, segment j), frame t = j*L + it % L, or a bundle of remainder
//     pieces (PcArgs), frame t = J*L + (it % L) % R of sequence (bundle * Q + (it % L) / R).
//     The first step of a slot / segment / piece that has no usable previous spectrum (j > 0, or the stream's first
//     call: prev := curr, src/FftMethod.cpp:1791-1793) transforms the previous frame itself (two_pass);
//     afterwards the previous spectrum is what the slot wrote one step earlier (ping-pong halves of
//     its scratch); the stream's last frame leaves its spectrum in the stream cache (:1872).
template <int P>
__device__ __forceinline__ void step_fill(const PcArgs& a, int it, int slot, int nslots, bool lockstep, StepDesc& d) {
  typedef PG<P> G;
  const size_t spec_f4 = G::SPEC_F4;
  d.flags = 0;
  d.img0 = d.img1 = d.next_img = nullptr;
  if (a.mode == PC_SEQ) {
    int ks, t, ph;   // sequence, frame, step parity of the slot's own spectrum chain
    bool fresh;      // the slot holds no spectrum of frame t - 1 of this sequence
    if (a.seq_C > 0) {   // tape cut: frames [slot * C, (slot + 1) * C) of the sequences laid end to end
      const int g = slot * a.seq_C + it;
      if (it >= a.seq_C || g >= a.seq_nseq * a.seq_T) return;
      ks = g / a.seq_T;
      t = g - ks * a.seq_T;
      ph = it;
      fresh = it == 0 || t == 0;
    } else {
      const int r = it / a.seq_L, tt = it - r * a.seq_L;
      const int item = slot + r * nslots;
      if (item >= a.n_sections) return;
      const int nmain = a.seq_nseq * a.seq_J;
      int j;
      if (item < nmain) {
        ks = item / a.seq_J;
        j = item - ks * a.seq_J;
        ph = tt;
      } else {   // a bundle: the remainder pieces of seq_Q sequences, one after the other
        const int pidx = tt / a.seq_R;
        ks = (item - nmain) * a.seq_Q + pidx;
        if (pidx >= a.seq_Q || ks >= a.seq_nseq) return;
        j = a.seq_J;
        ph = tt - pidx * a.seq_R;
      }
      t = j * a.seq_L + ph;
      if (t >= a.seq_T) return;
      fresh = ph == 0;
    }
    const int k = ks / a.S, s_ = ks - k * a.S;
    const int sj = s_ / a.sq, si = s_ - sj * a.sq;
    const size_t off = (size_t)k * a.stream_stride + (size_t)(sj * G::N) * a.row_stride + si * G::N;
    const unsigned fl = a.stream_ids ? a.first[k] : a.inline_flags[k];
    const int sid = a.stream_ids ? a.stream_ids[k] : a.inline_ids[k];
    const unsigned par = (fl >> 1) & 1u;
    f4* cache = reinterpret_cast<f4*>(a.spec_cache) + ((size_t)sid * a.S + s_) * 2 * spec_f4;
    f4* X = reinterpret_cast<f4*>(a.seq_scratch) + (size_t)slot * 2 * spec_f4;
    const bool first_call = (t == 0) && (fl & 1u);
    const bool two_pass = fresh && (t > 0 || first_call);
    d.img1 = a.curr + off + (size_t)t * a.frame_stride;
    d.img0 = (two_pass && !first_call) ? d.img1 - a.frame_stride : d.img1;
    d.spec0 = (fresh && !two_pass) ? cache + par * spec_f4 : X + ((ph & 1) ^ 1) * spec_f4;
    d.spec_out = (t == a.seq_T - 1) ? cache + (par ^ 1u) * spec_f4 : X + (ph & 1) * spec_f4;
    d.out_idx = (k * a.seq_T + t) * a.S + s_;
    d.s = s_;
    d.flags = SD_VALID | SD_CORRELATE | (two_pass ? SD_TWO_PASS : 0) | ((fresh && !two_pass) ? SD_PREV_GLOBAL : 0) |
              ((t == a.seq_T - 1) ? SD_OUT_GLOBAL : 0);
    return;
  }
  int sec = slot + it * nslots;
  if (sec >= a.n_sections) {
    if (!lockstep) return;
    sec = a.n_sections - 1;
  }
  const int f = sec / a.S, s_ = sec - f * a.S;
  const int j = s_ / a.sq, i = s_ - j * a.sq;
  const size_t off = (size_t)f * a.frame_stride + (size_t)(j * G::N) * a.row_stride + i * G::N;
  d.out_idx = sec;
  d.s = s_;
  if (a.mode == PC_PAIRS) {
    d.img0 = a.prev + off;
    d.img1 = a.curr + off;
    d.spec0 = reinterpret_cast<f4*>(a.scratch) + (size_t)slot * spec_f4;
    d.spec_out = nullptr;
    d.flags = SD_VALID | SD_TWO_PASS | SD_CORRELATE;
  } else {
    // stream cache: two halves per (stream, section); bit 1 of first[] = half holding the previous
    // frame's spectrum, bit 0 = first call of the stream (imPrev := imCurr, src/FftMethod.cpp:1791-1793)
    const unsigned fl = a.stream_ids ? a.first[f] : a.inline_flags[f];
    const int sid = a.stream_ids ? a.stream_ids[f] : a.inline_ids[f];
    const unsigned par = (fl >> 1) & 1u;
    f4* base = reinterpret_cast<f4*>(a.spec_cache) + ((size_t)sid * a.S + s_) * 2 * spec_f4;
    d.img0 = a.curr + off;
    d.img1 = d.img0;
    d.spec0 = base + par * spec_f4;
    d.spec_out = base + (par ^ 1u) * spec_f4;   // imPrev = imCurr.clone() (:1872)
    d.flags = SD_VALID | ((a.mode == PC_SPECTRUM || (fl & 1u)) ? SD_TWO_PASS : 0) |
              (a.mode != PC_SPECTRUM ? SD_CORRELATE : 0);
  }
}

// descriptor of step `it` into shared memory, with the first image of the step after it (one thread)
template <int P>
__device__ __noinline__ void make_step(const PcArgs& a, int it, int iters, int slot, int nslots, bool lockstep,
                                       StepDesc* out) {
  StepDesc d, dn;
  step_fill<P>(a, it, slot, nslots, lockstep, d);
  d.next_img = nullptr;   // no next step
  if (it + 1 < iters) {
    step_fill<P>(a, it + 1, slot, nslots, lockstep, dn);
    if (dn.flags & SD_VALID) d.next_img = dn.img0;
  }
  *out = d;
}

// ---- OCL-compatible numerics (PcArgs::ocl; cl/FftMethod.cl, "parity unpinned") ---------------------
// After the inverse row stage: scale the surface by 1/N^2 (cl/FftMethod.cl:733), zero every sample
// whose unshifted row or column index lies strictly between SEARCH_RADIUS = 55 and N - 55 (:737-747,
// :826-829), and redo the warp's maximum / first location on what is left (minmaxloc, :1164-1313).
template <int P>
__device__ __noinline__ void ocl_mask_scan(c32* A, int task, int lane, float& wm, int& wloc) {
  typedef PG<P> G;
  constexpr int R = 55;
  const float scale = 1.0f / ((float)G::N * (float)G::N);
  float best = -INFINITY;
  int loc = 0x7fffffff;
  for (int e = lane; e < G::RPT * G::N; e += 32) {
    const int qi = e / G::N, sl = e - qi * G::N;
    const int q = task * G::RPT + qi;
    const int ax = P == 16 ? sl / 15 : sl % P, bx = P == 16 ? sl % 15 : sl / P;
    const int x = crt<P>(ax, bx);
    const int ya = crt<P>(q / 15, q % 15);
    const int yb = ya >= G::M ? ya - G::M : ya + G::M;
    c32 v = A[q * G::SQ + sl];
    const bool mx = x > R && x < G::N - R;
    v.x = (mx || (ya > R && ya < G::N - R)) ? 0.0f : v.x * scale;
    v.y = (mx || (yb > R && yb < G::N - R)) ? 0.0f : v.y * scale;
    A[q * G::SQ + sl] = v;
    const int xs = x >= G::M ? x - G::M : x + G::M;
    const int ia = yb * G::N + xs, ib = ya * G::N + xs;   // fft-shifted row of row a is yb and vice versa
    if (v.x > best || (v.x == best && ia < loc)) { best = v.x; loc = ia; }
    if (v.y > best || (v.y == best && ib < loc)) { best = v.y; loc = ib; }
  }
  wm = warp_max(best);
  wloc = __reduce_min_sync(0xffffffffu, (best == wm && best > -INFINITY) ? loc : 0x7fffffff);
}

// coordinates of tap `tap` (0..48) of the radius-3 window around peak_idx; false outside the image
// (the reference clamps the window to the image, cl/FftMethod.cl:1333-1336)
template <int P>
__device__ __forceinline__ bool ocl_tap_coords(int peak_idx, int tap, int& y, int& x) {
  typedef PG<P> G;
  y = peak_idx / G::N + tap / 7 - 3;
  x = peak_idx % G::N + tap % 7 - 3;
  return !(y < 0 || y >= G::N || x < 0 || x >= G::N);
}

// refine (cl/FftMethod.cl:1315-1379) by ONE thread: positive samples only, FP32 accumulation of frame
// coordinates in row-major order, sum seeded with FLT_EPSILON; result = centroid - (origin + N/2).
// TAPS(t) yields the value of tap t.  peak_idx < 0 (no maximum: all-NaN surface) -> NaN ("LARGE SHIFT",
// src/FftMethod.cpp:461-464).
template <int P, typename TAPS>
__device__ __forceinline__ void ocl_refine(const PcArgs& a, int out_idx, int s_, int peak_idx, TAPS taps) {
  typedef PG<P> G;
  double xy[2] = {(double)NAN, (double)NAN};
  if (peak_idx >= 0) {
    const int x0 = (s_ % a.sq) * G::N, y0 = (s_ / a.sq) * G::N;
    float cx = 0.0f, cy = 0.0f, sum = FLT_EPSILON;
    for (int t = 0; t < 49; t++) {
      int y, x;
      if (!ocl_tap_coords<P>(peak_idx, t, y, x)) continue;
      const float v = taps(t, y, x);
      if (v > 0.0f) {
        cx = __fmaf_rn((float)(x0 + x), v, cx);
        cy = __fmaf_rn((float)(y0 + y), v, cy);
        sum += v;
      }
    }
    cx /= sum;
    cy /= sum;
    const double sx = (double)(cx - (float)(x0 + G::N / 2)), sy = (double)(cy - (float)(y0 + G::N / 2));
    bool valid = !(sx * sx + sy * sy > a.max_px_speed_sq || fabs(sx) > (double)G::N / 2.0 ||
                   fabs(sy) > (double)G::N / 2.0);   // src/FftMethod.cpp:1840-1856, applied to both branches
    if (sx != sx || sy != sy) valid = false;
    if (valid) { xy[0] = sx; xy[1] = sy; }
  }
  a.out[(size_t)out_idx * 2 + 0] = xy[0];
  a.out[(size_t)out_idx * 2 + 1] = xy[1];
}

// result of one section from the centroid sums (lane 0 of the finishing warp)
template <int P>
__device__ __forceinline__ void write_result(const PcArgs& a, int out_idx, double cx, double cy, double sum) {
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
  a.out[(size_t)out_idx * 2 + 0] = xy[0];
  a.out[(size_t)out_idx * 2 + 1] = xy[1];
}

// TM: pairs mode with the first image's spectrum of the regular units in tensor memory (N = 240 / 120)
// SEQTM: PC_SEQ with the spectrum chain of the regular units in tensor memory (N = 240; the correlation surface
// stays in shared memory then: its TMEM columns hold the spectrum across steps)
// EXTRAS: the OCL-compatible epilogue and the window option are compiled in (scratch variants only).  At N = 240
// their code costs the scratch variant 280 B of spills: the stream modes run an instantiation without them
// unless the context asks for either (4.59 -> 5.20 M frames/s in stream mode, profiles/r4q_ab_no_extras.txt).
template <int P, bool VEC, bool TM, bool SEQTM = false, bool EXTRAS = true>
__global__ void __launch_bounds__(pc_threads<P>(), KCfg<P>::MINB) pc_pfa_kernel(const __grid_constant__ PcArgs a) {
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
  StepDesc* s_desc = reinterpret_cast<StepDesc*>(s_scal + K::TEAMS * (SM_WMAX + SM_WLOC + SM_TAP + SM_PIDX)) + 2 * team;
  uint16_t* s_rowy = reinterpret_cast<uint16_t*>(reinterpret_cast<unsigned char*>(s_desc - 2 * team) + K::TEAMS * SM_DESC);
  // Teams of one CTA are independent, but they share the instruction cache.  In pairs mode every
  // team runs the same control flow, so the team barriers are made block barriers: all warps of the
  // SM then stay in the same stretch of this (large) kernel body.  A team without a section of its
  // own re-does the last section (identical result) to keep the barrier count uniform.
  const bool lockstep = K::TEAMS > 1 && a.mode == PC_PAIRS && ((SOLO ? MOF_LOCKSTEP_SOLO : MOF_LOCKSTEP) != 0);
  // N = 60: the 15 one-warp teams keep lock-step inside NG groups of consecutive warps (named barriers 1..NG) instead
  // of one block-wide group; group g starts one pixel fetch + row stage behind group g - 1 (barrier 8 + g, once), so
  // that the groups sit in different stages: while one waits for its pixels the others compute
  constexpr int NG = (SOLO && K::TEAMS == 15) ? MOF_SOLO_GROUPS : 0;
  constexpr bool GROUPS = NG > 1;
  constexpr int GW = GROUPS ? (15 + NG - 1) / NG : 15;           // warps per group (the last one may have fewer)
  const int grp = GROUPS ? (warp / GW < NG - 1 ? warp / GW : NG - 1) : 0;
  const int grp_threads = 32 * (grp == NG - 1 ? 15 - (NG - 1) * GW : GW);
  const int prev_threads = 32 * GW;                              // size of group grp - 1
  auto team_bar = [&]() {
    if (GROUPS && lockstep) asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(grp_threads) : "memory");
    else if (K::TEAMS == 1 || lockstep) __syncthreads();
    else if (SOLO) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "n"(WPT * 32) : "memory");
  };
  // between two stages of one warp task: a warp-level sync, or (N = 60 lock-step experiment) a block barrier
  auto stage_sync = [&]() {
    if (SOLO && MOF_SOLO_SYNC_ALL && lockstep) __syncthreads();
    else __syncwarp();
  };
  // column units: main warps own the regular units of their task, the aux role owns the packed unit;
  // the P-point stages treat both kinds alike, so a SOLO warp runs them over all M units at once
  const int ur0 = task * G::UPT;
  const int nur = (task == TASKS - 1) ? G::UPT - 1 : G::UPT;
  const int uy0 = SOLO ? 0 : (is_main ? ur0 : G::M - 1);
  const int nuy = SOLO ? G::M : (is_main ? nur : 1);

  const int slot = blockIdx.x * K::TEAMS + team;
  const int nslots = gridDim.x * K::TEAMS;
  const int iters = ((a.n_sections + nslots - 1) / nslots) * (a.mode == PC_SEQ ? a.seq_L : 1);
  for (int i = tid; i < G::N; i += (int)blockDim.x) fill_rowy<P>(s_rowy, i);
  if (!SOLO && is_aux && lane == 0) {
    make_step<P>(a, 0, iters, slot, nslots, lockstep, &s_desc[0]);
  }
  if (TM && P == 16 && MOF_SURF_TM && VEC && MOF_TMA_FETCH && tid < TASKS) {
    mbar_init(reinterpret_cast<uint64_t*>(s_tap + 32) + tid, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
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
  // N = 120: the staging area is the team's dedicated buffer: as soon as a warp has picked the pixel words of
  // one image it requests its rows of the NEXT image there (cp.async), a whole section pass ahead of their use
  constexpr bool XSTAGE = pc_xstage_bytes<P>() != 0;
  constexpr bool XSTAGE_ASYNC = XSTAGE && VEC;
  unsigned char* xstage_team = reinterpret_cast<unsigned char*>(s_rowy + G::N) + (size_t)team * pc_xstage_bytes<P>();
  const uint8_t* xstaged = nullptr;   // image whose rows are on their way into this warp's part of the buffer
  unsigned char* stage = XSTAGE ? xstage_team + task * 2 * G::RPT * G::N
                                : reinterpret_cast<unsigned char*>(A + task * G::RPT * G::SQ);

  // N = 240 pairs mode only (see stage_ix16_tm): surface in tensor memory
  constexpr bool SURF_TM = TM && P == 16 && (MOF_SURF_TM != 0) && !SEQTM;
  constexpr bool PV_EARLY = (TM && P == 16 && VEC && (MOF_PV_EARLY != 0)) || (SOLO && VEC && (MOF_SOLO_PV_EARLY & 1));
  constexpr bool PV_NEXT = SOLO && VEC && (MOF_SOLO_PV_EARLY & 2);   // N = 60: next section's first image before the inverse rows
  constexpr bool FETCH_ASYNC = P == 16 && (MOF_FETCH_ASYNC != 0);
  // second image of a pair parked in the columns pass 0 has already consumed (see park_issue)
  constexpr bool PARK2 = TM && P == 16 && VEC && !SEQTM && (MOF_PARK2 != 0) && !MOF_TMA_FETCH && !MOF_COL_MIX && !MOF_PV_EARLY;
  constexpr bool COL_MIX = TM && P == 16 && (MOF_COL_MIX != 0);   // see the column phase below
  // the aux warp as pixel producer: bulk row copies landing on one mbarrier per main warp (upper half of
  // the tap array: the 5x5 window needs 25 of its 64 floats; the 7x7 window never runs in this variant)
  constexpr bool TMA_FETCH = SURF_TM && VEC && (MOF_TMA_FETCH != 0);
  uint64_t* s_mbar = reinterpret_cast<uint64_t*>(s_tap + 32);
  unsigned mbar_phase = 0;
  bool aux_issued0 = false;   // aux warp: the first image of the coming step was requested behind IX16
  bool staged = false;    // the first image of this step is already on its way into the staging area (cp.async)
  bool pv_ready = false;  // pv[] holds the second image's pixels (requested during pass 0's last column stage)

  // the tensor-memory variants never run the OCL-compatible epilogue (launch_one sends it to the scratch
  // variants): its code stays out of the hot kernels' instruction footprint
  const int k_ocl = (TM || !EXTRAS) ? 0 : a.ocl;
  const float* const k_win = (TM || !EXTRAS) ? nullptr : a.win;   // same for the optional Hann window

  CLK_INIT;
  bool pend = false;   // aux warp (team mode): the previous section's centroid is still to be finished
  int pend_out = 0, pend_s = 0;
  // aux warp, after a team barrier: centroid of the previous section from the taps the main warps left
  auto finish_pending = [&]() {
    pend = false;
    if (k_ocl) {
      if (lane == 0) ocl_refine<P>(a, pend_out, pend_s, *s_pidx, [&](int t, int, int) { return s_tap[t]; });
      return;
    }
    double cx = 0.0, cy = 0.0, sum = 0.0;
    int y, x;
    if (lane < 25 && tap_coords<P>(*s_pidx, lane, y, x)) centroid_acc(y, x, s_tap[lane], cx, cy, sum);
    cx = warp_sum(cx);
    cy = warp_sum(cy);
    sum = warp_sum(sum);
    if (lane == 0) write_result<P>(a, pend_out, cx, cy, sum);
  };

  const bool tskew = !SOLO && K::TEAMS == 3 && (MOF_TEAM_SKEW != 0) && (a.mode == PC_PAIRS || MOF_TEAM_SKEW == 2) && !lockstep;
  bool tskew_pending = tskew;
  auto tskew_arrive = [&]() {   // this team's first row stage is done (or it has no work): the next team may start
    tskew_pending = false;
    if (team < K::TEAMS - 1) asm volatile("bar.arrive %0, %1;" ::"r"(9 + team), "n"(2 * WPT * 32) : "memory");
  };
  if (tskew && team) asm volatile("bar.sync %0, %1;" ::"r"(8 + team), "n"(2 * WPT * 32) : "memory");
  bool skew_pending = GROUPS && lockstep;
  // the team / group behind this one may start: called once, after this one's first row stage (MOF_SKEW_AT = 1) or
  // after its first whole pass (2: with three teams or groups they then sit about a third of a section apart)
  auto skew_arrive = [&]() {
    if (tskew_pending) tskew_arrive();
    if (GROUPS && skew_pending) {
      skew_pending = false;
      if (grp < NG - 1) {
        const int next_threads = 32 * (grp + 1 == NG - 1 ? 15 - (NG - 1) * GW : GW);
        asm volatile("bar.arrive %0, %1;" ::"r"(9 + grp), "r"(grp_threads + next_threads) : "memory");
      }
    }
  };
  if (GROUPS && lockstep && grp)   // group g starts a row stage behind group g - 1
    asm volatile("bar.sync %0, %1;" ::"r"(8 + grp), "r"(grp_threads + prev_threads) : "memory");
  // MOF_OPAQUE_STAGES (bit mask over P): every stage call gets the lane index through an empty asm, so that nothing
  // derived from it is hoisted out of the section loop and kept in registers (or spilled) through the other stages
  auto opaque_lane = [&]() { int l = lane; asm volatile("" : "+r"(l)); return l; };
#define OL ((((MOF_OPAQUE_STAGES) & P) != 0) ? opaque_lane() : lane)
  // pixel path of the tensor-memory (pairs) variants at N = 60 / 120: see MOF_OPAQUE_WORDC
#define OLW ((TM && (((MOF_OPAQUE_WORDC) & 12 & P) != 0)) ? opaque_lane() : lane)
  for (int it = 0; it < iters; it++) {
    if (GROUPS && lockstep) team_bar();
    else if (tskew) {}
    else if (K::TEAMS > 1 && (lockstep || MOF_LOCKSTEP != 2)) __syncthreads();   // (stream mode: at least one re-alignment per section)
    const StepDesc& d = s_desc[it & 1];
    if (SOLO) {   // a solo warp is its own aux: descriptor of this step, now
      if (lane == 0) make_step<P>(a, it, iters, slot, nslots, lockstep, &s_desc[it & 1]);
      __syncwarp();
    }
    const int dflags = d.flags;
    if (!(dflags & SD_VALID)) {   // never taken in lockstep mode
      if (K::TEAMS == 1 && a.mode != PC_SEQ) break;
      if (!SOLO) {   // keep the descriptor pipeline going: step it + 1 may have work (PC_SEQ)
        if (is_aux && lane == 0 && it + 1 < iters) make_step<P>(a, it + 1, iters, slot, nslots, lockstep, &s_desc[(it + 1) & 1]);
        team_bar();
      }
      if (tskew_pending) tskew_arrive();
      continue;
    }
    const bool d_correlate = dflags & SD_CORRELATE;
    if (TMA_FETCH && is_aux && !SOLO && !aux_issued0)   // first step: nobody asked for this image yet
      aux_fetch_image<P>(A, (dflags & SD_TWO_PASS) ? d.img0 : d.img1, a.row_stride, s_rowy, s_mbar, lane);

    uint32_t wa[15], wb[15];   // pixel words of the row stage (PARK2: those of pass 1 are picked at the end of pass 0)
    bool parked = false;
#pragma unroll 1
    for (int pass = (dflags & SD_TWO_PASS) ? 0 : 1; pass < 2; pass++) {
      CLK(0);
      // pass 0: spectrum of the first image -> spec0; pass 1: current image, correlate, invert columns
      if (is_main) {
        // rows: pixels -> staging (inside this warp's own rows of A) -> words -> FX15 -> FX16
        if (PARK2 && parked) {
          // the words were picked from the parked rows before the barrier that ended pass 0
        } else if (XSTAGE_ASYNC && xstaged == (pass ? d.img1 : d.img0)) {
          fx_stage_async_wait();            // requested right after the previous image's words were picked
        } else if (TMA_FETCH) {             // the aux warp's bulk copies land on this warp's mbarrier
          mbar_wait(&s_mbar[task], mbar_phase);
          mbar_phase ^= 1u;
        } else if (VEC) {
          if (SURF_TM && staged) {          // issued behind the previous section's IX16
            fx_stage_async_wait();
            staged = false;
          } else if ((PV_EARLY || PV_NEXT) && pv_ready) {  // requested one phase earlier
            fx_stage<P>(stage, OLW, pv);
            pv_ready = false;
          } else {
            // N = 240: cp.async straight into the staging area (no register round trip): 4.45 -> 4.67 M pairs/s;
            // N = 120 / 60 lose 2-3 % with it (measured, profiles/r2b_ab.txt) and keep the register path
            if (FETCH_ASYNC) {
              fx_stage_async<P>(stage, pass ? d.img1 : d.img0, a.row_stride, s_rowy, task, OLW);
              fx_stage_async_wait();
            } else {
              fx_prefetch<P>(pass ? d.img1 : d.img0, a.row_stride, s_rowy, task, OLW, pv);
              fx_stage<P>(stage, OLW, pv);
            }
          }
        }
        else fx_stage_bytes<P>(stage, pass ? d.img1 : d.img0, a.row_stride, s_rowy, task, lane);
        stage_sync();
        CLK(15);
        if (PARK2 && parked) parked = false;
        else if (XSTAGE) stage_fx15_read_from<P>(stage, task, OLW, wa, wb);
        else stage_fx15_read<P>(A, task, OLW, wa, wb);
        stage_sync();
        if (XSTAGE_ASYNC) {   // this warp's part of the buffer is free again: the image of the next row stage
          xstaged = (pass == 0 && d_correlate) ? d.img1 : d.next_img;
          if (xstaged) fx_stage_async<P>(stage, xstaged, a.row_stride, s_rowy, task, OLW);
        }
        CLK(16);
        stage_fx15<P>(A, task, OL, wa, wb, k_win, a.k23);
        stage_sync();
        CLK(1);
        stage_x16<P, 1, false>(A, task, OL, nullptr);
        CLK(2);
      }
      team_bar();   // rows done
      if (MOF_SKEW_AT == 1) skew_arrive();
      CLK(3);
      if (!SOLO && is_aux) {
        if (pend) finish_pending();
        // descriptor of the next step (once per step: in its last pass); its readers are two team barriers away
        if (pass == 1 || !d_correlate) {
          if (lane == 0 && it + 1 < iters) make_step<P>(a, it + 1, iters, slot, nslots, lockstep, &s_desc[(it + 1) & 1]);
          __syncwarp();
#if MOF_SPEC_PREFETCH
          // stream mode: the previous spectrum of the NEXT step comes from the stream cache in HBM (thousands of
          // streams x 230 KB do not stay in L2); ask for it a whole step ahead so that the cross-power stage
          // finds it in L2
          if (!TM && a.mode == PC_STREAM && it + 1 < iters) {
            const StepDesc& dn = s_desc[(it + 1) & 1];
            if ((dn.flags & (SD_VALID | SD_CORRELATE | SD_TWO_PASS)) == (SD_VALID | SD_CORRELATE)) {
              const char* sp = reinterpret_cast<const char*>(dn.spec0);
              for (int off = lane * 128; off < (int)(sizeof(f4) * G::SPEC_F4); off += 32 * 128)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(sp + off));
            }
          }
#endif
        }
      }
      // L2 hints for the next image one stage ahead: the second image of this section, or
      // (spectrum only) the first image of this team's next section
      // team mode: the aux warp (idle through the row stages, short column stage) issues the hints
      // for every task (+1.2 % at N = 240, +1.8 % at N = 120 over each main warp hinting its own
      // rows); a solo warp is its own aux
      // (SURF_TM requests the next section's first image behind IX16 already: its hints go out one
      // stage earlier, after the row stage of pass 1)
      if (XSTAGE_ASYNC) {
        // (N = 120: every main warp requests its own rows of the next image itself, see the row stage)
      } else if (VEC && is_aux && (pass == 0 || SURF_TM)) {
        const uint8_t* nx = (pass == 0 && d_correlate) ? d.img1 : d.next_img;
        if (nx) {
          if (SOLO) fx_prefetch_l2<P>(nx, a.row_stride, s_rowy, task, lane, a.mode == PC_PAIRS);
          else
            for (int t = 0; t < TASKS; t++) fx_prefetch_l2<P>(nx, a.row_stride, s_rowy, t, lane, a.mode == PC_PAIRS);
        }
      }
      if (SOLO && PV_EARLY && pass == 0 && d_correlate) {   // N = 60: flies through both column stages of pass 0
        fx_prefetch<P>(d.img1, a.row_stride, s_rowy, task, OLW, pv);
        pv_ready = true;
      }
      // column stages on this warp's units
      CLK(4);
      if (COL_MIX && is_main && ((wl >> 2) & 1)) {
        // Every second group of four main warps walks its units in two halves, each half through ALL of its
        // column stages before the next one starts (16-point, 15-point[, inverse 16-point]).  The 16-point
        // stages are bound by shared-memory wavefronts, the 15-point stages by the FP32 pipe; with all warps
        // in the same stage each pipe idles while the other limits.  Now two of an SM sub-partition's four
        // warps are in the other kind of stage for most of the column phase.  (Units own disjoint columns and
        // tensor-memory slots: the order inside a warp is free.)
#pragma unroll 1
        for (int h = 0; h < 2; h++) {
          const int hu0 = ur0 + 4 * h, hnu = h ? nur - 4 : 4;   // 15-point rounds take four units each
          stage_y16<P, false>(A, hu0, hnu, OL);
          __syncwarp();
          if (pass == 0) {
            stage_fy15_store_tm<P>(A, hu0, hnu, OL, tm_warp, h);
          } else {
            stage_fy15_corr_tm<P>(A, hu0, hnu, OL, tm_warp, 0, h);
            __syncwarp();
            stage_y16<P, true>(A, hu0, hnu, OL);
          }
          __syncwarp();
        }
        CLK(8);
      } else {
        stage_y16<P, false>(A, uy0, nuy, OL);
        stage_sync();
        CLK(5);
        if (pass == 0) {
          if (TM) {   // regular units -> tensor memory, packed unit -> global scratch
            if (!SOLO && PV_EARLY && is_main && d_correlate) {   // the loads fly during the 15-point stage and the barrier after it
              fx_prefetch<P>(d.img1, a.row_stride, s_rowy, task, OLW, pv);
              pv_ready = true;
            }
            if (PARK2) {
              // (every path ends with a definition of wa / wb: nothing of the row stage's words stays live through
              // the column stages for the sake of the loop-carried registers)
              if (is_main) {
                if (d_correlate) {
                  stage_fy15_store_tm<P>(A, ur0, nur, OL, tm_warp, 0, d.img1, a.row_stride, s_rowy, task);
                  fx_stage_async_wait();
                  __syncwarp();
                  park_read(A, ur0, lane, wa, wb);
                  parked = true;
                } else {
                  stage_fy15_store_tm<P>(A, ur0, nur, OL, tm_warp);
                  words_undef(wa, wb);
                }
              } else {
                stage_fy15_store<P>(A, G::M - 1, 1, OL, d.spec0);
                words_undef(wa, wb);
              }
            } else {
              if (is_main) stage_fy15_store_tm<P>(A, ur0, nur, OL, tm_warp);
              if (is_aux) stage_fy15_store<P>(A, G::M - 1, 1, OL, d.spec0);
            }
          } else {
            stage_fy15_store<P>(A, uy0, nuy, OL, d.spec0);
          }
          CLK(6);
        } else {
          if (SEQTM) {
            if (is_main)
              stage_fy15_corr_seqtm<P>(A, ur0, nur, OL, tm_warp, (dflags & SD_PREV_GLOBAL) ? d.spec0 : nullptr,
                                       (dflags & SD_OUT_GLOBAL) ? d.spec_out : nullptr);
          } else if (TM) {
            if (is_main) stage_fy15_corr_tm<P>(A, ur0, nur, OL, tm_warp, 0);
          } else if (is_main) {
            stage_fy15_corr<P>(A, ur0, nur, OL, d.spec0, d.spec_out, k_ocl);
          }
          if (is_aux) {
            sp_dft15<P, 1>(A, OL);
            __syncwarp();
            sp_cross_power<P>(A, OL, d.spec0, d.spec_out, k_ocl);
            __syncwarp();
            sp_dft15<P, -1>(A, OL);
          }
          stage_sync();
          CLK(7);
          stage_y16<P, true>(A, uy0, nuy, OL);
          CLK(8);
        }
      }
      team_bar();   // pass 0: A free again, spectrum visible to the team; pass 1: columns done
      if (MOF_SKEW_AT == 2) skew_arrive();
      if (TMA_FETCH && is_aux && !SOLO && pass == 0 && d_correlate)
        aux_fetch_image<P>(A, d.img1, a.row_stride, s_rowy, s_mbar, lane);
      CLK(9);
      if (!d_correlate) break;
    }
    if (!d_correlate) continue;

    // inverse rows with a running maximum, then each warp's own maximum and its first location in
    // fft-shifted row-major order (cv::minMaxLoc)
    float wm = -INFINITY;
    int wloc = 0x7fffffff;
    if (is_main) {
      MaxTrack mt;
      mt.vmax = -INFINITY;
      mt.which = 0;
      mt.multi = false;
      if (PV_NEXT && d.next_img) {
        fx_prefetch<P>(d.next_img, a.row_stride, s_rowy, task, OLW, pv);
        pv_ready = true;
      } else if (VEC && SOLO && d.next_img)
        fx_prefetch_l2<P>(d.next_img, a.row_stride, s_rowy, task, lane);
      CLK(10);
      stage_ix15<P>(A, task, OL);
      stage_sync();
      CLK(11);
      if (SURF_TM) {
        const uint8_t* nx = VEC ? d.next_img : nullptr;
        stage_ix16_tm<16>(A, task, OL, tm_warp, wm, wloc, nx, a.row_stride, s_rowy, TMA_FETCH ? nullptr : stage);
        staged = nx != nullptr;
        CLK(12);
      } else {
        stage_x16<P, -1, true>(A, task, OL, &mt);
        CLK(12);
        wm = warp_max(mt.vmax);
        __syncwarp();
        wloc = __reduce_min_sync(0xffffffffu, search_lane<P>(A, task, lane, mt, wm));
      }
      if (!SURF_TM && k_ocl) {
        __syncwarp();
        ocl_mask_scan<P>(A, task, lane, wm, wloc);
        __syncwarp();
      }
    }
    else if (TMA_FETCH) {   // aux warp: every main warp has consumed its first two row pairs -> next section's first image
      asm volatile("bar.sync 3, %0;" ::"n"(K::TEAMS * WPT * 32) : "memory");
      aux_issued0 = d.next_img != nullptr;
      if (aux_issued0) aux_fetch_image<P>(A, d.next_img, a.row_stride, s_rowy, s_mbar, lane);
    }
    else if (VEC && !SOLO && !SURF_TM) {   // aux warp: L2 hints for the first image of this team's next section
      const uint8_t* nx = d.next_img;
      if (nx)
        for (int t = 0; t < TASKS; t++) fx_prefetch_l2<P>(nx, a.row_stride, s_rowy, t, lane, a.mode == PC_PAIRS);
    }
    if (SOLO) {
      if (k_ocl) {
        if (lane == 0)
          ocl_refine<P>(a, d.out_idx, d.s, wloc == 0x7fffffff ? -1 : wloc,
                        [&](int, int y, int x) { return surf_at<P>(A, y, x); });
      } else {
        const int pidx = wloc == 0x7fffffff ? 0 : wloc;   // all-NaN surface: minMaxLoc leaves (0,0)
        double cx = 0.0, cy = 0.0, sum = 0.0;
        if (lane < 25) centroid_tap<P>(A, pidx, lane, cx, cy, sum);
        cx = warp_sum(cx);
        cy = warp_sum(cy);
        sum = warp_sum(sum);
        if (lane == 0) write_result<P>(a, d.out_idx, cx, cy, sum);
      }
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
        int y, x;
        if (SURF_TM) {
          if (pidx == 0x7fffffff) pidx = 0;
          taps_from_tm<16>(task, lane, tm_warp, pidx, s_tap, s_rowy);
        } else if (k_ocl) {   // 7 x 7 window; no maximum at all (NaN surface) is kept as -1
          if (pidx == 0x7fffffff) pidx = -1;
          for (int t = lane; t < 49 && pidx >= 0; t += 32)
            if (ocl_tap_coords<P>(pidx, t, y, x) && surf_rowpair<P>(y) / G::RPT == task) s_tap[t] = surf_at<P>(A, y, x);
        } else {
          if (pidx == 0x7fffffff) pidx = 0;
          if (lane < 25 && tap_coords<P>(pidx, lane, y, x) && surf_rowpair<P>(y) / G::RPT == task)
            s_tap[lane] = surf_at<P>(A, y, x);
        }
        if (wl == 0 && lane == 0) *s_pidx = pidx;
        __syncwarp();
      } else {
        pend = true;
        pend_out = d.out_idx;
        pend_s = d.s;
      }
    }
  }
  if (!SOLO) {
    team_bar();
    if (is_aux && pend) finish_pending();
  }
  if (TM) {   // every read of tensor memory precedes the barrier above (TM implies lockstep block barriers)
    __syncwarp();
    if (K::TEAMS > 1 && (!lockstep || GROUPS)) __syncthreads();   // free-running teams / groups: everybody is done with tensor memory
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
  if (P == 16) {
    e = cudaFuncSetAttribute(pc_pfa_kernel<16, true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)pc_smem_bytes<16>());
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(pc_pfa_kernel<16, true, false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)pc_smem_bytes<16>());
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(pc_pfa_kernel<16, false, false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)pc_smem_bytes<16>());
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
#if !defined(MOF_ONLY_P16)
  if ((e = setup_one<4>()) != cudaSuccess) return (int)e;
  if ((e = setup_one<8>()) != cudaSuccess) return (int)e;
#endif
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
#if !defined(MOF_ONLY_P16)
    case 60: return resident_one<4>(device);
    case 120: return resident_one<8>(device);
#endif
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
  const bool vec = (((size_t)args.curr | (size_t)args.prev | (size_t)args.row_stride | (size_t)args.frame_stride |
                     (args.mode == PC_SEQ ? args.stream_stride : (size_t)0)) & (size_t)(P - 1)) == 0;
  // (the OCL-compatible epilogue works on the surface in shared memory, and the window option is compiled out
  // of the tensor-memory variants to keep their instruction footprint small: scratch path for both)
  const bool tm = KCfg<P>::TM && vec && args.mode == PC_PAIRS && !args.ocl && !args.win && !getenv("MOF_NO_TMEM");
  // mof_fft_stream_push_many at N = 240: the spectrum chain of a sequence stays in tensor memory
  const bool seqtm = P == 16 && vec && args.mode == PC_SEQ && !args.ocl && !args.win && !getenv("MOF_NO_TMEM");
  if (seqtm) pc_pfa_kernel<16, true, true, true><<<grid, pc_threads<16>(), pc_smem_bytes<16>(), stream>>>(args);
  else if (tm) pc_pfa_kernel<P, true, KCfg<P>::TM><<<grid, pc_threads<P>(), pc_smem_bytes<P>(), stream>>>(args);
  else if (P == 16 && !args.ocl && !args.win && !getenv("MOF_EXTRAS_ALWAYS")) {   // scratch variants without the epilogue / window code
    if (vec) pc_pfa_kernel<16, true, false, false, false><<<grid, pc_threads<16>(), pc_smem_bytes<16>(), stream>>>(args);
    else pc_pfa_kernel<16, false, false, false, false><<<grid, pc_threads<16>(), pc_smem_bytes<16>(), stream>>>(args);
  }
  else if (vec) pc_pfa_kernel<P, true, false><<<grid, pc_threads<P>(), pc_smem_bytes<P>(), stream>>>(args);
  else pc_pfa_kernel<P, false, false><<<grid, pc_threads<P>(), pc_smem_bytes<P>(), stream>>>(args);
}

int pc_launch(int N, const PcArgs& args, int grid, cudaStream_t stream) {
  if (grid < 1) return (int)cudaErrorInvalidValue;
  switch (N) {
#if !defined(MOF_ONLY_P16)
    case 60: launch_one<4>(args, grid, stream); break;
    case 120: launch_one<8>(args, grid, stream); break;
#endif
    case 240: launch_one<16>(args, grid, stream); break;
    default: return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}

}  // namespace mof

// -------------------------------------------------------------------------------------------------
// mof_fft_pairs_noalloc_device: the pairs-mode launch for callers that cannot allocate (CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels as mof_fft_pairs_device (pc_pfa_kernel, pairs mode);
// the only difference is where the per-CTA spectrum scratch lives: in the file-scope array below instead
// of a context's cudaMalloc'ed buffer.  No context either (mof_create allocates): geometry comes as
// arguments.
//   * The array is MOF_NOALLOC_SCRATCH_BYTES = 36 MiB of device memory on every GPU that loads the
//     library, held until the process exits: one section spectrum (240 KB at N = 240, 60 KB at N = 120,
//     15 KB at N = 60) per resident CTA slot of a 148-SM B200 (148 / 444 / 2220 slots = 34.7 / 26.0 /
//     32.5 MiB).  A launch whose slots need more (a GPU with more SMs) is rejected, not shrunk.
//   * Every call on a device shares the array.  Nothing in it has to start at zero: a slot's spectrum is
//     written by pass 0 of a section before pass 1 reads it, so the previous call's contents are never
//     read.  Calls that OVERLAP on the GPU (two streams, two host threads) race on it: serialise them.
//   * In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the
//     kernel gets comes from cudaGetSymbolAddress, once per device, with that device current.
// -------------------------------------------------------------------------------------------------
#include <mutex>
#include <stdarg.h>
#include <stdio.h>

#include "../../include/mof.h"

static __device__ __align__(256) unsigned char g_noalloc_scratch[MOF_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int NOALLOC_MAX_DEVICES = 64;
struct NoallocDev {
  bool ready = false;
  void* scratch = nullptr;       // device address of g_noalloc_scratch on this device
  int slots[3] = {0, 0, 0};      // resident slots for N = 60 / 120 / 240
};
NoallocDev g_noalloc_dev[NOALLOC_MAX_DEVICES];
std::mutex g_noalloc_mu;

int noalloc_fail(char* err, size_t err_len, int code, const char* fmt, ...) {
  if (err && err_len) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, err_len, fmt, ap);
    va_end(ap);
  }
  return code;
}

}  // namespace

extern "C" int mof_fft_pairs_noalloc_device(int device, const uint8_t* d_curr, const uint8_t* d_prev,
                                            size_t pair_stride, int row_stride, int frame_size,
                                            int sample_point_size, double max_px_speed, int n_pairs,
                                            double* d_shifts_xy, void* cuda_stream, char* err, size_t err_len) {
  if (err && err_len) err[0] = 0;
  if (device < 0 || device >= NOALLOC_MAX_DEVICES)
    return noalloc_fail(err, err_len, MOF_ERR_INVALID, "device %d out of range (0..%d)", device, NOALLOC_MAX_DEVICES - 1);
  // geometry fix-ups of FftMethod::FftMethod (src/FftMethod.cpp:1706-1719), as mof_create does them
  int F = frame_size, N = sample_point_size;
  if (F <= 0 || N <= 0) return noalloc_fail(err, err_len, MOF_ERR_INVALID, "frame_size and sample_point_size must be positive");
  if (F % 2 == 1) F--;
  if (F % N != 0) N = F;
  const int ni = N == 60 ? 0 : (N == 120 ? 1 : (N == 240 ? 2 : -1));
  if (ni < 0)
    return noalloc_fail(err, err_len, MOF_ERR_INVALID,
                        "section size %d: the no-allocation entry point runs the prime-factor kernel only (60, 120 or 240); "
                        "other sizes need an N x N complex scratch per CTA that the %d-byte static array is not sized for",
                        N, MOF_NOALLOC_SCRATCH_BYTES);
  if (n_pairs < 0 || !d_curr || !d_prev || !d_shifts_xy || row_stride < F || pair_stride < (size_t)row_stride * (F - 1) + F)
    return noalloc_fail(err, err_len, MOF_ERR_INVALID, "bad argument (null pointer, negative count, or strides smaller than the frame)");
  const int sq = F / N, S = sq * sq;
  if ((long long)n_pairs * S > 0x7fffffffLL) return noalloc_fail(err, err_len, MOF_ERR_INVALID, "too many sections");
  if (n_pairs == 0) return MOF_OK;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return noalloc_fail(err, err_len, MOF_ERR_CUDA, "cudaSetDevice(%d) failed: %s", device, cudaGetErrorString(e));
  NoallocDev& d = g_noalloc_dev[device];
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!d.ready) {   // once per device, with that device current
      int rc = mof::pc_setup(device);
      if (rc) return noalloc_fail(err, err_len, MOF_ERR_CUDA, "kernel set-up failed: %s", cudaGetErrorString((cudaError_t)rc));
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
      if (e != cudaSuccess || !p)
        return noalloc_fail(err, err_len, MOF_ERR_CUDA, "cudaGetSymbolAddress(scratch) failed: %s", cudaGetErrorString(e));
      const int sizes[3] = {60, 120, 240};
      for (int i = 0; i < 3; i++) {
        d.slots[i] = mof::pc_resident_ctas(sizes[i], device);
        if (d.slots[i] <= 0) return noalloc_fail(err, err_len, MOF_ERR_CUDA, "occupancy query failed for N=%d", sizes[i]);
      }
      d.scratch = p;
      d.ready = true;
    }
  }
  const int n_sections = n_pairs * S;
  const int slots = n_sections < d.slots[ni] ? n_sections : d.slots[ni];
  const size_t need = (size_t)slots * mof::pc_spectrum_bytes(N);
  if (need > (size_t)MOF_NOALLOC_SCRATCH_BYTES)
    return noalloc_fail(err, err_len, MOF_ERR_INVALID,
                        "this launch needs %zu bytes of spectrum scratch (%d resident slots x %zu bytes at N=%d); the static "
                        "array holds %d: use mof_fft_pairs_device with a context on this GPU",
                        need, slots, mof::pc_spectrum_bytes(N), N, MOF_NOALLOC_SCRATCH_BYTES);
  mof::PcArgs a{};
  a.curr = d_curr; a.prev = d_prev; a.frame_stride = pair_stride; a.row_stride = row_stride;
  a.sq = sq; a.S = S; a.n_sections = n_sections; a.win = nullptr;
  a.scratch = d.scratch; a.out = d_shifts_xy;
  a.max_px_speed_sq = max_px_speed * max_px_speed;
  a.mode = mof::PC_PAIRS; a.post = mof::PC_POST_SHIFT; a.ocl = 0;
  const int rc = mof::pc_launch(N, a, slots, (cudaStream_t)cuda_stream);
  if (rc) return noalloc_fail(err, err_len, MOF_ERR_CUDA, "phase-correlation kernel launch failed: %s", cudaGetErrorString((cudaError_t)rc));
  return MOF_OK;
}
