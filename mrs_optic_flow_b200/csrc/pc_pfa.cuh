// pc_pfa.cuh -- twiddle-free (prime-factor) section pipeline for phase correlation, written as
// warp-task "stage" functions.  Host+device: tests/emu/emu_pfa.cpp runs the very same functions lane by
// lane on the CPU, so every index map is verified against the oracle without a GPU.
//
// What it computes (reference: FftMethod::processImage CPU branch, src/FftMethod.cpp:1829-1856,
// i.e. shift = -cv::phaseCorrelate(curr_roi, prev_roi) per N x N section):
//   2-D real FFT of both sections -> P = F_curr * conj(F_prev) -> C = P*|P| / (|P|^2 + FLT_EPSILON)
//   -> unscaled inverse -> fftshift -> first-maximum argmax -> clamped 5x5 FP64 centroid.
//
// Index maps.  N = 15*P with P in {4, 8, 16} coprime to 15, so a length-N DFT is EXACTLY a
// P x 15 two-dimensional DFT with no twiddle factors between the two stages (Good-Thomas):
//   space:      x <-> (ax, bx) = (x mod P, x mod 15)           x = (ax*EP + bx*16) mod N
//   frequency:  k <-> (ka, kb),  k = (15*ka + P*kb) mod N      -k <-> (-ka mod P, -kb mod 15)
//   X(ka, kb) = sum_{ax,bx} x(ax, bx) W_P^(ax ka) W_15^(bx kb)
// and the 15-point stage is itself a 3 x 5 prime-factor butterfly in registers (dft15).  The same
// maps are used along y.  DC is (0,0); the Nyquist frequency N/2 is (P/2, 0).
//
// Shared-memory section array A[M][SQ] of complex floats, M = N/2 row pairs, SQ = N+1:
//   row pair q = ayp*15 + by (ayp < P/2) packs image rows y(ayp, by) and y + N/2 = y(ayp + P/2, by)
//   as z = row_a + i*row_b;  column slot of (ax, bx) (space) or (ka, kb) (frequency) = slot<P>():
//   a*15 + b at N = 240, b*P + a at N = 120 / 60.
// Every stage is in place: an item writes exactly the slots it read, so stages of one warp task
// only need a warp-level sync between them.  Slot layout, odd row pitch and the per-size lane maps
// (*_item functions) are chosen so that the 16 lanes of a half warp hit 16 different 8-byte banks
// (tools/bank_check.py): at N = 240 a half warp walks the 16 butterflies of one row, at N = 120 /
// 60 it walks 15-16 different rows (or column units) at one fixed slot.
//
// Column units.  The half spectrum is processed in M "units": M-1 regular units own a column k
// and its mirror -k (slots kcol, mcol); the last unit packs the two purely real columns 0 and N/2.
// Regular units carry spectra scaled by 2 (the 1/2 of the real-pair un-packing is skipped); the
// normalised cross-power is invariant to that when FLT_EPSILON is scaled by 16, which it is.
#pragma once

#include <float.h>
#include <math.h>
#include <stdint.h>

#include "fft_regs.cuh"

#if defined(__CUDACC__)
#define MOF_ALIGN16 __align__(16)
#else
#define MOF_ALIGN16 alignas(16)
#endif

#if !defined(MOF_OPAQUE_WORDC)
#define MOF_OPAQUE_WORDC 12   // bit mask over P: lane-derived addresses of the pixel path recomputed per stage instead of hoisted and spilled
                              // (N = 60 / 120 pairs variants: +1 .. 4 %, profiles/r4l_*; the stream variants lose 4 % with it and keep the hoisting; 16 goes with MOF_PARK2, alone -1.2 %)
#endif
#if !defined(MOF_PREFETCH_L1)
#define MOF_PREFETCH_L1 8   // bit mask over P (4 | 8 | 16): the aux warp pulls the next image into L1 with real loads (N = 120, F = 480: +2.9 %, profiles/r4i_*)
#endif
#if !defined(MOF_Y16_PIPELINE)
#define MOF_Y16_PIPELINE 0   // (1: inverse, 2: both) same for the 16-point column stages
#endif
#if !defined(MOF_X16_PIPELINE)
#define MOF_X16_PIPELINE 0   // (measured: -2.6 %, profiles/r2h_ab.txt) N = 240: loads of the next round of a 16-point row stage issued before this round's butterfly
#endif

namespace mof {
namespace pfa {

template <int P>
struct PG {
  static constexpr int N = 15 * P;
  static constexpr int M = N / 2;
  static constexpr int HP = P / 2;
  static constexpr int SQ = N + 1;
  static constexpr int A_ELEMS = M * SQ;
  static constexpr int UNITS = M;
  static constexpr int ITEMS = UNITS * HP;      // (unit, j) items of the 15-point column stage
  static constexpr int SPEC_F4 = 16 * ITEMS;    // float4 per section spectrum: [half 2][t 8][item]
  static constexpr int PINV = (P == 16) ? 1 : (P == 8 ? 2 : 4);     // P^-1 mod 15
  static constexpr int EP = (P == 16) ? 225 : (P == 8 ? 105 : 45);  // 15 * (15^-1 mod P)
  static constexpr int G4 = P / 4;              // 4-byte pixel words per P pixels
  // warp-task geometry: a task = RPT row pairs (row stages) or UPT units (column stages)
  static constexpr int RPT = (P == 16) ? 8 : (P == 8 ? 15 : 30);
  static constexpr int UPT = RPT;
  static constexpr int TASKS = M / RPT;         // 15, 4, 1
};

// column slot of (a, b) inside a row of A
template <int P>
MOF_HD int slot(int a, int b) {
  return P == 16 ? a * 15 + b : b * P + a;
}
// distance between the slots of a and a+1 (fixed b) / of b and b+1 (fixed a)
template <int P>
struct SlotStep {
  static constexpr int A = (P == 16) ? 15 : 1;
  static constexpr int B = (P == 16) ? 1 : P;
};

// image row of (ay, by) / pixel column of (ax, bx)
template <int P>
MOF_HD int crt(int a, int b) {
  return (a * PG<P>::EP + b * 16) % PG<P>::N;
}

// permutation of the column index inside a unit group so that the lanes of a half warp hit
// distinct banks in the 15-point column stage
template <int P>
MOF_HD int ka_perm(int i) {
  if (P == 16) return ((i & 1) << 3) | (i >> 1);
  return i;  // N = 120 / 60: a half warp holds 15-16 different units, no permutation needed
}

// slots of unit u: regular u < M-1 -> (kcol, mcol) = slots of k and -k; u = M-1 -> (0, Nyquist)
template <int P>
MOF_HD void unit_cols(int u, int& kcol, int& mcol) {
  typedef PG<P> G;
  int ka, kb;
  if (u < 7 * P) {
    kb = 1 + u / P;
    ka = ka_perm<P>(u % P);
  } else if (u < G::M - 1) {
    kb = 0;
    ka = u - 7 * P + 1;
  } else {
    kcol = slot<P>(0, 0);
    mcol = slot<P>(G::HP, 0);
    return;
  }
  kcol = slot<P>(ka, kb);
  mcol = slot<P>((P - ka) % P, (15 - kb) % 15);
}

// butterfly index of (j, half) in the 15-point column stage: half 0 -> va = j, half 1 -> its
// partner -j (j = 0 pairs with P/2)
template <int P>
MOF_HD int va_of(int j, int half) {
  return half ? ((j == 0) ? PG<P>::HP : P - j) : j;
}
// element offset (in c32) of butterfly va of a unit, by = 0: rows (va mod P/2)*15 + by
template <int P>
MOF_HD int va_base(int va, int kcol, int mcol) {
  typedef PG<P> G;
  return ((va % G::HP) * 15) * G::SQ + ((va < G::HP) ? kcol : mcol);
}

// item index of (unit, butterfly pair j) in a section spectrum: the lanes of a half warp must land
// on consecutive items (N = 240: consecutive j of a unit; N = 120 / 60: consecutive units at one j)
template <int P>
MOF_HD int item_of(int u, int j) {
  return P == 16 ? u * PG<P>::HP + j : j * PG<P>::UNITS + u;
}
// spectrum (scratch / stream cache) addressing: c32 index of bin vb of an item, half
template <int P>
MOF_HD size_t spec_c32(int item, int half, int vb) {
  return 2 * ((size_t)(half * 8 + (vb >> 1)) * PG<P>::ITEMS + item) + (vb & 1);
}

MOF_HD int imin(int a, int b) { return a < b ? a : b; }

struct MOF_ALIGN16 f4 {
  float x, y, z, w;
};

// ------------------------------------------------------------------------------------------------
// FX: forward rows.  Pixels of the two image rows of a row pair are fetched as 4-byte words:
// word (g, t) covers pixels x = 4g + P*t + {0..3}, i.e. the four butterflies ax = 4g + j, and
// pixel j of the word has bx = (s + j) mod 15 with s = (4g + P*t) mod 15.  Words are kept in
// registers indexed by s, which makes every later register index a compile-time constant.
// ------------------------------------------------------------------------------------------------
// pixel-column offset of word s for word group g: 4g + P*t(s), t(s) = (s - 4g) * P^-1 mod 15
template <int P>
MOF_HD int word_off(int s, int c) {
  int t = (s * PG<P>::PINV) % 15 + c;
  t = t >= 15 ? t - 15 : t;
  return P * t;
}
template <int P>
MOF_HD int word_c(int g) {
  return (15 - (4 * g * PG<P>::PINV) % 15) % 15;
}

// Pixel path.  A warp task needs the 2*RPT image rows of its row pairs (N bytes each).  They are
// fetched from global memory one stage ahead with fully coalesced VW-byte vector loads (VW = P:
// 16/8/4 bytes, 15 vectors per row, element e = 32*i + lane of the task's [row][vector] list) and
// parked in registers; when the row stage starts they are dropped into the task's own (still
// unused) part of A, and every thread picks its 2 x 15 words from there.
template <int P>
struct LoadCfg {
  static constexpr int VW = P;                              // bytes per vector load
  static constexpr int W = VW / 4;                          // 32-bit registers per vector
  static constexpr int NE = 2 * PG<P>::RPT * 15;            // vectors per task
  static constexpr int NV = (NE + 31) / 32;                 // vectors per lane
  static constexpr int NREG = NV * W;
};

// rowy[2q + h] = image row of half h of row pair q
template <int P>
MOF_HD void fill_rowy(uint16_t* rowy, int idx) {
  typedef PG<P> G;
  const int q = idx >> 1;
  const int ya = crt<P>(q / 15, q % 15);
  rowy[idx] = (uint16_t)((idx & 1) ? (ya >= G::M ? ya - G::M : ya + G::M) : ya);
}

template <int P>
MOF_HD void fx_prefetch(const uint8_t* sec, int row_stride, const uint16_t* rowy, int task, int lane,
                        uint32_t (&v)[LoadCfg<P>::NREG]) {
  typedef LoadCfg<P> L;
#if defined(__CUDA_ARCH__) && MOF_OPAQUE_WORDC
  if (((MOF_OPAQUE_WORDC) & 16 & P) != 0) asm volatile("" : "+r"(lane));
#endif
#pragma unroll
  for (int i = 0; i < L::NV; i++) {
    const int e = i * 32 + lane;
    if (e < L::NE) {
      const int r = e / 15, c = e - 15 * r;
      const uint8_t* p = sec + (size_t)rowy[task * 2 * PG<P>::RPT + r] * row_stride + c * L::VW;
      const uint32_t* p32 = reinterpret_cast<const uint32_t*>(p);
      if (L::W == 4) {
#if defined(__CUDA_ARCH__)
        const uint4 t = *reinterpret_cast<const uint4*>(p);
        v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
#else
        for (int k = 0; k < 4; k++) v[4 * i + k] = p32[k];
#endif
      } else if (L::W == 2) {
#if defined(__CUDA_ARCH__)
        const uint2 t = *reinterpret_cast<const uint2*>(p);
        v[2 * i] = t.x; v[2 * i + 1] = t.y;
#else
        for (int k = 0; k < 2; k++) v[2 * i + k] = p32[k];
#endif
      } else {
        v[i] = p32[0];
      }
    } else {
      // full definition of the register set (keeps the previous contents from staying live)
#pragma unroll
      for (int k = 0; k < L::W; k++) v[L::W * i + k] = 0;
    }
  }
}

// variant of fx_prefetch that only warms L2 (no registers held): one hint per 128-byte line
template <int P>
MOF_HD void fx_prefetch_l2(const uint8_t* sec, int row_stride, const uint16_t* rowy, int task, int lane, bool l1 = false) {
#if defined(__CUDA_ARCH__)
  // 2*RPT rows of N bytes: two hints per row (N <= 240 < 256) cover it whatever the alignment
  for (int e = lane; e < 2 * PG<P>::RPT * 2; e += 32) {
    const int r = e >> 1;
    const uint8_t* p = sec + (size_t)rowy[task * 2 * PG<P>::RPT + r] * row_stride + (e & 1) * (PG<P>::N - 1);
#if MOF_PREFETCH_L1
    if (((MOF_PREFETCH_L1) & P) != 0 && l1) {   // bit mask over P = 4 / 8 / 16; pairs mode only (streams: -0.6 %)
      // a real load whose value nobody reads: unlike the hint (measured: no effect) it allocates the line in L1
      unsigned dummy;
      asm volatile("ld.global.ca.u8 %0, [%1];" : "=r"(dummy) : "l"(p));
    }
    else
#endif
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
  }
#else
  (void)sec; (void)row_stride; (void)rowy; (void)task; (void)lane; (void)l1;
#endif
}

// global -> staging area without a register round trip: cp.async (LDGSTS), VW bytes per lane and vector;
// the caller waits with fx_stage_async_wait() and a warp sync
template <int P>
MOF_HD void fx_stage_async(unsigned char* stage, const uint8_t* sec, int row_stride, const uint16_t* rowy, int task,
                           int lane) {
#if defined(__CUDA_ARCH__)
  typedef LoadCfg<P> L;
#if MOF_OPAQUE_WORDC
  // (bit mask over P) the row / vector split of the lane's elements is recomputed here instead of living in
  // registers for the whole kernel (nvcc hoists it out of the section loop and ptxas spills elsewhere)
  if (((MOF_OPAQUE_WORDC) & 16 & P) != 0) asm volatile("" : "+r"(lane));
#endif
#pragma unroll
  for (int i = 0; i < L::NV; i++) {
    const int e = i * 32 + lane;
    if (e < L::NE) {
      const int r = e / 15, c = e - 15 * r;
      const uint8_t* p = sec + (size_t)rowy[task * 2 * PG<P>::RPT + r] * row_stride + c * L::VW;
      const uint32_t d = (uint32_t)__cvta_generic_to_shared(stage + e * L::VW);
      if (L::VW == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(p) : "memory");
      else if (L::VW == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(p) : "memory");
      else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(p) : "memory");
    }
  }
#else
  (void)stage; (void)sec; (void)row_stride; (void)rowy; (void)task; (void)lane;
#endif
}
MOF_HD void fx_stage_async_wait() {
#if defined(__CUDA_ARCH__)
  asm volatile("cp.async.wait_all;" ::: "memory");
#endif
}

// registers -> staging area (the first 2*RPT*N bytes of the task's rows in A)
template <int P>
MOF_HD void fx_stage(unsigned char* stage, int lane, const uint32_t (&v)[LoadCfg<P>::NREG]) {
  typedef LoadCfg<P> L;
#if defined(__CUDA_ARCH__) && MOF_OPAQUE_WORDC
  if (((MOF_OPAQUE_WORDC) & 16 & P) != 0) asm volatile("" : "+r"(lane));
#endif
#pragma unroll
  for (int i = 0; i < L::NV; i++) {
    const int e = i * 32 + lane;
    if (e < L::NE) {
      uint32_t* o = reinterpret_cast<uint32_t*>(stage + e * L::VW);
#if defined(__CUDA_ARCH__)
      if (L::W == 4) *reinterpret_cast<uint4*>(o) = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      else if (L::W == 2) *reinterpret_cast<uint2*>(o) = make_uint2(v[2 * i], v[2 * i + 1]);
      else o[0] = v[i];
#else
      for (int k = 0; k < L::W; k++) o[k] = v[L::W * i + k];
#endif
    }
  }
}
// unaligned frames: synchronous byte copy global -> staging (no prefetch)
template <int P>
MOF_HD void fx_stage_bytes(unsigned char* stage, const uint8_t* sec, int row_stride, const uint16_t* rowy,
                           int task, int lane) {
  typedef PG<P> G;
  for (int e = lane; e < 2 * G::RPT * G::N; e += 32) {
    const int r = e / G::N, c = e - r * G::N;
    stage[e] = sec[(size_t)rowy[task * 2 * G::RPT + r] * row_stride + c];
  }
}

// staging -> this thread's 2 x 15 pixel words (rows a and b of row pair qi, word group g)
template <int P>
MOF_HD void fx_read_words(const unsigned char* stage, int qi, int g, uint32_t (&wa)[15], uint32_t (&wb)[15]) {
  typedef PG<P> G;
  const unsigned char* ra = stage + (2 * qi) * G::N + 4 * g;
  int c = word_c<P>(g);
#if defined(__CUDA_ARCH__) && MOF_OPAQUE_WORDC
  // keeps nvcc from hoisting the 15 lane-dependent word addresses out of the section loop (they spill)
  if (((MOF_OPAQUE_WORDC) & 16 & P) != 0) asm volatile("" : "+r"(c));
#endif
#pragma unroll
  for (int s = 0; s < 15; s++) {
    const int off = word_off<P>(s, c);
    wa[s] = *reinterpret_cast<const uint32_t*>(ra + off);
    wb[s] = *reinterpret_cast<const uint32_t*>(ra + G::N + off);
  }
}

// byte j of a word as float: PRMT into the mantissa of 2^23, minus 2^23 (two full-rate
// instructions instead of a quarter-rate I2F)
template <int J>
MOF_HD float byte_f32(uint32_t w) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650 | J)) - 8388608.0f;
#else
  return (float)((w >> (8 * J)) & 0xffu);
#endif
}

// bytes J of two words as one complex float: two PRMTs (selector immediate, the 2^23 exponent
// pattern comes in through `k23`, see below) and one packed add
template <int J>
MOF_HD c32 bytes_c32(uint32_t wa, uint32_t wb, uint32_t k23) {
#if defined(__CUDA_ARCH__)
  return cadd(mk(__uint_as_float(__byte_perm(wa, k23, 0x7650 | J)), __uint_as_float(__byte_perm(wb, k23, 0x7650 | J))),
              mk(-8388608.0f, -8388608.0f));
#else
  (void)k23;
  return mk((float)((wa >> (8 * J)) & 0xffu), (float)((wb >> (8 * J)) & 0xffu));
#endif
}
// PRMT takes ONE immediate.  With the 2^23 pattern as a compile-time constant ptxas makes IT the
// immediate and keeps the selector in a register that the instruction overwrites -- one extra MOV
// per converted byte.  The pattern therefore arrives as a kernel argument (PcArgs::k23, constant
// bank), which cannot be folded: PRMT Rd, Ra, 0x765J, c[0][k23].
// butterfly J of the thread's four: 15-point DFT over bx of z = row_a + i*row_b at ax = 4g + J
template <int P, int J>
MOF_HD void fx15_butterfly(c32 (&z)[15], const uint32_t (&wa)[15], const uint32_t (&wb)[15],
                           const float* win, int q, int g, uint32_t k23) {
  typedef PG<P> G;
#pragma unroll
  for (int bx = 0; bx < 15; bx++) {
    const int s = (bx - J + 15) % 15;
    z[bx] = bytes_c32<J>(wa[s], wb[s], k23);
  }
  if (win) {  // optional cv::createHanningWindow weights, N x N row-major
    const int ya = crt<P>(q / 15, q % 15);
    const int yb = ya >= G::M ? ya - G::M : ya + G::M;
    const int c = word_c<P>(g);
#pragma unroll
    for (int bx = 0; bx < 15; bx++) {
      const int s = (bx - J + 15) % 15;
      const int x = 4 * g + word_off<P>(s, c) + J;
      z[bx] = pmul(z[bx], mk(win[ya * G::N + x], win[yb * G::N + x]));
    }
  }
  dft15<1>(z);
}

template <int P, int J>
MOF_HD void fx15_store(c32* A, const c32 (&z)[15], int q, int g) {
  c32* o = A + q * PG<P>::SQ + slot<P>(4 * g + J, 0);
#pragma unroll
  for (int kb = 0; kb < 15; kb++) o[kb * SlotStep<P>::B] = z[kb];
}

// stage FX15 of warp task `task` for `lane`: (row pair, word group) = (lane / G4, lane % G4)
template <int P>
MOF_HD bool fx15_lane(int task, int lane, int& q, int& g) {
  typedef PG<P> G;
  // N = 240: (row pair, word group) = (lane / 4, lane % 4); N = 120: the two word groups of a row
  // pair sit in different half warps (15 rows each); N = 60: one row pair per lane
  const int qi = (P == 8) ? (lane & 15) : lane / G::G4;
  g = (P == 8) ? (lane >> 4) : lane % G::G4;
  q = task * G::RPT + qi;
  return qi < G::RPT;
}

// stage FX15 of a warp task, after the task's pixels were staged (fx_stage + warp sync): pick the
// words, warp sync (the staging area is about to be overwritten), four butterflies.
template <int P>
MOF_HD void stage_fx15_read(const c32* A, int task, int lane, uint32_t (&wa)[15], uint32_t (&wb)[15]) {
  int q, g;
  if (!fx15_lane<P>(task, lane, q, g)) return;
  fx_read_words<P>(reinterpret_cast<const unsigned char*>(A + task * PG<P>::RPT * PG<P>::SQ),
                   q - task * PG<P>::RPT, g, wa, wb);
}
// the same from a staging area that is not part of A (N = 120: a dedicated buffer per team)
template <int P>
MOF_HD void stage_fx15_read_from(const unsigned char* stage, int task, int lane, uint32_t (&wa)[15], uint32_t (&wb)[15]) {
  int q, g;
  if (!fx15_lane<P>(task, lane, q, g)) return;
  fx_read_words<P>(stage, q - task * PG<P>::RPT, g, wa, wb);
}
template <int P>
MOF_HD void stage_fx15(c32* A, int task, int lane, const uint32_t (&wa)[15],
                       const uint32_t (&wb)[15], const float* win, uint32_t k23 = 0x4B000000u) {
  int q, g;
  if (!fx15_lane<P>(task, lane, q, g)) return;
  c32 z[15];
  fx15_butterfly<P, 0>(z, wa, wb, win, q, g, k23);
  fx15_store<P, 0>(A, z, q, g);
  fx15_butterfly<P, 1>(z, wa, wb, win, q, g, k23);
  fx15_store<P, 1>(A, z, q, g);
  fx15_butterfly<P, 2>(z, wa, wb, win, q, g, k23);
  fx15_store<P, 2>(A, z, q, g);
  fx15_butterfly<P, 3>(z, wa, wb, win, q, g, k23);
  fx15_store<P, 3>(A, z, q, g);
}

// running maximum of the correlation surface seen by one thread: value, the round of the P-point
// stage whose item first reached it, and whether another item tied with it
struct MaxTrack {
  float vmax;
  int which;
  bool multi;
};

// Item maps of the row stages: round r of a warp task, lane -> (local row pair qi, b or ka).
//   N = 240: a half warp = the 15 b (16 ka) of one row;  N = 120: a half warp = the 15 rows of the
//   task at one b (ka), the two half warps take consecutive b (ka);  N = 60: lane = row.
template <int P>
struct RowMap {
  static constexpr int X16_ROUNDS = (P == 16) ? PG<P>::RPT / 2 : (P == 8 ? 8 : 15);
  static constexpr int X15_ROUNDS = 4;
  // P-point stage: (row, b)
  static MOF_HD bool x16(int r, int lane, int& qi, int& b) {
    if (P == 16) { qi = 2 * r + (lane >> 4); b = lane & 15; return b < 15; }
    if (P == 8) { qi = lane & 15; b = 2 * r + (lane >> 4); return qi < 15 && b < 15; }
    qi = lane; b = r; return lane < 30;
  }
  // 15-point stage: (row, ka)
  static MOF_HD bool x15(int r, int lane, int& qi, int& ka) {
    if (P == 16) { const int it = 32 * r + lane; qi = it >> 4; ka = it & 15; return true; }
    if (P == 8) { qi = lane & 15; ka = 2 * r + (lane >> 4); return qi < 15; }
    qi = lane; ka = r; return lane < 30;
  }
};

// stage FX16 / IX16: P-point DFT along ax <-> ka at fixed (row pair, bx).
// TRACK (inverse only): the outputs are the correlation surface (x = row a, y = row b); fold them
// into the thread's running maximum (its location is recovered afterwards by rescan_item); `which`
// remembers the round that first reached it.
template <int P, int DIR, bool TRACK>
MOF_HD void stage_x16(c32* A, int task, int lane, MaxTrack* mt) {
  typedef PG<P> G;
#if defined(__CUDA_ARCH__)
  constexpr bool PIPE = P == 16 && (MOF_X16_PIPELINE != 0);
#else
  constexpr bool PIPE = false;
#endif
  if (PIPE) {
    // N = 240: every lane has an item in every round (lanes 15 / 31 excepted), so the loads of round r + 1
    // are issued before the butterfly of round r: the shared-memory latency hides behind 80 packed
    // instructions instead of sitting at the top of every round (in place: rounds touch different rows)
    int qi, b;
    const bool act = RowMap<P>::x16(0, lane, qi, b);
    if (!act) return;
    c32 zn[P];
    {
      const c32* a0 = A + (task * G::RPT + qi) * G::SQ + slot<P>(0, b);
#pragma unroll
      for (int i = 0; i < P; i++) zn[i] = a0[SlotStep<P>::A * i];
    }
#pragma unroll 1
    for (int r = 0; r < RowMap<P>::X16_ROUNDS; r++) {
      RowMap<P>::x16(r, lane, qi, b);
      c32* a = A + (task * G::RPT + qi) * G::SQ + slot<P>(0, b);
      c32 z[P];
#pragma unroll
      for (int i = 0; i < P; i++) z[i] = zn[i];
      if (r + 1 < RowMap<P>::X16_ROUNDS) {
        const c32* an = a + 2 * G::SQ;   // round r + 1: two row pairs further
#pragma unroll
        for (int i = 0; i < P; i++) zn[i] = an[SlotStep<P>::A * i];
      }
      DftP<P, DIR>::run(z);
#pragma unroll
      for (int i = 0; i < P; i++) a[SlotStep<P>::A * i] = z[i];
      if (TRACK) {
        float m = fmaxf(z[0].x, z[0].y);
#pragma unroll
        for (int i = 1; i < P; i++) m = fmaxf(m, fmaxf(z[i].x, z[i].y));
        if (m > mt->vmax) {
          mt->vmax = m;
          mt->which = r;
          mt->multi = false;
        } else if (m == mt->vmax) {
          mt->multi = true;
        }
      }
    }
  } else {
#pragma unroll 1
  for (int r = 0; r < RowMap<P>::X16_ROUNDS; r++) {
    int qi, b;
    if (!RowMap<P>::x16(r, lane, qi, b)) continue;
    c32* a = A + (task * G::RPT + qi) * G::SQ + slot<P>(0, b);
    c32 z[P];
#pragma unroll
    for (int i = 0; i < P; i++) z[i] = a[SlotStep<P>::A * i];
    DftP<P, DIR>::run(z);
#pragma unroll
    for (int i = 0; i < P; i++) a[SlotStep<P>::A * i] = z[i];
    if (TRACK) {
      float m = fmaxf(z[0].x, z[0].y);
#pragma unroll
      for (int i = 1; i < P; i++) m = fmaxf(m, fmaxf(z[i].x, z[i].y));
      if (m > mt->vmax) {
        mt->vmax = m;
        mt->which = r;
        mt->multi = false;
      } else if (m == mt->vmax) {
        mt->multi = true;
      }
    }
  }
}
}

// stage IX15: 15-point inverse DFT along kb -> bx at fixed (row pair, ka)
template <int P>
MOF_HD void stage_ix15(c32* A, int task, int lane) {
  typedef PG<P> G;
#pragma unroll 1
  for (int r = 0; r < RowMap<P>::X15_ROUNDS; r++) {
    int qi, ka;
    if (!RowMap<P>::x15(r, lane, qi, ka)) continue;
    c32* a = A + (task * G::RPT + qi) * G::SQ + slot<P>(ka, 0);
    c32 z[15];
#pragma unroll
    for (int i = 0; i < 15; i++) z[i] = a[SlotStep<P>::B * i];
    dft15<-1>(z);
#pragma unroll
    for (int i = 0; i < 15; i++) a[SlotStep<P>::B * i] = z[i];
  }
}

// ------------------------------------------------------------------------------------------------
// FY: columns.  A task owns UPT units; all column stages of a task are in place on the task's own
// slots, so they chain with warp-level syncs only.
// ------------------------------------------------------------------------------------------------

// stage FY16: un-pack the row pairs (Ra[k] = (Z[k] + conj Z[-k])/2, Rb[k] = (Z[k] - conj Z[-k])/2i;
// the 1/2 is skipped for regular units) and P-point DFT along ay -> va at fixed (unit, by).
// Two units per round, 15 lanes each.  Output va lands in the slot butterfly va reads (va_base).
// PACKED is a template argument (not a test of u) so that the regular path has no control-flow
// merge with the packed one: the +-i / conjugate operand swizzles then fold into the butterflies.
template <int P, bool PACKED>
MOF_HD void fy16_item(c32* A, int u, int by) {
  typedef PG<P> G;
  int kcol, mcol;
  unit_cols<P>(u, kcol, mcol);
  c32* ak = A + by * G::SQ + kcol;
  c32* am = A + by * G::SQ + mcol;
  c32 z[P];
  if (PACKED) {  // packed real columns 0 and N/2: C = R[0] + i R[N/2]
#pragma unroll
    for (int m = 0; m < G::HP; m++) {
      const c32 za = ak[m * 15 * G::SQ], zb = am[m * 15 * G::SQ];
      z[m] = mk(za.x, zb.x);
      z[m + G::HP] = mk(za.y, zb.y);
    }
  } else {
#pragma unroll
    for (int m = 0; m < G::HP; m++) {
      const c32 za = ak[m * 15 * G::SQ], zb = am[m * 15 * G::SQ];
      // one packed FMA each with a (+1, -1) / (-1, +1) constant: written as adds of hand-built
      // (x, -y) pairs the compiler re-pairs the components through a dozen register moves
      z[m] = pfma(zb, mk(1.f, -1.f), za);                    // Z[k] + conj Z[-k]
      z[m + G::HP] = mul_mi<1>(pfma(zb, mk(-1.f, 1.f), za));  // (Z[k] - conj Z[-k]) / i
    }
  }
  DftP<P, 1>::run(z);
#pragma unroll
  for (int m = 0; m < G::HP; m++) {
    ak[m * 15 * G::SQ] = z[m];
    am[m * 15 * G::SQ] = z[m + G::HP];
  }
}

// stage IY16: P-point inverse DFT along va -> ay, then re-pack for the row transform:
// X[q][k] = Ga + i Gb, X[q][-k] = conj Ga + i conj Gb (Ga/Gb = rows a/b of the pair).
template <int P, bool PACKED>
MOF_HD void iy16_item(c32* A, int u, int by) {
  typedef PG<P> G;
  int kcol, mcol;
  unit_cols<P>(u, kcol, mcol);
  c32* ak = A + by * G::SQ + kcol;
  c32* am = A + by * G::SQ + mcol;
  c32 z[P];
#pragma unroll
  for (int m = 0; m < G::HP; m++) {
    z[m] = ak[m * 15 * G::SQ];
    z[m + G::HP] = am[m * 15 * G::SQ];
  }
  DftP<P, -1>::run(z);
  if (PACKED) {  // z[ay] = G[y,0] + i G[y,N/2], both real sequences
#pragma unroll
    for (int m = 0; m < G::HP; m++) {
      ak[m * 15 * G::SQ] = mk(z[m].x, z[m + G::HP].x);
      am[m * 15 * G::SQ] = mk(z[m].y, z[m + G::HP].y);
    }
  } else {
#pragma unroll
    for (int m = 0; m < G::HP; m++) {
      const c32 ga = z[m], gb = z[m + G::HP];
      ak[m * 15 * G::SQ] = cadd(ga, mul_mi<-1>(gb));                            // Ga + i Gb
      am[m * 15 * G::SQ] = pmul(cadd(ga, mul_mi<1>(gb)), mk(1.f, -1.f));        // conj(Ga - i Gb)
    }
  }
}

// FY16 / IY16 of a regular unit split into "load" and "transform + store" (software pipelining, N = 240)
template <int P>
MOF_HD void y16_load(const c32* A, int by, int kcol, int mcol, c32 (&zr)[P]) {
  typedef PG<P> G;
  const c32* ak = A + by * G::SQ + kcol;
  const c32* am = A + by * G::SQ + mcol;
#pragma unroll
  for (int m = 0; m < G::HP; m++) {
    zr[m] = ak[m * 15 * G::SQ];
    zr[m + G::HP] = am[m * 15 * G::SQ];
  }
}
template <int P, bool INV>
MOF_HD void y16_finish(c32* A, int by, int kcol, int mcol, const c32 (&zr)[P]) {
  typedef PG<P> G;
  c32* ak = A + by * G::SQ + kcol;
  c32* am = A + by * G::SQ + mcol;
  c32 z[P];
  if (INV) {
#pragma unroll
    for (int i = 0; i < P; i++) z[i] = zr[i];
    DftP<P, -1>::run(z);
#pragma unroll
    for (int m = 0; m < G::HP; m++) {
      const c32 ga = z[m], gb = z[m + G::HP];
      ak[m * 15 * G::SQ] = cadd(ga, mul_mi<-1>(gb));                            // Ga + i Gb
      am[m * 15 * G::SQ] = pmul(cadd(ga, mul_mi<1>(gb)), mk(1.f, -1.f));        // conj(Ga - i Gb)
    }
  } else {
#pragma unroll
    for (int m = 0; m < G::HP; m++) {
      const c32 za = zr[m], zb = zr[m + G::HP];
      z[m] = pfma(zb, mk(1.f, -1.f), za);                    // Z[k] + conj Z[-k]
      z[m + G::HP] = mul_mi<1>(pfma(zb, mk(-1.f, 1.f), za));  // (Z[k] - conj Z[-k]) / i
    }
    DftP<P, 1>::run(z);
#pragma unroll
    for (int m = 0; m < G::HP; m++) {
      ak[m * 15 * G::SQ] = z[m];
      am[m * 15 * G::SQ] = z[m + G::HP];
    }
  }
}

// round loop shared by FY16 / IY16: units [u0, u0 + nu), two per round
template <int P, bool INV>
MOF_HD void stage_y16(c32* A, int u0, int nu, int lane) {
  typedef PG<P> G;
  const int l = lane & 15;
  if (l == 15) return;
  const bool has_packed = (u0 + nu == G::M);   // the packed unit is the last one
  const int nreg = has_packed ? nu - 1 : nu;
#if defined(__CUDA_ARCH__) && MOF_Y16_PIPELINE
  if (P == 16 && (INV || MOF_Y16_PIPELINE > 1)) {   // loads of the next unit issued before this unit's butterfly (units own disjoint columns)
    int ui = lane >> 4;
    c32 zn[P];
    int kc = 0, mc = 0;
    if (ui < nreg) {
      unit_cols<P>(u0 + ui, kc, mc);
      y16_load<P>(A, l, kc, mc, zn);
    }
#pragma unroll 1
    for (; ui < nreg; ui += 2) {
      c32 zr[P];
#pragma unroll
      for (int i = 0; i < P; i++) zr[i] = zn[i];
      const int kcur = kc, mcur = mc;
      if (ui + 2 < nreg) {
        unit_cols<P>(u0 + ui + 2, kc, mc);
        y16_load<P>(A, l, kc, mc, zn);
      }
      y16_finish<P, INV>(A, l, kcur, mcur, zr);
    }
  } else
#endif
  {
#pragma unroll 1
    for (int ui = lane >> 4; ui < nreg; ui += 2) {
      if (INV) iy16_item<P, false>(A, u0 + ui, l);
      else fy16_item<P, false>(A, u0 + ui, l);
    }
  }
  if (has_packed && lane < 16) {
    if (INV) iy16_item<P, true>(A, G::M - 1, l);
    else fy16_item<P, true>(A, G::M - 1, l);
  }
}

// normalised cross-power, fast form: C = P * rsqrt(|P|^2), valid when |P|^2 + eps == |P|^2 in
// float (checked by the caller through the running minimum of |P|^2)
MOF_HD c32 cross_power_fast(c32 a, c32 b, float& m2min) {
  // P = a * conj(b) = (a.x, a.y) * b.x + (a.y, -a.x) * b.y, two packed instructions
  const c32 p = pfma(mk(a.y, -a.x), mk(b.y, b.y), pmul(a, mk(b.x, b.x)));
  const float m2 = fmaf_(p.y, p.y, p.x * p.x);
  m2min = fminf(m2min, m2);
#if defined(__CUDA_ARCH__)
  float s;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(m2));
#else
  const float s = 1.0f / sqrtf(m2);
#endif
  return pmul(p, mk(s, s));
}
// exact form (magSpectrums + divSpectrums, src/FftMethod.cpp:113-114, 1166-1172):
// C = P*|P| / (|P|^2 + eps);  eps = FLT_EPSILON * (scale of the spectra)^4
// ocl != 0: the OpenCL kernel's form instead, P * rsqrt(|P|^2 + eps) (cmulnormf, cl/FftMethod.cl:976-982)
MOF_HD c32 cross_power_exact(c32 a, c32 b, float eps, int ocl = 0) {
  const float pr = a.x * b.x + a.y * b.y;
  const float pi = a.y * b.x - a.x * b.y;
  if (ocl) {
    const float r = 1.0f / sqrtf(fmaf_(pr, pr, pi * pi + eps));
    return mk(pr * r, pi * r);
  }
  const float mag = sqrtf(pr * pr + pi * pi);
  const float s = mag / (mag * mag + eps);
  return mk(pr * s, pi * s);
}
// the four purely real bins: C = P / (|P| + eps)   (divSpectrums :1124-1126 with cv2-4.13's |P|)
// ocl != 0: the OpenCL kernel writes 1 / (a * b) into the four CCS corner bins (cl/FftMethod.cl:1031)
MOF_HD float cross_power_real(float a, float b, int ocl = 0) {
  const float p = a * b;
  if (ocl) return 1.0f / p;
  return p / (fabsf(p) + FLT_EPSILON);
}

#define MOF_CP_FAST_MIN 1024.0f  /* |P|^2 above which eps*16 is below half an ulp */

template <int P>
MOF_HD void spec_load15(const f4* spec, int item, int half, c32 (&f)[15]) {
  const f4* p = spec + (size_t)(half * 8) * PG<P>::ITEMS + item;
#pragma unroll
  for (int t = 0; t < 7; t++) {
    const f4 v = p[(size_t)t * PG<P>::ITEMS];
    f[2 * t] = mk(v.x, v.y);
    f[2 * t + 1] = mk(v.z, v.w);
  }
  const c32* last = reinterpret_cast<const c32*>(p + (size_t)7 * PG<P>::ITEMS);
  f[14] = last[0];
}
template <int P>
MOF_HD void spec_store15(f4* spec, int item, int half, const c32 (&f)[15]) {
  f4* p = spec + (size_t)(half * 8) * PG<P>::ITEMS + item;
#pragma unroll
  for (int t = 0; t < 7; t++) {
    f4 v;
    v.x = f[2 * t].x; v.y = f[2 * t].y; v.z = f[2 * t + 1].x; v.w = f[2 * t + 1].y;
    p[(size_t)t * PG<P>::ITEMS] = v;
  }
  c32* last = reinterpret_cast<c32*>(p + (size_t)7 * PG<P>::ITEMS);
  last[0] = f[14];
}

template <int P>
MOF_HD void col15_load(const c32* A, int base, c32 (&z)[15]) {
#pragma unroll
  for (int i = 0; i < 15; i++) z[i] = A[base + i * PG<P>::SQ];
}
template <int P>
MOF_HD void col15_store(c32* A, int base, const c32 (&z)[15]) {
#pragma unroll
  for (int i = 0; i < 15; i++) A[base + i * PG<P>::SQ] = z[i];
}

// stage FY15 (spectrum-only tail): 15-point DFT along by -> vb of butterflies (j, half 0/1) of a
// unit; result = the image's spectrum F[(va, vb), k], written to `spec`.  Works for the packed
// unit as well (its spectrum is kept packed).
template <int P>
MOF_HD void fy15_store_item(const c32* A, int u, int j, f4* spec) {
  typedef PG<P> G;
  int kcol, mcol;
  unit_cols<P>(u, kcol, mcol);
#pragma unroll 1
  for (int half = 0; half < 2; half++) {
    c32 z[15];
    col15_load<P>(A, va_base<P>(va_of<P>(j, half), kcol, mcol), z);
    dft15<1>(z);
    spec_store15<P>(spec, item_of<P>(u, j), half, z);
  }
}

// exact-path redo of one butterfly (rare: some |P|^2 below MOF_CP_FAST_MIN)
template <int P>
#if defined(__CUDACC__)
__device__ __host__ __noinline__
#endif
void fy15_corr_slow(c32* A, int base, int item, int half, const f4* spec_prev, int ocl = 0) {
  c32 z[15], fp[15];
  spec_load15<P>(spec_prev, item, half, fp);
  col15_load<P>(A, base, z);
  dft15<1>(z);
#pragma unroll
  for (int i = 0; i < 15; i++) z[i] = cross_power_exact(z[i], fp[i], 16.0f * FLT_EPSILON, ocl);
  dft15<-1>(z);
  col15_store<P>(A, base, z);
}

// stage FY15 + cross-power + IY15 of a REGULAR unit: forward 15-point DFT, C = normalised
// F_curr * conj(F_prev), inverse 15-point DFT, back in place.  spec_out (optional, must NOT alias
// spec_prev: the exact-path redo re-reads the previous spectrum) keeps the current spectrum.
template <int P>
MOF_HD void fy15_corr_item(c32* A, int u, int j, const f4* spec_prev, f4* spec_out, int ocl = 0) {
  typedef PG<P> G;
  int kcol, mcol;
  unit_cols<P>(u, kcol, mcol);
  const int item = item_of<P>(u, j);
#pragma unroll 1
  for (int half = 0; half < 2; half++) {
    const int base = va_base<P>(va_of<P>(j, half), kcol, mcol);
    c32 z[15], fp[15];
#if defined(MOF_FAKE_NOSPEC)   /* timing experiment only (wrong results): upper bound of keeping the spectra on chip */
    col15_load<P>(A, base, z);
    dft15<1>(z);
#pragma unroll
    for (int i = 0; i < 15; i++) fp[i] = z[(i + 1) % 15];
#else
    spec_load15<P>(spec_prev, item, half, fp);
    col15_load<P>(A, base, z);
    dft15<1>(z);
    if (spec_out) spec_store15<P>(spec_out, item, half, z);
#endif
    float m2min = 3.0e38f;
#pragma unroll
    for (int i = 0; i < 15; i++) z[i] = cross_power_fast(z[i], fp[i], m2min);
    if (m2min >= MOF_CP_FAST_MIN) {
      dft15<-1>(z);
      col15_store<P>(A, base, z);
    } else {
      fy15_corr_slow<P>(A, base, item, half, spec_prev, ocl);
    }
  }
}

// Item map of the 15-point column stages: round r, lane -> (local unit ul, butterfly pair j).
//   N = 240: four units x 8 j per round (a half warp = two units whose ka differ by 8);
//   N = 120: a half warp = the (up to 15) units of the task at one j;  N = 60: a half warp = 16
//   consecutive units at one j, both j in the same round.
template <int P>
struct ColMap {
  static MOF_HD int rounds(int nu) {
    if (P == 16) return (nu * PG<P>::HP + 31) / 32;
    if (P == 8) return 2;
    return (nu + 15) / 16;
  }
  static MOF_HD bool y15(int r, int lane, int nu, int& ul, int& j) {
    if (P == 16) { const int it = 32 * r + lane; ul = it >> 3; j = it & 7; }
    else if (P == 8) { ul = lane & 15; j = 2 * r + (lane >> 4); }
    else { ul = 16 * r + (lane & 15); j = lane >> 4; }
    return ul < nu;
  }
};

// round loops of the 15-point column stages over the units [u0, u0 + nu) of a task
template <int P>
MOF_HD void stage_fy15_store(const c32* A, int u0, int nu, int lane, f4* spec) {
#pragma unroll 1
  for (int r = 0; r < ColMap<P>::rounds(nu); r++) {
    int ul, j;
    if (ColMap<P>::y15(r, lane, nu, ul, j)) fy15_store_item<P>(A, u0 + ul, j, spec);
  }
}
template <int P>
MOF_HD void stage_fy15_corr(c32* A, int u0, int nu, int lane, const f4* spec_prev, f4* spec_out, int ocl = 0) {
#pragma unroll 1
  for (int r = 0; r < ColMap<P>::rounds(nu); r++) {
    int ul, j;
    if (ColMap<P>::y15(r, lane, nu, ul, j)) fy15_corr_item<P>(A, u0 + ul, j, spec_prev, spec_out, ocl);
  }
}

// ------------------------------------------------------------------------------------------------
// The packed unit (columns 0 and N/2), correlate path.  Its cross-power needs bins v and -v
// together, so it runs as four small warp-synchronous steps: SP1 forward 15-point DFTs in place,
// SP2 split / cross-power / re-pack per bin pair, SP3 inverse 15-point DFTs, then iy16_item.
// ------------------------------------------------------------------------------------------------
template <int P, int DIR>
MOF_HD void sp_dft15(c32* A, int lane) {
  typedef PG<P> G;
  if (lane >= P) return;
  const int base = va_base<P>(lane, slot<P>(0, 0), slot<P>(G::HP, 0));
  c32 z[15];
  col15_load<P>(A, base, z);
  dft15<DIR>(z);
  col15_store<P>(A, base, z);
}

template <int P>
MOF_HD void sp_cross_power(c32* A, int lane, const f4* spec_prev, f4* spec_out, int ocl = 0) {
  typedef PG<P> G;
  const c32* sp = reinterpret_cast<const c32*>(spec_prev);
  c32* so = reinterpret_cast<c32*>(spec_out);
#pragma unroll 1
  for (int id = lane; id < G::N; id += 32) {
    const int va = id / 15, vb = id % 15;
    const int vam = (P - va) % P, vbm = (15 - vb) % 15;
    const int idm = vam * 15 + vbm;
    if (id > idm) continue;
    const int ev = va_base<P>(va, slot<P>(0, 0), slot<P>(G::HP, 0)) + vb * G::SQ;
    const int em = va_base<P>(vam, slot<P>(0, 0), slot<P>(G::HP, 0)) + vbm * G::SQ;
    const size_t sv = spec_c32<P>(item_of<P>(G::M - 1, va <= G::HP ? va % G::HP : P - va), va >= G::HP, vb);
    const size_t sm = spec_c32<P>(item_of<P>(G::M - 1, vam <= G::HP ? vam % G::HP : P - vam), vam >= G::HP, vbm);
    const c32 wc = A[ev], wcm = A[em];
    const c32 wp = sp[sv], wpm = sp[sm];
    if (so) {
      so[sv] = wc;
      if (id != idm) so[sm] = wcm;
    }
    c32 r0, rM;
    if (id == idm) {  // v = -v: the four purely real bins
      r0 = mk(cross_power_real(wc.x, wp.x, ocl), 0.f);
      rM = mk(cross_power_real(wc.y, wp.y, ocl), 0.f);
    } else {
      const c32 fc0 = mk(0.5f * (wc.x + wcm.x), 0.5f * (wc.y - wcm.y));
      const c32 fcM = mk(0.5f * (wc.y + wcm.y), 0.5f * (wcm.x - wc.x));
      const c32 fp0 = mk(0.5f * (wp.x + wpm.x), 0.5f * (wp.y - wpm.y));
      const c32 fpM = mk(0.5f * (wp.y + wpm.y), 0.5f * (wpm.x - wp.x));
      r0 = cross_power_exact(fc0, fp0, FLT_EPSILON, ocl);
      rM = cross_power_exact(fcM, fpM, FLT_EPSILON, ocl);
    }
    A[ev] = mk(r0.x - rM.y, r0.y + rM.x);               // R[v,0] + i R[v,N/2]
    if (id != idm) A[em] = mk(r0.x + rM.y, rM.x - r0.y);  // conj R[v,0] + i conj R[v,N/2]
  }
}

// ------------------------------------------------------------------------------------------------
// Peak search and sub-pixel refinement on the surface left in A by the tracked IX16 stage.
// ------------------------------------------------------------------------------------------------
// first (fft-shifted, row-major) index at which item (row pair q, bx) equals vmax; INT_MAX if none
template <int P>
MOF_HD int rescan_item(const c32* A, int q, int bx, float vmax) {
  typedef PG<P> G;
  const int ya = crt<P>(q / 15, q % 15);            // row a; row b = ya +- M
  const int ysa = ya >= G::M ? ya - G::M : ya + G::M;  // shifted row of row a; shifted row of b = ya
  const c32* a = A + q * G::SQ + slot<P>(0, bx);
  int best = 0x7fffffff;
#pragma unroll
  for (int ax = 0; ax < P; ax++) {
    const c32 v = a[SlotStep<P>::A * ax];
    const int x = crt<P>(ax, bx);
    const int xs = x >= G::M ? x - G::M : x + G::M;
    if (v.x == vmax) best = imin(best, ysa * G::N + xs);
    if (v.y == vmax) best = imin(best, ya * G::N + xs);
  }
  return best;
}

// search of one thread after its warp's maximum is known (mt from the tracked IX16 stage)
template <int P>
MOF_HD int search_lane(const c32* A, int task, int lane, const MaxTrack& mt, float bmax) {
  typedef PG<P> G;
  if (!(mt.vmax == bmax)) return 0x7fffffff;
  int qi, b;
  if (!mt.multi) {
    if (!RowMap<P>::x16(mt.which, lane, qi, b)) return 0x7fffffff;
    return rescan_item<P>(A, task * G::RPT + qi, b, bmax);
  }
  int best = 0x7fffffff;
  for (int r = 0; r < RowMap<P>::X16_ROUNDS; r++)
    if (RowMap<P>::x16(r, lane, qi, b)) best = imin(best, rescan_item<P>(A, task * G::RPT + qi, b, bmax));
  return best;
}

// value of the fft-shifted correlation surface at (ys, xs)
template <int P>
MOF_HD float surf_at(const c32* A, int ys, int xs) {
  typedef PG<P> G;
  const int y = ys >= G::M ? ys - G::M : ys + G::M;
  const int x = xs >= G::M ? xs - G::M : xs + G::M;
  const int ay = y % P, by = y % 15;
  const c32 v = A[((ay % G::HP) * 15 + by) * G::SQ + slot<P>(x % P, x % 15)];
  return ay < G::HP ? v.x : v.y;
}

// row pair holding fft-shifted surface row ys (its owner in the row stages is task q / RPT)
template <int P>
MOF_HD int surf_rowpair(int ys) {
  typedef PG<P> G;
  const int y = ys >= G::M ? ys - G::M : ys + G::M;
  return ((y % P) % G::HP) * 15 + y % 15;
}

// coordinates of centroid tap `tap` (0..24) around peak_idx; false when it falls outside the image
template <int P>
MOF_HD bool tap_coords(int peak_idx, int tap, int& y, int& x) {
  typedef PG<P> G;
  y = peak_idx / G::N + tap / 5 - 2;
  x = peak_idx % G::N + tap % 5 - 2;
  return !(y < 0 || y >= G::N || x < 0 || x >= G::N);
}

// FP64 accumulation of one tap value (weightedCentroid, src/FftMethod.cpp:1330-1384)
MOF_HD void centroid_acc(int y, int x, float val, double& cx, double& cy, double& sum) {
  const double v = (double)val;
  cx += (double)x * v;
  cy += (double)y * v;
  sum += v;
}

// one tap of the 5x5 centroid window around the peak, clamped to the image, all values, FP64
// (weightedCentroid, src/FftMethod.cpp:1330-1384): tap = 0..24; adds into (cx, cy, sum)
template <int P>
MOF_HD void centroid_tap(const c32* A, int peak_idx, int tap, double& cx, double& cy, double& sum) {
  typedef PG<P> G;
  const int y = peak_idx / G::N + tap / 5 - 2, x = peak_idx % G::N + tap % 5 - 2;
  if (y < 0 || y >= G::N || x < 0 || x >= G::N) return;
  const double v = (double)surf_at<P>(A, y, x);
  cx += (double)x * v;
  cy += (double)y * v;
  sum += v;
}

// shift = t - N/2 and the validity rules of src/FftMethod.cpp:1840-1856
template <int P>
MOF_HD void centroid_finish(double cx, double cy, double sum, double max_px_speed_sq, double* out_xy) {
  typedef PG<P> G;
  sum += DBL_EPSILON;
  const double sx = cx / sum - (double)G::N / 2.0;  // -(center - t)
  const double sy = cy / sum - (double)G::N / 2.0;
  bool valid = !(sx * sx + sy * sy > max_px_speed_sq || fabs(sx) > (double)G::N / 2.0 ||
                 fabs(sy) > (double)G::N / 2.0);
  if (sx != sx || sy != sy) valid = false;
  const double qnan = (double)NAN;
  out_xy[0] = valid ? sx : qnan;
  out_xy[1] = valid ? sy : qnan;
}

}  // namespace pfa
}  // namespace mof
