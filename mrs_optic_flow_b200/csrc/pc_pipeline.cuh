// pc_pipeline.cuh -- per-section phase-correlation pipeline, written as "work items" that a CTA's
// threads execute between block barriers.  Host+device: the same item functions are run
// sequentially on the CPU by tests/emu/emu_pc.cpp to verify every index map without a GPU.
//
// What it computes (reference: FftMethod::processImage CPU branch, src/FftMethod.cpp:1829-1856,
// i.e. shift = -cv::phaseCorrelate(curr_roi, prev_roi) per N x N section):
//   2-D real FFT of both sections -> P = F_curr * conj(F_prev) -> C = P*|P| / (|P|^2 + FLT_EPSILON)
//   -> unscaled inverse -> fftshift -> first-maximum argmax -> clamped 5x5 FP64 centroid.
//
// Layout.  N = 15*P (P = 4, 8, 16 -> N = 60, 120, 240), M = N/2.  One section lives in shared
// memory as a [M][N+1] array of complex floats "A": row-pair q holds the packed transform of image
// rows q and q+M (z = row_q + i*row_{q+M}), one padding slot per row-pair makes the q-stride odd so
// that lanes<->q accesses are bank-conflict free; lanes<->k accesses are contiguous.
//   A size: N=240 -> 120*241*8 = 231,360 B (one CTA per SM), N=120 -> 58,080 B, N=60 -> 14,640 B.
// Every 1-D transform of length N is two register-resident butterflies (15 and P) with one
// exchange through A; all exchanges are in place (each item writes exactly the slots it read).
//
// Column k (1 <= k < M) of the half spectrum owns slots (q, k) and (q, N-k); the purely real
// columns 0 and M are packed into one complex column that owns slots (q, 0) and (q, M).
#pragma once

#include <float.h>
#include <math.h>
#include <stdint.h>

#include "fft_regs.cuh"

namespace mof {

// w_240^j = exp(-2 pi i j / 240); w_N^j = tw240[j * (240 / N)].
#if defined(__CUDACC__)
__constant__ c32 c_tw240[240];
#endif
static c32 h_tw240[240];
#if defined(__CUDA_ARCH__)
#define MOF_TW(i) c_tw240[(i)]
#else
#define MOF_TW(i) h_tw240[(i)]
#endif

// host-side table fill (the device copy is uploaded with cudaMemcpyToSymbol by the backend)
inline void fill_twiddles_host() {
  for (int j = 0; j < 240; j++) {
    double a = -2.0 * 3.14159265358979323846 * (double)j / 240.0;
    h_tw240[j].x = (float)cos(a);
    h_tw240[j].y = (float)sin(a);
  }
}

template <int P>
struct Geo {
  static constexpr int N = 15 * P;
  static constexpr int M = N / 2;
  static constexpr int SQ = N + 1;           // row-pair stride in c32 units (odd)
  static constexpr int H = P / 2;            // butterfly pairs per column / row-pair
  static constexpr int TWS = 16 / P;         // index scale into the 240-entry twiddle table
  static constexpr int A_ELEMS = M * SQ;     // c32 elements of the section array
  static constexpr int LD = (M + 31) / 32 * 32;  // lane dimension padded to whole warps
};

// slot of column k's private storage: (q, side) -> element index in A
template <int P>
MOF_HD int col_slot(int k, int q, int side) {
  typedef Geo<P> G;
  int kk = side ? (k == 0 ? G::M : G::N - k) : k;
  return q * G::SQ + kk;
}

// ------------------------------------------------------------------------------------------------
// P1a: forward rows, stage 1.  item (q, g): the P/4 adjacent butterflies n2 = NB*g .. NB*g+NB-1
// (NB = P/4) of row pair q.  Butterfly n2 is the 15-point DFT over x = P*n1 + n2 of
// z = row_q + i*row_{q+M}; its raw result (k1, n2) goes to A[q][k1 + 15 n2].  The NB adjacent pixels
// of both rows are fetched with one NB-byte load per (row, n1) when the addresses allow it
// (VEC = true), else byte by byte.  Optional window (cv::createHanningWindow layout).
// ------------------------------------------------------------------------------------------------
template <int NB>
MOF_HD unsigned load_px(const uint8_t* p, bool vec) {
  if (NB == 1) return p[0];
  if (vec) {
    if (NB == 2) return *reinterpret_cast<const uint16_t*>(p);
    return *reinterpret_cast<const uint32_t*>(p);
  }
  unsigned v = 0;
#pragma unroll
  for (int b = 0; b < NB; b++) v |= (unsigned)p[b] << (8 * b);
  return v;
}

template <int P>
MOF_HD void p1a_item(c32* A, int q, int g, const uint8_t* sec, int row_stride, const float* win,
                     bool vec) {
  typedef Geo<P> G;
  constexpr int NB = P / 4;
  const uint8_t* ra = sec + (size_t)q * row_stride + NB * g;
  const uint8_t* rb = sec + (size_t)(q + G::M) * row_stride + NB * g;
  unsigned wa[15], wb[15];
#pragma unroll
  for (int n1 = 0; n1 < 15; n1++) {
    wa[n1] = load_px<NB>(ra + P * n1, vec);
    wb[n1] = load_px<NB>(rb + P * n1, vec);
  }
#pragma unroll 1
  for (int b = 0; b < NB; b++) {
    const int n2 = NB * g + b, sh = 8 * b;
    c32 z[15];
#pragma unroll
    for (int n1 = 0; n1 < 15; n1++)
      z[n1] = mk((float)((wa[n1] >> sh) & 0xffu), (float)((wb[n1] >> sh) & 0xffu));
    if (win) {
      const float* va = win + q * G::N + n2;
      const float* vb = win + (q + G::M) * G::N + n2;
#pragma unroll
      for (int n1 = 0; n1 < 15; n1++) z[n1] = mk(va[P * n1] * z[n1].x, vb[P * n1] * z[n1].y);
    }
    dft15<1>(z);
    c32* o = A + q * G::SQ + 15 * n2;
#pragma unroll
    for (int k1 = 0; k1 < 15; k1++) o[k1] = z[k1];
  }
}

// P1b: forward rows, stage 2.  item (k1, q): twiddle + P-point DFT over n2, in place;
// afterwards A[q][k] = Z_q[k], k = k1 + 15 k2, natural order.
template <int P>
MOF_HD void p1b_item(c32* A, int k1, int q) {
  typedef Geo<P> G;
  c32* a = A + q * G::SQ + k1;
  c32 z[P];
#pragma unroll
  for (int n2 = 0; n2 < P; n2++) {
    c32 v = a[15 * n2];
    z[n2] = (n2 == 0) ? v : twmul<1>(v, MOF_TW(n2 * k1 * G::TWS));
  }
  DftP<P, 1>::run(z);
#pragma unroll
  for (int k2 = 0; k2 < P; k2++) a[15 * k2] = z[k2];
}

// P2: forward columns, stage 1.  item (n2, k): un-pack the row pairs on the fly
// (C_q = (Z[k] + conj Z[N-k])/2, C_{q+M} = (Z[k] - conj Z[N-k])/(2i); column 0 packs the two real
// columns 0 and M), P-point DFT over y = n2 + 15 n1, twiddle, store (k1, n2) in the column's slots.
template <int P>
MOF_HD void p2_item(c32* A, int n2, int k) {
  typedef Geo<P> G;
  const int kp = (k == 0) ? G::M : G::N - k;
  c32 z[P];
#pragma unroll
  for (int m = 0; m < G::H; m++) {
    const int q = n2 + 15 * m;
    c32 za = A[q * G::SQ + k], zb = A[q * G::SQ + kp];
    c32 ca, cb;
    if (k == 0) {
      ca = mk(za.x, zb.x);
      cb = mk(za.y, zb.y);
    } else {
      ca = mk(0.5f * (za.x + zb.x), 0.5f * (za.y - zb.y));
      cb = mk(0.5f * (za.y + zb.y), 0.5f * (zb.x - za.x));
    }
    z[m] = ca;
    z[m + G::H] = cb;
  }
  DftP<P, 1>::run(z);
#pragma unroll
  for (int k1 = 0; k1 < P; k1++) {
    c32 v = (k1 == 0) ? z[0] : twmul<1>(z[k1], MOF_TW(n2 * k1 * G::TWS));
    const int q = n2 + 15 * (k1 % G::H);
    A[q * G::SQ + ((k1 / G::H) ? kp : k)] = v;
  }
}

// normalised cross-power of one bin: C = P*|P| / (|P|^2 + eps), P = a * conj(b)
// (magSpectrums + divSpectrums, src/FftMethod.cpp:113-114, 1166-1172)
MOF_HD c32 cross_power(c32 a, c32 b) {
  float pr = a.x * b.x + a.y * b.y;
  float pi = a.y * b.x - a.x * b.y;
  float m2 = pr * pr + pi * pi;
  if (m2 >= 16.0f) {
    // |P|^2 + FLT_EPSILON == |P|^2 in float here, so C = P / |P|
#if defined(__CUDA_ARCH__)
    float s;  // m2 >= 16: no denormal / zero handling needed -> bare MUFU.RSQ
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(m2));
#else
    float s = 1.0f / sqrtf(m2);
#endif
    return mk(pr * s, pi * s);
  }
  float mag = sqrtf(m2);
  float s = mag / (mag * mag + FLT_EPSILON);
  return mk(pr * s, pi * s);
}
// the four purely real bins: C = P / (|P| + eps)   (divSpectrums :1124-1126 with cv2-4.13's |P|)
MOF_HD float cross_power_real(float a, float b) {
  float p = a * b;
  return p / (fabsf(p) + FLT_EPSILON);
}

// P3 load of one forward stage-2 butterfly (k1) of column k: 15 values over n2
template <int P>
MOF_HD void p3_load(const c32* A, int k, int k1, c32 (&z)[15]) {
  typedef Geo<P> G;
  const int kk = (k1 / G::H) ? ((k == 0) ? G::M : G::N - k) : k;
  const c32* a = A + (15 * (k1 % G::H)) * G::SQ + kk;
#pragma unroll
  for (int n2 = 0; n2 < 15; n2++) z[n2] = a[n2 * G::SQ];
}
template <int P>
MOF_HD void p3_store(c32* A, int k, int k1, const c32 (&z)[15]) {
  typedef Geo<P> G;
  const int kk = (k1 / G::H) ? ((k == 0) ? G::M : G::N - k) : k;
  c32* a = A + (15 * (k1 % G::H)) * G::SQ + kk;
#pragma unroll
  for (int n2 = 0; n2 < 15; n2++) a[n2 * G::SQ] = z[n2];
}

// P3 (spectrum-only tail): item (j, k): the two stage-2 butterflies k1a = j, k1b = partner(j),
// result F[v, k] (v = k1 + P k2) to global spec[v*M + k].  Column 0 is stored packed.
template <int P>
MOF_HD void p3_store_item(const c32* A, int j, int k, c32* spec) {
  typedef Geo<P> G;
#pragma unroll 1
  for (int half = 0; half < 2; half++) {
    const int k1 = half ? ((j == 0) ? G::H : P - j) : j;
    c32 z[15];
    p3_load<P>(A, k, k1, z);
    dft15<1>(z);
    c32* o = spec + (size_t)k1 * G::M + k;
#pragma unroll
    for (int k2 = 0; k2 < 15; k2++) o[(size_t)(P * k2) * G::M] = z[k2];
  }
}

// inverse-column stage 1 on a register butterfly: 15-point inverse DFT over k2, twiddle
// conj(w_N^(k1 y1)), store (y1, k1) in the column's slots (same slots the forward butterfly read).
template <int P>
MOF_HD void p3_inverse_store(c32* A, int k, int k1, c32 (&z)[15]) {
  typedef Geo<P> G;
  dft15<-1>(z);
#pragma unroll
  for (int y1 = 1; y1 < 15; y1++) z[y1] = twmul<-1>(z[y1], MOF_TW(k1 * (y1 * G::TWS)));
  p3_store<P>(A, k, k1, z);
}

// P3 (correlate tail).  item (j, k).  General column 1 <= k < M: forward stage 2, cross-power
// against the previous image's spectrum (prefetched before the butterfly), inverse stage 1.
// Packed column k = 0: forward stage 2 only, result back into its slots -- its cross-power needs
// the bins v and -v together and is done by P3Z / P3I below.
// spec_prev: previous spectrum (as written by p3_store_item); spec_out (optional, may alias
// spec_prev): where to keep this image's spectrum for the next frame of a stream.
template <int P>
MOF_HD void p3_corr_item(c32* A, int j, int k, const c32* spec_prev, c32* spec_out) {
  typedef Geo<P> G;
#pragma unroll 1
  for (int half = 0; half < 2; half++) {
    const int k1 = half ? ((j == 0) ? G::H : P - j) : j;
    c32 z[15];
    if (k == 0) {
      p3_load<P>(A, 0, k1, z);
      dft15<1>(z);
      p3_store<P>(A, 0, k1, z);
      continue;
    }
    const size_t g0 = (size_t)k1 * G::M + k;
    c32 fp[15];
#pragma unroll
    for (int k2 = 0; k2 < 15; k2++) fp[k2] = spec_prev[g0 + (size_t)(P * k2) * G::M];
    p3_load<P>(A, k, k1, z);
    dft15<1>(z);
    if (spec_out) {
#pragma unroll
      for (int k2 = 0; k2 < 15; k2++) spec_out[g0 + (size_t)(P * k2) * G::M] = z[k2];
    }
#pragma unroll
    for (int k2 = 0; k2 < 15; k2++) z[k2] = cross_power(z[k2], fp[k2]);
    p3_inverse_store<P>(A, k, k1, z);
  }
}

// slot of bin v of the packed column 0 after forward stage 2 (v = k1 + P k2 sits where (k1, n2=k2)
// was read from)
template <int P>
MOF_HD int col0_slot(int v) {
  typedef Geo<P> G;
  const int k1 = v % P, k2 = v / P;
  return (k2 + 15 * (k1 % G::H)) * G::SQ + ((k1 / G::H) ? G::M : 0);
}

// P3Z: packed column 0, one item per bin pair {v, -v}, v in [0, M]: split into F[v,0] and F[v,M]
// for both images, cross-power each, re-pack R[v,0] + i R[v,M] (and its mirror) in place.
template <int P>
MOF_HD void p3z_item(c32* A, int v, const c32* spec_prev, c32* spec_out) {
  typedef Geo<P> G;
  const int mv = (G::N - v) % G::N;
  const int sv = col0_slot<P>(v), sm = col0_slot<P>(mv);
  const c32 wc = A[sv], wcm = A[sm];
  const c32 wp = spec_prev[(size_t)v * G::M], wpm = spec_prev[(size_t)mv * G::M];
  if (spec_out) {  // after the loads: spec_out may alias spec_prev (in-place stream cache)
    spec_out[(size_t)v * G::M] = wc;
    if (v != mv) spec_out[(size_t)mv * G::M] = wcm;
  }
  c32 r0, rM;
  if (v == mv) {  // v = 0 or v = M: the four purely real bins
    r0 = mk(cross_power_real(wc.x, wp.x), 0.f);
    rM = mk(cross_power_real(wc.y, wp.y), 0.f);
  } else {
    c32 fc0 = mk(0.5f * (wc.x + wcm.x), 0.5f * (wc.y - wcm.y));
    c32 fcM = mk(0.5f * (wc.y + wcm.y), 0.5f * (wcm.x - wc.x));
    c32 fp0 = mk(0.5f * (wp.x + wpm.x), 0.5f * (wp.y - wpm.y));
    c32 fpM = mk(0.5f * (wp.y + wpm.y), 0.5f * (wpm.x - wp.x));
    r0 = cross_power(fc0, fp0);
    rM = cross_power(fcM, fpM);
  }
  A[sv] = mk(r0.x - rM.y, r0.y + rM.x);                // R[v,0] + i R[v,M]
  if (v != mv) A[sm] = mk(r0.x + rM.y, rM.x - r0.y);   // conj R[v,0] + i conj R[v,M]
}

// P3I: packed column 0, inverse stage 1: item k1 in [0, P)
template <int P>
MOF_HD void p3i_item(c32* A, int k1) {
  c32 z[15];
  p3_load<P>(A, 0, k1, z);
  p3_inverse_store<P>(A, 0, k1, z);
}

// P4: inverse columns, stage 2.  item (y1, k): P-point inverse DFT over k1; afterwards
// slot(k; q, 0) = G[q, k], slot(k; q, 1) = G[q + M, k]  (G = column-inverse of the cross-power).
template <int P>
MOF_HD void p4_item(c32* A, int y1, int k) {
  typedef Geo<P> G;
  const int kp = (k == 0) ? G::M : G::N - k;
  c32 z[P];
#pragma unroll
  for (int k1 = 0; k1 < P; k1++) z[k1] = A[(y1 + 15 * (k1 % G::H)) * G::SQ + ((k1 / G::H) ? kp : k)];
  DftP<P, -1>::run(z);
#pragma unroll
  for (int y2 = 0; y2 < P; y2++) A[(y1 + 15 * (y2 % G::H)) * G::SQ + ((y2 / G::H) ? kp : k)] = z[y2];
}

// P5a: inverse rows (C2R, rows q and q+M together), stage 1.  item (j, q): the two 15-point inverse
// butterflies n2a = j, n2b = partner(j) over k = P n1 + n2 of X = G[q,:] + i G[q+M,:] (Hermitian
// extension on the fly), twiddle, store (x1, n2) at A[q][P x1 + n2] (in place).
template <int P>
MOF_HD void p5a_item(c32* A, int j, int q) {
  typedef Geo<P> G;
  c32* a = A + q * G::SQ;
  c32 za[15], zb[15];
  const int n2a = j, n2b = (j == 0) ? G::H : P - j;
  if (j != 0) {
    // k = P n1 + n2a < M  <=>  n1 <= 7 ; mirror N - k = P (14 - n1) + n2b
#pragma unroll
    for (int n1 = 0; n1 < 15; n1++) {
      const int k = P * n1 + n2a;
      const int kk = (n1 <= 7) ? k : G::N - k;
      c32 ga = a[kk], gb = a[G::N - kk];
      c32 xk = mk(ga.x - gb.y, ga.y + gb.x);   // X[kk]   = G[q,kk] + i G[q+M,kk]
      c32 xm = mk(ga.x + gb.y, gb.x - ga.y);   // X[N-kk] = conj G[q,kk] + i conj G[q+M,kk]
      za[n1] = (n1 <= 7) ? xk : xm;
      zb[14 - n1] = (n1 <= 7) ? xm : xk;
    }
  } else {
    // n2a = 0: k = P n1, mirror P (15 - n1); k = 0 comes from the packed column
    // n2b = H: k = P n1 + H, mirror P (14 - n1) + H; k = M (n1 = 7) from the packed column
    c32 s0 = a[0], s1 = a[G::M];
    za[0] = mk(s0.x, s1.x);   // X[0] = G[q,0] + i G[q+M,0]
    zb[7] = mk(s0.y, s1.y);   // X[M] = G[q,M] + i G[q+M,M]
#pragma unroll
    for (int n1 = 1; n1 <= 7; n1++) {
      const int k = P * n1;
      c32 ga = a[k], gb = a[G::N - k];
      za[n1] = mk(ga.x - gb.y, ga.y + gb.x);
      za[15 - n1] = mk(ga.x + gb.y, gb.x - ga.y);
    }
#pragma unroll
    for (int n1 = 0; n1 <= 6; n1++) {
      const int k = P * n1 + G::H;
      c32 ga = a[k], gb = a[G::N - k];
      zb[n1] = mk(ga.x - gb.y, ga.y + gb.x);
      zb[14 - n1] = mk(ga.x + gb.y, gb.x - ga.y);
    }
  }
  dft15<-1>(za);
  dft15<-1>(zb);
#pragma unroll
  for (int x1 = 0; x1 < 15; x1++) {
    a[P * x1 + n2a] = (x1 == 0) ? za[0] : twmul<-1>(za[x1], MOF_TW(n2a * (x1 * G::TWS)));
    a[P * x1 + n2b] = (x1 == 0) ? zb[0] : twmul<-1>(zb[x1], MOF_TW(n2b * (x1 * G::TWS)));
  }
}

// P5b: inverse rows, stage 2.  item (x1, q): P-point inverse DFT over n2; result
// (corr[q][x], corr[q+M][x]) for x = x1 + 15 x2 stays in A[q][P x1 + x2]; the running maximum of
// the surface is folded in (its location is found afterwards by P6).
template <int P>
MOF_HD void p5b_item(c32* A, int x1, int q, float& vmax) {
  typedef Geo<P> G;
  c32* a = A + q * G::SQ + P * x1;
  c32 z[P];
#pragma unroll
  for (int n2 = 0; n2 < P; n2++) z[n2] = a[n2];
  DftP<P, -1>::run(z);
#pragma unroll
  for (int x2 = 0; x2 < P; x2++) {
    a[x2] = z[x2];
    vmax = fmaxf(vmax, fmaxf(z[x2].x, z[x2].y));
  }
}

// P6: location of the maximum, first occurrence in fft-shifted row-major order (cv::minMaxLoc).
// item = one slot (q, kslot); returns the shifted linear index if the slot holds the maximum,
// else INT_MAX-like sentinel.  Caller reduces with min.
template <int P>
MOF_HD int p6_item(const c32* A, int q, int kslot, float vmax) {
  typedef Geo<P> G;
  const c32 v = A[q * G::SQ + kslot];
  int best = 0x7fffffff;
  if (v.x == vmax || v.y == vmax) {
    const int x = (kslot / P) + 15 * (kslot % P);
    const int xs = (x < G::M) ? x + G::M : x - G::M;
    if (v.y == vmax) best = q * G::N + xs;                         // row q+M -> shifted row q
    else best = (q + G::M) * G::N + xs;                            // row q   -> shifted row q+M
    // both equal: the smaller index is the row-q+M one (shifted row q < q+M), already chosen
  }
  return best;
}

// value of the fft-shifted correlation surface at (ys, xs) after P5b
template <int P>
MOF_HD float surf_at(const c32* A, int ys, int xs) {
  typedef Geo<P> G;
  const int y = (ys < G::M) ? ys + G::M : ys - G::M;
  const int x = (xs < G::M) ? xs + G::M : xs - G::M;
  const int q = (y < G::M) ? y : y - G::M;
  c32 v = A[q * G::SQ + P * (x % 15) + x / 15];
  return (y < G::M) ? v.x : v.y;
}

// 5x5 centroid around the peak, clamped to the image, all values, FP64
// (weightedCentroid, src/FftMethod.cpp:1330-1384), then shift = t - N/2 and the validity rules of
// src/FftMethod.cpp:1840-1856.  Sequential version (one thread / host).
template <int P>
MOF_HD void centroid_shift(const c32* A, int peak_idx, double max_px_speed_sq, double* out_xy) {
  typedef Geo<P> G;
  const int py = peak_idx / G::N, px = peak_idx % G::N;
  int minr = py - 2, maxr = py + 2, minc = px - 2, maxc = px + 2;
  if (minr < 0) minr = 0;
  if (minc < 0) minc = 0;
  if (maxr > G::N - 1) maxr = G::N - 1;
  if (maxc > G::N - 1) maxc = G::N - 1;
  double cx = 0.0, cy = 0.0, sum = 0.0;
  for (int y = minr; y <= maxr; y++)
    for (int x = minc; x <= maxc; x++) {
      double v = (double)surf_at<P>(A, y, x);
      cx += (double)x * v;
      cy += (double)y * v;
      sum += v;
    }
  sum += DBL_EPSILON;
  double sx = cx / sum - (double)G::N / 2.0;  // -(center - t)
  double sy = cy / sum - (double)G::N / 2.0;
  bool valid = !(sx * sx + sy * sy > max_px_speed_sq || fabs(sx) > (double)G::N / 2.0 ||
                 fabs(sy) > (double)G::N / 2.0);
  if (sx != sx || sy != sy) valid = false;
  const double qnan = (double)NAN;
  out_xy[0] = valid ? sx : qnan;
  out_xy[1] = valid ? sy : qnan;
}

// ------------------------------------------------------------------------------------------------
// Phase driver shared by the CUDA kernel (tid/nthreads = thread index / block size, a block barrier
// between phases) and the CPU emulator (tid = 0, nthreads = 1).  Items of one phase are independent.
// ------------------------------------------------------------------------------------------------
enum { PH_1A = 0, PH_1B, PH_2, PH_3S, PH_3C, PH_3Z, PH_3I, PH_4, PH_5A, PH_5B };

template <int P>
struct SecCtx {
  c32* A;                 // shared-memory section array, Geo<P>::A_ELEMS
  const uint8_t* img;     // top-left pixel of the section being transformed
  int row_stride;         // bytes between image rows
  bool vec;               // pixel addresses allow P/4-byte loads
  const float* win;       // optional N*N window (cv::createHanningWindow layout) or nullptr
  const c32* spec_prev;   // PH_3C/3Z: previous image's spectrum [N][M]
  c32* spec_out;          // PH_3S: where the spectrum goes; PH_3C/3Z: optional copy for the next frame
  float vmax;             // PH_5B: this thread's running maximum
};

template <int P, int PH>
MOF_HD void run_phase(SecCtx<P>& c, int tid, int nthreads) {
  typedef Geo<P> G;
  if (PH == PH_1A) {
    for (int it = tid; it < G::M * 4; it += nthreads)
      p1a_item<P>(c.A, it / 4, it % 4, c.img, c.row_stride, c.win, c.vec);
    return;
  }
  if (PH == PH_3Z) {
    for (int it = tid; it <= G::M; it += nthreads) p3z_item<P>(c.A, it, c.spec_prev, c.spec_out);
    return;
  }
  if (PH == PH_3I) {
    for (int it = tid; it < P; it += nthreads) p3i_item<P>(c.A, it);
    return;
  }
  const int nb = (PH == PH_3S || PH == PH_3C || PH == PH_5A) ? G::H : 15;
  for (int it = tid; it < nb * G::LD; it += nthreads) {
    const int l = it % G::LD, b = it / G::LD;
    if (l >= G::M) continue;
    if (PH == PH_1B) p1b_item<P>(c.A, b, l);
    if (PH == PH_2) p2_item<P>(c.A, b, l);
    if (PH == PH_3S) p3_store_item<P>(c.A, b, l, c.spec_out);
    if (PH == PH_3C) p3_corr_item<P>(c.A, b, l, c.spec_prev, c.spec_out);
    if (PH == PH_4) p4_item<P>(c.A, b, l);
    if (PH == PH_5A) p5a_item<P>(c.A, b, l);
    if (PH == PH_5B) p5b_item<P>(c.A, b, l, c.vmax);
  }
}

// P6 driver: each thread scans its share of the surface slots for the maximum's first location.
template <int P>
MOF_HD int run_search(const c32* A, float vmax, int tid, int nthreads) {
  typedef Geo<P> G;
  int best = 0x7fffffff;
  for (int it = tid; it < G::M * G::N; it += nthreads) {
    const int b = p6_item<P>(A, it / G::N, it % G::N, vmax);
    best = b < best ? b : best;
  }
  return best;
}

}  // namespace mof
