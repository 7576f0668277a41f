// pg_generic.cuh -- phase correlation for section sizes outside the prime-factor plan (N != 60/120/240).
//
// FftMethod's constructor accepts any samplePointSize and falls back to N := frameSize when it does not
// divide the frame (src/FftMethod.cpp:1706-1719): config/default.yaml's 480 px frame can therefore ask
// for one 480 px section, and users pick 30 / 80 / 96 / 160.  cv::phaseCorrelate (the call at :1836)
// transforms such a section as it is whenever N is an "optimal DFT size" (2^a 3^b 5^c); this file covers
// exactly those sizes (N even, <= PG_MAX_N) with a mixed-radix Stockham transform:
//   * both images of a pair ride ONE complex transform, z = curr + i prev; their spectra are separated
//     inside the cross-power pass (A(k) = (Z(k) + conj Z(-k)) / 2, B(k) = (Z(k) - conj Z(-k)) / 2i);
//   * a line (row or column) is transformed by one warp in shared memory, radix 4/2/3/5 passes, twiddles
//     from one table W_N^j computed in double precision on the host;
//   * the section array lives in an L2/HBM scratch of N*N complex floats per resident CTA: 2 passes over
//     it per 2-D transform (rows, column tiles of 16) + 1 for the cross-power.
// The epilogue is the CPU branch's: first maximum in fft-shifted row-major order, 5x5 FP64 centroid,
// NaN rules (shared with the fast path: centroid_acc / the rules of src/FftMethod.cpp:1840-1856).
//
// Host+device like pc_pfa.cuh, so that tests/emu can run it lane by lane without a GPU.
#pragma once

#include <float.h>
#include <math.h>
#include <stdint.h>

#include "fft_regs.cuh"

namespace mof {
namespace pg {

constexpr int PG_MAX_N = 768;

struct PgPlan {
  int M;          // transform length = section size
  int nr;         // number of passes
  int radix[12];  // radix of pass k (4s first, then 2, 3s, 5s)
  // per pass: stride s (product of the earlier radices), s * m (m = n / r: distance between a butterfly's
  // inputs), M / n (twiddle step), and ceil(2^32 / s) so that b / s is one multiply-high for b < 2^16
  int s[12], sm[12], tstep[12];
  uint32_t magic[12];
};

// false when M is not an even 2-3-5-smooth size in range
inline bool pg_plan(int M, PgPlan* pl) {
  if (M < 4 || M > PG_MAX_N || (M & 1)) return false;
  int m = M, nr = 0;
  while (m % 4 == 0) { pl->radix[nr++] = 4; m /= 4; }
  while (m % 2 == 0) { pl->radix[nr++] = 2; m /= 2; }
  while (m % 3 == 0) { pl->radix[nr++] = 3; m /= 3; }
  while (m % 5 == 0) { pl->radix[nr++] = 5; m /= 5; }
  if (m != 1 || nr > 12) return false;
  pl->M = M;
  pl->nr = nr;
  int n = M, st = 1;
  for (int k = 0; k < nr; k++) {
    const int r = pl->radix[k];
    pl->s[k] = st;
    pl->sm[k] = st * (n / r);
    pl->tstep[k] = M / n;
    pl->magic[k] = st == 1 ? 0u : (uint32_t)(((uint64_t)1 << 32) / (uint64_t)st) + 1u;
    n /= r;
    st *= r;
  }
  return true;
}
MOF_HD int pg_div(int b, int s, uint32_t magic) {   // b / s for 0 <= b < 2^16, s <= 768
  if (s == 1) return b;
#if defined(__CUDA_ARCH__)
  return (int)__umulhi((uint32_t)b, magic);
#else
  return (int)(((uint64_t)(uint32_t)b * magic) >> 32);
#endif
}

// line buffer pitch (complex elements): two ping-pong halves of M, padded so that the 16 warps' buffers
// start in different banks when a tile row (16 consecutive columns) is scattered into them
MOF_HD int pg_half(int M) { return M + 1; }
MOF_HD int pg_pitch(int M) {
  int p = 2 * (M + 1);
  while ((p & 15) != 1) p++;
  return p;
}

MOF_HD c32 cmulc(c32 a, c32 w) { return mk(a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x); }
template <bool INV> MOF_HD c32 tw_of(const c32* tw, int j) {   // W^j forward, conj for the inverse kernel
  const c32 w = tw[j];
  return INV ? mk(w.x, -w.y) : w;
}
// multiplication by -i (forward) / +i (inverse)
template <bool INV> MOF_HD c32 rot(c32 a) { return INV ? mk(-a.y, a.x) : mk(a.y, -a.x); }

// One butterfly of one Stockham pass (decimation in frequency, autosort): sub-transform length n = r m,
// stride s, butterfly b = p s + q in [0, M / r):
//   y[q + s (r p + k)] = W_n^{p k} * sum_j x[q + s (p + m j)] w_r^{j k}
// xi = x + q + s p, yo = y + q + s r p, ts = (M / n) p (W_n^p = W_M^ts), sm = s m
template <bool INV, int R>
MOF_HD void pg_butterfly(const c32* xi, c32* yo, const c32* tw, int s, int sm, int ts) {
  if (R == 4) {
    const c32 a0 = xi[0], a1 = xi[sm], a2 = xi[2 * sm], a3 = xi[3 * sm];
    const c32 t0 = cadd(a0, a2), t1 = csub(a0, a2), t2 = cadd(a1, a3), t3 = rot<INV>(csub(a1, a3));
    yo[0] = cadd(t0, t2);
    yo[s] = cmulc(cadd(t1, t3), tw_of<INV>(tw, ts));
    yo[2 * s] = cmulc(csub(t0, t2), tw_of<INV>(tw, 2 * ts));
    yo[3 * s] = cmulc(csub(t1, t3), tw_of<INV>(tw, 3 * ts));
  } else if (R == 2) {
    const c32 a0 = xi[0], a1 = xi[sm];
    yo[0] = cadd(a0, a1);
    yo[s] = cmulc(csub(a0, a1), tw_of<INV>(tw, ts));
  } else if (R == 3) {
    const float S3 = 0.86602540378443864676f;
    const c32 a0 = xi[0], a1 = xi[sm], a2 = xi[2 * sm];
    const c32 t1 = cadd(a1, a2);
    const c32 t2 = mk(a0.x - 0.5f * t1.x, a0.y - 0.5f * t1.y);
    const c32 d = csub(a1, a2);
    const c32 t3 = rot<INV>(mk(S3 * d.x, S3 * d.y));   // -+ i sin(2 pi / 3) (a1 - a2)
    yo[0] = cadd(a0, t1);
    yo[s] = cmulc(cadd(t2, t3), tw_of<INV>(tw, ts));
    yo[2 * s] = cmulc(csub(t2, t3), tw_of<INV>(tw, 2 * ts));
  } else {   // R == 5
    const float C1 = 0.30901699437494742410f, C2 = -0.80901699437494742410f;
    const float S1 = 0.95105651629515357212f, S2 = 0.58778525229247312917f;
    const c32 a0 = xi[0], a1 = xi[sm], a2 = xi[2 * sm], a3 = xi[3 * sm], a4 = xi[4 * sm];
    const c32 t1 = cadd(a1, a4), t2 = cadd(a2, a3), t3 = csub(a1, a4), t4 = csub(a2, a3);
    const c32 m1 = mk(a0.x + C1 * t1.x + C2 * t2.x, a0.y + C1 * t1.y + C2 * t2.y);
    const c32 m2 = mk(a0.x + C2 * t1.x + C1 * t2.x, a0.y + C2 * t1.y + C1 * t2.y);
    const c32 n1 = rot<INV>(mk(S1 * t3.x + S2 * t4.x, S1 * t3.y + S2 * t4.y));
    const c32 n2 = rot<INV>(mk(S2 * t3.x - S1 * t4.x, S2 * t3.y - S1 * t4.y));
    yo[0] = cadd(a0, cadd(t1, t2));
    yo[s] = cmulc(cadd(m1, n1), tw_of<INV>(tw, ts));
    yo[2 * s] = cmulc(cadd(m2, n2), tw_of<INV>(tw, 2 * ts));
    yo[3 * s] = cmulc(csub(m2, n2), tw_of<INV>(tw, 3 * ts));
    yo[4 * s] = cmulc(csub(m1, n1), tw_of<INV>(tw, 4 * ts));
  }
}

template <bool INV, int R>
MOF_HD void pg_line_pass_r(const c32* x, c32* y, const c32* tw, const PgPlan& pl, int k, int lane) {
  const int s = pl.s[k], sm = pl.sm[k], tstep = pl.tstep[k], nb = pl.M / R;
  const uint32_t magic = pl.magic[k];
  for (int b = lane; b < nb; b += 32) {
    const int p = pg_div(b, s, magic), q = b - p * s;
    pg_butterfly<INV, R>(x + q + s * p, y + q + s * R * p, tw, s, sm, tstep * p);
  }
}

// this lane's butterflies of pass k on one line; buf = the line's two halves (pitch pg_half(M)).
// Input of pass k is half (k & 1), output the other one.  A __syncwarp() belongs between passes.
template <bool INV>
MOF_HD void pg_line_pass(c32* buf, const c32* tw, const PgPlan& pl, int k, int lane) {
  const c32* x = buf + (k & 1) * pg_half(pl.M);
  c32* y = buf + ((k & 1) ^ 1) * pg_half(pl.M);
  switch (pl.radix[k]) {
    case 4: pg_line_pass_r<INV, 4>(x, y, tw, pl, k, lane); break;
    case 2: pg_line_pass_r<INV, 2>(x, y, tw, pl, k, lane); break;
    case 3: pg_line_pass_r<INV, 3>(x, y, tw, pl, k, lane); break;
    default: pg_line_pass_r<INV, 5>(x, y, tw, pl, k, lane); break;
  }
}
// half of the line buffer that holds the result after all passes
MOF_HD int pg_result_half(const PgPlan& pl) { return pl.nr & 1; }

// Cross-power of the bin pair (k, -k) from the packed transform Z = FFT(curr + i prev):
//   A = FFT(curr), B = FFT(prev), P = A conj(B) (mulSpectrums, conjB = true), C = P |P| / (|P|^2 + eps)
// (magSpectrums + divSpectrums, src/FftMethod.cpp:113-114, 1166-1172).  zk = Z(k), zm = Z(-k);
// returns C(k); C(-k) is its conjugate.  self: k == -k (the four real bins): C = P / (|P| + eps).
MOF_HD c32 pg_cross(c32 zk, c32 zm, bool self) {
  // 2A = Z(k) + conj Z(-k);  2B = (Z(k) - conj Z(-k)) / i
  const float ax = zk.x + zm.x, ay = zk.y - zm.y;
  const float dx = zk.x - zm.x, dy = zk.y + zm.y;   // Z(k) - conj Z(-k) = (dx, dy)
  const float bx = dy, by = -dx;                    // divided by i
  if (self) {   // A, B real: ay == by == 0 exactly
    const float p = 0.25f * (ax * bx);
    return mk(p / (fabsf(p) + FLT_EPSILON), 0.f);
  }
  const float pr = 0.25f * (ax * bx + ay * by);
  const float pi = 0.25f * (ay * bx - ax * by);
  const float mag = sqrtf(pr * pr + pi * pi);
  const float sc = mag / (mag * mag + FLT_EPSILON);
  return mk(pr * sc, pi * sc);
}

// fft-shifted row-major index of surface element (y, x), M even (cv2's fftShift swaps the halves)
MOF_HD int pg_shifted_index(int y, int x, int M) {
  const int h = M >> 1;
  const int ys = y < h ? y + h : y - h, xs = x < h ? x + h : x - h;
  return ys * M + xs;
}
MOF_HD int pg_unshift(int v, int M) { const int h = M >> 1; return v < h ? v + h : v - h; }

// running "first maximum" (cv::minMaxLoc over the shifted surface): larger value, or equal value at a
// smaller shifted index
MOF_HD void pg_max_update(float v, int sidx, float& best, int& bidx) {
  if (v > best || (v == best && sidx < bidx)) { best = v; bidx = sidx; }
}

// FP64 accumulation of one centroid tap at shifted coordinates (weightedCentroid, src/FftMethod.cpp:1330-1384)
MOF_HD void pg_acc(int ys, int xs, float val, double& cx, double& cy, double& sum) {
  const double v = (double)val;
  cx += (double)xs * v;
  cy += (double)ys * v;
  sum += v;
}

// shift and validity rules from the centroid sums (src/FftMethod.cpp:1840-1856); N = section size
MOF_HD void pg_finish(double cx, double cy, double sum, int N, double max_px_speed_sq, double* out_xy) {
  sum += DBL_EPSILON;
  const double sx = cx / sum - (double)N / 2.0;
  const double sy = cy / sum - (double)N / 2.0;
  bool valid = !(sx * sx + sy * sy > max_px_speed_sq || fabs(sx) > (double)N / 2.0 || fabs(sy) > (double)N / 2.0);
  if (sx != sx || sy != sy) valid = false;
  out_xy[0] = valid ? sx : (double)NAN;
  out_xy[1] = valid ? sy : (double)NAN;
}

// scaleRotationEstimator::processImage tail (src/scaleRotationEstimator.cpp:117-126) on the shift of the
// log-polar pair: pt = -(shift); |pt.x| > res / 2 or NaN -> (1, 0) (or NaN, NaN when reject_nan)
MOF_HD void pg_finish_scalerot(double cx, double cy, double sum, int N, double sr_m, int reject_nan, double* out_xy) {
  double xy[2];
  pg_finish(cx, cy, sum, N, 1e300, xy);
  const double ptx = -xy[0], pty = -xy[1];
  if (fabs(ptx) > (double)(N / 2) || ptx != ptx || pty != pty) {
    out_xy[0] = reject_nan ? (double)NAN : 1.0;
    out_xy[1] = reject_nan ? (double)NAN : 0.0;
  } else {
    const double Ky = (double)N / 360.0;
    out_xy[0] = exp(ptx / sr_m);
    out_xy[1] = (pty / Ky) * (3.14159265358979323846 / 180.0);
  }
}

}  // namespace pg
}  // namespace mof
