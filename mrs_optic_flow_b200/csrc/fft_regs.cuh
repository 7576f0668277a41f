// fft_regs.cuh -- register-resident small DFTs (radix 2/3/4/5 primitives, 15 = 3x5 prime-factor,
// 8 and 16 = Cooley-Tukey over radix 4/2) used by the phase-correlation kernels.
//
// Everything here is host+device so the section pipeline can be executed thread-by-thread on the
// CPU by tests/emu (index-map verification without a GPU).  On the device the complex arithmetic
// uses Blackwell's packed FP32x2 instructions (FADD2 / FMUL2 / FFMA2: one instruction per complex
// add, the swap and sign flips of a multiplication by +-i ride on the operand modifiers, constants
// live in uniform registers), which halves the issue slots of every butterfly; the host path does
// the same IEEE operations one component at a time.  DIR = +1: forward kernel
// exp(-2 pi i nk/n) (cv::dft); DIR = -1: inverse kernel exp(+2 pi i nk/n), unscaled (cv::idft
// without DFT_SCALE, as cv::phaseCorrelate calls it -- reference call site src/FftMethod.cpp:1836).
#pragma once

#if defined(__CUDACC__)
#define MOF_HD __host__ __device__ __forceinline__
#define MOF_ALIGN8 __align__(8)
#else
#define MOF_HD inline
#define MOF_ALIGN8 alignas(8)
#endif

namespace mof {

struct MOF_ALIGN8 c32 {
  float x, y;
};

MOF_HD c32 mk(float x, float y) { c32 r; r.x = x; r.y = y; return r; }
#if defined(__CUDA_ARCH__)
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(c32 a) {
  u64 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a.x), "f"(a.y));
  return r;
}
__device__ __forceinline__ c32 upk(u64 v) {
  c32 r;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
  return r;
}
#endif
// per-component a + b, a - b, a * b, a * b + c (one packed instruction each on the device)
MOF_HD c32 cadd(c32 a, c32 b) {
#if defined(__CUDA_ARCH__)
  u64 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)));
  return upk(d);
#else
  return mk(a.x + b.x, a.y + b.y);
#endif
}
MOF_HD c32 csub(c32 a, c32 b) {
#if defined(__CUDA_ARCH__)
  u64 d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)));
  return upk(d);
#else
  return mk(a.x - b.x, a.y - b.y);
#endif
}
MOF_HD c32 pmul(c32 a, c32 b) {
#if defined(__CUDA_ARCH__)
  u64 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)));
  return upk(d);
#else
  return mk(a.x * b.x, a.y * b.y);
#endif
}
MOF_HD c32 pfma(c32 a, c32 b, c32 c) {
#if defined(__CUDA_ARCH__)
  u64 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)), "l"(pk(c)));
  return upk(d);
#else
  return mk(a.x * b.x + c.x, a.y * b.y + c.y);
#endif
}
MOF_HD c32 pswap(c32 a) { return mk(a.y, a.x); }
MOF_HD c32 pneg(c32 a) { return mk(-a.x, -a.y); }
MOF_HD c32 cmul(c32 a, c32 b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
MOF_HD c32 cconj(c32 a) { return mk(a.x, -a.y); }
// multiply by -i (DIR=+1) or +i (DIR=-1)
// (v, -v) / (-v, v) constant pairs of the butterflies.  As literals ptxas rebuilds each pair with two UMOV
// wherever it is used; MOF_CONST_BANK = 1 reads them from constant memory instead (one LDCU per pair or two,
// kept in uniform registers where the allocator can).
#if !defined(MOF_CONST_BANK)
#define MOF_CONST_BANK 0
#endif
#define MOF_K_S3 0.86602540378443864676f
#define MOF_K_S5 0.95105651629515357212f
#define MOF_K_S16 0.38268343236508978178f
#define MOF_K_R2 0.70710678118654752440f
#define MOF_K_C16 0.92387953251128673848f
#if defined(__CUDACC__) && MOF_CONST_BANK
static __constant__ float2 mof_kpn[10] = {
    {MOF_K_S3, -MOF_K_S3},   {-MOF_K_S3, MOF_K_S3},   {MOF_K_S5, -MOF_K_S5}, {-MOF_K_S5, MOF_K_S5},
    {MOF_K_S16, -MOF_K_S16}, {-MOF_K_S16, MOF_K_S16}, {MOF_K_R2, -MOF_K_R2}, {-MOF_K_R2, MOF_K_R2},
    {MOF_K_C16, -MOF_K_C16}, {-MOF_K_C16, MOF_K_C16}};
#endif
// the pair (v, -v) for v = +-{S3, S5, S16, R2, C16}[which]; `v` must be a compile-time constant after inlining
MOF_HD c32 kpn(int which, float v) {
#if defined(__CUDA_ARCH__) && MOF_CONST_BANK
  const float2 k = mof_kpn[2 * which + (v < 0.f ? 1 : 0)];
  return mk(k.x, k.y);
#else
  return mk(v, -v);
#endif
}

template <int DIR>
MOF_HD c32 mul_mi(c32 a) { return DIR > 0 ? mk(a.y, -a.x) : mk(-a.y, a.x); }
// multiply by a forward-table twiddle w (DIR=+1) or its conjugate (DIR=-1)
template <int DIR>
MOF_HD c32 twmul(c32 a, c32 w) {
  const float sy = DIR > 0 ? w.y : -w.y;
#if defined(__CUDA_ARCH__) && MOF_CONST_BANK
  const float m = sy < 0.f ? -sy : sy;   // dft8 / dft16 pass literals: the chain folds to one table entry
  if (m == MOF_K_S16) return pfma(a, mk(w.x, w.x), pmul(pswap(a), kpn(2, -sy)));
  if (m == MOF_K_R2) return pfma(a, mk(w.x, w.x), pmul(pswap(a), kpn(3, -sy)));
  if (m == MOF_K_C16) return pfma(a, mk(w.x, w.x), pmul(pswap(a), kpn(4, -sy)));
#endif
  return pfma(a, mk(w.x, w.x), pmul(pswap(a), mk(-sy, sy)));
}

template <int DIR>
MOF_HD void dft2(c32& a, c32& b) {
  c32 t = a;
  a = cadd(t, b);
  b = csub(t, b);
}

// fused-multiply-add helpers: a + s*b and a - s*b on one component
MOF_HD float fmaf_(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(a, b, c);
#else
  return a * b + c;
#endif
}

// 3-point DFT, 6 packed instructions (2 add + 1 add + 3 fma): the sqrt(3)/2 rotation is folded
// into the output fmas.
template <int DIR>
MOF_HD void dft3(c32& x0, c32& x1, c32& x2) {
  const float S = DIR > 0 ? 0.86602540378443864676f : -0.86602540378443864676f;
  const c32 s = cadd(x1, x2), d = csub(x1, x2);
  const c32 t = pfma(s, mk(-0.5f, -0.5f), x0);
  x0 = cadd(x0, s);
  // forward: X1 = t - i*S*d, X2 = t + i*S*d
  const c32 sd = pswap(d);
  x1 = pfma(sd, kpn(0, S), t);
  x2 = pfma(sd, kpn(0, -S), t);
}

template <int DIR>
MOF_HD void dft4(c32& x0, c32& x1, c32& x2, c32& x3) {
  c32 a = cadd(x0, x2), b = csub(x0, x2), c = cadd(x1, x3), d = csub(x1, x3);
  c32 e = mul_mi<DIR>(d);
  x0 = cadd(a, c);
  x1 = cadd(b, e);
  x2 = csub(a, c);
  x3 = csub(b, e);
}

// 5-point DFT, 16 packed instructions (6 add + 10 fma): v1 = S1 (b1 + r b2), v2 = S1 (r b1 - b2)
// with r = S2/S1, and the S1 scaling folded into the output fmas.
template <int DIR>
MOF_HD void dft5(c32& x0, c32& x1, c32& x2, c32& x3, c32& x4) {
  const float C1 = 0.30901699437494742410f, C2 = -0.80901699437494742410f;
  const float S1 = DIR > 0 ? 0.95105651629515357212f : -0.95105651629515357212f;
  const float R = 0.61803398874989484820f;  // S2 / S1
  const c32 a1 = cadd(x1, x4), b1 = csub(x1, x4), a2 = cadd(x2, x3), b2 = csub(x2, x3);
  const c32 t1 = pfma(a2, mk(C2, C2), pfma(a1, mk(C1, C1), x0));
  const c32 t2 = pfma(a2, mk(C1, C1), pfma(a1, mk(C2, C2), x0));
  const c32 w1 = pfma(b2, mk(R, R), b1);
  const c32 w2 = pfma(b1, mk(R, R), pneg(b2));
  x0 = cadd(cadd(x0, a1), a2);
  // forward: X1 = t1 - i S1 w1, X4 = t1 + i S1 w1, X2 = t2 - i S1 w2, X3 = t2 + i S1 w2
  const c32 s1 = pswap(w1), s2 = pswap(w2);
  x1 = pfma(s1, kpn(1, S1), t1);
  x4 = pfma(s1, kpn(1, -S1), t1);
  x2 = pfma(s2, kpn(1, S1), t2);
  x3 = pfma(s2, kpn(1, -S1), t2);
}

// 15-point DFT, Good-Thomas 3x5 (no internal twiddles).
// input  n = (5 n1 + 3 n2) mod 15, output k = (10 k1 + 6 k2) mod 15.
template <int DIR>
MOF_HD void dft15(c32 (&x)[15]) {
  c32 a[5][3];
#pragma unroll
  for (int n2 = 0; n2 < 5; n2++) {
#pragma unroll
    for (int n1 = 0; n1 < 3; n1++) a[n2][n1] = x[(5 * n1 + 3 * n2) % 15];
    dft3<DIR>(a[n2][0], a[n2][1], a[n2][2]);
  }
#pragma unroll
  for (int k1 = 0; k1 < 3; k1++) {
    dft5<DIR>(a[0][k1], a[1][k1], a[2][k1], a[3][k1], a[4][k1]);
#pragma unroll
    for (int k2 = 0; k2 < 5; k2++) x[(10 * k1 + 6 * k2) % 15] = a[k2][k1];
  }
}

// W16^k = exp(-2 pi i k/16) table entries used by dft8/dft16 (k = 1, 2, 3 and products)
#define MOF_C1_16 0.92387953251128673848f /* cos(pi/8) */
#define MOF_S1_16 0.38268343236508978178f /* sin(pi/8) */
#define MOF_R2 0.70710678118654752440f    /* cos(pi/4) */

template <int DIR>
MOF_HD c32 mul_w8_1(c32 a) {  // * exp(-+ i pi/4)
  return twmul<DIR>(a, mk(MOF_R2, -MOF_R2));
}
template <int DIR>
MOF_HD c32 mul_w8_3(c32 a) {  // * exp(-+ i 3pi/4)
  return twmul<DIR>(a, mk(-MOF_R2, -MOF_R2));
}

// 8-point DFT in natural order: n = 2 n1 + n2 (n1<4, n2<2), k = k1 + 4 k2.
template <int DIR>
MOF_HD void dft8(c32 (&x)[8]) {
  // two 4-point DFTs over even / odd samples
  dft4<DIR>(x[0], x[2], x[4], x[6]);
  dft4<DIR>(x[1], x[3], x[5], x[7]);
  // twiddle odd branch by W8^k1
  x[3] = mul_w8_1<DIR>(x[3]);
  x[5] = mul_mi<DIR>(x[5]);
  x[7] = mul_w8_3<DIR>(x[7]);
  // radix-2 combine: X[k1] = E[k1] + O[k1], X[k1+4] = E[k1] - O[k1]
  c32 e0 = x[0], e1 = x[2], e2 = x[4], e3 = x[6];
  c32 o0 = x[1], o1 = x[3], o2 = x[5], o3 = x[7];
  x[0] = cadd(e0, o0); x[4] = csub(e0, o0);
  x[1] = cadd(e1, o1); x[5] = csub(e1, o1);
  x[2] = cadd(e2, o2); x[6] = csub(e2, o2);
  x[3] = cadd(e3, o3); x[7] = csub(e3, o3);
}

// 16-point DFT in natural order: n = 4 n1 + n2, k = k1 + 4 k2 (4x4 Cooley-Tukey).
template <int DIR>
MOF_HD void dft16(c32 (&x)[16]) {
  // stage 1: for each n2, 4-point DFT over n1 (inputs x[4 n1 + n2]) -> index k1
#pragma unroll
  for (int n2 = 0; n2 < 4; n2++) dft4<DIR>(x[n2], x[4 + n2], x[8 + n2], x[12 + n2]);
  // now x[4 k1 + n2] holds stage-1 output (k1, n2); twiddle by W16^(n2 k1)
  const c32 w1 = mk(MOF_C1_16, -MOF_S1_16);   // W16^1
  const c32 w2 = mk(MOF_R2, -MOF_R2);         // W16^2
  const c32 w3 = mk(MOF_S1_16, -MOF_C1_16);   // W16^3
  const c32 w6 = mk(-MOF_R2, -MOF_R2);        // W16^6
  const c32 w9 = mk(-MOF_C1_16, MOF_S1_16);   // W16^9
  x[4 * 1 + 1] = twmul<DIR>(x[4 * 1 + 1], w1);
  x[4 * 1 + 2] = twmul<DIR>(x[4 * 1 + 2], w2);
  x[4 * 1 + 3] = twmul<DIR>(x[4 * 1 + 3], w3);
  x[4 * 2 + 1] = twmul<DIR>(x[4 * 2 + 1], w2);
  x[4 * 2 + 2] = mul_mi<DIR>(x[4 * 2 + 2]);   // W16^4 = -i
  x[4 * 2 + 3] = twmul<DIR>(x[4 * 2 + 3], w6);
  x[4 * 3 + 1] = twmul<DIR>(x[4 * 3 + 1], w3);
  x[4 * 3 + 2] = twmul<DIR>(x[4 * 3 + 2], w6);
  x[4 * 3 + 3] = twmul<DIR>(x[4 * 3 + 3], w9);
  // stage 2: for each k1, 4-point DFT over n2 -> k2 ; result X[k1 + 4 k2] lands in x[4 k1 + k2]
#pragma unroll
  for (int k1 = 0; k1 < 4; k1++) dft4<DIR>(x[4 * k1], x[4 * k1 + 1], x[4 * k1 + 2], x[4 * k1 + 3]);
  // transpose 4x4 so that x[k] is natural order: X[k1 + 4 k2] = x[4 k1 + k2]
#pragma unroll
  for (int k1 = 0; k1 < 4; k1++)
#pragma unroll
    for (int k2 = k1 + 1; k2 < 4; k2++) {
      c32 t = x[4 * k1 + k2];
      x[4 * k1 + k2] = x[4 * k2 + k1];
      x[4 * k2 + k1] = t;
    }
}

template <int P, int DIR>
struct DftP;
template <int DIR>
struct DftP<4, DIR> {
  static MOF_HD void run(c32 (&x)[4]) { dft4<DIR>(x[0], x[1], x[2], x[3]); }
};
template <int DIR>
struct DftP<8, DIR> {
  static MOF_HD void run(c32 (&x)[8]) { dft8<DIR>(x); }
};
template <int DIR>
struct DftP<16, DIR> {
  static MOF_HD void run(c32 (&x)[16]) { dft16<DIR>(x); }
};

}  // namespace mof
