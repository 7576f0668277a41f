/*
 * mof_oracle.c -- CPU ORACLE (test infrastructure, see mof_oracle.h for scope and citations).
 *
 * Part 1: float32 mixed-radix FFT + cv::phaseCorrelate restatement + FftMethod CPU branch.
 * Part 2: FastSpacedBM SAD search + histogram vote.
 * Part 3: cv::logPolar / cv::remap fixed-point restatement + scaleRotationEstimator.
 *
 * Nothing in here is copied from /root/reference or from OpenCV; it restates the published
 * algorithms (file:line citations inline) with its own data layout (SoA complex planes,
 * Stockham autosort over interleaved sequences).
 */
#include "mof_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <pthread.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* Minimal pthread parallel-for: `nthreads` workers pull indices from a shared counter.
 * (libgomp's spec file is not reachable from this image's gcc wrapper, so no OpenMP.) */
typedef void (*pf_body)(int index, void* thread_state, void* ctx);
typedef void* (*pf_state_new)(void* ctx);
typedef void (*pf_state_free)(void* thread_state);
typedef struct {
  int n;
  int next;
  pthread_mutex_t mu;
  pf_body body;
  pf_state_new snew;
  pf_state_free sfree;
  void* ctx;
  int fail;
} pf_job;

static void* pf_worker(void* arg) {
  pf_job* job = (pf_job*)arg;
  void* st = job->snew ? job->snew(job->ctx) : NULL;
  if (job->snew && !st) {
    pthread_mutex_lock(&job->mu);
    job->fail = 1;
    pthread_mutex_unlock(&job->mu);
    return NULL;
  }
  for (;;) {
    pthread_mutex_lock(&job->mu);
    int i = job->next++;
    pthread_mutex_unlock(&job->mu);
    if (i >= job->n) break;
    job->body(i, st, job->ctx);
  }
  if (job->sfree && st) job->sfree(st);
  return NULL;
}

static int parallel_for(int n, int nthreads, pf_body body, pf_state_new snew, pf_state_free sfree,
                        void* ctx) {
  pf_job job;
  job.n = n; job.next = 0; job.body = body; job.snew = snew; job.sfree = sfree; job.ctx = ctx;
  job.fail = 0;
  pthread_mutex_init(&job.mu, NULL);
  if (nthreads < 1) nthreads = 1;
  if (nthreads > n) nthreads = n > 0 ? n : 1;
  if (nthreads == 1) {
    pf_worker(&job);
  } else {
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    int started = 0;
    for (int t = 0; t < nthreads; t++)
      if (pthread_create(&th[t], NULL, pf_worker, &job) == 0) started++; else break;
    if (started == 0) pf_worker(&job);
    for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
    free(th);
  }
  pthread_mutex_destroy(&job.mu);
  return job.fail ? -2 : 0;
}

/* ------------------------------------------------------------------------------------------ */
/* Part 1a: Stockham FFT over `s0` interleaved sequences of length n (SoA float planes).       */
/* ------------------------------------------------------------------------------------------ */

typedef struct {
  int n;
  int nr;
  int radix[32];
  float* tw_re; /* w_n^k = exp(-2 pi i k / n), k in [0,n) */
  float* tw_im;
} fft_plan;

static int plan_init(fft_plan* pl, int n) {
  pl->n = n;
  pl->nr = 0;
  int m = n;
  while (m % 4 == 0) { pl->radix[pl->nr++] = 4; m /= 4; }
  while (m % 2 == 0) { pl->radix[pl->nr++] = 2; m /= 2; }
  while (m % 3 == 0) { pl->radix[pl->nr++] = 3; m /= 3; }
  while (m % 5 == 0) { pl->radix[pl->nr++] = 5; m /= 5; }
  if (m != 1) return -1; /* not 2-3-5 smooth: cv::phaseCorrelate would zero-pad; unsupported */
  pl->tw_re = (float*)malloc(sizeof(float) * (size_t)n);
  pl->tw_im = (float*)malloc(sizeof(float) * (size_t)n);
  if (!pl->tw_re || !pl->tw_im) return -2;
  for (int k = 0; k < n; k++) {
    double a = -2.0 * M_PI * (double)k / (double)n;
    pl->tw_re[k] = (float)cos(a);
    pl->tw_im[k] = (float)sin(a);
  }
  return 0;
}

static void plan_free(fft_plan* pl) {
  free(pl->tw_re);
  free(pl->tw_im);
  pl->tw_re = pl->tw_im = NULL;
}

/* One DIF Stockham pass of radix r.  x -> y.  l = current sub-length, s = current stride. */
static void pass_r2(int l, int s, int n, const fft_plan* pl, int inv, const float* xr,
                    const float* xi, float* yr, float* yi) {
  const int m = l / 2, ts = n / l;
  for (int p = 0; p < m; p++) {
    float wr = pl->tw_re[(p * ts) % n], wi = pl->tw_im[(p * ts) % n];
    if (inv) wi = -wi;
    const float *a_r = xr + (size_t)s * p, *a_i = xi + (size_t)s * p;
    const float *b_r = xr + (size_t)s * (p + m), *b_i = xi + (size_t)s * (p + m);
    float *o0r = yr + (size_t)s * (2 * p), *o0i = yi + (size_t)s * (2 * p);
    float *o1r = yr + (size_t)s * (2 * p + 1), *o1i = yi + (size_t)s * (2 * p + 1);
    for (int q = 0; q < s; q++) {
      float ar = a_r[q], ai = a_i[q], br = b_r[q], bi = b_i[q];
      o0r[q] = ar + br;
      o0i[q] = ai + bi;
      float dr = ar - br, di = ai - bi;
      o1r[q] = dr * wr - di * wi;
      o1i[q] = dr * wi + di * wr;
    }
  }
}

static void pass_r4(int l, int s, int n, const fft_plan* pl, int inv, const float* xr,
                    const float* xi, float* yr, float* yi) {
  const int m = l / 4, ts = n / l;
  for (int p = 0; p < m; p++) {
    float w1r = pl->tw_re[(p * ts) % n], w1i = pl->tw_im[(p * ts) % n];
    float w2r = pl->tw_re[(2 * p * ts) % n], w2i = pl->tw_im[(2 * p * ts) % n];
    float w3r = pl->tw_re[(3 * p * ts) % n], w3i = pl->tw_im[(3 * p * ts) % n];
    if (inv) { w1i = -w1i; w2i = -w2i; w3i = -w3i; }
    const float sg = inv ? -1.f : 1.f; /* multiply by -i (fwd) or +i (inv) */
    const float *x0r = xr + (size_t)s * p, *x0i = xi + (size_t)s * p;
    const float *x1r = xr + (size_t)s * (p + m), *x1i = xi + (size_t)s * (p + m);
    const float *x2r = xr + (size_t)s * (p + 2 * m), *x2i = xi + (size_t)s * (p + 2 * m);
    const float *x3r = xr + (size_t)s * (p + 3 * m), *x3i = xi + (size_t)s * (p + 3 * m);
    float *o0r = yr + (size_t)s * (4 * p), *o0i = yi + (size_t)s * (4 * p);
    float *o1r = o0r + s, *o1i = o0i + s, *o2r = o0r + 2 * s, *o2i = o0i + 2 * s;
    float *o3r = o0r + 3 * s, *o3i = o0i + 3 * s;
    for (int q = 0; q < s; q++) {
      float ar = x0r[q] + x2r[q], ai = x0i[q] + x2i[q];
      float br = x0r[q] - x2r[q], bi = x0i[q] - x2i[q];
      float cr = x1r[q] + x3r[q], ci = x1i[q] + x3i[q];
      float dr = x1r[q] - x3r[q], di = x1i[q] - x3i[q];
      /* -i*d (fwd) = (di, -dr); +i*d (inv) = (-di, dr) */
      float er = sg * di, ei = -sg * dr;
      o0r[q] = ar + cr;
      o0i[q] = ai + ci;
      float t1r = br + er, t1i = bi + ei;
      float t2r = ar - cr, t2i = ai - ci;
      float t3r = br - er, t3i = bi - ei;
      o1r[q] = t1r * w1r - t1i * w1i;
      o1i[q] = t1r * w1i + t1i * w1r;
      o2r[q] = t2r * w2r - t2i * w2i;
      o2i[q] = t2r * w2i + t2i * w2r;
      o3r[q] = t3r * w3r - t3i * w3i;
      o3i[q] = t3r * w3i + t3i * w3r;
    }
  }
}

static void pass_r3(int l, int s, int n, const fft_plan* pl, int inv, const float* xr,
                    const float* xi, float* yr, float* yi) {
  const int m = l / 3, ts = n / l;
  const float c = -0.5f;
  const float sn = inv ? 0.86602540378443864676f : -0.86602540378443864676f;
  for (int p = 0; p < m; p++) {
    float w1r = pl->tw_re[(p * ts) % n], w1i = pl->tw_im[(p * ts) % n];
    float w2r = pl->tw_re[(2 * p * ts) % n], w2i = pl->tw_im[(2 * p * ts) % n];
    if (inv) { w1i = -w1i; w2i = -w2i; }
    const float *x0r = xr + (size_t)s * p, *x0i = xi + (size_t)s * p;
    const float *x1r = xr + (size_t)s * (p + m), *x1i = xi + (size_t)s * (p + m);
    const float *x2r = xr + (size_t)s * (p + 2 * m), *x2i = xi + (size_t)s * (p + 2 * m);
    float *o0r = yr + (size_t)s * (3 * p), *o0i = yi + (size_t)s * (3 * p);
    float *o1r = o0r + s, *o1i = o0i + s, *o2r = o0r + 2 * s, *o2i = o0i + 2 * s;
    for (int q = 0; q < s; q++) {
      float sr = x1r[q] + x2r[q], si = x1i[q] + x2i[q];
      float dr = x1r[q] - x2r[q], di = x1i[q] - x2i[q];
      o0r[q] = x0r[q] + sr;
      o0i[q] = x0i[q] + si;
      float tr = x0r[q] + c * sr, ti = x0i[q] + c * si;
      /* i*sn*d = (-sn*di, sn*dr) */
      float ur = -sn * di, ui = sn * dr;
      float t1r = tr + ur, t1i = ti + ui;
      float t2r = tr - ur, t2i = ti - ui;
      o1r[q] = t1r * w1r - t1i * w1i;
      o1i[q] = t1r * w1i + t1i * w1r;
      o2r[q] = t2r * w2r - t2i * w2i;
      o2i[q] = t2r * w2i + t2i * w2r;
    }
  }
}

static void pass_r5(int l, int s, int n, const fft_plan* pl, int inv, const float* xr,
                    const float* xi, float* yr, float* yi) {
  const int m = l / 5, ts = n / l;
  const float c1 = 0.30901699437494742410f, c2 = -0.80901699437494742410f;
  const float s1 = inv ? 0.95105651629515357212f : -0.95105651629515357212f;
  const float s2 = inv ? 0.58778525229247312917f : -0.58778525229247312917f;
  for (int p = 0; p < m; p++) {
    float wr[5], wi[5];
    for (int j = 1; j < 5; j++) {
      wr[j] = pl->tw_re[(j * p * ts) % n];
      wi[j] = pl->tw_im[(j * p * ts) % n];
      if (inv) wi[j] = -wi[j];
    }
    const float* xr_[5];
    const float* xi_[5];
    for (int j = 0; j < 5; j++) {
      xr_[j] = xr + (size_t)s * (p + j * m);
      xi_[j] = xi + (size_t)s * (p + j * m);
    }
    float* or_ = yr + (size_t)s * (5 * p);
    float* oi_ = yi + (size_t)s * (5 * p);
    for (int q = 0; q < s; q++) {
      float a1r = xr_[1][q] + xr_[4][q], a1i = xi_[1][q] + xi_[4][q];
      float b1r = xr_[1][q] - xr_[4][q], b1i = xi_[1][q] - xi_[4][q];
      float a2r = xr_[2][q] + xr_[3][q], a2i = xi_[2][q] + xi_[3][q];
      float b2r = xr_[2][q] - xr_[3][q], b2i = xi_[2][q] - xi_[3][q];
      float x0r = xr_[0][q], x0i = xi_[0][q];
      or_[q] = x0r + a1r + a2r;
      oi_[q] = x0i + a1i + a2i;
      float t1r = x0r + c1 * a1r + c2 * a2r, t1i = x0i + c1 * a1i + c2 * a2i;
      float t2r = x0r + c2 * a1r + c1 * a2r, t2i = x0i + c2 * a1i + c1 * a2i;
      /* i*(s1*b1 + s2*b2), i*(s2*b1 - s1*b2) */
      float u1r = -(s1 * b1i + s2 * b2i), u1i = (s1 * b1r + s2 * b2r);
      float u2r = -(s2 * b1i - s1 * b2i), u2i = (s2 * b1r - s1 * b2r);
      float y1r = t1r + u1r, y1i = t1i + u1i;
      float y4r = t1r - u1r, y4i = t1i - u1i;
      float y2r = t2r + u2r, y2i = t2i + u2i;
      float y3r = t2r - u2r, y3i = t2i - u2i;
      or_[q + s] = y1r * wr[1] - y1i * wi[1];
      oi_[q + s] = y1r * wi[1] + y1i * wr[1];
      or_[q + 2 * s] = y2r * wr[2] - y2i * wi[2];
      oi_[q + 2 * s] = y2r * wi[2] + y2i * wr[2];
      or_[q + 3 * s] = y3r * wr[3] - y3i * wi[3];
      oi_[q + 3 * s] = y3r * wi[3] + y3i * wr[3];
      or_[q + 4 * s] = y4r * wr[4] - y4i * wi[4];
      oi_[q + 4 * s] = y4r * wi[4] + y4i * wr[4];
    }
  }
}

/* Transform the s0 interleaved sequences x[q + s0*p], p in [0,n).  Result lands in *res_re /
 * *res_im (either the a- or the b-buffer).  Unscaled in both directions. */
static void fft_interleaved(const fft_plan* pl, int s0, int inv, float* ar, float* ai, float* br,
                            float* bi, float** res_re, float** res_im) {
  int l = pl->n, s = s0;
  float *xr = ar, *xi = ai, *yr = br, *yi = bi;
  for (int k = 0; k < pl->nr; k++) {
    int r = pl->radix[k];
    switch (r) {
      case 2: pass_r2(l, s, pl->n, pl, inv, xr, xi, yr, yi); break;
      case 3: pass_r3(l, s, pl->n, pl, inv, xr, xi, yr, yi); break;
      case 4: pass_r4(l, s, pl->n, pl, inv, xr, xi, yr, yi); break;
      default: pass_r5(l, s, pl->n, pl, inv, xr, xi, yr, yi); break;
    }
    float* t;
    t = xr; xr = yr; yr = t;
    t = xi; xi = yi; yi = t;
    l /= r;
    s *= r;
  }
  *res_re = xr;
  *res_im = xi;
}

static void transpose_sq(int n, const float* src, float* dst) {
  const int T = 16;
  for (int i0 = 0; i0 < n; i0 += T)
    for (int j0 = 0; j0 < n; j0 += T) {
      int i1 = i0 + T < n ? i0 + T : n, j1 = j0 + T < n ? j0 + T : n;
      for (int i = i0; i < i1; i++)
        for (int j = j0; j < j1; j++) dst[(size_t)j * n + i] = src[(size_t)i * n + j];
    }
}

/* ------------------------------------------------------------------------------------------ */
/* Part 1b: cv::phaseCorrelate restated.                                                        */
/* ------------------------------------------------------------------------------------------ */

typedef struct {
  int n;
  fft_plan plan;
  float *ar, *ai, *br, *bi; /* n*n planes, ping-pong */
  float* surf;              /* fft-shifted correlation surface */
  float *f1r, *f1i;         /* spectrum of the first image */
} pc_ws;

static int ws_init(pc_ws* w, int n) {
  memset(w, 0, sizeof(*w));
  w->n = n;
  int rc = plan_init(&w->plan, n);
  if (rc) return rc;
  size_t nn = (size_t)n * n;
  w->ar = (float*)malloc(sizeof(float) * nn);
  w->ai = (float*)malloc(sizeof(float) * nn);
  w->br = (float*)malloc(sizeof(float) * nn);
  w->bi = (float*)malloc(sizeof(float) * nn);
  w->surf = (float*)malloc(sizeof(float) * nn);
  w->f1r = (float*)malloc(sizeof(float) * nn);
  w->f1i = (float*)malloc(sizeof(float) * nn);
  if (!w->ar || !w->ai || !w->br || !w->bi || !w->surf || !w->f1r || !w->f1i) return -2;
  return 0;
}

static void ws_free(pc_ws* w) {
  plan_free(&w->plan);
  free(w->ar); free(w->ai); free(w->br); free(w->bi); free(w->surf); free(w->f1r); free(w->f1i);
  memset(w, 0, sizeof(*w));
}

/* Full 2-D transform of the complex n x n array held in (ar, ai): columns, transpose,
 * columns.  The result is indexed [u][v] (u = x-frequency, v = y-frequency) when going
 * forward from [y][x]; applying it again with inv=1 on [u][v] returns [y][x]. */
static void fft2d(pc_ws* w, int inv, float** pr, float** pi) {
  const int n = w->n;
  float *rr, *ri;
  fft_interleaved(&w->plan, n, inv, w->ar, w->ai, w->br, w->bi, &rr, &ri);
  float* orr = (rr == w->ar) ? w->br : w->ar;
  float* ori = (ri == w->ai) ? w->bi : w->ai;
  transpose_sq(n, rr, orr);
  transpose_sq(n, ri, ori);
  /* second pass starts from (orr, ori); make them the "a" side */
  float *xr = orr, *xi = ori, *yr = rr, *yi = ri;
  float *r2, *i2;
  fft_interleaved(&w->plan, n, inv, xr, xi, yr, yi, &r2, &i2);
  *pr = r2;
  *pi = i2;
}

/* flags: bit 0 = old in-tree magSpectrums on the four real bins; bit 1 = OCL-compatible numerics
 * ("parity unpinned": a restatement of cl/FftMethod.cl that no OpenCL device here can run), with
 * (ocl_x0, ocl_y0) = frame coordinates of the section origin (the OCL centroid accumulates frame
 * coordinates in FP32, cl/FftMethod.cl:1333-1352). */
static int pc_core_ex(pc_ws* w, const float* src1, int stride1, const float* src2, int stride2,
                      const float* window, int flags, int ocl_x0, int ocl_y0, double* out_xy, double* response) {
  const int n = w->n;
  const int h = n / 2;
  const int ocl = (flags & 2) != 0;
  /* cv::phaseCorrelate transforms the two images SEPARATELY (dft(src1), dft(src2)); so does this
   * restatement.  (Packing them as src1 + i*src2 into one complex transform would be cheaper, but
   * the split leaks rounding noise of the stronger image into the spectrum of the weaker one -- a
   * black second frame must give an exactly zero spectrum, as it does in cv2.)  The window, when
   * given, multiplies both first. */
  float *Zr = NULL, *Zi = NULL;
  double ocl_rb[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};   /* exact real bins (0,0), (h,0), (0,h), (h,h) of src1 / src2 */
  if (ocl)
    for (int im = 0; im < 2; im++) {
      const float* src = im ? src2 : src1;
      const int stride = im ? stride2 : stride1;
      for (int y = 0; y < n; y++)
        for (int x = 0; x < n; x++) {
          const double pv = (double)src[(size_t)y * stride + x];
          const double sx = (x & 1) ? -pv : pv, sy = (y & 1) ? -pv : pv;
          ocl_rb[im][0] += pv;
          ocl_rb[im][1] += sx;
          ocl_rb[im][2] += sy;
          ocl_rb[im][3] += ((x ^ y) & 1) ? -pv : pv;
        }
    }
  for (int pass = 0; pass < 2; pass++) {
    const float* src = pass ? src2 : src1;
    const int stride = pass ? stride2 : stride1;
    for (int y = 0; y < n; y++) {
      const float* a = src + (size_t)y * stride;
      float* zr = w->ar + (size_t)y * n;
      float* zi = w->ai + (size_t)y * n;
      if (window) {
        const float* wv = window + (size_t)y * n;
        for (int x = 0; x < n; x++) { zr[x] = wv[x] * a[x]; zi[x] = 0.f; }
      } else {
        for (int x = 0; x < n; x++) { zr[x] = a[x]; zi[x] = 0.f; }
      }
    }
    fft2d(w, 0, &Zr, &Zi);
    if (pass == 0) {
      memcpy(w->f1r, Zr, sizeof(float) * (size_t)n * n);
      memcpy(w->f1i, Zi, sizeof(float) * (size_t)n * n);
    }
  }
  /* cross-power spectrum into the other buffer pair, indexed [u][v] */
  float* Rr = (Zr == w->ar) ? w->br : w->ar;
  float* Ri = (Zi == w->ai) ? w->bi : w->ai;
  const float eps = FLT_EPSILON;
  for (int u = 0; u < n; u++) {
    const int mu = (n - u) % n;
    for (int v = 0; v <= h; v++) {
      const int mv = (n - v) % n;
      const size_t i0 = (size_t)u * n + v, i1 = (size_t)mu * n + mv;
      const float f1r = w->f1r[i0], f1i = w->f1i[i0];
      const float f2r = Zr[i0], f2i = Zi[i0];
      /* mulSpectrums(FFT1, FFT2, P, 0, conjB=true): P = F1 * conj(F2) */
      float pr = f1r * f2r + f1i * f2i;
      float pim = f1i * f2r - f1r * f2i;
      int real_bin = (u == 0 || u == h) && (v == 0 || v == h);
      float cr, ci;
      if (ocl && real_bin) {
        /* mulAndNormalizeSpectrums, CCS corner bins: 1.0f / (temp1 * temp2)  (cl/FftMethod.cl:1031).
         * These four bins are plain sums of +-pixels: integers below 2^24 for N <= 240, i.e. EXACT in
         * float32 for any transform that adds them without twiddle products (the CUDA kernel's
         * prime-factor transform does).  This restatement's Stockham passes leave ~1e-5 of rounding
         * noise where the exact sum is 0 (about one section in ten at N = 60: blurred textures have
         * next to no energy at the Nyquist frequency), which would turn the reference's 1/0 = inf (no
         * maximum -> NaN, "LARGE SHIFT") into a huge finite spike; so the bins are taken exactly. */
        const int k = (u == h ? 1 : 0) + (v == h ? 2 : 0);
        cr = 1.0f / ((float)ocl_rb[0][k] * (float)ocl_rb[1][k]);
        ci = 0.f;
      } else if (ocl) {
        /* cmulnormf (cl/FftMethod.cl:976-982): mul * rsqrt(mad(mul.x, mul.x, mul.y*mul.y + FLT_EPSILON)) */
        float denom = 1.0f / sqrtf(fmaf(pr, pr, pim * pim + eps));
        cr = pr * denom;
        ci = pim * denom;
      } else if (real_bin) {
        /* CCS stores only the real part of these four bins (imag is exactly 0 in exact
         * arithmetic).  magSpectrums: cv2 4.13 -> |P|; the old in-tree copy
         * (src/FftMethod.cpp:107-109) -> P*P.  divSpectrums: A / (B + eps)
         * (src/FftMethod.cpp:1124-1126). */
        float mag = (flags & 1) ? pr * pr : fabsf(pr);
        cr = pr / (mag + eps);
        ci = 0.f;
      } else {
        /* magSpectrums (src/FftMethod.cpp:113-114, :130): sqrt in double, stored float;
         * divSpectrums conjB=false with B = (mag, 0) (src/FftMethod.cpp:1166-1172) */
        float mag = (float)sqrt((double)pr * pr + (double)pim * pim);
        double denom = (double)(mag * mag + eps);
        double re = (double)(pr * mag);
        double im = (double)(pim * mag);
        cr = (float)(re / denom);
        ci = (float)(im / denom);
      }
      Rr[i0] = cr;
      Ri[i0] = ci;
      if (v != 0 && v != h) { /* Hermitian mirror (what idft assumes for CCS input) */
        Rr[i1] = cr;
        Ri[i1] = -ci;
      }
    }
  }
  /* columns v=0 and v=h: the CCS layout stores them as packed Hermitian columns, i.e. only
   * u in [0,h] is independent; enforce the same symmetry */
  for (int u = h + 1; u < n; u++) {
    const int mu = n - u;
    Rr[(size_t)u * n] = Rr[(size_t)mu * n];
    Ri[(size_t)u * n] = -Ri[(size_t)mu * n];
    Rr[(size_t)u * n + h] = Rr[(size_t)mu * n + h];
    Ri[(size_t)u * n + h] = -Ri[(size_t)mu * n + h];
  }
  /* idft(C, C) without DFT_SCALE.  fft2d works on (ar,ai) as input: move R there if needed. */
  if (Rr != w->ar) {
    float* t;
    t = w->ar; w->ar = w->br; w->br = t;
    t = w->ai; w->ai = w->bi; w->bi = t;
  }
  float *cr_, *ci_;
  fft2d(w, 1, &cr_, &ci_);
  /* fftShift (src/FftMethod.cpp:1257-1323, even sizes: quadrant swap) */
  for (int y = 0; y < n; y++) {
    const float* srow = cr_ + (size_t)((y + h) % n) * n;
    float* drow = w->surf + (size_t)y * n;
    memcpy(drow, srow + h, sizeof(float) * (size_t)h);
    memcpy(drow + h, srow, sizeof(float) * (size_t)h);
  }
  if (ocl) {
    /* ifft_multi_radix_rows (cl/FftMethod.cl:724-846): output scaled by 1/N^2 (:733), written
     * fft-shifted, and every sample whose UNSHIFTED row or column index lies strictly between
     * SEARCH_RADIUS = 55 (src/FftMethod.cpp:819-820) and N - 55 is written as 0 (:737-747, :826-829). */
    const int R = 55;
    const float scale = 1.0f / ((float)n * (float)n);
    for (int ys = 0; ys < n; ys++) {
      const int y = (ys + h) % n;   /* unshifted index of shifted row ys */
      for (int xs = 0; xs < n; xs++) {
        const int x = (xs + h) % n;
        float* v = w->surf + (size_t)ys * n + xs;
        if ((y > R && y < n - R) || (x > R && x < n - R)) *v = 0.0f;
        else *v = *v * scale;
      }
    }
    /* minmaxloc (cl/FftMethod.cl:1164-1313): maximum with the lowest index among equals inside a row
     * (:1262-1266) and the first row among equal rows (strict >, :1302); an all-NaN surface leaves
     * INDEX_MAX, which the host turns into "LARGE SHIFT" -> NaN (src/FftMethod.cpp:461-464). */
    int py = -1, px = -1;
    float best = -FLT_MAX;
    for (int y = 0; y < n; y++)
      for (int x = 0; x < n; x++) {
        float vv = w->surf[(size_t)y * n + x];
        if (py < 0 ? (vv >= best) : (vv > best)) { best = vv; py = y; px = x; }
      }
    if (py < 0) { out_xy[0] = NAN; out_xy[1] = NAN; if (response) *response = 0.0; return 0; }
    /* refine (cl/FftMethod.cl:1315-1379), radius 3: clamped 7x7 window, positive samples only, FP32
     * accumulation of FRAME coordinates in row-major order, sum seeded with FLT_EPSILON */
    int xmin = ocl_x0 + (px - 3 >= 0 ? px - 3 : 0), xmax = ocl_x0 + (px + 3 < n ? px + 3 : n - 1);
    int ymin = ocl_y0 + (py - 3 >= 0 ? py - 3 : 0), ymax = ocl_y0 + (py + 3 < n ? py + 3 : n - 1);
    float cxf = 0.0f, cyf = 0.0f, sumf = FLT_EPSILON;
    for (int y = ymin; y <= ymax; y++)
      for (int x = xmin; x <= xmax; x++) {
        const float vv = w->surf[(size_t)(y - ocl_y0) * n + (x - ocl_x0)];
        if (vv > 0.0f) {
          cxf += (float)x * vv;
          cyf += (float)y * vv;
          sumf += vv;
        }
      }
    cxf /= sumf;
    cyf /= sumf;
    /* the kernel writes centroid - (origin + N/2) (:1368-1369); the caller negates like it does for
     * cv::phaseCorrelate, so hand back the negative */
    out_xy[0] = -(double)(cxf - (float)(ocl_x0 + (n >> 1)));
    out_xy[1] = -(double)(cyf - (float)(ocl_y0 + (n >> 1)));
    if (response) *response = (double)best;
    return 0;
  }
  /* minMaxLoc: first maximum in row-major order */
  int py = 0, px = 0;
  float best = w->surf[0];
  for (int y = 0; y < n; y++)
    for (int x = 0; x < n; x++) {
      float vv = w->surf[(size_t)y * n + x];
      if (vv > best) { best = vv; py = y; px = x; }
    }
  /* weightedCentroid 5x5, clamped to the image, all values, FP64
   * (src/FftMethod.cpp:1330-1384) */
  int minr = py - 2, maxr = py + 2, minc = px - 2, maxc = px + 2;
  if (minr < 0) minr = 0;
  if (minc < 0) minc = 0;
  if (maxr > n - 1) maxr = n - 1;
  if (maxc > n - 1) maxc = n - 1;
  double cx = 0.0, cy = 0.0, sum = 0.0;
  for (int y = minr; y <= maxr; y++)
    for (int x = minc; x <= maxc; x++) {
      double vv = (double)w->surf[(size_t)y * n + x];
      cx += (double)x * vv;
      cy += (double)y * vv;
      sum += vv;
    }
  if (response) *response = sum / ((double)n * (double)n);
  sum += DBL_EPSILON;
  cx /= sum;
  cy /= sum;
  out_xy[0] = (double)n / 2.0 - cx;
  out_xy[1] = (double)n / 2.0 - cy;
  return 0;
}

static int pc_core(pc_ws* w, const float* src1, int stride1, const float* src2, int stride2,
                   const float* window, int flags, double* out_xy, double* response) {
  return pc_core_ex(w, src1, stride1, src2, stride2, window, flags, 0, 0, out_xy, response);
}

int mof_oracle_phase_correlate_surface(const float* src1, int stride1, const float* src2,
                                       int stride2, int n, const float* window, int flags,
                                       double* out_xy, double* response, float* surf) {
  if (n < 2 || (n & 1)) return -1;
  pc_ws w;
  int rc = ws_init(&w, n);
  if (rc) { ws_free(&w); return rc; }
  rc = pc_core(&w, src1, stride1, src2, stride2, window, flags, out_xy, response);
  if (surf) memcpy(surf, w.surf, sizeof(float) * (size_t)n * n);
  ws_free(&w);
  return rc;
}

int mof_oracle_phase_correlate(const float* src1, int stride1, const float* src2, int stride2,
                               int n, const float* window, int flags, double* out_xy,
                               double* response) {
  return mof_oracle_phase_correlate_surface(src1, stride1, src2, stride2, n, window, flags,
                                            out_xy, response, NULL);
}

/* cv::createHanningWindow: w(y,x) = sqrt(wr[y]*wc[x])?  No -- OpenCV computes
 * wc[j] = 0.5*(1-cos(2 pi j/(cols-1))), wr[i] likewise, dst(i,j) = (float)(wr[i]*wc[j]) and
 * then takes cv::sqrt of the whole matrix (imgproc/src/phasecorr.cpp createHanningWindow). */
void mof_oracle_hann(int n, float* win) {
  double* wc = (double*)malloc(sizeof(double) * (size_t)n);
  double coeff = 2.0 * M_PI / (double)(n - 1);
  for (int j = 0; j < n; j++) wc[j] = 0.5 * (1.0 - cos(coeff * j));
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) win[(size_t)i * n + j] = sqrtf((float)(wc[i] * wc[j]));
  free(wc);
}

typedef struct {
  const uint8_t *curr, *prev;
  size_t pair_stride;
  int row_stride, F, N, sq;
  double max_sq;
  const float* win;
  double* shifts;
  int ocl;
} fftp_ctx;
typedef struct {
  pc_ws w;
  float *cf, *pf;
} fftp_state;

static void fftp_state_free(void* p) {
  fftp_state* st = (fftp_state*)p;
  if (!st) return;
  ws_free(&st->w);
  free(st->cf);
  free(st->pf);
  free(st);
}
static void* fftp_state_new(void* c) {
  fftp_ctx* ctx = (fftp_ctx*)c;
  fftp_state* st = (fftp_state*)calloc(1, sizeof(fftp_state));
  if (!st) return NULL;
  int rc = ws_init(&st->w, ctx->N);
  st->cf = (float*)malloc(sizeof(float) * (size_t)ctx->F * ctx->F);
  st->pf = (float*)malloc(sizeof(float) * (size_t)ctx->F * ctx->F);
  if (rc || !st->cf || !st->pf) { fftp_state_free(st); return NULL; }
  return st;
}
static void fftp_body(int p, void* s_, void* c_) {
  fftp_state* st = (fftp_state*)s_;
  fftp_ctx* ctx = (fftp_ctx*)c_;
  const int F = ctx->F, N = ctx->N, sq = ctx->sq, S = sq * sq;
  const uint8_t* c = ctx->curr + (size_t)p * ctx->pair_stride;
  const uint8_t* q = ctx->prev + (size_t)p * ctx->pair_stride;
  /* imCurr.convertTo(imCurrF, CV_32FC1) on the whole frame (src/FftMethod.cpp:1805-1806) */
  for (int y = 0; y < F; y++)
    for (int x = 0; x < F; x++) {
      st->cf[(size_t)y * F + x] = (float)c[(size_t)y * ctx->row_stride + x];
      st->pf[(size_t)y * F + x] = (float)q[(size_t)y * ctx->row_stride + x];
    }
  for (int j = 0; j < sq; j++)
    for (int i = 0; i < sq; i++) {
      double t[2];
      pc_core_ex(&st->w, st->cf + (size_t)(j * N) * F + i * N, F,
                 st->pf + (size_t)(j * N) * F + i * N, F, ctx->win, ctx->ocl ? 2 : 0, i * N, j * N, t, NULL);
      /* shift = -cv::phaseCorrelate(curr, prev) (src/FftMethod.cpp:1836) */
      double sx = -t[0], sy = -t[1];
      int valid = 1;
      if (sx * sx + sy * sy > ctx->max_sq || fabs(sx) > (double)N / 2 || fabs(sy) > (double)N / 2)
        valid = 0; /* src/FftMethod.cpp:1841-1845 */
      if (isnan(sx) || isnan(sy)) valid = 0; /* :1846-1849 */
      double* o = ctx->shifts + ((size_t)p * S + (size_t)(i + j * sq)) * 2;
      o[0] = valid ? sx : NAN; /* :1853 */
      o[1] = valid ? sy : NAN;
    }
}

int mof_oracle_fft_pairs(const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                         int row_stride, int F, int N, double max_px_speed, int hann,
                         int n_pairs, double* shifts_xy, int nthreads) {
  return mof_oracle_fft_pairs_ex(curr, prev, pair_stride, row_stride, F, N, max_px_speed, hann ? 1 : 0, n_pairs,
                                 shifts_xy, nthreads);
}

/* flags: bit 0 = Hann window, bit 1 = OCL-compatible numerics (useOCL = true branch,
 * src/FftMethod.cpp:1824-1825, :1833-1834: the kernel's output is used as the shift, un-negated). */
int mof_oracle_fft_pairs_ex(const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                            int row_stride, int F, int N, double max_px_speed, int flags,
                            int n_pairs, double* shifts_xy, int nthreads) {
  const int hann = flags & 1;
  /* geometry fix-ups of FftMethod::FftMethod (src/FftMethod.cpp:1706-1719) */
  if (F % 2 == 1) F--;
  if (F % N != 0) N = F;
  fftp_ctx ctx;
  ctx.curr = curr; ctx.prev = prev; ctx.pair_stride = pair_stride; ctx.row_stride = row_stride;
  ctx.F = F; ctx.N = N; ctx.sq = F / N; ctx.max_sq = max_px_speed * max_px_speed;
  ctx.shifts = shifts_xy;
  ctx.ocl = (flags & 2) != 0;
  float* win = NULL;
  if (hann) {
    win = (float*)malloc(sizeof(float) * (size_t)N * N);
    if (!win) return -2;
    mof_oracle_hann(N, win);
  }
  ctx.win = win;
  int rc = parallel_for(n_pairs, nthreads, fftp_body, fftp_state_new, fftp_state_free, &ctx);
  free(win);
  return rc;
}

/* ------------------------------------------------------------------------------------------ */
/* Part 2: FastSpacedBM (src/FastSpacedBMMethod.cl, src/FastSpacedBMMethod_OCL.cpp).            */
/* ------------------------------------------------------------------------------------------ */

static int bm_core(const uint8_t* curr, const uint8_t* prev, int W, int H, int row_stride, int B,
                   int step, int R, int no_gate, int8_t* flow_x, int8_t* flow_y, float* voted_xy,
                   int32_t* sad_out) {
  const int D = 2 * R + 1;
  if (D > 50 || B <= 0 || step < 0 || R < 0) return -1; /* arraySize 50, FastSpacedBMMethod.cl:1 */
  /* grid = (cols - 2R)/(B+step) x (rows - 2R)/(B+step)  (FastSpacedBMMethod_OCL.cpp:81-90) */
  const int pitch = B + step;
  const int gx = (W - 2 * R) / pitch, gy = (H - 2 * R) / pitch;
  if (gx <= 0 || gy <= 0) return 0;
  int32_t* sad = (int32_t*)malloc(sizeof(int32_t) * (size_t)D * D);
  int histx[50], histy[50];
  memset(histx, 0, sizeof(histx));
  memset(histy, 0, sizeof(histy));
  for (int by = 0; by < gy; by++)
    for (int bx = 0; bx < gx; bx++) {
      const int x0 = bx * pitch, y0 = by * pitch;
      /* SAD(sx,sy) = sum_{i,j<B} |curr[y0+R+j][x0+R+i] - prev[y0+j+sy][x0+i+sx]|
       * (FastSpacedBMMethod.cl:28-35; input_1 = curr, input_2 = prev per the host's argument
       * order, FastSpacedBMMethod_OCL.cpp:108-109).  NB: rows/cols beyond the image are never
       * touched because x0+B+2R <= W by the grid formula only when step >= 0 ... it is the
       * caller's job to pass frames with W - 2R >= gx*pitch (true by construction). */
      for (int sy = 0; sy < D; sy++)
        for (int sx = 0; sx < D; sx++) {
          int32_t acc = 0;
          for (int j = 0; j < B; j++) {
            const uint8_t* c = curr + (size_t)(y0 + R + j) * row_stride + x0 + R;
            const uint8_t* p = prev + (size_t)(y0 + j + sy) * row_stride + x0 + sx;
            for (int i = 0; i < B; i++) acc += abs((int)c[i] - (int)p[i]);
          }
          sad[sy * D + sx] = acc;
        }
      if (sad_out) memcpy(sad_out + (size_t)(by * gx + bx) * D * D, sad, sizeof(int32_t) * D * D);
      /* first minimum scanning sy outer, sx inner (strict '>' at :53 and :69) */
      int32_t mn = sad[0];
      int mx_ = 0, my_ = 0;
      for (int sy = 0; sy < D; sy++)
        for (int sx = 0; sx < D; sx++)
          if (sad[sy * D + sx] < mn) { mn = sad[sy * D + sx]; mx_ = sx; my_ = sy; }
      int fx = mx_ - R, fy = my_ - R;
      /* noise gate: (SAD(R,R) - min) <= R*R*0.2 -> (0,0)   (:2, :77-82; the reference writes
       * the zero through imgSrcOffset instead of imgDstOffset -- both are 0 for the
       * continuous matrices it allocates, so the intended semantics is restated) */
      if (!no_gate && (double)(sad[R * D + R] - mn) <= (double)(R * R) * 0.2) { fx = 0; fy = 0; }
      flow_x[by * gx + bx] = (int8_t)fx;
      flow_y[by * gx + bx] = (int8_t)fy;
      histx[fx + R]++;
      histy[fy + R]++;
    }
  /* Histogram_C1_D0 (:86-169): per-axis histogram, stable descending sort -> index 0 is the
   * mode, ties resolved to the lowest shift.  (The 1x1-grid Y-sort omission of the reference,
   * :140-151, is a launch-geometry bug and is not reproduced.) */
  int bxm = 0, bym = 0;
  for (int k = 1; k < D; k++) {
    if (histx[k] > histx[bxm]) bxm = k;
    if (histy[k] > histy[bym]) bym = k;
  }
  if (voted_xy) {
    voted_xy[0] = (float)(bxm - R);
    voted_xy[1] = (float)(bym - R);
  }
  free(sad);
  return gx * gy;
}

int mof_oracle_bm(const uint8_t* curr, const uint8_t* prev, int W, int H, int row_stride, int B,
                  int step, int R, int8_t* flow_x, int8_t* flow_y, float* voted_xy,
                  int32_t* sad_out) {
  return bm_core(curr, prev, W, H, row_stride, B, step, R, 0, flow_x, flow_y, voted_xy, sad_out);
}

/* Integer stage of BlockMethod::processImage (src/BlockMethod.cpp:25-75): blocks of B x B at
 * (n*B + R, m*B + R), n, m < (F - 2R)/B (:11, :45); SAD against prev displaced by (j, i) in
 * [-R, R]^2 (:52-59); cv::minMaxLoc first minimum in row-major (i outer, j inner) order (:66);
 * per-axis histograms and std::max_element, i.e. ties -> lowest shift (:68-69, :79-80).  The same
 * search as FastSpacedBM with step = 0 and without its noise gate.  Refine (:96-147) is not
 * restated. */
int mof_oracle_blockmethod(const uint8_t* curr, const uint8_t* prev, int F, int row_stride, int B, int R,
                           int8_t* flow_x, int8_t* flow_y, float* voted_xy) {
  return bm_core(curr, prev, F, F, row_stride, B, 0, R, 1, flow_x, flow_y, voted_xy, NULL);
}

typedef struct {
  const uint8_t *curr, *prev;
  size_t pair_stride;
  int W, H, row_stride, B, step, R, nb;
  int8_t* flow;
  float* voted;
} bmp_ctx;
static void bmp_body(int p, void* s_, void* c_) {
  (void)s_;
  bmp_ctx* c = (bmp_ctx*)c_;
  /* flow_xy layout per pair: nb x-shifts followed by nb y-shifts */
  mof_oracle_bm(c->curr + (size_t)p * c->pair_stride, c->prev + (size_t)p * c->pair_stride, c->W,
                c->H, c->row_stride, c->B, c->step, c->R, c->flow + (size_t)p * 2 * c->nb,
                c->flow + (size_t)p * 2 * c->nb + c->nb, c->voted + (size_t)p * 2, NULL);
}

int mof_oracle_bm_pairs(const uint8_t* curr, const uint8_t* prev, size_t pair_stride, int W, int H,
                        int row_stride, int B, int step, int R, int n_pairs, int8_t* flow_xy,
                        float* voted_xy, int nthreads) {
  const int pitch = B + step;
  if (pitch <= 0) return -1;
  const int gx = (W - 2 * R) / pitch, gy = (H - 2 * R) / pitch;
  if (2 * R + 1 > 50 || gx <= 0 || gy <= 0) return -1;
  bmp_ctx c;
  c.curr = curr; c.prev = prev; c.pair_stride = pair_stride; c.W = W; c.H = H;
  c.row_stride = row_stride; c.B = B; c.step = step; c.R = R; c.nb = gx * gy;
  c.flow = flow_xy; c.voted = voted_xy;
  int rc = parallel_for(n_pairs, nthreads, bmp_body, NULL, NULL, &c);
  return rc ? rc : c.nb;
}

/* ------------------------------------------------------------------------------------------ */
/* Part 3: cv::logPolar -> cv::warpPolar -> cv::remap (CV_8UC1) restated, then the             */
/*         scaleRotationEstimator pair function (src/scaleRotationEstimator.cpp:34-148).        */
/* ------------------------------------------------------------------------------------------ */

#define TAB_BITS 5
#define TAB_SIZE 32
#define COEF_BITS 15
#define COEF_SCALE (1 << COEF_BITS)

static int16_t g_tab_cubic[TAB_SIZE * TAB_SIZE * 16];
static int16_t g_tab_lanczos[TAB_SIZE * TAB_SIZE * 64];
static int g_tabs_ready = 0;

static void coeffs_cubic(float x, float* c) {
  const float A = -0.75f;
  c[0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
  c[1] = ((A + 2) * x - (A + 3)) * x * x + 1;
  c[2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
  c[3] = 1.f - c[0] - c[1] - c[2];
}

static void coeffs_lanczos4(float x, float* c) {
  static const double s45 = 0.70710678118654752440084436210485;
  static const double cs[8][2] = {{1, 0},  {-s45, -s45}, {0, 1},  {s45, -s45},
                                  {-1, 0}, {s45, s45},   {0, -1}, {-s45, s45}};
  if (x < FLT_EPSILON) {
    for (int i = 0; i < 8; i++) c[i] = 0.f;
    c[3] = 1.f;
    return;
  }
  float sum = 0.f;
  double y0 = -(x + 3) * M_PI * 0.25, s0 = sin(y0), c0 = cos(y0);
  for (int i = 0; i < 8; i++) {
    double y = -(x + 3 - i) * M_PI * 0.25;
    c[i] = (float)((cs[i][0] * s0 + cs[i][1] * c0) / (y * y));
    sum += c[i];
  }
  sum = 1.f / sum;
  for (int i = 0; i < 8; i++) c[i] *= sum;
}

static int16_t sat_short(float v) {
  long r = lrintf(v); /* round-half-even like cvRound */
  if (r > 32767) r = 32767;
  if (r < -32768) r = -32768;
  return (int16_t)r;
}

static void build_tab2d(int ks, int16_t* itab) {
  float tab1[TAB_SIZE * 8];
  for (int i = 0; i < TAB_SIZE; i++) {
    float x = (float)i * (1.f / TAB_SIZE);
    if (ks == 4) coeffs_cubic(x, tab1 + i * ks);
    else coeffs_lanczos4(x, tab1 + i * ks);
  }
  for (int i = 0; i < TAB_SIZE; i++)       /* fractional y */
    for (int j = 0; j < TAB_SIZE; j++) {   /* fractional x */
      int16_t* it = itab + (size_t)(i * TAB_SIZE + j) * ks * ks;
      int isum = 0;
      for (int k1 = 0; k1 < ks; k1++) {
        float vy = tab1[i * ks + k1];
        for (int k2 = 0; k2 < ks; k2++) {
          float v = vy * tab1[j * ks + k2];
          it[k1 * ks + k2] = sat_short(v * COEF_SCALE);
          isum += it[k1 * ks + k2];
        }
      }
      if (isum != COEF_SCALE) {
        int diff = isum - COEF_SCALE;
        int ks2 = ks / 2, Mk1 = ks2, Mk2 = ks2, mk1 = ks2, mk2 = ks2;
        for (int k1 = ks2; k1 < ks2 + 2; k1++)
          for (int k2 = ks2; k2 < ks2 + 2; k2++) {
            if (it[k1 * ks + k2] < it[mk1 * ks + mk2]) { mk1 = k1; mk2 = k2; }
            else if (it[k1 * ks + k2] > it[Mk1 * ks + Mk2]) { Mk1 = k1; Mk2 = k2; }
          }
        if (diff < 0) it[Mk1 * ks + Mk2] = (int16_t)(it[Mk1 * ks + Mk2] - diff);
        else it[mk1 * ks + mk2] = (int16_t)(it[mk1 * ks + mk2] - diff);
      }
    }
}

static pthread_once_t g_tabs_once = PTHREAD_ONCE_INIT;
static void build_tabs(void) {
  build_tab2d(4, g_tab_cubic);
  build_tab2d(8, g_tab_lanczos);
  g_tabs_ready = 1;
}
static void ensure_tabs(void) { pthread_once(&g_tabs_once, build_tabs); }

const int16_t* mof_oracle_interp_table(int kind) {
  ensure_tabs();
  return kind == 2 ? g_tab_cubic : (kind == 4 ? g_tab_lanczos : NULL);
}

static int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) {
    if (p < 0) p = -p;
    else p = 2 * (len - 1) - p;
  }
  return p;
}

int mof_oracle_logpolar(const uint8_t* src, int src_stride, int res, double M, int interp,
                        uint8_t* dst, int dst_stride) {
  if (interp != 2 && interp != 4) return -1;
  ensure_tabs();
  const int ks = interp == 2 ? 4 : 8;
  const int off = ks / 2 - 1; /* taps start at floor(coord) - off */
  const int16_t* tab = interp == 2 ? g_tab_cubic : g_tab_lanczos;
  /* cv::logPolar: maxRadius = exp(width / M); warpPolar(..., WARP_POLAR_LOG):
   * Kangle = 2 pi / height, Kmag = log(maxRadius) / width, rho_k = (float)(exp(k*Kmag) - 1),
   * map = (float)(rho*cos + cx), (float)(rho*sin + cy), centre (res/2, res/2) as Point2f
   * (src/scaleRotationEstimator.cpp:25) */
  const double maxRadius = M > 0 ? exp((double)res / M) : 1.0;
  const double Kangle = 2.0 * M_PI / (double)res;
  const double Kmag = log(maxRadius) / (double)res;
  const float cxy = (float)(res / 2);
  float* rhos = (float*)malloc(sizeof(float) * (size_t)res);
  for (int r = 0; r < res; r++) rhos[r] = (float)(exp(r * Kmag) - 1.0);
  for (int phi = 0; phi < res; phi++) {
    double a = Kangle * phi, cp = cos(a), sp = sin(a);
    for (int r = 0; r < res; r++) {
      float mx = (float)((double)rhos[r] * cp + (double)cxy);
      float my = (float)((double)rhos[r] * sp + (double)cxy);
      /* remap: fixed-point coordinates, 1/32 px */
      int sx = (int)lrintf(mx * (float)TAB_SIZE);
      int sy = (int)lrintf(my * (float)TAB_SIZE);
      int fxy = (sy & (TAB_SIZE - 1)) * TAB_SIZE + (sx & (TAB_SIZE - 1));
      int ix = sx >> TAB_BITS, iy = sy >> TAB_BITS;
      if (ix > 32767) ix = 32767;
      if (ix < -32768) ix = -32768;
      if (iy > 32767) iy = 32767;
      if (iy < -32768) iy = -32768;
      /* BORDER_TRANSPARENT: skip when the centre tap is outside the source */
      if ((unsigned)ix >= (unsigned)res || (unsigned)iy >= (unsigned)res) continue;
      const int16_t* w = tab + (size_t)fxy * ks * ks;
      int x0 = ix - off, y0 = iy - off;
      int sum = 0;
      for (int k1 = 0; k1 < ks; k1++) {
        int yy = reflect101(y0 + k1, res);
        const uint8_t* row = src + (size_t)yy * src_stride;
        for (int k2 = 0; k2 < ks; k2++) {
          int xx = reflect101(x0 + k2, res);
          sum += (int)row[xx] * (int)w[k1 * ks + k2];
        }
      }
      int v = (sum + (1 << (COEF_BITS - 1))) >> COEF_BITS;
      if (v < 0) v = 0;
      if (v > 255) v = 255;
      dst[(size_t)phi * dst_stride + r] = (uint8_t)v;
    }
  }
  free(rhos);
  return 0;
}

typedef struct {
  const uint8_t *curr, *prev;
  size_t pair_stride;
  int row_stride, res;
  double M;
  double* out;
} srp_ctx;
typedef struct {
  pc_ws w;
  uint8_t* lp;
  float *pf, *cf;
} srp_state;
static void srp_state_free(void* p) {
  srp_state* st = (srp_state*)p;
  if (!st) return;
  ws_free(&st->w);
  free(st->lp); free(st->pf); free(st->cf);
  free(st);
}
static void* srp_state_new(void* c_) {
  srp_ctx* c = (srp_ctx*)c_;
  srp_state* st = (srp_state*)calloc(1, sizeof(srp_state));
  if (!st) return NULL;
  size_t nn = (size_t)c->res * c->res;
  int rc = ws_init(&st->w, c->res);
  st->lp = (uint8_t*)malloc(nn);
  st->pf = (float*)malloc(sizeof(float) * nn);
  st->cf = (float*)malloc(sizeof(float) * nn);
  if (rc || !st->lp || !st->pf || !st->cf) { srp_state_free(st); return NULL; }
  return st;
}
static void srp_body(int p, void* s_, void* c_) {
  srp_state* st = (srp_state*)s_;
  srp_ctx* c = (srp_ctx*)c_;
  const int res = c->res;
  const size_t nn = (size_t)res * res;
  /* first call: tempIm = zeros; logPolar(INTER_CUBIC) -> prevIm_F32  (:27, :45, :48) */
  memset(st->lp, 0, nn);
  mof_oracle_logpolar(c->prev + (size_t)p * c->pair_stride, c->row_stride, res, c->M, 2, st->lp,
                      res);
  for (size_t k = 0; k < nn; k++) st->pf[k] = (float)st->lp[k];
  /* later call: same persistent tempIm, logPolar(INTER_LANCZOS4)  (:112, :115) */
  mof_oracle_logpolar(c->curr + (size_t)p * c->pair_stride, c->row_stride, res, c->M, 4, st->lp,
                      res);
  for (size_t k = 0; k < nn; k++) st->cf[k] = (float)st->lp[k];
  double pt[2];
  pc_core(&st->w, st->cf, res, st->pf, res, NULL, 0, pt, NULL); /* phaseCorrelate(cur, prev) :117 */
  double* o = c->out + (size_t)p * 2;
  /* :119-121: both clauses test pt.x */
  if (fabs(pt[0]) > (double)(res / 2)) {
    o[0] = 1.0;
    o[1] = 0.0;
  } else {
    const double Ky = (double)res / 360.0; /* :26 */
    o[0] = exp(pt[0] / c->M);              /* :123 */
    o[1] = (pt[1] / Ky) * (M_PI / 180.0);  /* :124 */
  }
}

int mof_oracle_scalerot_pairs(const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                              int row_stride, int res, double M, int n_pairs, double* scale_rot,
                              int nthreads) {
  ensure_tabs();
  srp_ctx c;
  c.curr = curr; c.prev = prev; c.pair_stride = pair_stride; c.row_stride = row_stride;
  c.res = res; c.M = M; c.out = scale_rot;
  return parallel_for(n_pairs, nthreads, srp_body, srp_state_new, srp_state_free, &c);
}


/* ------------------------------------------------------------------------------------------------
 * Camera front end (src/optic_flow.cpp:1603-1622) and the long-range down-scale
 * (src/FftMethod.cpp:1931-1932), restated from what cv2 4.13 computes for them:
 *   cv::resize INTER_LINEAR with an exact even integer ratio k: the sample point k*dx + (k-1)/2
 *   lies half way between source pixels k*dx + k/2 - 1 and k*dx + k/2 on both axes, and the
 *   fixed-point path reduces to (a + b + c + d + 2) >> 2;  k = 1 is a copy;
 *   cv::cvtColor(CV_RGB2GRAY) on 8-bit: (9798*c0 + 19235*c1 + 3735*c2 + 16384) >> 15.
 * Pinned by tests/golden/frontend_cv2.npz.
 * ---------------------------------------------------------------------------------------------- */
static unsigned fe_sample(const uint8_t* img, int row_stride, int ch, int k, int xs, int ys, int c) {
  if (k == 1) return img[(size_t)ys * row_stride + (size_t)xs * ch + c];
  const uint8_t* p = img + (size_t)(ys * k + k / 2 - 1) * row_stride + (size_t)(xs * k + k / 2 - 1) * ch + c;
  return ((unsigned)p[0] + p[ch] + p[row_stride] + p[row_stride + ch] + 2u) >> 2;
}

int mof_oracle_frontend(const uint8_t* raw, int width, int height, int channels, int row_stride,
                        int scale_factor, int cx, int F, uint8_t* gray) {
  const int k = scale_factor <= 0 ? 1 : scale_factor;
  if ((channels != 1 && channels != 3) || (k != 1 && k != 2 && k != 4) || width % k || height % k) return -1;
  const int ws = width / k, hs = height / k;
  const int x0 = cx - F / 2, y0 = hs / 2 - F / 2;   /* :1610-1617 */
  if (x0 < 0 || y0 < 0 || x0 + F > ws || y0 + F > hs) return -2;
  for (int y = 0; y < F; y++)
    for (int x = 0; x < F; x++) {
      unsigned v;
      if (channels == 1) {
        v = fe_sample(raw, row_stride, 1, k, x0 + x, y0 + y, 0);
      } else {
        const unsigned c0 = fe_sample(raw, row_stride, 3, k, x0 + x, y0 + y, 0);
        const unsigned c1 = fe_sample(raw, row_stride, 3, k, x0 + x, y0 + y, 1);
        const unsigned c2 = fe_sample(raw, row_stride, 3, k, x0 + x, y0 + y, 2);
        v = (9798u * c0 + 19235u * c1 + 3735u * c2 + 16384u) >> 15;   /* :1622 */
      }
      gray[(size_t)y * F + x] = (uint8_t)v;
    }
  return 0;
}

/* cv::resize(im, Size(), 1/4, 1/4) of a gray F x F frame (F % 4 == 0) -> F/4 x F/4, contiguous */
int mof_oracle_quarter(const uint8_t* src, int row_stride, int F, uint8_t* dst) {
  if (F % 4) return -1;
  const int Fl = F / 4;
  for (int y = 0; y < Fl; y++)
    for (int x = 0; x < Fl; x++) dst[(size_t)y * Fl + x] = (uint8_t)fe_sample(src, row_stride, 1, 4, x, y, 0);
  return 0;
}


/* ---- BlockMethod::Refine ---------------------------------------------------------------------- */
/* index pair and 11-bit weights of output sample d of a 2x up-scale of n samples (cv::resize
 * INTER_LINEAR: src = (d + 0.5) / 2 - 0.5).  Columns zero the fraction at the borders, rows clamp
 * the indices and keep it (resizeGeneric_ / VResizeLinear as cv2 4.13 behaves; pinned by
 * tests/golden/refine_cv2.npz). */
static void up2_coef(int d, int n, int cols, int* s0, int* s1, int* c0, int* c1) {
  int s = (d & 1) ? (d - 1) / 2 : d / 2 - 1;
  int f = (d & 1) ? 512 : 1536;
  if (cols) {
    if (s < 0) { s = 0; f = 0; }
    if (s >= n - 1) { s = n - 1; f = 0; }
    *s0 = s;
    *s1 = s + 1 < n ? s + 1 : n - 1;
  } else {
    *s0 = s < 0 ? 0 : (s > n - 1 ? n - 1 : s);
    *s1 = s + 1 < 0 ? 0 : (s + 1 > n - 1 ? n - 1 : s + 1);
  }
  *c1 = f;
  *c0 = 2048 - f;
}

void mof_oracle_up2(const uint8_t* src, int w, int h, int row_stride, uint8_t* dst) {
  const int W2 = 2 * w, H2 = 2 * h;
  for (int dy = 0; dy < H2; dy++) {
    int y0, y1, b0, b1;
    up2_coef(dy, h, 0, &y0, &y1, &b0, &b1);
    const uint8_t* r0 = src + (size_t)y0 * row_stride;
    const uint8_t* r1 = src + (size_t)y1 * row_stride;
    for (int dx = 0; dx < W2; dx++) {
      int x0, x1, a0, a1;
      up2_coef(dx, w, 1, &x0, &x1, &a0, &a1);
      const int h0 = r0[x0] * a0 + r0[x1] * a1;
      const int h1 = r1[x0] * a0 + r1[x1] * a1;
      dst[(size_t)dy * W2 + dx] = (uint8_t)((((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2);
    }
  }
}

int mof_oracle_refine(const uint8_t* curr, int w, int h, int row_stride, int full_x, int full_y, int passes,
                      double* refined_xy) {
  const int W2 = 2 * w, H2 = 2 * h;
  uint8_t* up = (uint8_t*)malloc((size_t)W2 * H2);
  if (!up) return -2;
  /* pass 1 up-scales imCurr into imPrev2x and imCurr2x; later passes "resize" 2F -> 2F (a copy) */
  mof_oracle_up2(curr, w, h, row_stride, up);
  int ox = full_x, oy = full_y, scale = 1, rc = 0;
  for (int p = 1; p <= passes; p++) {
    scale *= 2;
    ox *= 2;
    oy *= 2;
    const int sx = ox < 0 ? -ox + 1 : 1, sy = oy < 0 ? -oy + 1 : 1;      /* :113-121 */
    const int cw = W2 - (abs(ox) + 2), ch = H2 - (abs(oy) + 2);          /* :123 */
    if (cw <= 0 || ch <= 0) { rc = -1; break; }
    long long best = -1;
    int bn = 0, bm = 0;
    for (int m = -1; m <= 1; m++)
      for (int n = -1; n <= 1; n++) {                                     /* :127-137 */
        long long sad = 0;
        for (int y = 0; y < ch; y++) {
          const uint8_t* pa = up + (size_t)(1 + y) * W2 + 1;
          const uint8_t* pb = up + (size_t)(sy + m + y) * W2 + sx + n;
          for (int x = 0; x < cw; x++) sad += abs((int)pa[x] - (int)pb[x]);
        }
        if (best < 0 || sad < best) { best = sad; bn = n; bm = m; }       /* first minimum, :142 */
      }
    ox += bn;
    oy += bm;                                                             /* :144 */
  }
  free(up);
  if (rc) {
    refined_xy[0] = refined_xy[1] = NAN;
    return rc;
  }
  refined_xy[0] = (double)((float)ox / (float)scale);                     /* :146 */
  refined_xy[1] = (double)((float)oy / (float)scale);
  return 0;
}
