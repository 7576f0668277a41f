// mof_capi.cu -- the C-ABI boundary (include/mof.h) and the C++ host backend behind it: context,
// device buffers, staging/overlap of host<->device copies, stream state.  This replaces the
// OpenCL context/queue/buffer code of the reference (FftMethod::FftMethod src/FftMethod.cpp:1680-1766,
// phaseCorrelateField :1391-1575, OCL_FftPlan::enqueueTransform :276-475) with CUDA streams.
// No CPU fallback: every entry point fails with MOF_ERR_CUDA when the device is not usable.
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#if !defined(MOF_PIPE_CHUNKS_DEFAULT)
#define MOF_PIPE_CHUNKS_DEFAULT 8
#endif
#include <string>
#include <vector>

#include "../../include/mof.h"
#include "mof_internal.h"

namespace {

thread_local std::string g_create_error;

struct TimedLaunch {
  cudaEvent_t a, b;
};

}  // namespace

struct mof_ctx {
  mof_config cfg{};
  int F = 0, N = 0, sq = 0, S = 0;
  int device = 0;
  int sms = 0;
  int pc_grid = 0;      // resident CTAs for section size N
  int sr_grid = 0;      // resident CTAs for section size sr_resolution
  cudaStream_t stream[2] = {nullptr, nullptr};
  cudaEvent_t last_done = nullptr;
  // staging for host-pointer calls (two slots, double buffered)
  uint8_t* d_curr[2] = {nullptr, nullptr};
  uint8_t* d_prev[2] = {nullptr, nullptr};
  double* d_out[2] = {nullptr, nullptr};
  void* d_scratch[2] = {nullptr, nullptr};  // per-CTA spectrum scratch (pairs mode)
  size_t scratch_bytes = 0;
  float* d_win = nullptr;
  // streams
  void* d_spec_cache = nullptr;
  std::vector<uint8_t> stream_first;  // host: 1 = next push is the first call of that stream
  std::vector<uint8_t> stream_parity; // host: which half of the stream's cache holds the previous spectrum
  // per-push metadata (stream ids + first/parity flags): a ring of META_SLOTS pinned/device blocks so
  // that a push does not have to wait for the previous push's kernel before it can be queued
  static constexpr int META_SLOTS = 4;
  uint8_t* d_meta = nullptr;    // META_SLOTS x [n_streams ints | n_streams flag bytes, padded to 8*n_streams]
  uint8_t* h_meta = nullptr;    // pinned, same layout
  cudaEvent_t meta_done[META_SLOTS] = {nullptr, nullptr, nullptr, nullptr};
  int meta_slot = 0;
  void* d_seq_scratch = nullptr;   // mof_fft_stream_push_many: resident slots x 2 spectra (lazily allocated)
  // block matching staging (lazily sized)
  uint8_t* d_bm_curr = nullptr;
  uint8_t* d_bm_prev = nullptr;
  int8_t* d_bm_flow = nullptr;
  float* d_bm_voted = nullptr;
  size_t bm_cap_bytes = 0, bm_cap_flow = 0;
  int bm_cap_pairs = 0;
  // scale / rotation
  uint32_t* d_lp_pk = nullptr;   // packed log-polar map (LpMap::pk)
  int16_t *d_tab_cubic = nullptr, *d_tab_lanczos = nullptr;
  uint8_t* d_lp_img = nullptr;   // 2 * max_batch log-polar images (prev block then curr block)
  void* d_sr_scratch = nullptr;
  double* d_sr_out = nullptr;
  bool sr_ready = false;
  bool fft_ready = false;
  bool fft_generic = false;     // N outside {60, 120, 240}: pg_kernels.cu (streams cache the previous FRAME)
  float* d_pg_tw = nullptr;     // its twiddle table
  float* d_sr_tw = nullptr;     // scale / rotation at a resolution outside the plan: twiddles of the generic kernel
  bool sr_generic = false;
  uint8_t* d_frame_cache = nullptr;   // generic path: n_streams previous frames, F x F bytes each
  uint8_t* d_sr_temp = nullptr;  // persistent tempIm of the stateful estimator
  uint8_t* d_sr_prev = nullptr;  // previous accepted log-polar image
  bool sr_first = true;
  // camera front end / long-range staging (lazily sized, grow-only)
  uint8_t* d_fe_raw = nullptr;
  size_t fe_raw_cap = 0;
  cudaEvent_t fe_copied[2] = {nullptr, nullptr};     // host push_many_raw pipeline: H2D of a staging half done
  cudaEvent_t fe_consumed[2] = {nullptr, nullptr};   //   front end has read that half
  uint8_t* d_fe_gray = nullptr;
  size_t fe_gray_cap = 0;
  uint8_t* d_lr = nullptr;       // n down-scaled curr frames then n down-scaled prev frames
  size_t lr_cap = 0;
  double* d_lr_out = nullptr;
  size_t lr_out_cap = 0;
  // BlockMethod::Refine staging (grow-only): 2x images, host-API frames / offsets / results
  uint8_t* d_rf_up = nullptr;
  size_t rf_up_cap = 0;
  uint8_t* d_rf_io = nullptr;
  size_t rf_io_cap = 0;
  // bookkeeping
  std::string err;
  std::atomic_flag busy = ATOMIC_FLAG_INIT;
  long long launches = 0;
  bool timing = false;
  std::vector<TimedLaunch> timed;
};

namespace {

int fail(mof_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (c) c->err = buf;
  else g_create_error = buf;
  return code;
}

#define CU(c, call)                                                                       \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return fail((c), MOF_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                  __FILE__, __LINE__);                                                    \
  } while (0)

struct BusyGuard {
  mof_ctx* c;
  bool ok;
  explicit BusyGuard(mof_ctx* ctx) : c(ctx), ok(!ctx->busy.test_and_set(std::memory_order_acquire)) {}
  ~BusyGuard() {
    if (ok) c->busy.clear(std::memory_order_release);
  }
};

// ---- host tables for cv::remap / cv::warpPolar (computed like OpenCV does, in double/float) ------
void coeffs_cubic(float x, float* c) {
  const float A = -0.75f;
  c[0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
  c[1] = ((A + 2) * x - (A + 3)) * x * x + 1;
  c[2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
  c[3] = 1.f - c[0] - c[1] - c[2];
}

void coeffs_lanczos4(float x, float* c) {
  const double s45 = 0.70710678118654752440084436210485;
  const double cs[8][2] = {{1, 0},  {-s45, -s45}, {0, 1},  {s45, -s45},
                           {-1, 0}, {s45, s45},   {0, -1}, {-s45, s45}};
  if (x < 1.1920929e-07f) {
    for (int i = 0; i < 8; i++) c[i] = 0.f;
    c[3] = 1.f;
    return;
  }
  float sum = 0.f;
  const double pi = 3.14159265358979323846;
  double y0 = -(x + 3) * pi * 0.25, s0 = std::sin(y0), c0 = std::cos(y0);
  for (int i = 0; i < 8; i++) {
    double y = -(x + 3 - i) * pi * 0.25;
    c[i] = (float)((cs[i][0] * s0 + cs[i][1] * c0) / (y * y));
    sum += c[i];
  }
  sum = 1.f / sum;
  for (int i = 0; i < 8; i++) c[i] *= sum;
}

std::vector<int16_t> build_weight_table(int ks) {
  std::vector<float> t1(32 * ks);
  for (int i = 0; i < 32; i++) {
    float x = (float)i * (1.f / 32.f);
    if (ks == 4) coeffs_cubic(x, &t1[i * ks]);
    else coeffs_lanczos4(x, &t1[i * ks]);
  }
  std::vector<int16_t> tab((size_t)32 * 32 * ks * ks);
  for (int i = 0; i < 32; i++)
    for (int j = 0; j < 32; j++) {
      int16_t* it = &tab[(size_t)(i * 32 + j) * ks * ks];
      int isum = 0;
      for (int k1 = 0; k1 < ks; k1++)
        for (int k2 = 0; k2 < ks; k2++) {
          float v = t1[i * ks + k1] * t1[j * ks + k2];
          long r = std::lrintf(v * 32768.f);
          r = r > 32767 ? 32767 : (r < -32768 ? -32768 : r);
          it[k1 * ks + k2] = (int16_t)r;
          isum += (int)r;
        }
      if (isum != 32768) {  // push the rounding residue into the largest / smallest central tap
        int diff = isum - 32768, h = ks / 2, Mk1 = h, Mk2 = h, mk1 = h, mk2 = h;
        for (int k1 = h; k1 < h + 2; k1++)
          for (int k2 = h; k2 < h + 2; k2++) {
            if (it[k1 * ks + k2] < it[mk1 * ks + mk2]) { mk1 = k1; mk2 = k2; }
            else if (it[k1 * ks + k2] > it[Mk1 * ks + Mk2]) { Mk1 = k1; Mk2 = k2; }
          }
        if (diff < 0) it[Mk1 * ks + Mk2] = (int16_t)(it[Mk1 * ks + Mk2] - diff);
        else it[mk1 * ks + mk2] = (int16_t)(it[mk1 * ks + mk2] - diff);
      }
    }
  return tab;
}

// cv::logPolar(center = (res/2, res/2), M) -> warpPolar map -> remap's fixed-point coordinates
void build_logpolar_map(int res, double M, std::vector<int16_t>& ix, std::vector<int16_t>& iy,
                        std::vector<uint16_t>& fxy) {
  const double pi = 3.14159265358979323846;
  const double maxRadius = M > 0 ? std::exp((double)res / M) : 1.0;
  const double Kangle = 2.0 * pi / (double)res;
  const double Kmag = std::log(maxRadius) / (double)res;
  const float c = (float)(res / 2);
  std::vector<float> rho(res);
  for (int r = 0; r < res; r++) rho[r] = (float)(std::exp(r * Kmag) - 1.0);
  ix.resize((size_t)res * res);
  iy.resize((size_t)res * res);
  fxy.resize((size_t)res * res);
  for (int phi = 0; phi < res; phi++) {
    const double a = Kangle * phi, cp = std::cos(a), sp = std::sin(a);
    for (int r = 0; r < res; r++) {
      float mx = (float)((double)rho[r] * cp + (double)c);
      float my = (float)((double)rho[r] * sp + (double)c);
      int sx = (int)std::lrintf(mx * 32.f), sy = (int)std::lrintf(my * 32.f);
      int x = sx >> 5, y = sy >> 5;
      x = x > 32767 ? 32767 : (x < -32768 ? -32768 : x);
      y = y > 32767 ? 32767 : (y < -32768 ? -32768 : y);
      const size_t p = (size_t)phi * res + r;
      ix[p] = (int16_t)x;
      iy[p] = (int16_t)y;
      fxy[p] = (uint16_t)((sy & 31) * 32 + (sx & 31));
    }
  }
}

// cv::createHanningWindow(Size(n,n), CV_32F)
std::vector<float> build_hann(int n) {
  std::vector<double> wc(n);
  const double coeff = 2.0 * 3.14159265358979323846 / (double)(n - 1);
  for (int j = 0; j < n; j++) wc[j] = 0.5 * (1.0 - std::cos(coeff * j));
  std::vector<float> w((size_t)n * n);
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) w[(size_t)i * n + j] = std::sqrt((float)(wc[i] * wc[j]));
  return w;
}

bool supported_n(int n) { return n == 60 || n == 120 || n == 240; }

void free_ctx(mof_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  for (auto& t : c->timed) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
  for (int i = 0; i < 2; i++) {
    cudaFree(c->d_curr[i]); cudaFree(c->d_prev[i]); cudaFree(c->d_out[i]); cudaFree(c->d_scratch[i]);
    if (c->stream[i]) cudaStreamDestroy(c->stream[i]);
  }
  if (c->last_done) cudaEventDestroy(c->last_done);
  cudaFree(c->d_win); cudaFree(c->d_spec_cache); cudaFree(c->d_meta); cudaFree(c->d_seq_scratch);
  cudaFreeHost(c->h_meta);
  for (auto& e : c->meta_done) if (e) cudaEventDestroy(e);
  for (int i = 0; i < 2; i++) {
    if (c->fe_copied[i]) cudaEventDestroy(c->fe_copied[i]);
    if (c->fe_consumed[i]) cudaEventDestroy(c->fe_consumed[i]);
  }
  cudaFree(c->d_bm_curr); cudaFree(c->d_bm_prev); cudaFree(c->d_bm_flow); cudaFree(c->d_bm_voted);
  cudaFree(c->d_lp_pk);
  cudaFree(c->d_tab_cubic); cudaFree(c->d_tab_lanczos); cudaFree(c->d_lp_img);
  cudaFree(c->d_sr_scratch); cudaFree(c->d_sr_out); cudaFree(c->d_sr_temp); cudaFree(c->d_sr_prev);
  cudaFree(c->d_pg_tw); cudaFree(c->d_sr_tw); cudaFree(c->d_frame_cache);
  cudaFree(c->d_fe_raw); cudaFree(c->d_fe_gray); cudaFree(c->d_lr); cudaFree(c->d_lr_out);
  cudaFree(c->d_rf_up); cudaFree(c->d_rf_io);
  delete c;
}

int launch_pc(mof_ctx* c, int N, mof::PcArgs& a, int grid_cap, cudaStream_t st) {
  int grid = a.n_sections < grid_cap ? a.n_sections : grid_cap;
  if (grid < 1) return MOF_OK;
  TimedLaunch t{};
  const bool timed = c->timing && c->timed.size() < 8192;
  if (timed) {
    cudaEventCreate(&t.a);
    cudaEventCreate(&t.b);
    cudaEventRecord(t.a, st);
  }
  int rc;
  if (a.post == mof::PC_POST_SCALEROT && c->sr_generic)
    rc = mof::pg_launch(N, a, c->d_sr_tw, a.scratch, nullptr, a.S, grid, st);
  else if (a.post != mof::PC_POST_SCALEROT && c->fft_generic)
    rc = mof::pg_launch(N, a, c->d_pg_tw, a.scratch, c->d_frame_cache, a.seq_T > 0 ? a.seq_T * a.S : a.S, grid, st);
  else
    rc = mof::pc_launch(N, a, grid, st);
  if (timed) {
    cudaEventRecord(t.b, st);
    c->timed.push_back(t);
  }
  c->launches++;
  if (rc) return fail(c, MOF_ERR_CUDA, "phase-correlation kernel launch failed: %s",
                      cudaGetErrorString((cudaError_t)rc));
  return MOF_OK;
}

// copy n frames (F x F used) from a strided host layout into the compact device staging
int h2d_frames(mof_ctx* c, uint8_t* dst, const uint8_t* src, size_t frame_stride, int row_stride,
               int w, int h, int n, cudaStream_t st) {
  if (row_stride == w && ((size_t)w * h == frame_stride || n == 1)) {
    // fully contiguous: one linear DMA (2-D copies with 240-byte rows run an order of magnitude slower)
    CU(c, cudaMemcpyAsync(dst, src, (size_t)w * h * n, cudaMemcpyHostToDevice, st));
  } else if ((size_t)row_stride * h == frame_stride) {
    CU(c, cudaMemcpy2DAsync(dst, w, src, row_stride, w, (size_t)h * n, cudaMemcpyHostToDevice, st));
  } else {
    for (int k = 0; k < n; k++)
      CU(c, cudaMemcpy2DAsync(dst + (size_t)k * w * h, w, src + (size_t)k * frame_stride, row_stride,
                              w, h, cudaMemcpyHostToDevice, st));
  }
  return MOF_OK;
}

}  // namespace

extern "C" {

const char* mof_version(void) { return "mof-b200 0.1 (sm_100a)"; }

void mof_default_config(mof_config* cfg) {
  if (!cfg) return;
  std::memset(cfg, 0, sizeof(*cfg));
  cfg->frame_size = 480;
  cfg->sample_point_size = 120;
  cfg->max_px_speed = 80.0;
  cfg->max_batch = 64;
  cfg->bm_block = 120;
  cfg->bm_step = 24;
  cfg->bm_radius = 21;
  cfg->sr_resolution = 240;
  cfg->sr_m = 49.9;
}

const char* mof_last_error(const mof_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int mof_create(mof_ctx** out, const mof_config* cfg) {
  if (!out || !cfg) return fail(nullptr, MOF_ERR_INVALID, "mof_create: null argument");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(nullptr, MOF_ERR_CUDA, "no usable CUDA device (%s); this backend has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= ndev)
    return fail(nullptr, MOF_ERR_INVALID, "device %d out of range (have %d)", cfg->device, ndev);
  mof_ctx* c = new mof_ctx();
  c->cfg = *cfg;
  c->device = cfg->device;
  // geometry fix-ups of FftMethod::FftMethod (src/FftMethod.cpp:1706-1719)
  int F = cfg->frame_size, N = cfg->sample_point_size;
  if (F <= 0 || N <= 0) { free_ctx(c); return fail(nullptr, MOF_ERR_INVALID, "frame_size and sample_point_size must be positive"); }
  if (F % 2 == 1) F--;
  if (F % N != 0) N = F;
  // the FFT entry points need a section size of the radix plan; block matching and scale/rotation
  // contexts may use any frame size
  c->fft_generic = !supported_n(N) && mof::pg_supported(N);
  c->fft_ready = supported_n(N) || c->fft_generic;
  if (c->fft_generic && cfg->ocl_compat) {
    free_ctx(c);
    return fail(nullptr, MOF_ERR_INVALID, "ocl_compat needs a section size of 60, 120 or 240 (got %d)", N);
  }
  c->F = F; c->N = N; c->sq = F / N; c->S = c->sq * c->sq;
  if (c->cfg.max_batch <= 0) c->cfg.max_batch = 64;
#define CUC(call)                                                                              \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess) {                                                                   \
      int rc_ = fail(nullptr, MOF_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_));   \
      free_ctx(c);                                                                             \
      return rc_;                                                                              \
    }                                                                                          \
  } while (0)
  CUC(cudaSetDevice(c->device));
  CUC(cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, c->device));
  CUC((cudaError_t)mof::pc_setup(c->device));
  CUC((cudaError_t)mof::bm_setup(c->device));
  CUC((cudaError_t)mof::lp_setup(c->device));
  CUC((cudaError_t)mof::rf_setup(c->device));
  if (c->fft_ready) {
    c->pc_grid = c->fft_generic ? mof::pg_resident_ctas(N, c->device) : mof::pc_resident_ctas(N, c->device);
    if (c->pc_grid <= 0) { free_ctx(c); return fail(nullptr, MOF_ERR_CUDA, "occupancy query failed for N=%d", N); }
  }
  if (c->fft_generic) {
    std::vector<float> tw((size_t)2 * N);
    mof::pg_twiddles(N, tw.data());
    CUC(cudaMalloc(&c->d_pg_tw, tw.size() * sizeof(float)));
    CUC(cudaMemcpy(c->d_pg_tw, tw.data(), tw.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  for (int i = 0; i < 2; i++) CUC(cudaStreamCreateWithFlags(&c->stream[i], cudaStreamNonBlocking));
  CUC(cudaEventCreateWithFlags(&c->last_done, cudaEventDisableTiming));
  for (int i = 0; i < 2; i++) {
    CUC(cudaEventCreateWithFlags(&c->fe_copied[i], cudaEventDisableTiming));
    CUC(cudaEventCreateWithFlags(&c->fe_consumed[i], cudaEventDisableTiming));
  }
  const size_t fbytes = (size_t)F * F;
  c->scratch_bytes = !c->fft_ready ? 0
                     : c->fft_generic ? (size_t)c->pc_grid * mof::pg_scratch_bytes_per_cta(N)
                                      : (size_t)c->pc_grid * mof::pc_spectrum_bytes(N);
  for (int i = 0; i < 2; i++) {
    CUC(cudaMalloc(&c->d_curr[i], fbytes * c->cfg.max_batch));
    CUC(cudaMalloc(&c->d_prev[i], fbytes * c->cfg.max_batch));
    CUC(cudaMalloc(&c->d_out[i], sizeof(double) * 2 * c->S * c->cfg.max_batch));
    if (c->scratch_bytes) CUC(cudaMalloc(&c->d_scratch[i], c->scratch_bytes));
  }
  if (cfg->hann && c->fft_ready) {
    std::vector<float> w = build_hann(N);
    CUC(cudaMalloc(&c->d_win, w.size() * sizeof(float)));
    CUC(cudaMemcpy(c->d_win, w.data(), w.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  if (cfg->n_streams > 0 && c->fft_generic) {
    CUC(cudaMalloc(&c->d_frame_cache, fbytes * cfg->n_streams));
    c->stream_first.assign(cfg->n_streams, 1);
    c->stream_parity.assign(cfg->n_streams, 0);
  } else if (cfg->n_streams > 0 && c->fft_ready) {
    const size_t spec = (size_t)cfg->n_streams * c->S * 2 * mof::pc_spectrum_bytes(N);  // two halves per stream
    CUC(cudaMalloc(&c->d_spec_cache, spec));
    c->stream_first.assign(cfg->n_streams, 1);
    c->stream_parity.assign(cfg->n_streams, 0);
    CUC(cudaMalloc(&c->d_meta, (size_t)mof_ctx::META_SLOTS * 8 * cfg->n_streams));
    CUC(cudaMallocHost(&c->h_meta, (size_t)mof_ctx::META_SLOTS * 8 * cfg->n_streams));
    for (auto& e : c->meta_done) CUC(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  // scale / rotation: the log-polar kernel stages a (res + 8)^2 source image in shared memory (res <= 240, res % 4 == 0);
  // the correlation runs on the prime-factor kernel (60 / 120 / 240) or the generic one (other even 2^a 3^b 5^c)
  const bool sr_size_ok = cfg->sr_resolution >= 8 && cfg->sr_resolution <= 240 && cfg->sr_resolution % 4 == 0 &&
                          (supported_n(cfg->sr_resolution) || mof::pg_supported(cfg->sr_resolution));
  if (cfg->sr_resolution > 0 && sr_size_ok && cfg->sr_m > 0) {
    const int res = cfg->sr_resolution;
    std::vector<int16_t> ix, iy;
    std::vector<uint16_t> fxy;
    build_logpolar_map(res, cfg->sr_m, ix, iy, fxy);
    std::vector<int16_t> tc = build_weight_table(4), tl = build_weight_table(8);
    const size_t npx = (size_t)res * res;
    std::vector<uint32_t> pk(npx);
    for (size_t p = 0; p < npx; p++) {
      const int x = ix[p], y = iy[p];
      pk[p] = (x < 0 || x >= res || y < 0 || y >= res)
                  ? 0xffffffffu
                  : ((uint32_t)((y + 4) * (res + 8) + x + 4) << 10) | (uint32_t)fxy[p];
    }
    CUC(cudaMalloc(&c->d_lp_pk, npx * 4));
    CUC(cudaMemcpy(c->d_lp_pk, pk.data(), npx * 4, cudaMemcpyHostToDevice));
    CUC(cudaMalloc(&c->d_tab_cubic, tc.size() * 2));
    CUC(cudaMalloc(&c->d_tab_lanczos, tl.size() * 2));
    CUC(cudaMemcpy(c->d_tab_cubic, tc.data(), tc.size() * 2, cudaMemcpyHostToDevice));
    CUC(cudaMemcpy(c->d_tab_lanczos, tl.data(), tl.size() * 2, cudaMemcpyHostToDevice));
    CUC(cudaMalloc(&c->d_lp_img, npx * 2 * c->cfg.max_batch));
    c->sr_generic = !supported_n(res);
    if (c->sr_generic) {
      c->sr_grid = mof::pg_resident_ctas(res, c->device);
      if (c->sr_grid <= 0) { free_ctx(c); return fail(nullptr, MOF_ERR_CUDA, "occupancy query failed for sr_resolution=%d", res); }
      std::vector<float> tw((size_t)2 * res);
      mof::pg_twiddles(res, tw.data());
      CUC(cudaMalloc(&c->d_sr_tw, tw.size() * sizeof(float)));
      CUC(cudaMemcpy(c->d_sr_tw, tw.data(), tw.size() * sizeof(float), cudaMemcpyHostToDevice));
      CUC(cudaMalloc(&c->d_sr_scratch, (size_t)c->sr_grid * mof::pg_scratch_bytes_per_cta(res)));
    } else {
      c->sr_grid = mof::pc_resident_ctas(res, c->device);
      CUC(cudaMalloc(&c->d_sr_scratch, (size_t)c->sr_grid * mof::pc_spectrum_bytes(res)));
    }
    CUC(cudaMalloc(&c->d_sr_out, sizeof(double) * 2 * c->cfg.max_batch));
    CUC(cudaMalloc(&c->d_sr_temp, npx));
    CUC(cudaMalloc(&c->d_sr_prev, npx));
    c->sr_ready = true;
  }
#undef CUC
  *out = c;
  return MOF_OK;
}

void mof_destroy(mof_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  free_ctx(ctx);
}

int mof_get_geometry(const mof_ctx* c, int* F, int* N, int* S) {
  if (!c) return MOF_ERR_INVALID;
  if (F) *F = c->F;
  if (N) *N = c->N;
  if (S) *S = c->S;
  return MOF_OK;
}

int mof_sync(mof_ctx* c) {
  if (!c) return MOF_ERR_INVALID;
  CU(c, cudaSetDevice(c->device));
  for (int i = 0; i < 2; i++) CU(c, cudaStreamSynchronize(c->stream[i]));
  return MOF_OK;
}

long long mof_kernel_launches(const mof_ctx* c) { return c ? c->launches : 0; }

int mof_enable_kernel_timing(mof_ctx* c, int on) {
  if (!c) return MOF_ERR_INVALID;
  c->timing = on != 0;
  return MOF_OK;
}

int mof_fft_kernel_time(mof_ctx* c, double* avg_ms, long long* n, int reset) {
  if (!c) return MOF_ERR_INVALID;
  CU(c, cudaSetDevice(c->device));
  double tot = 0;
  long long cnt = 0;
  for (auto& t : c->timed) {
    if (cudaEventSynchronize(t.b) != cudaSuccess) continue;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) { tot += ms; cnt++; }
  }
  if (avg_ms) *avg_ms = cnt ? tot / (double)cnt : 0.0;
  if (n) *n = cnt;
  if (reset) {
    for (auto& t : c->timed) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
    c->timed.clear();
  }
  return MOF_OK;
}

// ---- FFT pairs ------------------------------------------------------------------------------------
static int fft_pairs_device_impl(mof_ctx* c, const uint8_t* d_curr, const uint8_t* d_prev,
                                 size_t pair_stride, int row_stride, int n_pairs, double* d_out,
                                 void* scratch, cudaStream_t st) {
  mof::PcArgs a{};
  a.curr = d_curr; a.prev = d_prev; a.frame_stride = pair_stride; a.row_stride = row_stride;
  a.sq = c->sq; a.S = c->S; a.n_sections = n_pairs * c->S; a.win = c->d_win;
  a.scratch = scratch; a.out = d_out;
  a.max_px_speed_sq = c->cfg.max_px_speed * c->cfg.max_px_speed;
  a.mode = mof::PC_PAIRS; a.post = mof::PC_POST_SHIFT; a.ocl = c->cfg.ocl_compat ? 1 : 0;
  return launch_pc(c, c->N, a, c->pc_grid, st);
}

int mof_fft_pairs_device(mof_ctx* c, const uint8_t* d_curr, const uint8_t* d_prev, size_t pair_stride,
                         int row_stride, int n_pairs, double* d_out, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (n_pairs == 0) return MOF_OK;
  if (!d_curr || !d_prev || !d_out || n_pairs < 0 || row_stride < c->F)
    return fail(c, MOF_ERR_INVALID, "mof_fft_pairs_device: bad argument");
  if ((long long)n_pairs * c->S > 0x7fffffffLL) return fail(c, MOF_ERR_INVALID, "too many sections");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  // the per-CTA scratch belongs to the context: order this launch after the previous one
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  int rc = fft_pairs_device_impl(c, d_curr, d_prev, pair_stride, row_stride, n_pairs, d_out,
                                 c->d_scratch[0], st);
  if (rc) return rc;
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

int mof_fft_pairs(mof_ctx* c, const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                  int row_stride, int n_pairs, double* shifts) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (n_pairs == 0) return MOF_OK;
  if (!curr || !prev || !shifts || n_pairs < 0 || row_stride < c->F)
    return fail(c, MOF_ERR_INVALID, "mof_fft_pairs: bad argument");
  CU(c, cudaSetDevice(c->device));
  const int F = c->F, mb = c->cfg.max_batch;
  const size_t fbytes = (size_t)F * F;
  // both streams must be ordered after earlier device-API work that used scratch slot 0
  CU(c, cudaStreamWaitEvent(c->stream[0], c->last_done, 0));
  int slot = 0;
  for (int p0 = 0; p0 < n_pairs; p0 += mb, slot ^= 1) {
    const int n = n_pairs - p0 < mb ? n_pairs - p0 : mb;
    cudaStream_t st = c->stream[slot];
    int rc = h2d_frames(c, c->d_curr[slot], curr + (size_t)p0 * pair_stride, pair_stride, row_stride,
                        F, F, n, st);
    if (rc) return rc;
    rc = h2d_frames(c, c->d_prev[slot], prev + (size_t)p0 * pair_stride, pair_stride, row_stride, F,
                    F, n, st);
    if (rc) return rc;
    rc = fft_pairs_device_impl(c, c->d_curr[slot], c->d_prev[slot], fbytes, F, n, c->d_out[slot],
                               c->d_scratch[slot], st);
    if (rc) return rc;
    CU(c, cudaMemcpyAsync(shifts + (size_t)p0 * c->S * 2, c->d_out[slot],
                          sizeof(double) * 2 * c->S * n, cudaMemcpyDeviceToHost, st));
  }
  CU(c, cudaStreamSynchronize(c->stream[0]));
  CU(c, cudaStreamSynchronize(c->stream[1]));
  CU(c, cudaEventRecord(c->last_done, c->stream[0]));
  return MOF_OK;
}

// ---- FFT streams ----------------------------------------------------------------------------------
// validates the ids of one push and stages (ids | first/parity flags) through the pinned ring; returns the
// device pointers.  The slot was last used META_SLOTS pushes ago.
static int stage_stream_meta(mof_ctx* c, int n, const int* ids, cudaStream_t st, int* slot_out, int** d_ids,
                             uint8_t** d_first, mof::PcArgs* inl) {
  if (!c->d_spec_cache) return fail(c, MOF_ERR_INVALID, "context was created with n_streams = 0");
  if (n < 0 || n > c->cfg.n_streams) return fail(c, MOF_ERR_INVALID, "n out of range");
  std::vector<char> seen(c->cfg.n_streams, 0);
  for (int k = 0; k < n; k++) {
    if (ids[k] < 0 || ids[k] >= c->cfg.n_streams) return fail(c, MOF_ERR_INVALID, "stream id %d out of range", ids[k]);
    if (seen[ids[k]]) return fail(c, MOF_ERR_INVALID, "stream id %d appears twice in one push", ids[k]);
    seen[ids[k]] = 1;
  }
  if (n <= mof::PcArgs::PC_INLINE && !getenv("MOF_NO_INLINE_IDS")) {   // in the kernel parameters: no copy, no ring slot
    for (int k = 0; k < n; k++) {
      inl->inline_ids[k] = ids[k];
      inl->inline_flags[k] = (uint8_t)(c->stream_first[ids[k]] | (c->stream_parity[ids[k]] << 1));
    }
    *d_ids = nullptr;
    *d_first = nullptr;
    *slot_out = -1;
    return MOF_OK;
  }
  const int slot = c->meta_slot;
  c->meta_slot = (slot + 1) % mof_ctx::META_SLOTS;
  CU(c, cudaEventSynchronize(c->meta_done[slot]));
  const size_t blk = (size_t)8 * c->cfg.n_streams;
  int* h_ids = reinterpret_cast<int*>(c->h_meta + slot * blk);
  uint8_t* h_first = c->h_meta + slot * blk + (size_t)4 * c->cfg.n_streams;
  *d_ids = reinterpret_cast<int*>(c->d_meta + slot * blk);
  *d_first = c->d_meta + slot * blk + (size_t)4 * c->cfg.n_streams;
  for (int k = 0; k < n; k++) {
    h_ids[k] = ids[k];
    h_first[k] = (uint8_t)(c->stream_first[ids[k]] | (c->stream_parity[ids[k]] << 1));
  }
  // one copy covers both arrays (ids, gap, flags)
  CU(c, cudaMemcpyAsync(*d_ids, h_ids, (size_t)4 * c->cfg.n_streams + n, cudaMemcpyHostToDevice, st));
  *slot_out = slot;
  return MOF_OK;
}

// Generic section sizes (pg_kernels.cu): a push is a pairs launch whose previous frames come from the streams'
// frame cache (first call: the frame itself), then the frames of the push replace them.  Batches of
// PC_INLINE frames (ids and flags travel in the kernel parameters).  T > 0: the results are frame t of a
// [n][T][S] array whose element [0][t] is d_out.
static int generic_stream_push(mof_ctx* c, int n, const int* ids, const uint8_t* d_frames, size_t frame_stride,
                               int row_stride, double* d_out, int T, cudaStream_t st, int mode) {
  if (!c->d_frame_cache) return fail(c, MOF_ERR_INVALID, "context was created with n_streams = 0");
  if (n < 0 || n > c->cfg.n_streams) return fail(c, MOF_ERR_INVALID, "n out of range");
  std::vector<char> seen(c->cfg.n_streams, 0);
  for (int k = 0; k < n; k++) {
    if (ids[k] < 0 || ids[k] >= c->cfg.n_streams) return fail(c, MOF_ERR_INVALID, "stream id %d out of range", ids[k]);
    if (seen[ids[k]]) return fail(c, MOF_ERR_INVALID, "stream id %d appears twice in one push", ids[k]);
    seen[ids[k]] = 1;
  }
  const int out_fs = T > 0 ? T * c->S : c->S;
  for (int k0 = 0; k0 < n; k0 += mof::PcArgs::PC_INLINE) {
    const int nb = n - k0 < mof::PcArgs::PC_INLINE ? n - k0 : mof::PcArgs::PC_INLINE;
    mof::PcArgs a{};
    a.curr = d_frames + (size_t)k0 * frame_stride; a.prev = nullptr; a.frame_stride = frame_stride; a.row_stride = row_stride;
    a.sq = c->sq; a.S = c->S; a.n_sections = nb * c->S; a.win = c->d_win; a.scratch = c->d_scratch[0];
    a.out = d_out + (size_t)k0 * out_fs * 2; a.seq_T = T;
    a.max_px_speed_sq = c->cfg.max_px_speed * c->cfg.max_px_speed;
    a.mode = mof::PC_STREAM; a.post = mof::PC_POST_SHIFT;
    for (int k = 0; k < nb; k++) {
      a.inline_ids[k] = ids[k0 + k];
      a.inline_flags[k] = c->stream_first[ids[k0 + k]];
    }
    if (mode != mof::PC_SPECTRUM) {
      int rc = launch_pc(c, c->N, a, c->pc_grid, st);
      if (rc) return rc;
    }
    int rc = mof::pg_cache_update(a, nb, c->d_frame_cache, c->F, st);
    c->launches++;
    if (rc) return fail(c, MOF_ERR_CUDA, "frame cache update failed: %s", cudaGetErrorString((cudaError_t)rc));
  }
  for (int k = 0; k < n; k++) c->stream_first[ids[k]] = 0;
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

static int stream_push_impl(mof_ctx* c, int n, const int* ids, const uint8_t* d_frames,
                            size_t frame_stride, int row_stride, double* d_out, cudaStream_t st,
                            int mode) {
  if (c->fft_generic) return generic_stream_push(c, n, ids, d_frames, frame_stride, row_stride, d_out, 0, st, mode);
  int slot = 0;
  int* d_ids = nullptr;
  uint8_t* d_first = nullptr;
  mof::PcArgs a{};
  int rc = stage_stream_meta(c, n, ids, st, &slot, &d_ids, &d_first, &a);
  if (rc) return rc;
  a.curr = d_frames; a.prev = nullptr; a.frame_stride = frame_stride; a.row_stride = row_stride;
  a.sq = c->sq; a.S = c->S; a.n_sections = n * c->S; a.win = c->d_win;
  a.spec_cache = c->d_spec_cache; a.stream_ids = d_ids; a.first = d_first;
  a.out = d_out;
  a.max_px_speed_sq = c->cfg.max_px_speed * c->cfg.max_px_speed;
  a.mode = mode; a.post = mof::PC_POST_SHIFT; a.ocl = c->cfg.ocl_compat ? 1 : 0;
  rc = launch_pc(c, c->N, a, c->pc_grid, st);
  if (rc) return rc;
  for (int k = 0; k < n; k++) {
    c->stream_first[ids[k]] = 0;
    if (mode == mof::PC_STREAM) c->stream_parity[ids[k]] ^= 1;  // the kernel wrote the other half
  }
  if (slot >= 0) CU(c, cudaEventRecord(c->meta_done[slot], st));
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

// Chunks of a host-frame push_many call (chunk i + 1 crosses PCIe while chunk i runs).  The copies are the
// bottleneck and run back to back, so what a call pays on top of them is a fixed cost per copy plus the LAST
// chunk's kernel: few chunks (MOF_PIPE_CHUNKS overrides, 1..64).
static int pipe_chunks() {
  static const int v = [] {
    const char* e = getenv("MOF_PIPE_CHUNKS");
    const int k = e ? atoi(e) : 0;
    return k >= 1 && k <= 64 ? k : MOF_PIPE_CHUNKS_DEFAULT;
  }();
  return v;
}

// T frames per stream in one launch (PC_SEQ).  The (stream, section) sequences are cut into J time
// segments so that the work items fill the resident slots: J minimises
//   rounds(J) * (frames per segment + the extra forward transform a later segment starts with).
static int stream_push_many_impl(mof_ctx* c, int n, const int* ids, int T, const uint8_t* d_frames,
                                 size_t stream_stride, size_t frame_stride, int row_stride, double* d_out,
                                 cudaStream_t st) {
  if (T == 1) return stream_push_impl(c, n, ids, d_frames, stream_stride, row_stride, d_out, st, mof::PC_STREAM);
  if (c->fft_generic) {   // frame by frame: the previous frame of every stream sits in the frame cache
    for (int t = 0; t < T; t++) {
      int rc = generic_stream_push(c, n, ids, d_frames + (size_t)t * frame_stride, stream_stride, row_stride,
                                   d_out + (size_t)t * c->S * 2, T, st, mof::PC_STREAM);
      if (rc) return rc;
    }
    return MOF_OK;
  }
  int slot = 0;
  int* d_ids = nullptr;
  uint8_t* d_first = nullptr;
  mof::PcArgs a{};
  int rc = stage_stream_meta(c, n, ids, st, &slot, &d_ids, &d_first, &a);
  if (rc) return rc;
  if (!c->d_seq_scratch) {
    if (cudaMalloc(&c->d_seq_scratch, (size_t)c->pc_grid * 2 * mof::pc_spectrum_bytes(c->N)) != cudaSuccess)
      return fail(c, MOF_ERR_NOMEM, "cudaMalloc of the sequence scratch failed");
  }
  // segment length L: T = J * L + R; the J whole segments of every sequence are work items of L steps, the
  // remainder pieces (R frames, first step with the extra transform: R + 0.6) are bundled Q to a work item so
  // that a bundle costs no more than a segment.  L = ceil(T / J) with Q = 1 is the uniform cut of before; e.g.
  // 128 sequences of 64 frames on 444 slots: L = 20, J = 3, R = 4, Q = 4 -> 416 items, 20 steps instead of 22.
  const long long items0 = (long long)n * c->S;
  int best_L = T, best_J = 1, best_R = 0, best_Q = 1;
  long long best_items = items0;
  double best_cost = 1e300;
  const bool uniform_only = getenv("MOF_SEQ_UNIFORM") != nullptr;   // A/B: the cut of before
  for (int L = T; L >= 1; L--) {
    const int J = T / L, R = T - J * L;
    if (J > 64) break;
    int Q = R ? (int)(((double)L + 0.6) / ((double)R + 0.6)) : 1;
    if (Q < 1 || uniform_only) Q = 1;
    if (uniform_only && L != (T + (J + (R ? 1 : 0)) - 1) / (J + (R ? 1 : 0))) continue;
    const long long items = items0 * J + (R ? (items0 + Q - 1) / Q : 0);
    const long long rounds = (items + c->pc_grid - 1) / c->pc_grid;
    const double cost = (double)rounds * ((double)L + ((J > 1 || R) ? 0.6 : 0.0));
    if (cost < best_cost - 1e-9) { best_cost = cost; best_L = L; best_J = J; best_R = R; best_Q = Q; best_items = items; }
  }
  if (best_items > 0x7fffffffLL || (long long)n * T * c->S > 0x3fffffffLL)
    return fail(c, MOF_ERR_INVALID, "too many sections");
  // The tape cut (default): all sequences end to end, every resident slot walks C = ceil(frames / slots)
  // consecutive frames -- across sequence boundaries, where the chain restarts from the stream cache -- so the
  // slots finish together whatever T and the number of sequences are: cost C + 0.6 against rounds * (L + 0.6).
  // MOF_SEQ_CUT=bundles (or MOF_SEQ_UNIFORM=1) keeps the segment cuts above for A/B.
  const char* cut = getenv("MOF_SEQ_CUT");
  int tape_C = 0;
  if (!uniform_only && !(cut && !strcmp(cut, "bundles"))) {
    const long long total = items0 * T;
    tape_C = (int)((total + c->pc_grid - 1) / c->pc_grid);
    best_L = tape_C;
    best_items = (total + tape_C - 1) / tape_C;   // slots in use
  }
  const int L = best_L;
  a.curr = d_frames; a.prev = nullptr; a.frame_stride = frame_stride; a.row_stride = row_stride;
  a.stream_stride = stream_stride;
  a.sq = c->sq; a.S = c->S; a.n_sections = (int)best_items; a.win = c->d_win;
  a.spec_cache = c->d_spec_cache; a.stream_ids = d_ids; a.first = d_first;
  a.seq_T = T; a.seq_L = L; a.seq_J = best_J; a.seq_R = best_R; a.seq_Q = best_Q; a.seq_nseq = (int)items0;
  a.seq_C = tape_C;
  a.seq_scratch = c->d_seq_scratch;
  a.out = d_out;
  a.max_px_speed_sq = c->cfg.max_px_speed * c->cfg.max_px_speed;
  a.mode = mof::PC_SEQ; a.post = mof::PC_POST_SHIFT; a.ocl = c->cfg.ocl_compat ? 1 : 0;
  rc = launch_pc(c, c->N, a, c->pc_grid, st);
  if (rc) return rc;
  for (int k = 0; k < n; k++) {
    c->stream_first[ids[k]] = 0;
    c->stream_parity[ids[k]] ^= 1;   // frame T-1 went into the other half of the stream's cache
  }
  if (slot >= 0) CU(c, cudaEventRecord(c->meta_done[slot], st));
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

int mof_fft_stream_push_many_device(mof_ctx* c, int n, const int* ids, int T, const uint8_t* d_frames,
                                    size_t stream_stride, size_t frame_stride, int row_stride, double* d_out,
                                    void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!ids || !d_frames || !d_out || n < 0 || T < 1 || row_stride < c->F)
    return fail(c, MOF_ERR_INVALID, "mof_fft_stream_push_many_device: bad argument");
  if (n == 0) return MOF_OK;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  return stream_push_many_impl(c, n, ids, T, d_frames, stream_stride, frame_stride, row_stride, d_out, st);
}

int mof_fft_stream_push_many(mof_ctx* c, int n, const int* ids, int T, const uint8_t* frames, size_t stream_stride,
                             size_t frame_stride, int row_stride, double* shifts) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!ids || !frames || !shifts || n < 0 || T < 1 || row_stride < c->F)
    return fail(c, MOF_ERR_INVALID, "mof_fft_stream_push_many: bad argument");
  if (n > c->cfg.max_batch) return fail(c, MOF_ERR_INVALID, "n = %d exceeds max_batch = %d", n, c->cfg.max_batch);
  if (n == 0) return MOF_OK;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  const int F = c->F;
  const size_t fb = (size_t)F * F;
  // chunks of Tc frames per stream; chunk i + 1 is copied (copy stream, the other staging slot) while chunk
  // i runs on stream[0] (the stream state chains from chunk to chunk, so launches and result copies stay
  // in order there)
  const int Tcap = c->cfg.max_batch / n;   // frames per stream per chunk that fit one staging slot
  int Tc = Tcap < T ? Tcap : T;
  if (T >= 16 && Tc > (T + pipe_chunks() - 1) / pipe_chunks()) Tc = (T + pipe_chunks() - 1) / pipe_chunks();
  const bool piped = Tc < T;
  cudaStream_t cp = piped ? c->stream[1] : st;
  if (piped) CU(c, cudaStreamWaitEvent(cp, c->last_done, 0));
  const size_t row = sizeof(double) * 2 * c->S;
  for (int i = 0, t0 = 0; t0 < T; i++, t0 += Tc) {
    const int tc = T - t0 < Tc ? T - t0 : Tc;
    const int b = piped ? (i & 1) : 0;
    if (i >= 2) CU(c, cudaStreamWaitEvent(cp, c->fe_consumed[b], 0));
    if (row_stride == F && frame_stride == fb) {
      // a stream's frames are contiguous: the chunk of ALL streams is one 2-D copy (row = tc frames of one
      // stream, pitch = stream_stride); per-stream copies of ~100 KB are API-call bound (17 vs 50 GB/s)
      CU(c, cudaMemcpy2DAsync(c->d_curr[b], (size_t)tc * fb, frames + (size_t)t0 * fb, stream_stride, (size_t)tc * fb, n,
                              cudaMemcpyHostToDevice, cp));
    } else {
      for (int k = 0; k < n; k++) {
        int rc = h2d_frames(c, c->d_curr[b] + (size_t)k * tc * fb, frames + (size_t)k * stream_stride + (size_t)t0 * frame_stride,
                            frame_stride, row_stride, F, F, tc, cp);
        if (rc) return rc;
      }
    }
    if (piped) {
      CU(c, cudaEventRecord(c->fe_copied[b], cp));
      CU(c, cudaStreamWaitEvent(st, c->fe_copied[b], 0));
    }
    int rc = stream_push_many_impl(c, n, ids, tc, c->d_curr[b], (size_t)tc * fb, fb, F, c->d_out[0], st);
    if (rc) return rc;
    if (piped) CU(c, cudaEventRecord(c->fe_consumed[b], st));
    CU(c, cudaMemcpy2DAsync(shifts + (size_t)t0 * c->S * 2, row * T, c->d_out[0], row * tc, row * tc, n,
                            cudaMemcpyDeviceToHost, st));
  }
  CU(c, cudaStreamSynchronize(st));
  return MOF_OK;
}

int mof_fft_stream_push_device(mof_ctx* c, int n, const int* ids, const uint8_t* d_frames,
                               size_t frame_stride, int row_stride, double* d_out, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!ids || !d_frames || !d_out || n < 0 || row_stride < c->F) return fail(c, MOF_ERR_INVALID, "mof_fft_stream_push_device: bad argument");
  if (n == 0) return MOF_OK;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  return stream_push_impl(c, n, ids, d_frames, frame_stride, row_stride, d_out, st, mof::PC_STREAM);
}

int mof_fft_stream_push(mof_ctx* c, int n, const int* ids, const uint8_t* frames, size_t frame_stride,
                        int row_stride, double* shifts) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!ids || !frames || !shifts || n < 0 || row_stride < c->F) return fail(c, MOF_ERR_INVALID, "mof_fft_stream_push: bad argument");
  if (n > c->cfg.max_batch) return fail(c, MOF_ERR_INVALID, "n = %d exceeds max_batch = %d", n, c->cfg.max_batch);
  if (n == 0) return MOF_OK;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));   // device-API pushes on caller streams share the cache
  const int F = c->F;
  int rc = h2d_frames(c, c->d_curr[0], frames, frame_stride, row_stride, F, F, n, st);
  if (rc) return rc;
  rc = stream_push_impl(c, n, ids, c->d_curr[0], (size_t)F * F, F, c->d_out[0], st, mof::PC_STREAM);
  if (rc) return rc;
  CU(c, cudaMemcpyAsync(shifts, c->d_out[0], sizeof(double) * 2 * c->S * n, cudaMemcpyDeviceToHost, st));
  CU(c, cudaStreamSynchronize(st));
  return MOF_OK;
}

int mof_fft_stream_set_prev(mof_ctx* c, int id, const uint8_t* frame, int row_stride) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!frame || row_stride < c->F) return fail(c, MOF_ERR_INVALID, "mof_fft_stream_set_prev: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  const int F = c->F;
  int rc = h2d_frames(c, c->d_curr[0], frame, (size_t)row_stride * F, row_stride, F, F, 1, st);
  if (rc) return rc;
  rc = stream_push_impl(c, 1, &id, c->d_curr[0], (size_t)F * F, F, c->d_out[0], st, mof::PC_SPECTRUM);
  if (rc) return rc;
  CU(c, cudaStreamSynchronize(st));
  return MOF_OK;
}

int mof_fft_stream_reset(mof_ctx* c, int id) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (c->stream_first.empty()) return fail(c, MOF_ERR_INVALID, "context was created with n_streams = 0");
  if (id < 0 || id >= c->cfg.n_streams) return fail(c, MOF_ERR_INVALID, "stream id out of range");
  c->stream_first[id] = 1;
  return MOF_OK;
}

// ---- block matching ---------------------------------------------------------------------------------
int mof_bm_geometry(const mof_ctx* c, int width, int height, int* gx, int* gy) {
  if (!c) return MOF_ERR_INVALID;
  const int pitch = c->cfg.bm_block + c->cfg.bm_step;
  if (pitch <= 0 || c->cfg.bm_block <= 0 || c->cfg.bm_radius < 0 || 2 * c->cfg.bm_radius + 1 > 50)
    return MOF_ERR_INVALID;
  // grid = (cols - 2R)/(B+step) x (rows - 2R)/(B+step)  (FastSpacedBMMethod_OCL.cpp:81-90)
  int x = (width - 2 * c->cfg.bm_radius) / pitch, y = (height - 2 * c->cfg.bm_radius) / pitch;
  if (gx) *gx = x > 0 ? x : 0;
  if (gy) *gy = y > 0 ? y : 0;
  return MOF_OK;
}

static int bm_device_impl(mof_ctx* c, const uint8_t* d_curr, const uint8_t* d_prev, size_t pair_stride,
                          int width, int height, int row_stride, int n_pairs, int8_t* d_flow,
                          float* d_voted, cudaStream_t st) {
  int gx = 0, gy = 0;
  if (mof_bm_geometry(c, width, height, &gx, &gy) != MOF_OK || gx <= 0 || gy <= 0)
    return fail(c, MOF_ERR_INVALID, "block-matching geometry invalid (B=%d step=%d R=%d on %dx%d)",
                c->cfg.bm_block, c->cfg.bm_step, c->cfg.bm_radius, width, height);
  mof::BmArgs a{};
  a.curr = d_curr; a.prev = d_prev; a.pair_stride = pair_stride; a.row_stride = row_stride;
  a.width = width; a.height = height; a.B = c->cfg.bm_block; a.step = c->cfg.bm_step;
  a.R = c->cfg.bm_radius; a.gx = gx; a.gy = gy; a.n_pairs = n_pairs; a.flow_xy = d_flow;
  a.voted_xy = d_voted;
  a.no_gate = c->cfg.bm_no_gate ? 1 : 0;
  if (n_pairs == 0) return MOF_OK;
  int nl = 0;
  int rc = mof::bm_launch(a, st, &nl);
  c->launches += nl;
  if (rc) return fail(c, MOF_ERR_CUDA, "block-matching launch failed: %s", cudaGetErrorString((cudaError_t)rc));
  return MOF_OK;
}

int mof_bm_pairs_device(mof_ctx* c, const uint8_t* d_curr, const uint8_t* d_prev, size_t pair_stride,
                        int width, int height, int row_stride, int n_pairs, int8_t* d_flow,
                        float* d_voted, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (n_pairs == 0) return MOF_OK;
  if (!d_curr || !d_prev || !d_flow || !d_voted || n_pairs < 0 || row_stride < width)
    return fail(c, MOF_ERR_INVALID, "mof_bm_pairs_device: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  return bm_device_impl(c, d_curr, d_prev, pair_stride, width, height, row_stride, n_pairs, d_flow,
                        d_voted, st);
}

int mof_bm_pairs(mof_ctx* c, const uint8_t* curr, const uint8_t* prev, size_t pair_stride, int width,
                 int height, int row_stride, int n_pairs, int8_t* flow, float* voted) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (n_pairs == 0) return MOF_OK;
  if (!curr || !prev || !flow || !voted || n_pairs < 0 || row_stride < width)
    return fail(c, MOF_ERR_INVALID, "mof_bm_pairs: bad argument");
  int gx = 0, gy = 0;
  if (mof_bm_geometry(c, width, height, &gx, &gy) != MOF_OK || gx <= 0 || gy <= 0)
    return fail(c, MOF_ERR_INVALID, "block-matching geometry invalid");
  CU(c, cudaSetDevice(c->device));
  const size_t fbytes = (size_t)width * height, nb = (size_t)gx * gy;
  const int chunk = c->cfg.max_batch;
  if (c->bm_cap_bytes < fbytes * chunk || c->bm_cap_flow < nb * 2 * chunk) {
    cudaFree(c->d_bm_curr); cudaFree(c->d_bm_prev); cudaFree(c->d_bm_flow); cudaFree(c->d_bm_voted);
    c->d_bm_curr = c->d_bm_prev = nullptr; c->d_bm_flow = nullptr; c->d_bm_voted = nullptr;
    c->bm_cap_bytes = c->bm_cap_flow = 0;
    CU(c, cudaMalloc(&c->d_bm_curr, fbytes * chunk));
    CU(c, cudaMalloc(&c->d_bm_prev, fbytes * chunk));
    CU(c, cudaMalloc(&c->d_bm_flow, nb * 2 * chunk));
    CU(c, cudaMalloc(&c->d_bm_voted, sizeof(float) * 2 * chunk));
    c->bm_cap_bytes = fbytes * chunk;
    c->bm_cap_flow = nb * 2 * chunk;
  }
  cudaStream_t st = c->stream[0];
  for (int p0 = 0; p0 < n_pairs; p0 += chunk) {
    const int n = n_pairs - p0 < chunk ? n_pairs - p0 : chunk;
    int rc = h2d_frames(c, c->d_bm_curr, curr + (size_t)p0 * pair_stride, pair_stride, row_stride,
                        width, height, n, st);
    if (rc) return rc;
    rc = h2d_frames(c, c->d_bm_prev, prev + (size_t)p0 * pair_stride, pair_stride, row_stride, width,
                    height, n, st);
    if (rc) return rc;
    rc = bm_device_impl(c, c->d_bm_curr, c->d_bm_prev, fbytes, width, height, width, n, c->d_bm_flow,
                        c->d_bm_voted, st);
    if (rc) return rc;
    CU(c, cudaMemcpyAsync(flow + (size_t)p0 * 2 * nb, c->d_bm_flow, nb * 2 * n, cudaMemcpyDeviceToHost, st));
    CU(c, cudaMemcpyAsync(voted + (size_t)p0 * 2, c->d_bm_voted, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, st));
    CU(c, cudaStreamSynchronize(st));
  }
  return MOF_OK;
}

// ---- scale / rotation -------------------------------------------------------------------------------
static int scalerot_device_impl(mof_ctx* c, const uint8_t* d_curr, const uint8_t* d_prev,
                                size_t pair_stride, int row_stride, int n, double* d_out,
                                cudaStream_t st) {
  const int res = c->cfg.sr_resolution;
  const size_t npx = (size_t)res * res;
  mof::LpMap m{c->d_lp_pk, c->d_tab_cubic, c->d_tab_lanczos, res};
  uint8_t* lp_prev = c->d_lp_img;
  uint8_t* lp_curr = c->d_lp_img + npx * c->cfg.max_batch;
  // first call: tempIm = zeros, logPolar(INTER_CUBIC)  (scaleRotationEstimator.cpp:27, :45)
  CU(c, cudaMemsetAsync(lp_prev, 0, npx * n, st));
  mof::LpArgs la{d_prev, pair_stride, row_stride, lp_prev, npx, res, n, 2};
  int rc = mof::lp_launch(m, la, st);
  c->launches++;
  if (rc) return fail(c, MOF_ERR_CUDA, "log-polar launch failed: %s", cudaGetErrorString((cudaError_t)rc));
  // later call: the same persistent tempIm is overwritten with INTER_LANCZOS4 (:112); transparent
  // pixels keep what the first remap left there (bg = lp_prev)
  mof::LpArgs lb{d_curr, pair_stride, row_stride, lp_curr, npx, res, n, 4, lp_prev, npx, res};
  rc = mof::lp_launch(m, lb, st);
  c->launches++;
  if (rc) return fail(c, MOF_ERR_CUDA, "log-polar launch failed: %s", cudaGetErrorString((cudaError_t)rc));
  mof::PcArgs a{};
  a.curr = lp_curr; a.prev = lp_prev; a.frame_stride = npx; a.row_stride = res; a.sq = 1; a.S = 1;
  a.n_sections = n; a.win = nullptr; a.scratch = c->d_sr_scratch; a.out = d_out;
  a.max_px_speed_sq = 1e300; a.sr_m = c->cfg.sr_m; a.mode = mof::PC_PAIRS; a.post = mof::PC_POST_SCALEROT;
  return launch_pc(c, res, a, c->sr_grid, st);
}

int mof_scalerot_pairs_device(mof_ctx* c, const uint8_t* d_curr, const uint8_t* d_prev, size_t pair_stride,
                              int row_stride, int n_pairs, double* d_out, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->sr_ready) return fail(c, MOF_ERR_INVALID, "scale/rotation not configured (sr_resolution: a multiple of 4 up to 240 of the form 2^a 3^b 5^c, sr_m > 0)");
  if (n_pairs == 0) return MOF_OK;
  if (!d_curr || !d_prev || !d_out || n_pairs < 0 || row_stride < c->cfg.sr_resolution)
    return fail(c, MOF_ERR_INVALID, "mof_scalerot_pairs_device: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  for (int p0 = 0; p0 < n_pairs; p0 += c->cfg.max_batch) {
    const int n = n_pairs - p0 < c->cfg.max_batch ? n_pairs - p0 : c->cfg.max_batch;
    int rc = scalerot_device_impl(c, d_curr + (size_t)p0 * pair_stride, d_prev + (size_t)p0 * pair_stride,
                                  pair_stride, row_stride, n, d_out + (size_t)p0 * 2, st);
    if (rc) return rc;
  }
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

int mof_scalerot_pairs(mof_ctx* c, const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                       int row_stride, int n_pairs, double* scale_rot) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->sr_ready) return fail(c, MOF_ERR_INVALID, "scale/rotation not configured (sr_resolution: a multiple of 4 up to 240 of the form 2^a 3^b 5^c, sr_m > 0)");
  const int res = c->cfg.sr_resolution;
  if (n_pairs == 0) return MOF_OK;
  if (!curr || !prev || !scale_rot || n_pairs < 0 || row_stride < res)
    return fail(c, MOF_ERR_INVALID, "mof_scalerot_pairs: bad argument");
  if ((size_t)res * res > (size_t)c->F * c->F)
    return fail(c, MOF_ERR_INVALID, "sr_resolution larger than frame_size: staging too small");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  const size_t npx = (size_t)res * res;
  for (int p0 = 0; p0 < n_pairs; p0 += c->cfg.max_batch) {
    const int n = n_pairs - p0 < c->cfg.max_batch ? n_pairs - p0 : c->cfg.max_batch;
    int rc = h2d_frames(c, c->d_curr[0], curr + (size_t)p0 * pair_stride, pair_stride, row_stride, res, res, n, st);
    if (rc) return rc;
    rc = h2d_frames(c, c->d_prev[0], prev + (size_t)p0 * pair_stride, pair_stride, row_stride, res, res, n, st);
    if (rc) return rc;
    rc = scalerot_device_impl(c, c->d_curr[0], c->d_prev[0], npx, res, n, c->d_sr_out, st);
    if (rc) return rc;
    CU(c, cudaMemcpyAsync(scale_rot + (size_t)p0 * 2, c->d_sr_out, sizeof(double) * 2 * n, cudaMemcpyDeviceToHost, st));
    CU(c, cudaStreamSynchronize(st));
  }
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

int mof_scalerot_reset(mof_ctx* c) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  c->sr_first = true;
  return MOF_OK;
}

int mof_scalerot_push(mof_ctx* c, const uint8_t* frame, int row_stride, double* scale_rot) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->sr_ready) return fail(c, MOF_ERR_INVALID, "scale/rotation not configured");
  const int res = c->cfg.sr_resolution;
  if (!frame || !scale_rot || row_stride < res) return fail(c, MOF_ERR_INVALID, "mof_scalerot_push: bad argument");
  if ((size_t)res * res > (size_t)c->F * c->F) return fail(c, MOF_ERR_INVALID, "staging too small");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  const size_t npx = (size_t)res * res;
  mof::LpMap m{c->d_lp_pk, c->d_tab_cubic, c->d_tab_lanczos, res};
  int rc = h2d_frames(c, c->d_curr[0], frame, (size_t)row_stride * res, row_stride, res, res, 1, st);
  if (rc) return rc;
  if (c->sr_first) {
    // tempIm = zeros (ctor, :27); logPolar(INTER_CUBIC) (:45); prevIm_F32 = tempIm (:48); return (1,0)
    CU(c, cudaMemsetAsync(c->d_sr_temp, 0, npx, st));
    mof::LpArgs la{c->d_curr[0], npx, res, c->d_sr_temp, npx, res, 1, 2};
    rc = mof::lp_launch(m, la, st);
    c->launches++;
    if (rc) return fail(c, MOF_ERR_CUDA, "log-polar launch failed: %s", cudaGetErrorString((cudaError_t)rc));
    CU(c, cudaMemcpyAsync(c->d_sr_prev, c->d_sr_temp, npx, cudaMemcpyDeviceToDevice, st));
    CU(c, cudaStreamSynchronize(st));
    c->sr_first = false;
    scale_rot[0] = 1.0;
    scale_rot[1] = 0.0;
    return MOF_OK;
  }
  mof::LpArgs la{c->d_curr[0], npx, res, c->d_sr_temp, npx, res, 1, 4};  // :112
  rc = mof::lp_launch(m, la, st);
  c->launches++;
  if (rc) return fail(c, MOF_ERR_CUDA, "log-polar launch failed: %s", cudaGetErrorString((cudaError_t)rc));
  mof::PcArgs a{};
  a.curr = c->d_sr_temp; a.prev = c->d_sr_prev; a.frame_stride = npx; a.row_stride = res; a.sq = 1; a.S = 1;
  a.n_sections = 1; a.scratch = c->d_sr_scratch; a.out = c->d_sr_out; a.max_px_speed_sq = 1e300;
  a.sr_m = c->cfg.sr_m; a.mode = mof::PC_PAIRS; a.post = mof::PC_POST_SCALEROT; a.sr_reject_nan = 1;
  rc = launch_pc(c, res, a, c->sr_grid, st);
  if (rc) return rc;
  double out[2];
  CU(c, cudaMemcpyAsync(out, c->d_sr_out, sizeof(out), cudaMemcpyDeviceToHost, st));
  CU(c, cudaStreamSynchronize(st));
  if (out[0] != out[0]) {  // rejected (:119-121): previous image is kept
    scale_rot[0] = 1.0;
    scale_rot[1] = 0.0;
  } else {
    scale_rot[0] = out[0];
    scale_rot[1] = out[1];
    CU(c, cudaMemcpyAsync(c->d_sr_prev, c->d_sr_temp, npx, cudaMemcpyDeviceToDevice, st));  // :128
    CU(c, cudaStreamSynchronize(st));
  }
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

int mof_logpolar(mof_ctx* c, const uint8_t* src, int src_stride, int interp, uint8_t* dst, int dst_stride,
                 int n_images, size_t src_image_stride, size_t dst_image_stride) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->sr_ready) return fail(c, MOF_ERR_INVALID, "scale/rotation not configured");
  const int res = c->cfg.sr_resolution;
  if (!src || !dst || n_images < 0 || src_stride < res || dst_stride < res || (interp != 2 && interp != 4))
    return fail(c, MOF_ERR_INVALID, "mof_logpolar: bad argument");
  if ((size_t)res * res > (size_t)c->F * c->F) return fail(c, MOF_ERR_INVALID, "staging too small");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  const size_t npx = (size_t)res * res;
  mof::LpMap m{c->d_lp_pk, c->d_tab_cubic, c->d_tab_lanczos, res};
  for (int p0 = 0; p0 < n_images; p0 += c->cfg.max_batch) {
    const int n = n_images - p0 < c->cfg.max_batch ? n_images - p0 : c->cfg.max_batch;
    int rc = h2d_frames(c, c->d_curr[0], src + (size_t)p0 * src_image_stride, src_image_stride, src_stride, res, res, n, st);
    if (rc) return rc;
    rc = h2d_frames(c, c->d_lp_img, dst + (size_t)p0 * dst_image_stride, dst_image_stride, dst_stride, res, res, n, st);
    if (rc) return rc;
    mof::LpArgs la{c->d_curr[0], npx, res, c->d_lp_img, npx, res, n, interp};
    rc = mof::lp_launch(m, la, st);
    c->launches++;
    if (rc) return fail(c, MOF_ERR_CUDA, "log-polar launch failed: %s", cudaGetErrorString((cudaError_t)rc));
    for (int k = 0; k < n; k++)
      CU(c, cudaMemcpy2DAsync(dst + (size_t)(p0 + k) * dst_image_stride, dst_stride, c->d_lp_img + (size_t)k * npx,
                              res, res, res, cudaMemcpyDeviceToHost, st));
    CU(c, cudaStreamSynchronize(st));
  }
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

// ---- camera front end --------------------------------------------------------------------------------
static int grow(mof_ctx* c, uint8_t** buf, size_t* cap, size_t need) {
  if (need <= *cap) return MOF_OK;
  CU(c, cudaDeviceSynchronize());
  cudaFree(*buf);
  *buf = nullptr;
  *cap = 0;
  if (cudaMalloc(buf, need) != cudaSuccess) return fail(c, MOF_ERR_NOMEM, "cudaMalloc of %zu bytes failed", need);
  *cap = need;
  return MOF_OK;
}

// geometry of src/optic_flow.cpp:1603-1617 -> kernel arguments
static int fe_args(mof_ctx* c, const mof_frontend_desc* d, int n, const uint8_t* src, uint8_t* dst,
                   size_t dst_image_stride, int dst_row_stride, mof::FeArgs* out) {
  if (!d || d->width <= 0 || d->height <= 0 || (d->channels != 1 && d->channels != 3))
    return fail(c, MOF_ERR_INVALID, "front end: bad descriptor (channels must be 1 or 3)");
  const int k = d->scale_factor <= 0 ? 1 : d->scale_factor;
  if ((k != 1 && k != 2 && k != 4) || d->width % k || d->height % k)
    return fail(c, MOF_ERR_INVALID, "front end: scale_factor %d unsupported (1, 2 or 4, dividing %d x %d)", k,
                d->width, d->height);
  if (d->row_stride < d->width * d->channels) return fail(c, MOF_ERR_INVALID, "front end: row_stride too small");
  const int ws = d->width / k, hs = d->height / k, F = c->F;
  const int x0 = d->cx - F / 2, y0 = hs / 2 - F / 2;
  if (x0 < 0 || y0 < 0 || x0 + F > ws || y0 + F > hs)
    return fail(c, MOF_ERR_INVALID, "front end: %d x %d crop at (%d, %d) leaves the %d x %d scaled frame", F, F,
                x0, y0, ws, hs);
  mof::FeArgs a{};
  a.src = src; a.src_image_stride = d->image_stride; a.src_row_stride = d->row_stride;
  a.channels = d->channels; a.k = k; a.x0 = x0; a.y0 = y0; a.out_w = F; a.out_h = F;
  a.dst = dst; a.dst_image_stride = dst_image_stride; a.dst_row_stride = dst_row_stride; a.n_images = n;
  *out = a;
  return MOF_OK;
}

static int fe_run(mof_ctx* c, const mof::FeArgs& a, cudaStream_t st) {
  int rc = mof::fe_launch(a, c->sms, st);
  c->launches++;
  if (rc) return fail(c, MOF_ERR_CUDA, "front-end kernel launch failed: %s", cudaGetErrorString((cudaError_t)rc));
  return MOF_OK;
}

// raw HOST frames -> device staging.  Only the window the front end reads crosses PCIe: rows
// [y0 k, (y0 + F) k) and byte columns [x0 k ch, (x0 + F) k ch) of every frame (cfg5: 1440 of 2256 bytes per
// row), landing as a tight (F k) x (F k) frame whose descriptor is *dd.  FE_SLACK bytes behind the last
// frame belong to the staging buffer (fe_bgr16_kernel reads up to 31 bytes past a crop row).
// MOF_FE_FULL_H2D=1 stages whole frames (A/B).
constexpr int FE_SLACK = 64;
static void fe_crop_desc(const mof_ctx* c, const mof_frontend_desc* d, mof_frontend_desc* dd, size_t* src_off,
                         bool* whole) {
  const int k = d->scale_factor <= 0 ? 1 : d->scale_factor;
  *whole = getenv("MOF_FE_FULL_H2D") != nullptr;
  *dd = *d;
  if (*whole) {
    dd->row_stride = d->width * d->channels;
    dd->image_stride = (size_t)dd->row_stride * d->height;
    *src_off = 0;
    return;
  }
  const int x0 = d->cx - c->F / 2, y0 = (d->height / k) / 2 - c->F / 2;
  dd->width = c->F * k;
  dd->height = c->F * k;
  dd->cx = c->F / 2;
  dd->row_stride = dd->width * d->channels;
  dd->image_stride = (size_t)dd->row_stride * dd->height;
  *src_off = (size_t)y0 * k * d->row_stride + (size_t)x0 * k * d->channels;
}

// n frames from `raw` (frame pitch d->image_stride) into dst (tight); d was validated by fe_args
static int fe_stage_frames(mof_ctx* c, const mof_frontend_desc* d, const mof_frontend_desc* dd, size_t src_off,
                           bool whole, const uint8_t* raw, int n, uint8_t* dst, cudaStream_t st) {
  if (whole || dd->height == d->height)   // every row of the frame is in the window: frames chain into one 2-D copy
    return h2d_frames(c, dst, raw + src_off, d->image_stride, d->row_stride, dd->row_stride, d->height, n, st);
  for (int k = 0; k < n; k++)
    CU(c, cudaMemcpy2DAsync(dst + (size_t)k * dd->image_stride, dd->row_stride,
                            raw + (size_t)k * d->image_stride + src_off, d->row_stride, dd->row_stride, dd->height,
                            cudaMemcpyHostToDevice, st));
  return MOF_OK;
}

static int fe_stage_raw(mof_ctx* c, const mof_frontend_desc* d, const uint8_t* raw, int n, cudaStream_t st,
                        mof_frontend_desc* dd) {
  size_t off;
  bool whole;
  fe_crop_desc(c, d, dd, &off, &whole);
  int rc = grow(c, &c->d_fe_raw, &c->fe_raw_cap, dd->image_stride * n + FE_SLACK);
  if (rc) return rc;
  return fe_stage_frames(c, d, dd, off, whole, raw, n, c->d_fe_raw, st);
}

int mof_frontend_device(mof_ctx* c, const mof_frontend_desc* d, const uint8_t* d_raw, int n, uint8_t* d_gray,
                        size_t gray_image_stride, int gray_row_stride, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!d_raw || !d_gray || n < 0 || gray_row_stride < c->F) return fail(c, MOF_ERR_INVALID, "mof_frontend_device: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  mof::FeArgs a;
  int rc = fe_args(c, d, n, d_raw, d_gray, gray_image_stride, gray_row_stride, &a);
  if (rc) return rc;
  return fe_run(c, a, st);
}

int mof_frontend(mof_ctx* c, const mof_frontend_desc* d, const uint8_t* raw, int n, uint8_t* gray) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!raw || !gray || n < 0) return fail(c, MOF_ERR_INVALID, "mof_frontend: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  mof::FeArgs a;
  const size_t fb = (size_t)c->F * c->F;
  int rc = fe_args(c, d, n, raw, nullptr, fb, c->F, &a);   // validates before anything is staged
  if (rc) return rc;
  mof_frontend_desc dd;
  if ((rc = fe_stage_raw(c, d, raw, n, st, &dd))) return rc;
  if ((rc = grow(c, &c->d_fe_gray, &c->fe_gray_cap, fb * n))) return rc;
  if ((rc = fe_args(c, &dd, n, c->d_fe_raw, c->d_fe_gray, fb, c->F, &a))) return rc;
  a.src_slack = FE_SLACK;
  if ((rc = fe_run(c, a, st))) return rc;
  CU(c, cudaMemcpyAsync(gray, c->d_fe_gray, fb * n, cudaMemcpyDeviceToHost, st));
  CU(c, cudaStreamSynchronize(st));
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

static int push_raw_impl(mof_ctx* c, int n, const int* ids, const mof_frontend_desc* d, const uint8_t* d_raw,
                         double* d_out, cudaStream_t st, int slack = 0) {
  const size_t fb = (size_t)c->F * c->F;
  int rc = grow(c, &c->d_fe_gray, &c->fe_gray_cap, fb * (size_t)(n > 0 ? n : 1));
  if (rc) return rc;
  mof::FeArgs a;
  if ((rc = fe_args(c, d, n, d_raw, c->d_fe_gray, fb, c->F, &a))) return rc;
  a.src_slack = slack;
  if ((rc = fe_run(c, a, st))) return rc;
  return stream_push_impl(c, n, ids, c->d_fe_gray, fb, c->F, d_out, st, mof::PC_STREAM);
}

int mof_fft_stream_push_raw_device(mof_ctx* c, int n, const int* ids, const mof_frontend_desc* d,
                                   const uint8_t* d_raw, double* d_out, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!ids || !d_raw || !d_out) return fail(c, MOF_ERR_INVALID, "mof_fft_stream_push_raw_device: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  return push_raw_impl(c, n, ids, d, d_raw, d_out, st);
}

int mof_fft_stream_push_raw(mof_ctx* c, int n, const int* ids, const mof_frontend_desc* d, const uint8_t* raw,
                            double* shifts) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!ids || !raw || !shifts || n < 0) return fail(c, MOF_ERR_INVALID, "mof_fft_stream_push_raw: bad argument");
  if (n > c->cfg.max_batch) return fail(c, MOF_ERR_INVALID, "n = %d exceeds max_batch = %d", n, c->cfg.max_batch);
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  mof::FeArgs a;
  int rc = fe_args(c, d, n, raw, nullptr, 0, c->F, &a);
  if (rc) return rc;
  mof_frontend_desc dd;
  if ((rc = fe_stage_raw(c, d, raw, n, st, &dd))) return rc;
  if ((rc = push_raw_impl(c, n, ids, &dd, c->d_fe_raw, c->d_out[0], st, FE_SLACK))) return rc;
  CU(c, cudaMemcpyAsync(shifts, c->d_out[0], sizeof(double) * 2 * c->S * n, cudaMemcpyDeviceToHost, st));
  CU(c, cudaStreamSynchronize(st));
  return MOF_OK;
}

static int push_many_raw_impl(mof_ctx* c, int n, const int* ids, int T, const mof_frontend_desc* d,
                              const uint8_t* d_raw, double* d_out, cudaStream_t st, int slack = 0,
                              cudaEvent_t raw_consumed = nullptr) {
  const size_t fb = (size_t)c->F * c->F;
  int rc = grow(c, &c->d_fe_gray, &c->fe_gray_cap, fb * (size_t)n * T);
  if (rc) return rc;
  mof::FeArgs a;
  if ((rc = fe_args(c, d, n * T, d_raw, c->d_fe_gray, fb, c->F, &a))) return rc;
  a.src_slack = slack;
  if ((rc = fe_run(c, a, st))) return rc;
  if (raw_consumed) CU(c, cudaEventRecord(raw_consumed, st));   // the raw staging may be refilled from here on
  return stream_push_many_impl(c, n, ids, T, c->d_fe_gray, (size_t)T * fb, fb, c->F, d_out, st);
}

int mof_fft_stream_push_many_raw_device(mof_ctx* c, int n, const int* ids, int T, const mof_frontend_desc* d,
                                        const uint8_t* d_raw, double* d_out, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!ids || !d_raw || !d_out || n < 0 || T < 1) return fail(c, MOF_ERR_INVALID, "mof_fft_stream_push_many_raw_device: bad argument");
  if (n == 0) return MOF_OK;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  return push_many_raw_impl(c, n, ids, T, d, d_raw, d_out, st);
}

int mof_fft_stream_push_many_raw(mof_ctx* c, int n, const int* ids, int T, const mof_frontend_desc* d,
                                 const uint8_t* raw, double* shifts) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (!ids || !raw || !shifts || n < 0 || T < 1) return fail(c, MOF_ERR_INVALID, "mof_fft_stream_push_many_raw: bad argument");
  if (n > c->cfg.max_batch) return fail(c, MOF_ERR_INVALID, "n = %d exceeds max_batch = %d", n, c->cfg.max_batch);
  if (n == 0) return MOF_OK;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  mof::FeArgs fa;
  int rc = fe_args(c, d, n, raw, nullptr, 0, c->F, &fa);   // validates before anything is staged
  if (rc) return rc;
  // The sequence is cut into chunks of tc frames per stream; chunk i + 1 crosses PCIe (copy stream, the
  // other half of the staging buffer) while chunk i runs (front end + PC_SEQ launch + result copy on
  // stream[0], in order: the stream state chains from chunk to chunk).
  mof_frontend_desc dd;
  size_t off;
  bool whole;
  fe_crop_desc(c, d, &dd, &off, &whole);
  const int Tcap = c->cfg.max_batch / n;
  int Tc = Tcap < T ? Tcap : T;
  if (T >= 16 && Tc > (T + pipe_chunks() - 1) / pipe_chunks()) Tc = (T + pipe_chunks() - 1) / pipe_chunks();
  const int n_chunks = (T + Tc - 1) / Tc;
  const size_t half = (dd.image_stride * (size_t)n * Tc + FE_SLACK + 255) & ~(size_t)255;
  if ((rc = grow(c, &c->d_fe_raw, &c->fe_raw_cap, half * (n_chunks > 1 ? 2 : 1)))) return rc;
  cudaStream_t cp = n_chunks > 1 ? c->stream[1] : st;
  if (cp != st) CU(c, cudaStreamWaitEvent(cp, c->last_done, 0));
  const size_t row = sizeof(double) * 2 * c->S;
  const bool one_3d = n > 1 && (whole || dd.height == d->height) && d->image_stride == (size_t)d->row_stride * d->height &&
                      dd.image_stride == (size_t)dd.row_stride * d->height && !getenv("MOF_FE_H2D_PER_STREAM");
  for (int i = 0, t0 = 0; t0 < T; i++, t0 += Tc) {
    const int tc = T - t0 < Tc ? T - t0 : Tc;
    const int b = i & 1;
    uint8_t* stage = c->d_fe_raw + (size_t)b * half;
    if (i >= 2) CU(c, cudaStreamWaitEvent(cp, c->fe_consumed[b], 0));
    // every row of every frame is in the window and frames follow each other row by row: the chunk of ALL
    // streams is one 3-D copy (x: the window's bytes of a row, y: the rows of tc frames, z: the streams)
    // instead of one 2-D copy per stream
    if (one_3d) {
      cudaMemcpy3DParms p3{};
      p3.srcPtr = make_cudaPitchedPtr(const_cast<uint8_t*>(raw) + (size_t)t0 * d->image_stride + off, (size_t)d->row_stride,
                                      (size_t)d->row_stride, (size_t)T * d->height);
      p3.dstPtr = make_cudaPitchedPtr(stage, (size_t)dd.row_stride, (size_t)dd.row_stride, (size_t)tc * d->height);
      p3.extent = make_cudaExtent((size_t)dd.row_stride, (size_t)tc * d->height, (size_t)n);
      p3.kind = cudaMemcpyHostToDevice;
      CU(c, cudaMemcpy3DAsync(&p3, cp));
    } else {
      for (int k = 0; k < n; k++)
        if ((rc = fe_stage_frames(c, d, &dd, off, whole, raw + ((size_t)k * T + t0) * d->image_stride, tc,
                                  stage + (size_t)k * tc * dd.image_stride, cp)))
          return rc;
    }
    if (cp != st) {
      CU(c, cudaEventRecord(c->fe_copied[b], cp));
      CU(c, cudaStreamWaitEvent(st, c->fe_copied[b], 0));
    }
    if ((rc = push_many_raw_impl(c, n, ids, tc, &dd, stage, c->d_out[0], st, FE_SLACK, c->fe_consumed[b]))) return rc;
    CU(c, cudaMemcpy2DAsync(shifts + (size_t)t0 * c->S * 2, row * T, c->d_out[0], row * tc, row * tc, n,
                            cudaMemcpyDeviceToHost, st));
  }
  CU(c, cudaStreamSynchronize(st));
  return MOF_OK;
}

// ---- long-range mode -----------------------------------------------------------------------------------
int mof_longrange_geometry(const mof_ctx* c, int* F_lr, int* S_lr) {
  if (!c) return MOF_ERR_INVALID;
  const int sq_lr = c->sq / 4;   // sqNum / LONG_RANGE_RATIO (src/FftMethod.cpp:1720)
  if (F_lr) *F_lr = c->F / 4;
  if (S_lr) *S_lr = sq_lr * sq_lr;
  return MOF_OK;
}

static int longrange_device_impl(mof_ctx* c, const uint8_t* d_curr, const uint8_t* d_prev, size_t pair_stride,
                                 int row_stride, int n, double* d_out, cudaStream_t st) {
  const int Fl = c->F / 4, sq_lr = c->sq / 4;
  if (c->F % 4) return fail(c, MOF_ERR_INVALID, "long range: frame size %d is not a multiple of 4", c->F);
  if (sq_lr == 0 || n == 0) return MOF_OK;   // the reference returns an empty vector
  const size_t fl = (size_t)Fl * Fl;
  int rc = grow(c, &c->d_lr, &c->lr_cap, 2 * fl * n);
  if (rc) return rc;
  // cv::resize(im, 1/4) of both frames (src/FftMethod.cpp:1931-1932)
  for (int h = 0; h < 2; h++) {
    mof::FeArgs a{};
    a.src = h ? d_prev : d_curr; a.src_image_stride = pair_stride; a.src_row_stride = row_stride;
    a.channels = 1; a.k = 4; a.x0 = 0; a.y0 = 0; a.out_w = Fl; a.out_h = Fl;
    a.dst = c->d_lr + (size_t)h * fl * n; a.dst_image_stride = fl; a.dst_row_stride = Fl; a.n_images = n;
    if ((rc = fe_run(c, a, st))) return rc;
  }
  mof::PcArgs a{};
  a.curr = c->d_lr; a.prev = c->d_lr + fl * n; a.frame_stride = fl; a.row_stride = Fl;
  a.sq = sq_lr; a.S = sq_lr * sq_lr; a.n_sections = n * a.S; a.win = c->d_win;
  a.scratch = c->d_scratch[0]; a.out = d_out;
  a.max_px_speed_sq = c->cfg.max_px_speed * c->cfg.max_px_speed;   // max_px_speed_lr = 1 * max_px_speed (:1687)
  a.mode = mof::PC_PAIRS; a.post = mof::PC_POST_SHIFT; a.ocl = c->cfg.ocl_compat ? 1 : 0;
  return launch_pc(c, c->N, a, c->pc_grid, st);
}

int mof_fft_pairs_longrange_device(mof_ctx* c, const uint8_t* d_curr, const uint8_t* d_prev, size_t pair_stride,
                                   int row_stride, int n_pairs, double* d_out, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (n_pairs == 0) return MOF_OK;
  if (!d_curr || !d_prev || !d_out || n_pairs < 0 || row_stride < c->F)
    return fail(c, MOF_ERR_INVALID, "mof_fft_pairs_longrange_device: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  int rc = longrange_device_impl(c, d_curr, d_prev, pair_stride, row_stride, n_pairs, d_out, st);
  if (rc) return rc;
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

int mof_fft_pairs_longrange(mof_ctx* c, const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                            int row_stride, int n_pairs, double* shifts) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (!c->fft_ready)
    return fail(c, MOF_ERR_INVALID, "section size %d unsupported by the FFT path (60, 120, 240, or any even 2^a 3^b 5^c up to 768)", c->N);
  if (n_pairs == 0) return MOF_OK;
  if (!curr || !prev || !shifts || n_pairs < 0 || row_stride < c->F)
    return fail(c, MOF_ERR_INVALID, "mof_fft_pairs_longrange: bad argument");
  CU(c, cudaSetDevice(c->device));
  const int F = c->F, mb = c->cfg.max_batch, sq_lr = c->sq / 4, S_lr = sq_lr * sq_lr;
  if (S_lr == 0) return MOF_OK;
  const size_t fbytes = (size_t)F * F;
  cudaStream_t st = c->stream[0];
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  if (c->lr_out_cap < (size_t)mb * S_lr * 2) {
    CU(c, cudaDeviceSynchronize());
    cudaFree(c->d_lr_out);
    c->d_lr_out = nullptr;
    c->lr_out_cap = 0;
    if (cudaMalloc(&c->d_lr_out, sizeof(double) * 2 * S_lr * mb) != cudaSuccess)
      return fail(c, MOF_ERR_NOMEM, "cudaMalloc failed");
    c->lr_out_cap = (size_t)mb * S_lr * 2;
  }
  for (int p0 = 0; p0 < n_pairs; p0 += mb) {
    const int n = n_pairs - p0 < mb ? n_pairs - p0 : mb;
    int rc = h2d_frames(c, c->d_curr[0], curr + (size_t)p0 * pair_stride, pair_stride, row_stride, F, F, n, st);
    if (rc) return rc;
    if ((rc = h2d_frames(c, c->d_prev[0], prev + (size_t)p0 * pair_stride, pair_stride, row_stride, F, F, n, st))) return rc;
    if ((rc = longrange_device_impl(c, c->d_curr[0], c->d_prev[0], fbytes, F, n, c->d_lr_out, st))) return rc;
    CU(c, cudaMemcpyAsync(shifts + (size_t)p0 * S_lr * 2, c->d_lr_out, sizeof(double) * 2 * S_lr * n,
                          cudaMemcpyDeviceToHost, st));
    CU(c, cudaStreamSynchronize(st));
  }
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

// ---- point pairs for getRT / get2DT ---------------------------------------------------------------------
int mof_point_pairs_device(mof_ctx* c, const double* d_shifts, int n_frames, int sq, int N, double* d_initial,
                           double* d_shifted, int* d_counts, void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (n_frames == 0) return MOF_OK;
  if (!d_shifts || !d_initial || !d_shifted || !d_counts || n_frames < 0 || sq <= 0 || N <= 0)
    return fail(c, MOF_ERR_INVALID, "mof_point_pairs_device: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  int rc = mof::pp_launch(d_shifts, n_frames, sq, N, d_initial, d_shifted, d_counts, st);
  c->launches++;
  if (rc) return fail(c, MOF_ERR_CUDA, "point-pair kernel launch failed: %s", cudaGetErrorString((cudaError_t)rc));
  return MOF_OK;
}

int mof_point_pairs(mof_ctx* c, const double* shifts, int n_frames, int sq, int N, double* initial,
                    double* shifted, int* counts) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (n_frames == 0) return MOF_OK;
  if (!shifts || !initial || !shifted || !counts || n_frames < 0 || sq <= 0 || N <= 0)
    return fail(c, MOF_ERR_INVALID, "mof_point_pairs: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream[0];
  const size_t nb = sizeof(double) * 2 * (size_t)sq * sq * n_frames;
  double *d_in = nullptr, *d_a = nullptr, *d_b = nullptr;
  int* d_cnt = nullptr;
  if (cudaMalloc(&d_in, nb) != cudaSuccess || cudaMalloc(&d_a, nb) != cudaSuccess ||
      cudaMalloc(&d_b, nb) != cudaSuccess || cudaMalloc(&d_cnt, sizeof(int) * n_frames) != cudaSuccess) {
    cudaFree(d_in); cudaFree(d_a); cudaFree(d_b); cudaFree(d_cnt);
    return fail(c, MOF_ERR_NOMEM, "cudaMalloc failed");
  }
  int rc = MOF_OK;
  cudaError_t e = cudaMemcpyAsync(d_in, shifts, nb, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) {
    int lr = mof::pp_launch(d_in, n_frames, sq, N, d_a, d_b, d_cnt, st);
    c->launches++;
    e = (cudaError_t)lr;
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(initial, d_a, nb, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(shifted, d_b, nb, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(counts, d_cnt, sizeof(int) * n_frames, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) rc = fail(c, MOF_ERR_CUDA, "mof_point_pairs: %s", cudaGetErrorString(e));
  cudaFree(d_in); cudaFree(d_a); cudaFree(d_b); cudaFree(d_cnt);
  return rc;
}

// ---- BlockMethod::Refine (src/BlockMethod.cpp:96-147) ----------------------------------------------------
static int refine_device_impl(mof_ctx* c, const uint8_t* d_curr, size_t frame_stride, int row_stride, int width,
                              int height, int n_frames, const int* d_fullpix_xy, int passes, double* d_refined_xy,
                              cudaStream_t st) {
  // (the fused variant keeps the 2x image of a frame in shared memory: no scratch)
  int rc = mof::rf_fused(width, height) ? 0 : grow(c, &c->d_rf_up, &c->rf_up_cap, (size_t)4 * width * height * n_frames + 64);   // word reads overrun
  if (rc) return rc;
  // the 2x working image belongs to the context: order this use after the previous one
  CU(c, cudaStreamWaitEvent(st, c->last_done, 0));
  mof::RfArgs a;
  a.curr = d_curr; a.frame_stride = frame_stride; a.row_stride = row_stride;
  a.w = width; a.h = height; a.n_frames = n_frames;
  a.fullpix = d_fullpix_xy; a.up = c->d_rf_up; a.out = d_refined_xy; a.passes = passes;
  int launches = 0;
  int lr = mof::rf_launch(a, c->sms, st, &launches);
  c->launches += launches;
  if (lr) return fail(c, MOF_ERR_CUDA, "refine kernel launch failed: %s", cudaGetErrorString((cudaError_t)lr));
  CU(c, cudaEventRecord(c->last_done, st));
  return MOF_OK;
}

int mof_bm_refine_device(mof_ctx* c, const uint8_t* d_curr, size_t frame_stride, int row_stride, int width,
                         int height, int n_frames, const int* d_fullpix_xy, int passes, double* d_refined_xy,
                         void* cuda_stream) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);
  if (!g.ok) return MOF_ERR_BUSY;
  if (n_frames == 0) return MOF_OK;
  if (!d_curr || !d_fullpix_xy || !d_refined_xy || n_frames < 0 || width < 2 || height < 2 ||
      row_stride < width || passes < 1 || passes > 8)
    return fail(c, MOF_ERR_INVALID, "mof_bm_refine_device: bad argument");
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream[0];
  return refine_device_impl(c, d_curr, frame_stride, row_stride, width, height, n_frames, d_fullpix_xy, passes,
                            d_refined_xy, st);
}

int mof_bm_refine(mof_ctx* c, const uint8_t* curr, size_t frame_stride, int row_stride, int width, int height,
                  int n_frames, const int* fullpix_xy, int passes, double* refined_xy) {
  if (!c) return MOF_ERR_INVALID;
  BusyGuard g(c);   // one guard for staging, kernels and the copy back
  if (!g.ok) return MOF_ERR_BUSY;
  if (n_frames == 0) return MOF_OK;
  if (!curr || !fullpix_xy || !refined_xy || n_frames < 0 || width < 2 || height < 2 || row_stride < width ||
      passes < 1 || passes > 8)
    return fail(c, MOF_ERR_INVALID, "mof_bm_refine: bad argument");
  size_t fb = (size_t)width * height;
  size_t off_xy = (fb * n_frames + 15) & ~(size_t)15;
  size_t off_out = off_xy + (((size_t)8 * n_frames + 15) & ~(size_t)15);
  size_t total = off_out + (size_t)16 * n_frames;
  CU(c, cudaSetDevice(c->device));
  int rc = grow(c, &c->d_rf_io, &c->rf_io_cap, total);
  if (rc) return rc;
  cudaStream_t st = c->stream[0];
  if (frame_stride == (size_t)row_stride * height || n_frames == 1) {
    CU(c, cudaMemcpy2DAsync(c->d_rf_io, width, curr, row_stride, width, (size_t)height * n_frames,
                            cudaMemcpyHostToDevice, st));
  } else {
    for (int f = 0; f < n_frames; f++)
      CU(c, cudaMemcpy2DAsync(c->d_rf_io + fb * f, width, curr + frame_stride * f, row_stride, width, height,
                              cudaMemcpyHostToDevice, st));
  }
  CU(c, cudaMemcpyAsync(c->d_rf_io + off_xy, fullpix_xy, (size_t)8 * n_frames, cudaMemcpyHostToDevice, st));
  rc = refine_device_impl(c, c->d_rf_io, fb, width, width, height, n_frames,
                          reinterpret_cast<const int*>(c->d_rf_io + off_xy), passes,
                          reinterpret_cast<double*>(c->d_rf_io + off_out), st);
  if (rc) return rc;
  CU(c, cudaMemcpyAsync(refined_xy, c->d_rf_io + off_out, (size_t)16 * n_frames, cudaMemcpyDeviceToHost, st));
  CU(c, cudaStreamSynchronize(st));
  return MOF_OK;
}

}  // extern "C"
