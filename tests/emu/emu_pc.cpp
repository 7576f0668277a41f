// CPU emulator of the CUDA section pipeline (tests only): runs the exact item functions of
// mrs_optic_flow_b200/csrc/pc_pipeline.cuh sequentially, phase by phase, so that every index map
// of the kernel can be verified against the oracle without a GPU.  `order` permutes the item
// execution order inside each phase (0 forward, 1 reversed, 2 strided) to expose accidental
// intra-phase dependencies that a real GPU would turn into races.
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "../../mrs_optic_flow_b200/csrc/pc_pipeline.cuh"

using namespace mof;

template <int P, int PH>
static void phase_all(SecCtx<P>& c, int order) {
  typedef Geo<P> G;
  // emulate a block of T "threads", each running run_phase with its tid; the order in which the
  // threads run is permuted.
  const int T = 4 * G::LD;
  for (int i = 0; i < T; i++) {
    int tid = order == 0 ? i : (order == 1 ? T - 1 - i : (i * 37) % T);
    run_phase<P, PH>(c, tid, T);
  }
}

template <int P>
static int run(const uint8_t* curr, const uint8_t* prev, int row_stride, const float* win,
               double max_sq, int order, double* out_xy, float* surf, int stream_mode) {
  typedef Geo<P> G;
  std::vector<c32> A(G::A_ELEMS), spec((size_t)G::N * G::M), spec2((size_t)G::N * G::M);
  for (auto& v : A) v = mk(std::numeric_limits<float>::quiet_NaN(), 0.f);
  SecCtx<P> c;
  c.A = A.data();
  c.row_stride = row_stride;
  c.win = win;
  c.vmax = -std::numeric_limits<float>::infinity();
  // exercise both pixel-load paths: vector loads when the addresses allow it
  c.vec = ((((size_t)curr | (size_t)prev | (size_t)row_stride) & (P / 4 - 1)) == 0) && (order != 1);
  // previous image -> spectrum
  c.img = prev;
  c.spec_prev = nullptr;
  c.spec_out = spec.data();
  phase_all<P, PH_1A>(c, order);
  phase_all<P, PH_1B>(c, order);
  phase_all<P, PH_2>(c, order);
  phase_all<P, PH_3S>(c, order);
  // current image -> correlate
  c.img = curr;
  c.spec_prev = spec.data();
  c.spec_out = stream_mode ? spec2.data() : nullptr;
  phase_all<P, PH_1A>(c, order);
  phase_all<P, PH_1B>(c, order);
  phase_all<P, PH_2>(c, order);
  phase_all<P, PH_3C>(c, order);
  phase_all<P, PH_3Z>(c, order);
  phase_all<P, PH_3I>(c, order);
  phase_all<P, PH_4>(c, order);
  phase_all<P, PH_5A>(c, order);
  phase_all<P, PH_5B>(c, order);
  const int pidx = run_search<P>(c.A, c.vmax, 0, 1);
  centroid_shift<P>(c.A, pidx, max_sq, out_xy);
  if (surf)
    for (int y = 0; y < G::N; y++)
      for (int x = 0; x < G::N; x++) surf[y * G::N + x] = surf_at<P>(c.A, y, x);
  if (stream_mode) {
    // the spectrum kept for the next frame must equal what a spectrum-only pass would store
    c.img = curr;
    c.spec_out = spec.data();
    phase_all<P, PH_1A>(c, order);
    phase_all<P, PH_1B>(c, order);
    phase_all<P, PH_2>(c, order);
    phase_all<P, PH_3S>(c, order);
    if (std::memcmp(spec.data(), spec2.data(), sizeof(c32) * spec.size()) != 0) return 7;
  }
  return 0;
}

extern "C" int emu_pc(int N, const uint8_t* curr, const uint8_t* prev, int row_stride,
                      const float* win, double max_sq, int order, double* out_xy, float* surf,
                      int stream_mode) {
  fill_twiddles_host();
  switch (N) {
    case 60: return run<4>(curr, prev, row_stride, win, max_sq, order, out_xy, surf, stream_mode);
    case 120: return run<8>(curr, prev, row_stride, win, max_sq, order, out_xy, surf, stream_mode);
    case 240: return run<16>(curr, prev, row_stride, win, max_sq, order, out_xy, surf, stream_mode);
  }
  return -1;
}

// naive DFT check of the register butterflies
#include <cmath>
template <int R, int DIR, typename F>
static double check_dft(F f) {
  c32 x[R], y[R];
  double worst = 0;
  for (int t = 0; t < 4; t++) {
    for (int i = 0; i < R; i++) x[i] = y[i] = mk((float)std::sin(1.3 * i + t), (float)std::cos(0.7 * i * i + t));
    f(y);
    for (int k = 0; k < R; k++) {
      double re = 0, im = 0;
      for (int n = 0; n < R; n++) {
        double a = -DIR * 2.0 * 3.14159265358979323846 * n * k / R;
        re += x[n].x * std::cos(a) - x[n].y * std::sin(a);
        im += x[n].x * std::sin(a) + x[n].y * std::cos(a);
      }
      worst = std::fmax(worst, std::fmax(std::fabs(re - y[k].x), std::fabs(im - y[k].y)));
    }
  }
  return worst;
}
extern "C" double emu_check_butterflies() {
  double w = 0;
  w = std::fmax(w, check_dft<15, 1>([](c32(&z)[15]) { dft15<1>(z); }));
  w = std::fmax(w, check_dft<15, -1>([](c32(&z)[15]) { dft15<-1>(z); }));
  w = std::fmax(w, check_dft<4, 1>([](c32(&z)[4]) { DftP<4, 1>::run(z); }));
  w = std::fmax(w, check_dft<4, -1>([](c32(&z)[4]) { DftP<4, -1>::run(z); }));
  w = std::fmax(w, check_dft<8, 1>([](c32(&z)[8]) { DftP<8, 1>::run(z); }));
  w = std::fmax(w, check_dft<8, -1>([](c32(&z)[8]) { DftP<8, -1>::run(z); }));
  w = std::fmax(w, check_dft<16, 1>([](c32(&z)[16]) { DftP<16, 1>::run(z); }));
  w = std::fmax(w, check_dft<16, -1>([](c32(&z)[16]) { DftP<16, -1>::run(z); }));
  return w;
}
