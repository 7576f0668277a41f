// CPU emulator of the twiddle-free section pipeline (tests only): runs the stage functions of
// mrs_optic_flow_b200/csrc/pc_pfa.cuh warp task by warp task, lane by lane, with a full "barrier"
// between stages, so that every index map of the CUDA kernel is verified against the oracle without
// a GPU.  `order` permutes the lane/task execution order inside a stage (0 forward, 1 reversed,
// 2 strided) to expose intra-stage dependencies that a GPU would turn into races.
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "../../mrs_optic_flow_b200/csrc/pc_pfa.cuh"

using namespace mof;
using namespace mof::pfa;

template <typename F>
static void for_lanes(int ntasks, int order, F f) {
  const int T = ntasks * 32;
  for (int i = 0; i < T; i++) {
    const int id = order == 0 ? i : (order == 1 ? T - 1 - i : (i * 37) % T);
    f(id / 32, id % 32);
  }
}

template <int P>
struct Emu {
  typedef PG<P> G;
  std::vector<c32> A;
  int row_stride;
  const float* win;
  int order;
  Emu() : A(G::A_ELEMS) {
    for (auto& v : A) v = mk(std::numeric_limits<float>::quiet_NaN(), 0.f);
  }
  static int nu_regular(int task) { return task == G::TASKS - 1 ? G::UPT - 1 : G::UPT; }

  void forward_rows(const uint8_t* img) {
    typedef LoadCfg<P> L;
    const bool vec = ((((size_t)img | (size_t)row_stride) & (size_t)(P - 1)) == 0) && order != 1;
    std::vector<uint16_t> rowy(G::N);
    for (int i = 0; i < G::N; i++) fill_rowy<P>(rowy.data(), i);
    // global -> registers (one stage ahead in the kernel) -> staging inside the task's rows of A
    std::vector<uint32_t> regs((size_t)G::TASKS * 32 * L::NREG);
    auto stage_of = [&](int task) { return reinterpret_cast<unsigned char*>(A.data() + task * G::RPT * G::SQ); };
    if (vec) {
      for_lanes(G::TASKS, order, [&](int task, int lane) {
        fx_prefetch<P>(img, row_stride, rowy.data(), task, lane,
                       *reinterpret_cast<uint32_t(*)[L::NREG]>(&regs[((size_t)task * 32 + lane) * L::NREG]));
      });
      for_lanes(G::TASKS, order, [&](int task, int lane) {
        fx_stage<P>(stage_of(task), lane,
                    *reinterpret_cast<uint32_t(*)[L::NREG]>(&regs[((size_t)task * 32 + lane) * L::NREG]));
      });
    } else {
      for_lanes(G::TASKS, order, [&](int task, int lane) {
        fx_stage_bytes<P>(stage_of(task), img, row_stride, rowy.data(), task, lane);
      });
    }
    std::vector<uint32_t> words((size_t)G::TASKS * 32 * 30);
    for_lanes(G::TASKS, order, [&](int task, int lane) {
      uint32_t* w = &words[((size_t)task * 32 + lane) * 30];
      stage_fx15_read<P>(A.data(), task, lane, *reinterpret_cast<uint32_t(*)[15]>(w),
                         *reinterpret_cast<uint32_t(*)[15]>(w + 15));
    });
    for_lanes(G::TASKS, order, [&](int task, int lane) {
      uint32_t* w = &words[((size_t)task * 32 + lane) * 30];
      stage_fx15<P>(A.data(), task, lane, *reinterpret_cast<uint32_t(*)[15]>(w),
                    *reinterpret_cast<uint32_t(*)[15]>(w + 15), win);
    });
    for_lanes(G::TASKS, order, [&](int task, int lane) { stage_x16<P, 1, false>(A.data(), task, lane, nullptr); });
  }
  void spectrum(const uint8_t* img, f4* spec) {
    forward_rows(img);
    for_lanes(G::TASKS, order, [&](int task, int lane) { stage_y16<P, false>(A.data(), task * G::UPT, G::UPT, lane); });
    for_lanes(G::TASKS, order, [&](int task, int lane) {
      stage_fy15_store<P>(A.data(), task * G::UPT, nu_regular(task), lane, spec);
      if (task == G::TASKS - 1 && lane < G::HP) fy15_store_item<P>(A.data(), G::M - 1, lane, spec);
    });
  }
  void correlate(const uint8_t* img, const f4* spec_prev, f4* spec_out, double max_sq, double* out_xy,
                 float* surf) {
    forward_rows(img);
    for_lanes(G::TASKS, order, [&](int task, int lane) { stage_y16<P, false>(A.data(), task * G::UPT, G::UPT, lane); });
    for_lanes(G::TASKS, order, [&](int task, int lane) {
      stage_fy15_corr<P>(A.data(), task * G::UPT, nu_regular(task), lane, spec_prev, spec_out);
    });
    for_lanes(1, order, [&](int, int lane) { sp_dft15<P, 1>(A.data(), lane); });
    for_lanes(1, order, [&](int, int lane) { sp_cross_power<P>(A.data(), lane, spec_prev, spec_out); });
    for_lanes(1, order, [&](int, int lane) { sp_dft15<P, -1>(A.data(), lane); });
    for_lanes(G::TASKS, order, [&](int task, int lane) { stage_y16<P, true>(A.data(), task * G::UPT, G::UPT, lane); });
    for_lanes(G::TASKS, order, [&](int task, int lane) { stage_ix15<P>(A.data(), task, lane); });
    std::vector<MaxTrack> mt(G::TASKS * 32);
    for (auto& m : mt) { m.vmax = -std::numeric_limits<float>::infinity(); m.which = 0; m.multi = false; }
    for_lanes(G::TASKS, order, [&](int task, int lane) {
      stage_x16<P, -1, true>(A.data(), task, lane, &mt[task * 32 + lane]);
    });
    float bmax = -std::numeric_limits<float>::infinity();
    for (auto& m : mt) bmax = fmaxf(bmax, m.vmax);
    int peak = 0x7fffffff;
    for_lanes(G::TASKS, order, [&](int task, int lane) {
      peak = imin(peak, search_lane<P>(A.data(), task, lane, mt[task * 32 + lane], bmax));
    });
    if (peak == 0x7fffffff) peak = 0;  // all-NaN surface: cv::minMaxLoc leaves the location at (0,0)
    double cx = 0, cy = 0, sum = 0;
    for (int tap = 0; tap < 25; tap++) centroid_tap<P>(A.data(), peak, tap, cx, cy, sum);
    centroid_finish<P>(cx, cy, sum, max_sq, out_xy);
    if (surf)
      for (int y = 0; y < G::N; y++)
        for (int x = 0; x < G::N; x++) surf[y * G::N + x] = surf_at<P>(A.data(), y, x);
  }
};

template <int P>
static int run(const uint8_t* curr, const uint8_t* prev, int row_stride, const float* win, double max_sq,
               int order, double* out_xy, float* surf, int stream_mode) {
  typedef PG<P> G;
  Emu<P> e;
  e.row_stride = row_stride;
  e.win = win;
  e.order = order;
  std::vector<f4> spec(G::SPEC_F4), spec2(G::SPEC_F4), spec3(G::SPEC_F4);
  std::memset(spec.data(), 0xff, sizeof(f4) * spec.size());
  std::memset(spec2.data(), 0, sizeof(f4) * spec2.size());
  std::memset(spec3.data(), 0, sizeof(f4) * spec3.size());
  e.spectrum(prev, spec.data());
  e.correlate(curr, spec.data(), stream_mode ? spec2.data() : nullptr, max_sq, out_xy, surf);
  if (stream_mode) {
    // the spectrum kept for the next frame must equal what a spectrum-only pass would store
    e.spectrum(curr, spec3.data());
    if (std::memcmp(spec2.data(), spec3.data(), sizeof(f4) * spec2.size()) != 0) return 7;
  }
  return 0;
}

extern "C" int emu_pfa(int N, const uint8_t* curr, const uint8_t* prev, int row_stride, const float* win,
                       double max_sq, int order, double* out_xy, float* surf, int stream_mode) {
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
