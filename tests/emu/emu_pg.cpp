// CPU emulator of the generic-size phase correlation (tests only): runs the host+device functions of
// mrs_optic_flow_b200/csrc/pg_generic.cuh line by line and lane by lane in the order pg_kernel uses them,
// so that the Stockham index maps, the spectrum separation and the epilogue are checked without a GPU.
#include <cmath>
#include <cstdint>
#include <vector>

#include "../../mrs_optic_flow_b200/csrc/pg_generic.cuh"

using namespace mof;
using namespace mof::pg;

template <bool INV>
static void line_fft(std::vector<c32>& buf, const std::vector<c32>& tw, const PgPlan& pl) {
  for (int k = 0; k < pl.nr; k++)
    for (int lane = 31; lane >= 0; lane--) pg_line_pass<INV>(buf.data(), tw.data(), pl, k, lane);
}

extern "C" int emu_pg_supported(int N) {
  PgPlan pl;
  return pg_plan(N, &pl) ? 1 : 0;
}

// one section: curr/prev N x N uint8 (row_stride bytes between rows) -> out_xy[2]
extern "C" int emu_pg(int N, const uint8_t* curr, const uint8_t* prev, int row_stride, double max_px_speed_sq,
                      double* out_xy) {
  PgPlan pl;
  if (!pg_plan(N, &pl)) return -1;
  const int M = pl.M, H = pg_half(M);
  std::vector<c32> tw(M);
  for (int j = 0; j < M; j++) {
    const double a = -2.0 * M_PI * (double)j / (double)M;
    tw[j] = mk((float)cos(a), (float)sin(a));
  }
  std::vector<c32> Z((size_t)M * M), buf(pg_pitch(M));
  const int res = pg_result_half(pl) * H;
  for (int y = 0; y < M; y++) {   // forward rows of z = curr + i prev
    for (int x = 0; x < M; x++) buf[x] = mk((float)curr[y * row_stride + x], (float)prev[y * row_stride + x]);
    line_fft<false>(buf, tw, pl);
    for (int x = 0; x < M; x++) Z[(size_t)y * M + x] = buf[res + x];
  }
  for (int x = 0; x < M; x++) {   // forward columns
    for (int y = 0; y < M; y++) buf[y] = Z[(size_t)y * M + x];
    line_fft<false>(buf, tw, pl);
    for (int y = 0; y < M; y++) Z[(size_t)y * M + x] = buf[res + y];
  }
  for (int ky = 0; ky < M; ky++)   // cross-power, one visit per pair (k, -k)
    for (int kx = 0; kx <= M / 2; kx++) {
      const int my = (M - ky) % M, mx = (M - kx) % M;
      if (mx == kx && my < ky) continue;
      const bool self = mx == kx && my == ky;
      const c32 c = pg_cross(Z[(size_t)ky * M + kx], Z[(size_t)my * M + mx], self);
      Z[(size_t)ky * M + kx] = c;
      if (!self) Z[(size_t)my * M + mx] = mk(c.x, -c.y);
    }
  for (int x = 0; x < M; x++) {   // inverse columns
    for (int y = 0; y < M; y++) buf[y] = Z[(size_t)y * M + x];
    line_fft<true>(buf, tw, pl);
    for (int y = 0; y < M; y++) Z[(size_t)y * M + x] = buf[res + y];
  }
  std::vector<float> surf((size_t)M * M);
  float best = -INFINITY;
  int bidx = 0x7fffffff;
  for (int y = 0; y < M; y++) {   // inverse rows, real part, running first maximum
    for (int x = 0; x < M; x++) buf[x] = Z[(size_t)y * M + x];
    line_fft<true>(buf, tw, pl);
    for (int x = 0; x < M; x++) {
      const float v = buf[res + x].x;
      surf[(size_t)y * M + x] = v;
      pg_max_update(v, pg_shifted_index(y, x, M), best, bidx);
    }
  }
  double cx = 0, cy = 0, sum = 0;
  if (bidx != 0x7fffffff) {
    const int py = bidx / M, px = bidx % M;   // shifted coordinates of the peak
    for (int t = 0; t < 25; t++) {
      const int ys = py + t / 5 - 2, xs = px + t % 5 - 2;
      if (ys < 0 || ys >= M || xs < 0 || xs >= M) continue;
      pg_acc(ys, xs, surf[(size_t)pg_unshift(ys, M) * M + pg_unshift(xs, M)], cx, cy, sum);
    }
  }
  pg_finish(cx, cy, sum, N, max_px_speed_sq, out_xy);
  return 0;
}
