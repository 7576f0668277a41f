// mof_internal.h -- launch interfaces between the C-ABI layer (mof_capi.cu) and the kernel
// translation units.  Not part of the public boundary (that is include/mof.h).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace mof {

struct c32;

// ---- phase correlation (pc_pfa_kernel.cu) ---------------------------------------------------------
enum PcMode {
  PC_PAIRS = 0,     // prev and curr images given; spectrum of prev goes through per-CTA scratch
  PC_STREAM = 1,    // curr image + cached spectrum of the stream's previous frame (double buffered)
  PC_SPECTRUM = 2,  // spectrum of `curr` only -> cache (OpticFlowCalc::setImPrev)
  PC_SEQ = 3        // T consecutive frames per stream in one launch (mof_fft_stream_push_many): a work
                    // item = (stream, section, time segment); the resident slot that owns it walks the
                    // segment's frames in order and keeps the previous spectrum in its own scratch
};
enum PcPost {
  PC_POST_SHIFT = 0,     // FftMethod: shift = -phaseCorrelate, NaN validation
  PC_POST_SCALEROT = 1   // scaleRotationEstimator: (exp(pt.x/M), pt.y*360/res*pi/180), reject -> (1,0)
};

struct PcArgs {
  const uint8_t* curr;
  const uint8_t* prev;        // PC_PAIRS only
  size_t frame_stride;        // bytes between consecutive frames
  int row_stride;             // bytes between rows
  int sq;                     // sections per frame side
  int S;                      // sq*sq
  int n_sections;             // frames * S
  const float* win;           // N*N window or nullptr
  void* scratch;              // PC_PAIRS: gridDim.x section spectra (pc_spectrum_bytes each)
  void* spec_cache;           // PC_STREAM / PC_SPECTRUM: slots * 2 section spectra (double buffered)
  const int* stream_ids;      // per frame (device)
  const uint8_t* first;       // per frame (device): bit 0 = first call of that stream, bit 1 = which
                              // half of the stream's cache holds the previous frame's spectrum
  double* out;                // n_sections * 2
  double max_px_speed_sq;
  double sr_m;                // PC_POST_SCALEROT
  int mode;
  int post;
  int sr_reject_nan;          // PC_POST_SCALEROT: write (NaN, NaN) instead of (1, 0) on reject
  uint32_t k23;               // 0x4B000000 (2^23 as float bits); set by pc_launch, see pc_pfa.cuh
  int ocl;                    // OCL-compatible numerics (cl/FftMethod.cl): radius-55 mask, 7x7 positive-only
                              // FP32 centroid, rsqrt normalisation, no maximum -> NaN
  // PC_SEQ: a (stream, section) sequence of seq_T frames = seq_J segments of seq_L frames + a remainder piece of
  // seq_R < seq_L frames.  Work items (n_sections of them, seq_L steps each): the seq_nseq * seq_J segments first,
  // then bundles of seq_Q remainder pieces (of seq_Q different sequences) walked one after the other.
  int seq_T;                  // frames per stream in this launch
  int seq_L;                  // frames per time segment
  int seq_J;                  // whole segments per (stream, section): seq_T / seq_L
  int seq_R;                  // frames of the remainder piece: seq_T - seq_J * seq_L (0: none)
  int seq_Q;                  // remainder pieces per bundle (seq_Q * seq_R <= seq_L)
  int seq_nseq;               // sequences = streams * S
  int seq_C;                  // > 0: the "tape" cut instead -- all sequences laid end to end (sequence ks, frame t at
                              // ks * seq_T + t), slot i walks frames [i * seq_C, (i + 1) * seq_C) in seq_C steps;
                              // n_sections = slots in use, seq_L = seq_C
  size_t stream_stride;       // bytes between the first frames of consecutive streams
  void* seq_scratch;          // resident slots * 2 section spectra (previous / current, ping-pong)
  // stream_ids == nullptr in a stream mode: ids and flags of up to PC_INLINE frames travel in the kernel
  // parameters (no metadata copy in front of the launch: the camera-callback case of a few streams per call)
  static constexpr int PC_INLINE = 64;
  int inline_ids[PC_INLINE];
  uint8_t inline_flags[PC_INLINE];
};

// One-time per-device setup (twiddle upload, shared-memory opt-in).  Returns cudaError_t as int.
int pc_setup(int device);
// Resident CTAs the persistent grid should use for section size N on this device.
int pc_resident_ctas(int N, int device);
size_t pc_spectrum_bytes(int N);   // bytes of one section spectrum in the kernel's own layout
int pc_launch(int N, const PcArgs& args, int grid, cudaStream_t stream);

// ---- generic section sizes (pg_kernels.cu): even N = 2^a 3^b 5^c <= 768 outside the plan above ----------
bool pg_supported(int N);
int pg_resident_ctas(int N, int device);          // also opts the kernel into its shared memory
size_t pg_scratch_bytes_per_cta(int N);           // N * N complex floats
void pg_twiddles(int N, float* out);              // 2 * N floats, exp(-2 pi i j / N)
// PC_PAIRS, or PC_STREAM with ids / first flags inline (<= PcArgs::PC_INLINE frames) and the previous FRAMES in
// `cache` (F x F bytes per stream); out_fs = sections between the results of consecutive frames
int pg_launch(int N, const PcArgs& args, const float* d_tw, void* scratch, const uint8_t* cache, int out_fs, int grid,
              cudaStream_t stream);
int pg_cache_update(const PcArgs& args, int n_frames, uint8_t* cache, int F, cudaStream_t stream);

// ---- FastSpacedBM (bm_kernels.cu) ---------------------------------------------------------------
struct BmArgs {
  const uint8_t* curr;
  const uint8_t* prev;
  size_t pair_stride;
  int row_stride;
  int width, height;
  int B, step, R;
  int gx, gy;
  int n_pairs;
  int8_t* flow_xy;   // n_pairs * 2 * nb
  float* voted_xy;   // n_pairs * 2
  int no_gate;       // 1: no noise gate (BlockMethod semantics)
  int pair0;         // first pair of this launch inside the call (TMA coordinates address the whole call)
};
int bm_setup(int device);
int bm_launch(const BmArgs& args, cudaStream_t stream, int* launches);

// ---- log-polar resample (lp_kernels.cu) ---------------------------------------------------------
struct LpMap {
  // per output pixel: ((iy + 4) * (res + 8) + ix + 4) << 10 | fxy -- byte offset of the centre tap in
  // the kernel's padded source image and the 1/32-px fraction index; 0xffffffff = outside the source
  const uint32_t* pk;
  const int16_t* tab_cubic;    // 32*32*16
  const int16_t* tab_lanczos;  // 32*32*64
  int res;
};
struct LpArgs {
  const uint8_t* src;
  size_t src_image_stride;
  int src_row_stride;
  uint8_t* dst;
  size_t dst_image_stride;
  int dst_row_stride;
  int n_images;
  int interp;  // 2 cubic, 4 lanczos4
  // BORDER_TRANSPARENT pixels: left untouched when bg == nullptr, else copied from bg (the image the
  // reference's persistent tempIm held before this remap, src/scaleRotationEstimator.cpp:27, :112)
  const uint8_t* bg;
  size_t bg_image_stride;
  int bg_row_stride;
};
int lp_setup(int device);
int lp_launch(const LpMap& map, const LpArgs& args, cudaStream_t stream);

// ---- camera front end (fe_kernels.cu) -------------------------------------------------------------
struct FeArgs {
  const uint8_t* src;        // n raw frames, channels interleaved
  size_t src_image_stride;
  int src_row_stride;
  int channels;              // 1 or 3
  int k;                     // integer down-scale factor: 1, 2 or 4
  int x0, y0;                // crop origin in the scaled image
  int out_w, out_h;
  uint8_t* dst;              // n gray frames
  size_t dst_image_stride;
  int dst_row_stride;
  int n_images;
  int src_slack;             // bytes known to be readable after the last source image (0 if unknown)
};
int fe_launch(const FeArgs& args, int sms, cudaStream_t stream);
// point pairs for getRT / get2DT (fe_kernels.cu)
int pp_launch(const double* shifts, int n_frames, int sq, int N, double* initial, double* shifted, int* counts,
              cudaStream_t stream);

// ---- BlockMethod::Refine (rf_kernels.cu) ----------------------------------------------------------
struct RfArgs {
  const uint8_t* curr;     // n frames, w x h, 8-bit gray
  size_t frame_stride;
  int row_stride;
  int w, h;
  int n_frames;
  const int* fullpix;      // device: (x, y) integer flow per frame
  uint8_t* up;             // device scratch: n x (2h) x (2w) bytes
  double* out;             // device: refined (x, y) per frame
  int passes;              // 2 in the reference (src/BlockMethod.cpp:78)
};
int rf_setup(int device);
bool rf_fused(int w, int h);
int rf_launch(const RfArgs& args, int sms, cudaStream_t stream, int* launches);

}  // namespace mof
