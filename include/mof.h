/*
 * mof.h -- C ABI of the B200 (sm_100a) backend for mrs_optic_flow's per-section image-shift
 * estimation.  This is the drop-in boundary: plain pointers and sizes, `int` status codes
 * (0 = ok), NaN in-band for invalid sections, no C++ exceptions and no torch types cross it.
 *
 * Each entry point names the reference interface it replaces (paths relative to the reference
 * repository root).  INTEGRATION.md shows the C++ adapter (`FftMethodCuda : OpticFlowCalc`) and the
 * ctypes binding that sit on top.
 *
 * Threading: one call at a time per context (the reference serialises callers the same way:
 * `mutex_process_`, src/optic_flow.cpp:1683, and the `running` flag, src/FftMethod.cpp:1775-1777).
 * A second call entering while one is running returns MOF_ERR_BUSY and writes nothing -- the C++
 * adapter turns that into the reference's empty vector.
 *
 * There is no CPU fallback: every compute entry point fails with MOF_ERR_CUDA if the CUDA device
 * is missing.
 */
#ifndef MOF_H
#define MOF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MOF_OK 0
#define MOF_ERR_INVALID 1  /* bad argument / unsupported geometry (see mof_last_error) */
#define MOF_ERR_CUDA 2     /* CUDA runtime failure */
#define MOF_ERR_BUSY 3     /* re-entrant call while the context is running */
#define MOF_ERR_NOMEM 4

typedef struct mof_ctx mof_ctx;

/* Mirrors the constructor arguments of the three reference classes:
 *   FftMethod(frameSize, samplePointSize, max_px_speed, ...)            include/FftMethod.h:434-435
 *   FastSpacedBMMethod(samplePointSize, scanRadius, stepSize, ...)      include/FastSpacedBMMethod_OCL.h:39-40
 *   scaleRotationEstimator(resolution, m, ...)                          include/scaleRotationEstimator.h:15
 * Zero-initialise, then fill what you use (mof_default_config does it with config/default.yaml). */
typedef struct mof_config {
  int device;            /* CUDA device ordinal */
  int frame_size;        /* F: frames are F x F uint8 (odd F is decremented, FftMethod.cpp:1706-1708) */
  int sample_point_size; /* N: section side; F % N != 0 forces N = F (FftMethod.cpp:1709-1716).
                            N after the fix-up: 60, 120, 240 run the prime-factor kernel (15 x {4,8,16});
                            any other even 2^a 3^b 5^c up to 768 (30, 80, 96, 160, 480 ...) the generic
                            mixed-radix kernel; other sizes: FFT entry points return MOF_ERR_INVALID */
  double max_px_speed;   /* sections with |shift| above this become NaN (FftMethod.cpp:1841) */
  int hann;              /* 1: multiply sections by cv::createHanningWindow first (default 0: the
                            reference passes no window, FftMethod.cpp:1836) */
  int max_batch;         /* capacity (frame pairs) of the internal staging for host-pointer calls */
  int n_streams;         /* number of stateful camera streams (mof_fft_stream_push); 0 = none */
  int bm_block;          /* B  = samplePointSize of FastSpacedBMMethod */
  int bm_step;           /* stepSize */
  int bm_radius;         /* R  = scanRadius, 2R+1 <= 50 (arraySize, FastSpacedBMMethod.cl:1) */
  int sr_resolution;     /* res of scaleRotationEstimator (square frames): a multiple of 4 up to 240 of the form
                            2^a 3^b 5^c (60 / 120 / 240 correlate on the prime-factor kernel, the rest on the generic one) */
  double sr_m;           /* log-polar magnitude M (config/default.yaml:13) */
  int bm_no_gate;        /* 1: skip FastSpacedBM's noise gate -- with bm_step = 0 this is the integer
                            stage of BlockMethod::processImage (src/BlockMethod.cpp:25-75): blocks
                            at n*B + R, first-minimum cv::minMaxLoc, per-axis histogram argmax */
  int ocl_compat;        /* 1: the FFT entry points reproduce the numerics of the reference's OpenCL branch
                            (useOCL = true, the deployed default, launch/optic_flow.launch:40-41) where it
                            differs from the CPU branch: P * rsqrt(|P|^2 + FLT_EPSILON) and 1/(a*b) on the four
                            real bins (cl/FftMethod.cl:976-982, :1031), surface scaled by 1/N^2 and zeroed beyond
                            SEARCH_RADIUS = 55 (:737-747, :826-829; src/FftMethod.cpp:819-820), radius-3
                            positive-only FP32 centroid of frame coordinates seeded with FLT_EPSILON
                            (:1315-1379), no maximum -> NaN ("LARGE SHIFT", src/FftMethod.cpp:461-464).
                            Default 0 = CPU branch (cv::phaseCorrelate), the parity target.  PARITY UNPINNED:
                            no OpenCL device exists on either box; the oracle is a restatement of the .cl */
  int reserved[6];
} mof_config;

/* frame_size 480, sample_point_size 120, max_px_speed 80, bm 120/24/21, sr 240/49.9
 * (config/default.yaml:13,28-32,63), max_batch 64, everything else 0. */
void mof_default_config(mof_config* cfg);

/* Replaces FftMethod::FftMethod (src/FftMethod.cpp:1680-1766): geometry fix-ups, device buffers. */
int mof_create(mof_ctx** out, const mof_config* cfg);
void mof_destroy(mof_ctx* ctx);
/* Last error text of this context (or of the failed mof_create when ctx == NULL). */
const char* mof_last_error(const mof_ctx* ctx);
/* Geometry after the constructor fix-ups: F, N, sections per frame S = (F/N)^2. */
int mof_get_geometry(const mof_ctx* ctx, int* frame_size, int* sample_point_size, int* sections);

/* ---- FftMethod::processImage, CPU-branch semantics (src/FftMethod.cpp:1772-1903, :1829-1856) ----
 * For every frame pair and every section (index i + j*sqNum, i = column tile):
 *   shift = -cv::phaseCorrelate(curr_roi, prev_roi), NaN when invalid.
 * curr/prev: HOST pointers; pair p at base + p*pair_stride, rows row_stride bytes apart.
 * shifts_xy: HOST, n_pairs * S * 2 doubles (x, y).  Blocking (returns with results on the host,
 * like k_phase_corr.run(..., sync=true), src/FftMethod.cpp:398).  n_pairs may exceed max_batch
 * (processed in chunks). */
int mof_fft_pairs(mof_ctx* ctx, const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                  int row_stride, int n_pairs, double* shifts_xy);

/* Same computation on DEVICE pointers, asynchronous on `cuda_stream` (a cudaStream_t, NULL =
 * the context's own stream); call mof_sync or synchronise the stream yourself. */
int mof_fft_pairs_device(mof_ctx* ctx, const uint8_t* d_curr, const uint8_t* d_prev,
                         size_t pair_stride, int row_stride, int n_pairs, double* d_shifts_xy,
                         void* cuda_stream);

/* Stateful form = what OpticFlowCalc::processImage does for one camera (imPrev semantics,
 * src/FftMethod.cpp:1791-1793 first call prev := curr, :1872 imPrev = imCurr.clone()).  The
 * previous frame is kept on the device as its spectrum, so each frame costs one forward transform.
 * frames: HOST, n frames, frame k belongs to stream_ids[k] (distinct ids within one call);
 * shifts_xy: HOST, n * S * 2 doubles.  Blocking. */
int mof_fft_stream_push(mof_ctx* ctx, int n, const int* stream_ids, const uint8_t* frames,
                        size_t frame_stride, int row_stride, double* shifts_xy);
/* Device-pointer, asynchronous variant (stream_ids is still a HOST array). */
int mof_fft_stream_push_device(mof_ctx* ctx, int n, const int* stream_ids, const uint8_t* d_frames,
                               size_t frame_stride, int row_stride, double* d_shifts_xy,
                               void* cuda_stream);
/* T consecutive frames per stream in ONE call and ONE launch (BASELINE config 5: 64 streams x 64 frames):
 * frame t of stream_ids[k] is at frames + k*stream_stride + t*frame_stride; the result of that frame --
 * what T successive mof_fft_stream_push calls would return, frame t against frame t-1, frame 0 against
 * the stream's cached previous frame or itself on a first call (src/FftMethod.cpp:1791-1793, :1872) --
 * lands at shifts_xy + ((k*T + t)*S)*2.  A resident CTA keeps walking the frames of one (stream, section)
 * and holds the previous spectrum on chip, so a frame costs one forward transform and nothing travels
 * through HBM between frames; long sequences are cut into time segments to fill the GPU.  Afterwards the
 * stream's previous frame is frame T-1.  Host variant: n * ceil(...) chunks of max_batch frames. */
int mof_fft_stream_push_many(mof_ctx* ctx, int n, const int* stream_ids, int T, const uint8_t* frames,
                             size_t stream_stride, size_t frame_stride, int row_stride,
                             double* shifts_xy);
int mof_fft_stream_push_many_device(mof_ctx* ctx, int n, const int* stream_ids, int T,
                                    const uint8_t* d_frames, size_t stream_stride, size_t frame_stride,
                                    int row_stride, double* d_shifts_xy, void* cuda_stream);
/* OpticFlowCalc::setImPrev (include/OpticFlowCalc.h:14-16): seed a stream's previous frame. */
int mof_fft_stream_set_prev(mof_ctx* ctx, int stream_id, const uint8_t* frame, int row_stride);
/* Forget a stream's previous frame (next push behaves like the first call). */
int mof_fft_stream_reset(mof_ctx* ctx, int stream_id);

/* ---- FastSpacedBMMethod::processImage (src/FastSpacedBMMethod_OCL.cpp:71-184) -------------------
 * SAD block matching (OptFlow_C1_D0, src/FastSpacedBMMethod.cl:4-84) + per-axis histogram vote
 * (Histogram_C1_D0, :86-169).  Frames are width x height uint8 (HOST).  Per pair:
 *   flow_xy : 2*nb int8, nb x-shifts then nb y-shifts, block index by*gx + bx,
 *             gx = (width-2R)/(B+step), gy = (height-2R)/(B+step)  (:90)
 *   voted_xy: 2 floats, the Point2f the reference returns (:172-175). */
int mof_bm_geometry(const mof_ctx* ctx, int width, int height, int* gx, int* gy);
int mof_bm_pairs(mof_ctx* ctx, const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                 int width, int height, int row_stride, int n_pairs, int8_t* flow_xy,
                 float* voted_xy);
int mof_bm_pairs_device(mof_ctx* ctx, const uint8_t* d_curr, const uint8_t* d_prev,
                        size_t pair_stride, int width, int height, int row_stride, int n_pairs,
                        int8_t* d_flow_xy, float* d_voted_xy, void* cuda_stream);

/* ---- scaleRotationEstimator::processImage (src/scaleRotationEstimator.cpp:34-148) ---------------
 * Pair form: prev goes through the first-call branch (cv::logPolar INTER_CUBIC into a zeroed
 * uint8 image, :45), curr through the later-call branch (INTER_LANCZOS4 into the same image, :112),
 * then cv::phaseCorrelate(cur, prev) (:117) -> scale_rot[2p] = exp(pt.x/M), scale_rot[2p+1] =
 * pt.y*360/res*pi/180 (:123-124); (1, 0) when rejected (:119-121).  res x res uint8 HOST frames. */
int mof_scalerot_pairs(mof_ctx* ctx, const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                       int row_stride, int n_pairs, double* scale_rot);
int mof_scalerot_pairs_device(mof_ctx* ctx, const uint8_t* d_curr, const uint8_t* d_prev,
                              size_t pair_stride, int row_stride, int n_pairs,
                              double* d_scale_rot, void* cuda_stream);
/* Stateful form = scaleRotationEstimator::processImage as the nodelet would call it for one camera
 * (src/scaleRotationEstimator.cpp:34-148): first call -> INTER_CUBIC into the zeroed persistent
 * tempIm, returns (1, 0); later calls -> INTER_LANCZOS4 into the same tempIm, correlate against the
 * previous log-polar image, which is replaced only when the estimate is accepted (:119-128).
 * frame: res x res uint8 HOST; scale_rot: 2 doubles.  Blocking. */
int mof_scalerot_push(mof_ctx* ctx, const uint8_t* frame, int row_stride, double* scale_rot);
int mof_scalerot_reset(mof_ctx* ctx);

/* The log-polar resample alone (cv::logPolar restated; interp 2 = INTER_CUBIC, 4 = INTER_LANCZOS4).
 * dst is read-modify-write: BORDER_TRANSPARENT leaves untouched pixels as they were.  HOST. */
int mof_logpolar(mof_ctx* ctx, const uint8_t* src, int src_stride, int interp, uint8_t* dst,
                 int dst_stride, int n_images, size_t src_image_stride, size_t dst_image_stride);

/* ---- camera front end (src/optic_flow.cpp:1603-1622) ---------------------------------------------
 * What the nodelet does to each camera frame before FftMethod::processImage, moved to the device:
 *   cv::resize to (width/scale_factor, height/scale_factor)  (:1603-1604; INTER_LINEAR)
 *   crop F x F at (cx - F/2, height_scaled/2 - F/2)           (:1610-1617; cx = principal point)
 *   cv::cvtColor(CV_RGB2GRAY)                                  (:1622)
 * raw: width x height frames, `channels` (1 or 3) interleaved bytes per pixel; scale_factor 1, 2
 * or 4 with width and height divisible by it (other ratios: MOF_ERR_INVALID); the crop must lie
 * inside the scaled frame.  gray: n frames of F x F bytes.  Bit-exact against cv2 4.13. */
typedef struct mof_frontend_desc {
  int width, height;     /* raw frame size in pixels */
  int channels;          /* 1 (MONO8) or 3 (BGR8 as cv_bridge delivers, src/optic_flow.cpp:1465) */
  int row_stride;        /* bytes between raw rows */
  size_t image_stride;   /* bytes between raw frames */
  int scale_factor;      /* mrs_optic_flow/scale_factor (config/default.yaml:39) */
  int cx;                /* crop centre x in the scaled frame (cx_, src/optic_flow.cpp:1611) */
} mof_frontend_desc;
int mof_frontend(mof_ctx* ctx, const mof_frontend_desc* desc, const uint8_t* raw, int n,
                 uint8_t* gray);
int mof_frontend_device(mof_ctx* ctx, const mof_frontend_desc* desc, const uint8_t* d_raw, int n,
                        uint8_t* d_gray, size_t gray_image_stride, int gray_row_stride,
                        void* cuda_stream);
/* Front end + stateful FftMethod::processImage in one call: raw camera frames in, per-section
 * shifts out (same semantics as mof_fft_stream_push).  HOST / DEVICE pointer variants. */
int mof_fft_stream_push_raw(mof_ctx* ctx, int n, const int* stream_ids,
                            const mof_frontend_desc* desc, const uint8_t* raw, double* shifts_xy);
int mof_fft_stream_push_raw_device(mof_ctx* ctx, int n, const int* stream_ids,
                                   const mof_frontend_desc* desc, const uint8_t* d_raw,
                                   double* d_shifts_xy, void* cuda_stream);

/* Raw camera frames, T per stream: frame t of stream k at raw + (k*T + t)*desc->image_stride. */
int mof_fft_stream_push_many_raw(mof_ctx* ctx, int n, const int* stream_ids, int T,
                                 const mof_frontend_desc* desc, const uint8_t* raw, double* shifts_xy);
int mof_fft_stream_push_many_raw_device(mof_ctx* ctx, int n, const int* stream_ids, int T,
                                        const mof_frontend_desc* desc, const uint8_t* d_raw,
                                        double* d_shifts_xy, void* cuda_stream);

/* ---- FftMethod::processImageLongRange, CPU branch (src/FftMethod.cpp:1905-1990) ------------------
 * Both frames are down-scaled by LONG_RANGE_RATIO = 4 (cv::resize, :1931-1932) and correlated in
 * sections of samplePointSize_lr = N (:1685) on a sqNum_lr = (F/N)/4 grid (:1720); validity uses
 * max_px_speed_lr = max_px_speed (:1687) and N/2.  S_lr = sqNum_lr^2 results per pair (0 when the
 * section grid is smaller than 4 x 4: the reference then returns an empty vector). */
int mof_longrange_geometry(const mof_ctx* ctx, int* frame_size_lr, int* sections_lr);
int mof_fft_pairs_longrange(mof_ctx* ctx, const uint8_t* curr, const uint8_t* prev,
                            size_t pair_stride, int row_stride, int n_pairs, double* shifts_xy);
int mof_fft_pairs_longrange_device(mof_ctx* ctx, const uint8_t* d_curr, const uint8_t* d_prev,
                                   size_t pair_stride, int row_stride, int n_pairs,
                                   double* d_shifts_xy, void* cuda_stream);

/* ---- point pairs for OpticFlow::getRT / get2DT (src/optic_flow.cpp:527-550, :402-416) -------------
 * Per frame, the sections whose shift is finite, in section order, as initial = section centre
 * (i*N + N/2, j*N + N/2) and shifted = initial + shift; counts[f] = how many (the caller compares it
 * with shifted_pts_thr, :552).  shifts_xy: n_frames * sq*sq * 2 doubles as the FFT entry points
 * write them; initial_xy / shifted_xy: same size, entries past counts[f] are NaN.  sq and N are
 * arguments so that the long-range grid (sqNum_lr) can be used as well. */
int mof_point_pairs(mof_ctx* ctx, const double* shifts_xy, int n_frames, int sq, int section_size,
                    double* initial_xy, double* shifted_xy, int* counts);
int mof_point_pairs_device(mof_ctx* ctx, const double* d_shifts_xy, int n_frames, int sq,
                           int section_size, double* d_initial_xy, double* d_shifted_xy,
                           int* d_counts, void* cuda_stream);

/* ---- BlockMethod::Refine (src/BlockMethod.cpp:96-147) -----------------------------------------------
 * Sub-pixel refinement of an integer flow, as the reference computes it: `passes` (2 at :78)
 * doublings of the offset, each followed by a 3x3 SAD search on the 2x up-scaled (cv::resize,
 * INTER_LINEAR) CURRENT frame against itself -- the reference up-scales imCurr into both working
 * images (:110-111), so imPrev does not enter.  curr: n_frames 8-bit gray frames width x height;
 * fullpix_xy: n_frames (x, y) ints (e.g. mof_bm_pairs' voted flow); refined_xy: n_frames (x, y)
 * doubles = offset / 2^passes.  An offset too large for the cut-out (the reference's cv::Rect would
 * throw) gives NaN. */
int mof_bm_refine(mof_ctx* ctx, const uint8_t* curr, size_t frame_stride, int row_stride, int width,
                  int height, int n_frames, const int* fullpix_xy, int passes, double* refined_xy);
int mof_bm_refine_device(mof_ctx* ctx, const uint8_t* d_curr, size_t frame_stride, int row_stride,
                         int width, int height, int n_frames, const int* d_fullpix_xy, int passes,
                         double* d_refined_xy, void* cuda_stream);

/* ---- several GPUs behind one handle (north_star (c)) -------------------------------------
 * One mof_ctx and one host worker thread per device inside this library; no NCCL, no peer access.
 * Bags of pairs are cut into contiguous slices, one per device; camera stream s lives on device
 * s % G (its previous-frame spectrum stays there).  Every device copies its results D2H straight into
 * its own disjoint slice of the caller's array.  cfg->device is ignored, cfg->n_streams is the TOTAL
 * number of streams, cfg->max_batch applies per device.  All pointers are HOST pointers; calls block. */
typedef struct mof_multi mof_multi;
int mof_multi_create(mof_multi** out, const mof_config* cfg, const int* devices, int n_devices);
void mof_multi_destroy(mof_multi* m);
int mof_multi_device_count(const mof_multi* m);
const char* mof_multi_last_error(const mof_multi* m);
int mof_multi_get_geometry(const mof_multi* m, int* frame_size, int* sample_point_size, int* sections);
int mof_multi_fft_pairs(mof_multi* m, const uint8_t* curr, const uint8_t* prev, size_t pair_stride,
                        int row_stride, int n_pairs, double* shifts_xy);
int mof_multi_fft_stream_push(mof_multi* m, int n, const int* stream_ids, const uint8_t* frames,
                              size_t frame_stride, int row_stride, double* shifts_xy);
int mof_multi_fft_stream_push_many(mof_multi* m, int n, const int* stream_ids, int T,
                                   const uint8_t* frames, size_t stream_stride, size_t frame_stride,
                                   int row_stride, double* shifts_xy);
int mof_multi_fft_stream_push_many_raw(mof_multi* m, int n, const int* stream_ids, int T,
                                       const mof_frontend_desc* desc, const uint8_t* raw,
                                       double* shifts_xy);
long long mof_multi_kernel_launches(const mof_multi* m);

/* ---- pairs without any allocation ------------------------------------------------------------------
 * No counterpart in the reference.  For callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the
 * computation of mof_fft_pairs_device -- same kernels, same results -- without a context.  The entry point
 * calls no cudaMalloc / cudaFree / allocating library routine: the per-CTA spectrum scratch of the kernel
 * is a file-scope static __device__ array of MOF_NOALLOC_SCRATCH_BYTES (36 MiB, sized for the resident
 * grid of a 148-SM B200) that every GPU loading the library holds until the process exits.
 *   frame_size / sample_point_size: the fields of the same name in struct mof_config, same fix-ups; after them the section size must be
 *     60, 120 or 240 (the prime-factor kernel); other sizes, and launches whose resident slots would need
 *     more scratch than the array (a GPU with more SMs), return MOF_ERR_INVALID with a message in `err`.
 *   d_curr / d_prev / d_shifts_xy: DEVICE pointers on `device`; n_pairs frames of frame_size^2 bytes at
 *     pair_stride / row_stride; result [n_pairs][S][2] doubles.  No Hann window, no ocl_compat.
 *   cuda_stream: a cudaStream_t, NULL = the default stream; the launch is asynchronous on it.
 *   err / err_len: optional buffer for the error text (there is no context to ask).
 * The first call on a device sets the kernels up (function attributes, cudaGetSymbolAddress, occupancy
 * queries; no allocation) -- make it outside a graph capture.  ALL calls on a device share the scratch
 * array; it keeps the previous call's contents, none of which is ever read (every slot is written before
 * it is read within a section), so nothing is cleared.  Calls that overlap on the GPU -- different streams
 * or host threads without an ordering between the launches -- race on it: serialise them. */
#define MOF_NOALLOC_SCRATCH_BYTES (36 * 1024 * 1024)
int mof_fft_pairs_noalloc_device(int device, const uint8_t* d_curr, const uint8_t* d_prev,
                                 size_t pair_stride, int row_stride, int frame_size,
                                 int sample_point_size, double max_px_speed, int n_pairs,
                                 double* d_shifts_xy, void* cuda_stream, char* err, size_t err_len);

/* Wait for everything queued on the context's own stream. */
int mof_sync(mof_ctx* ctx);

/* Instrumentation for bench.py: number of kernels this context has launched so far, and the
 * average device time [ms] of the dominant (phase-correlation) kernel over the launches since the
 * last call with reset != 0 (CUDA events recorded on the launching stream). */
long long mof_kernel_launches(const mof_ctx* ctx);
int mof_fft_kernel_time(mof_ctx* ctx, double* avg_ms, long long* n_launches, int reset);
int mof_enable_kernel_timing(mof_ctx* ctx, int on);

/* Library / build identification, e.g. "mof-b200 0.1 sm_100a". */
const char* mof_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MOF_H */
