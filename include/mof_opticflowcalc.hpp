// mof_opticflowcalc.hpp -- header-only C++ adapters that put the B200 backend (include/mof.h,
// libmof_b200.so) behind the reference's own plugin interface, so that the nodelet
// (src/optic_flow.cpp:1001-1002 construction, :1685-1690 call) can switch backends by changing one
// `new`.  Needs the reference's include/OpticFlowCalc.h and OpenCV core headers on the include
// path; nothing here touches CUDA types.
//
//   FftMethodCuda              : OpticFlowCalc   <- FftMethod            (include/FftMethod.h:434-441)
//   FastSpacedBMMethodCuda                       <- FastSpacedBMMethod   (include/FastSpacedBMMethod_OCL.h:39-42)
//   scaleRotationEstimatorCuda                   <- scaleRotationEstimator (include/scaleRotationEstimator.h:15-16)
//   BlockMethodCuda                              <- BlockMethod          (include/BlockMethod.h:43-45)
//
// Semantics kept from the reference: first call returns zero shifts (prev := curr,
// src/FftMethod.cpp:1791-1793); invalid sections are (NaN, NaN) (:1840-1856); a re-entrant call
// returns an empty vector (:1775-1776); `gui`, `debug`, `midPoint_t`, `yaw_angle`, `rot_center`,
// `i_fx`, `i_fy` are accepted and unused exactly as in the CPU branch; `raw_output` is never
// written (the reference does not write it either).  Video dump / preview are not provided.
#ifndef MOF_OPTICFLOWCALC_HPP
#define MOF_OPTICFLOWCALC_HPP

#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "OpticFlowCalc.h"  // the reference's abstract interface, unchanged
#include "mof.h"

namespace mof_adapter {

inline void throw_on(int rc, const mof_ctx* ctx, const char* what) {
  if (rc != MOF_OK && rc != MOF_ERR_BUSY)
    throw std::runtime_error(std::string(what) + ": " + mof_last_error(ctx));
}

// the backend wants 8-bit single-channel rows; cv::Mat gives data + step for any ROI view.
// A frame larger than frameSize x frameSize is accepted and its top-left frameSize x frameSize is
// used: that is what the reference's section loop reads too (ROIs i*N, j*N with i, j < sqNum,
// src/FftMethod.cpp:1829-1836, whatever the size of the Mat the caller passes).
inline const uint8_t* gray8(const cv::Mat& m, int min_side) {
  if (m.type() != CV_8UC1 || m.rows < min_side || m.cols < min_side)
    throw std::invalid_argument("mof: expected a CV_8UC1 image of at least frameSize x frameSize");
  return m.ptr<uint8_t>(0);
}

}  // namespace mof_adapter

class FftMethodCuda : public OpticFlowCalc {
public:
  // same argument list as FftMethod::FftMethod; storeVideo/videoPath/videoFPS/cl_file_name/useOCL
  // are ignored (no video dump, no OpenCL).  `device` selects the GPU.
  FftMethodCuda(int i_frameSize, int i_samplePointSize, double max_px_speed_t, bool /*i_storeVideo*/,
                bool /*i_raw_enable*/, bool /*i_rot_corr_enable*/, bool /*i_tilt_corr_enable*/,
                std::string* /*videoPath*/, int /*videoFPS*/, std::string /*i_cl_file_name*/,
                bool /*i_useOCL*/, int device = 0) {
    mof_config cfg;
    mof_default_config(&cfg);
    cfg.device = device;
    cfg.frame_size = i_frameSize;
    cfg.sample_point_size = i_samplePointSize;
    cfg.max_px_speed = max_px_speed_t;
    cfg.n_streams = 1;
    cfg.max_batch = 1;
    cfg.sr_resolution = 0;   // no scale/rotation tables or scratch in this context
    max_px_speed_sq = max_px_speed_t * max_px_speed_t;
    int rc = mof_create(&ctx_, &cfg);
    if (rc != MOF_OK) throw std::runtime_error(std::string("mof_create: ") + mof_last_error(nullptr));
    mof_get_geometry(ctx_, &frameSize_, &samplePointSize_, &sections_);
  }
  ~FftMethodCuda() { mof_destroy(ctx_); }
  FftMethodCuda(const FftMethodCuda&) = delete;
  FftMethodCuda& operator=(const FftMethodCuda&) = delete;

  std::vector<cv::Point2d> processImage(cv::Mat imCurr_t, bool /*gui*/, bool /*debug*/,
                                        cv::Point /*midPoint_t*/, double /*yaw_angle*/,
                                        cv::Point2d /*rot_center*/,
                                        std::vector<cv::Point2d>& /*raw_output*/,
                                        double /*i_fx*/ = 300, double /*i_fy*/ = 300) override {
    // One `first` flag and one imPrev for both modes, as in the reference (src/FftMethod.cpp:1761,
    // :1791-1793, :1900).  On the first call the reference overwrites whatever setImPrev stored
    // (the nodelet seeds a black frame right after construction, src/optic_flow.cpp:1016-1018):
    // prev := curr, zero shifts.  Later, an imPrev that is not the clone this adapter made itself
    // (setImPrev was called again, or long-range mode replaced it) is re-transformed on the device.
    if (first_) {
      mof_adapter::throw_on(mof_fft_stream_reset(ctx_, 0), ctx_, "mof_fft_stream_reset");
    } else if (imPrev.empty() || imPrev.ptr<uint8_t>(0) != cached_prev_) {
      int rc0 = mof_fft_stream_set_prev(ctx_, 0, mof_adapter::gray8(imPrev, frameSize_), (int)imPrev.step);
      if (rc0 == MOF_ERR_BUSY) return std::vector<cv::Point2d>();
      mof_adapter::throw_on(rc0, ctx_, "mof_fft_stream_set_prev");
    }
    std::vector<double> xy(2 * (size_t)sections_);
    const int stream_id = 0;
    const uint8_t* px = mof_adapter::gray8(imCurr_t, frameSize_);
    int rc = mof_fft_stream_push(ctx_, 1, &stream_id, px, (size_t)imCurr_t.step * (size_t)frameSize_,
                                 (int)imCurr_t.step, xy.data());
    if (rc == MOF_ERR_BUSY) return std::vector<cv::Point2d>();  // src/FftMethod.cpp:1775-1776
    mof_adapter::throw_on(rc, ctx_, "mof_fft_stream_push");
    imPrev = imCurr_t.clone();   // src/FftMethod.cpp:1872
    cached_prev_ = imPrev.ptr<uint8_t>(0);   // the device cache now holds exactly this frame's spectrum
    first_ = false;              // :1900
    std::vector<cv::Point2d> speeds((size_t)sections_);
    for (int s = 0; s < sections_; s++) speeds[s] = cv::Point2d(xy[2 * s], xy[2 * s + 1]);
    return speeds;
  }

  // FftMethod::processImageLongRange (include/FftMethod.h:440-441, src/FftMethod.cpp:1905-1990):
  // both frames down-scaled by 4 on the device, (sqNum/4)^2 sections of samplePointSize.  The
  // previous full-resolution frame is kept as the reference keeps it (imPrev, a deep copy).
  std::vector<cv::Point2d> processImageLongRange(cv::Mat imCurr_t, bool /*gui*/, bool /*debug*/,
                                                 cv::Point /*midPoint_t*/, double /*yaw_angle*/,
                                                 cv::Point2d /*rot_center*/,
                                                 std::vector<cv::Point2d>& /*raw_output*/,
                                                 double /*i_fx*/ = 300, double /*i_fy*/ = 300) {
    int f_lr = 0, s_lr = 0;
    mof_longrange_geometry(ctx_, &f_lr, &s_lr);
    mof_adapter::gray8(imCurr_t, frameSize_);   // type / size check
    if (first_ || imPrev.empty()) imPrev = imCurr_t.clone();   // first call (:1919-1922), also after setImPrev
    std::vector<double> xy(2 * (size_t)s_lr + 2);
    cv::Mat cur_copy = imCurr_t.clone();
    // one pair; both images share the row stride of their own deep copies
    int rc = MOF_OK;
    if (s_lr > 0) {
      if (cur_copy.step != imPrev.step) throw std::invalid_argument("mof: frame geometry changed between calls");
      // pair_stride is irrelevant for n_pairs = 1
      rc = mof_fft_pairs_longrange(ctx_, cur_copy.ptr<uint8_t>(0), mof_adapter::gray8(imPrev, frameSize_), 0,
                                   (int)cur_copy.step, 1, xy.data());
    }
    if (rc == MOF_ERR_BUSY) return std::vector<cv::Point2d>();
    mof_adapter::throw_on(rc, ctx_, "mof_fft_pairs_longrange");
    imPrev = cur_copy;   // imPrev = imCurr.clone()
    cached_prev_ = nullptr;   // the cached spectrum of the normal mode no longer matches imPrev
    first_ = false;           // :2004
    std::vector<cv::Point2d> speeds((size_t)s_lr);
    for (int s = 0; s < s_lr; s++) speeds[s] = cv::Point2d(xy[2 * s], xy[2 * s + 1]);
    return speeds;
  }

  int sections() const { return sections_; }

private:
  mof_ctx* ctx_ = nullptr;
  int frameSize_ = 0, samplePointSize_ = 0, sections_ = 0;
  bool first_ = true;                     // FftMethod::first (src/FftMethod.cpp:1761)
  const uint8_t* cached_prev_ = nullptr;  // pixels of the imPrev whose spectrum the device cache holds
};

class FastSpacedBMMethodCuda {
public:
  FastSpacedBMMethodCuda(int i_samplePointSize, int i_scanRadius, int i_stepSize, int /*cx*/, int /*cy*/,
                         int /*fx*/, int /*fy*/, int /*k1*/, int /*k2*/, int /*k3*/, int /*p1*/, int /*p2*/,
                         bool /*i_storeVideo*/, std::string* /*videoPath*/, int device = 0) {
    mof_config cfg;
    mof_default_config(&cfg);
    cfg.device = device;
    cfg.bm_block = i_samplePointSize;
    cfg.bm_radius = i_scanRadius;
    cfg.bm_step = i_stepSize;
    cfg.max_batch = 1;
    cfg.sr_resolution = 0;
    int rc = mof_create(&ctx_, &cfg);
    if (rc != MOF_OK) throw std::runtime_error(std::string("mof_create: ") + mof_last_error(nullptr));
  }
  ~FastSpacedBMMethodCuda() { mof_destroy(ctx_); }
  void setImPrev(cv::Mat imPrev_t) { imPrev_ = imPrev_t.clone(); }

  // src/FastSpacedBMMethod_OCL.cpp:71-184: one voted Point2f; first call -> prev := curr
  std::vector<cv::Point2f> processImage(cv::Mat imCurr_t, bool /*gui*/, bool /*debug*/,
                                        cv::Point /*midPoint_t*/, double /*yaw_angle*/,
                                        cv::Point2d /*tiltCorr*/) {
    if (imPrev_.empty()) imPrev_ = imCurr_t.clone();
    cv::Mat c = imCurr_t.isContinuous() ? imCurr_t : imCurr_t.clone();
    int gx = 0, gy = 0;
    mof_adapter::throw_on(mof_bm_geometry(ctx_, c.cols, c.rows, &gx, &gy), ctx_, "mof_bm_geometry");
    std::vector<int8_t> flow(2 * (size_t)gx * gy);
    float voted[2] = {0.f, 0.f};
    mof_adapter::throw_on(mof_bm_pairs(ctx_, c.ptr<uint8_t>(0), imPrev_.ptr<uint8_t>(0), 0, c.cols, c.rows,
                                       (int)c.step, 1, flow.data(), voted), ctx_, "mof_bm_pairs");
    imPrev_ = c.clone();  // :181
    return std::vector<cv::Point2f>(1, cv::Point2f(voted[0], voted[1]));
  }

private:
  mof_ctx* ctx_ = nullptr;
  cv::Mat imPrev_;
};

// BlockMethod (include/BlockMethod.h:43-45, src/BlockMethod.cpp:3-94): SAD search at stride
// samplePointSize without a noise gate, histogram vote, then Refine(imCurr, imPrev, vote, 2) (:78).
// imPrev starts as a black frame (:19-20).  Not derived from OpticFlowCalc here because the
// reference class overrides nothing of it either (its processImage has a different signature).
class BlockMethodCuda {
public:
  BlockMethodCuda(int i_frameSize, int i_samplePointSize, int i_scanRadius, int /*i_scanDiameter*/,
                  int /*i_scanCount*/, int /*i_stepSize*/, int device = 0)
      : frameSize_(i_frameSize) {
    mof_config cfg;
    mof_default_config(&cfg);
    cfg.device = device;
    cfg.frame_size = i_frameSize;
    cfg.sample_point_size = i_frameSize;
    cfg.bm_block = i_samplePointSize;
    cfg.bm_radius = i_scanRadius;
    cfg.bm_step = 0;
    cfg.bm_no_gate = 1;
    cfg.max_batch = 1;
    cfg.sr_resolution = 0;
    int rc = mof_create(&ctx_, &cfg);
    if (rc != MOF_OK) throw std::runtime_error(std::string("mof_create: ") + mof_last_error(nullptr));
    imPrev_ = cv::Mat(i_frameSize, i_frameSize, CV_8UC1);
    for (int r = 0; r < i_frameSize; r++) std::memset(imPrev_.ptr<uint8_t>(r), 0, (size_t)i_frameSize);
  }
  ~BlockMethodCuda() { mof_destroy(ctx_); }

  std::vector<cv::Point2d> processImage(cv::Mat imCurr_t, bool /*gui*/, bool /*debug*/,
                                        cv::Point /*midPoint_t*/, double /*yaw_angle*/,
                                        cv::Point2d /*tiltCorr*/) {
    mof_adapter::gray8(imCurr_t, frameSize_);
    cv::Mat c = imCurr_t.isContinuous() ? imCurr_t : imCurr_t.clone();
    int gx = 0, gy = 0;
    mof_adapter::throw_on(mof_bm_geometry(ctx_, frameSize_, frameSize_, &gx, &gy), ctx_, "mof_bm_geometry");
    std::vector<int8_t> flow(2 * (size_t)gx * gy);
    float voted[2] = {0.f, 0.f};
    mof_adapter::throw_on(mof_bm_pairs(ctx_, c.ptr<uint8_t>(0), imPrev_.ptr<uint8_t>(0), 0, frameSize_, frameSize_,
                                       (int)c.step, 1, flow.data(), voted), ctx_, "mof_bm_pairs");
    const int full[2] = {(int)voted[0], (int)voted[1]};
    double refined[2] = {0.0, 0.0};
    mof_adapter::throw_on(mof_bm_refine(ctx_, c.ptr<uint8_t>(0), 0, (int)c.step, frameSize_, frameSize_, 1, full,
                                        2, refined), ctx_, "mof_bm_refine");
    imPrev_ = c.clone();  // :86
    return std::vector<cv::Point2d>(1, cv::Point2d(refined[0], refined[1]));
  }

private:
  mof_ctx* ctx_ = nullptr;
  int frameSize_ = 0;
  cv::Mat imPrev_;
};

class scaleRotationEstimatorCuda {
public:
  scaleRotationEstimatorCuda(int resolution, double m, bool /*i_storeVideo*/, std::string* /*videoPath*/,
                             int /*videoFPS*/, int device = 0) {
    mof_config cfg;
    mof_default_config(&cfg);
    cfg.device = device;
    cfg.frame_size = resolution;
    cfg.sample_point_size = resolution;
    cfg.sr_resolution = resolution;
    cfg.sr_m = m;
    cfg.max_batch = 1;
    res_ = resolution;
    int rc = mof_create(&ctx_, &cfg);
    if (rc != MOF_OK) throw std::runtime_error(std::string("mof_create: ") + mof_last_error(nullptr));
  }
  ~scaleRotationEstimatorCuda() { mof_destroy(ctx_); }

  // src/scaleRotationEstimator.cpp:34-148: Point2d(scale, rotation[rad]); first call -> (1, 0)
  cv::Point2d processImage(cv::Mat imCurr, bool /*gui*/, bool /*debug*/) {
    double sr[2] = {1.0, 0.0};
    mof_adapter::throw_on(mof_scalerot_push(ctx_, mof_adapter::gray8(imCurr, res_), (int)imCurr.step, sr),
                          ctx_, "mof_scalerot_push");
    return cv::Point2d(sr[0], sr[1]);
  }

private:
  mof_ctx* ctx_ = nullptr;
  int res_ = 0;
};

#endif  // MOF_OPTICFLOWCALC_HPP
