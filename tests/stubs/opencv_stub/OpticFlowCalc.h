// TEST STAND-IN (not product code): the smallest cv:: surface + the abstract plugin interface that
// include/mof_opticflowcalc.hpp needs, so that the adapter can be syntax- and link-checked in an
// image that has neither OpenCV C++ headers nor ROS.  The real build uses the reference's own
// include/OpticFlowCalc.h (same virtual signature, include/OpticFlowCalc.h:9-12) and <opencv2/core.hpp>.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8UC1 0
namespace cv {
template <typename T> struct Point_ { T x, y; Point_() : x(0), y(0) {} Point_(T a, T b) : x(a), y(b) {} };
typedef Point_<int> Point; typedef Point_<int> Point2i; typedef Point_<double> Point2d; typedef Point_<float> Point2f;
struct Mat {
  int rows = 0, cols = 0; size_t step = 0; std::shared_ptr<std::vector<uint8_t>> buf; uint8_t* data = nullptr;
  Mat() {}
  Mat(int r, int c, int /*type*/) : rows(r), cols(c), step((size_t)c), buf(new std::vector<uint8_t>((size_t)r * c)) { data = buf->data(); }
  int type() const { return CV_8UC1; }
  bool empty() const { return data == nullptr; }
  bool isContinuous() const { return step == (size_t)cols; }
  template <typename T> T* ptr(int r) { return (T*)(data + (size_t)r * step); }
  template <typename T> const T* ptr(int r) const { return (const T*)(data + (size_t)r * step); }
  Mat clone() const { Mat m(rows, cols, 0); for (int r = 0; r < rows; r++) std::memcpy(m.data + (size_t)r * m.step, data + (size_t)r * step, (size_t)cols); return m; }
};
}  // namespace cv

class OpticFlowCalc {
public:
  virtual std::vector<cv::Point2d> processImage(cv::Mat imCurr, bool gui, bool debug, cv::Point midPoint_t, double yaw_angle,
                                                cv::Point2d rot_center, std::vector<cv::Point2d>& output_vectors_raw,
                                                double i_fx = 300, double i_fy = 300) = 0;
  virtual ~OpticFlowCalc() {}
  void setImPrev(cv::Mat imPrev_t) { imPrev = imPrev_t; }
protected:
  cv::Mat imPrev, imCurr, imView;
  cv::Point2i midPoint;
  double max_px_speed_sq;
};
