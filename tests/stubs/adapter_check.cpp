// Builds the C++ adapter against the stand-in headers and drives it through the C ABI.
// usage: adapter_check            -> expects construction to FAIL loudly when no CUDA device exists
//        adapter_check gpu        -> runs FftMethodCuda on two synthetic frames, prints the shifts
#include <cstdio>
#include <cstring>
#include "mof_opticflowcalc.hpp"

int main(int argc, char** argv) {
  const bool gpu = argc > 1 && !std::strcmp(argv[1], "gpu");
  try {
    FftMethodCuda fm(240, 120, 80.0, false, false, false, false, nullptr, 30, "", false);
    OpticFlowCalc* calc = &fm;
    cv::Mat a(240, 240, CV_8UC1), b(240, 240, CV_8UC1);
    unsigned s = 12345;
    for (int y = 0; y < 240; y++)
      for (int x = 0; x < 240; x++) { s = s * 1664525u + 1013904223u; a.ptr<uint8_t>(y)[x] = (uint8_t)(s >> 24); }
    for (int y = 0; y < 240; y++)   // b = a circularly shifted by (+3, -2)
      for (int x = 0; x < 240; x++) b.ptr<uint8_t>(y)[x] = a.ptr<uint8_t>((y + 2) % 240)[(x + 240 - 3) % 240];
    std::vector<cv::Point2d> raw;
    // the nodelet's own sequence (src/optic_flow.cpp:1016-1018): a black imPrev right after construction;
    // the reference's first processImage overwrites it with the current frame (src/FftMethod.cpp:1791-1793)
    cv::Mat black(240, 240, CV_8UC1);
    calc->setImPrev(black);
    std::vector<cv::Point2d> first = calc->processImage(a, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    std::vector<cv::Point2d> second = calc->processImage(b, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    std::printf("sections %zu\n", first.size());
    for (size_t i = 0; i < first.size(); i++) std::printf("first %.9f %.9f\n", first[i].x, first[i].y);
    for (size_t i = 0; i < second.size(); i++) std::printf("second %.9f %.9f\n", second[i].x, second[i].y);
    // setImPrev after the first call is honoured: a against b gives the opposite shift
    calc->setImPrev(b);
    std::vector<cv::Point2d> third = calc->processImage(a, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    for (size_t i = 0; i < third.size(); i++) std::printf("third %.9f %.9f\n", third[i].x, third[i].y);
    // and the frame after that is correlated against `a` again (imPrev = imCurr.clone(), :1872)
    std::vector<cv::Point2d> fourth = calc->processImage(b, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    for (size_t i = 0; i < fourth.size(); i++) std::printf("fourth %.9f %.9f\n", fourth[i].x, fourth[i].y);
    // long-range mode on a 2 x 2 section grid: sqNum / 4 = 0 -> empty result (src/FftMethod.cpp:1720, :1947)
    std::vector<cv::Point2d> lr = fm.processImageLongRange(b, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    std::printf("longrange %zu\n", lr.size());
    // 4 x 4 grid of 60 px sections -> one long-range section of the 60 px down-scaled frame
    FftMethodCuda fm60(240, 60, 80.0, false, false, false, false, nullptr, 30, "", false);
    fm60.processImageLongRange(a, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    lr = fm60.processImageLongRange(b, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    for (size_t i = 0; i < lr.size(); i++) std::printf("lr60 %.9f %.9f\n", lr[i].x, lr[i].y);
    // samplePointSize 100 does not divide 240: the constructor falls back to one 240 px section
    // (src/FftMethod.cpp:1709-1716); 80 px sections run the generic-size kernel
    FftMethodCuda fmf(240, 100, 80.0, false, false, false, false, nullptr, 30, "", false);
    fmf.processImage(a, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    std::vector<cv::Point2d> fb = fmf.processImage(b, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    for (size_t i = 0; i < fb.size(); i++) std::printf("fallback240 %.9f %.9f\n", fb[i].x, fb[i].y);
    FftMethodCuda fm80(240, 80, 80.0, false, false, false, false, nullptr, 30, "", false);
    fm80.processImage(a, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    std::vector<cv::Point2d> g80 = fm80.processImage(b, false, false, cv::Point(), 0.0, cv::Point2d(), raw);
    for (size_t i = 0; i < g80.size(); i++) std::printf("generic80 %.9f %.9f\n", g80[i].x, g80[i].y);
    // BlockMethod: black imPrev on the first call, then b against a; vote refined by Refine (:78)
    BlockMethodCuda bmm(240, 60, 8, 17, 0, 0);
    std::vector<cv::Point2d> bm1 = bmm.processImage(a, false, false, cv::Point(), 0.0, cv::Point2d());
    std::vector<cv::Point2d> bm2 = bmm.processImage(b, false, false, cv::Point(), 0.0, cv::Point2d());
    std::printf("blockmethod %.9f %.9f %.9f %.9f\n", bm1[0].x, bm1[0].y, bm2[0].x, bm2[0].y);
    scaleRotationEstimatorCuda sre(240, 49.9, false, nullptr, 30);
    cv::Point2d sr = sre.processImage(a, false, false);
    std::printf("scalerot first (%.3f, %.3f)\n", sr.x, sr.y);
    return gpu ? 0 : 3;  // on a CPU-only box construction must have thrown
  } catch (const std::exception& e) {
    std::printf("exception: %s\n", e.what());
    return gpu ? 2 : 0;
  }
}
