// TEST INFRASTRUCTURE (CPU oracle) -- not part of the shipped product path.
//
// CPU restatement of the six OpenCV primitives the reference's ORB extractor calls
// (/root/reference/src/ORBextractor.cc:103 fastAtan2, :809/:814 FAST, :1086 GaussianBlur,
//  :1120 resize, :1122/:1127 copyMakeBorder, cvRound/cvFloor/cvCeil throughout).
// OpenCV is NOT vendored in the reference and its version is not pinned there
// (CMakeLists.txt:33-39), so this arithmetic is pinned to cv2 4.13.0 -- the only OpenCV in
// the build container -- and checked bit-for-bit against it by tests/test_oracle_vs_cv2.py
// (run in the container) and against the committed fixtures under tests/golden/.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace orbo {

// cvRound: round-half-to-even (SSE cvtsd2si / lrint in OpenCV).
int cv_round(double v);
int cv_floor(double v);
int cv_ceil(double v);

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) on CV_8UC1 (fixed point, 11-bit coefficients).
void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride,
                      uint8_t* dst, int dw, int dh, size_t dstride);

// Per-axis coefficient table of the resize above (exposed so host code of the product can be
// checked against the same table in tests).  ofs[d] = first source tap, a0/a1 = 11-bit weights.
void resize_axis_table(int S, int D, int* ofs, int16_t* a0, int16_t* a1);

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) on CV_8UC1.
// src and dst may alias.
void gaussian_blur7_u8(const uint8_t* src, int w, int h, size_t sstride,
                       uint8_t* dst, size_t dstride);

// cv::copyMakeBorder(..., BORDER_REFLECT_101) on CV_8UC1: dst is (w+l+r) x (h+t+b).
void copy_make_border_reflect101_u8(const uint8_t* src, int w, int h, size_t sstride,
                                    uint8_t* dst, size_t dstride, int top, int bottom,
                                    int left, int right);

struct FastPoint { int x, y, score; };
// cv::FAST(img, kps, threshold, /*nonmaxSuppression=*/true), FAST-9/16, row-major emission.
// score == KeyPoint::response (largest threshold at which the pixel is still a corner).
void fast9_nms(const uint8_t* img, int w, int h, size_t stride, int threshold,
               std::vector<FastPoint>& out);
// threshold-independent arc score: max over the 16 contiguous 9-arcs of
// max(min_arc(d), -max_arc(d)); pixel is a corner at t iff value > t.
int fast9_best(const uint8_t* p, size_t stride);

// cv::fastAtan2(y, x): degrees in [0,360), 7th-order odd polynomial.
float fast_atan2_deg(float y, float x);

}  // namespace orbo
