// TEST INFRASTRUCTURE.  Implementation of the OpenCV stand-in (opencv2/core/core.hpp): every
// image primitive forwards to oracle/cvprim.{h,cc} (arithmetic pinned to cv2 4.13.0).
#include "opencv2/core/core.hpp"

#include "cvprim.h"

int cvRound(double v) { return orbo::cv_round(v); }
int cvFloor(double v) { return orbo::cv_floor(v); }
int cvCeil(double v) { return orbo::cv_ceil(v); }

namespace cv {

void resize(InputArray src_, OutputArray dst_, Size dsize, double, double, int interpolation) {
  assert(interpolation == INTER_LINEAR);
  Mat src = src_.getMat();
  dst_.create(dsize.height, dsize.width, CV_8UC1);  // no-op for a correctly sized ROI
  Mat dst = dst_.getMat();
  orbo::resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right,
                    int borderType) {
  assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
  Mat src = src_.getMat();
  dst_.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
  Mat dst = dst_.getMat();
  // src may be the centre ROI of dst (ORBextractor.cc:1122): build into a scratch copy first.
  std::vector<uchar> tmp((size_t)dst.rows * dst.cols);
  orbo::copy_make_border_reflect101_u8(src.data, src.cols, src.rows, src.step, tmp.data(), dst.cols,
                                       top, bottom, left, right);
  for (int y = 0; y < dst.rows; ++y) std::memcpy(dst.ptr(y), &tmp[(size_t)y * dst.cols], dst.cols);
}

void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY, int borderType) {
  assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
  Mat src = src_.getMat();
  dst_.create(src.rows, src.cols, CV_8UC1);
  Mat dst = dst_.getMat();
  orbo::gaussian_blur7_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.step);
}

void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
  assert(nonmaxSuppression);
  Mat img = image.getMat();
  std::vector<orbo::FastPoint> pts;
  orbo::fast9_nms(img.data, img.cols, img.rows, img.step, threshold, pts);
  keypoints.clear();
  for (const orbo::FastPoint& p : pts)
    keypoints.push_back(KeyPoint((float)p.x, (float)p.y, 7.f, -1.f, (float)p.score));
}

float fastAtan2(float y, float x) { return orbo::fast_atan2_deg(y, x); }

}  // namespace cv
