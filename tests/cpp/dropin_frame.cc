// TEST INFRASTRUCTURE.  The reference's own Frame constructors (src/Frame.cc, compiled VERBATIM from /root/reference
// with its own include/Frame.h) running on top of the PRODUCT's drop-in ORBextractor.h
// (c_orb_slam_b200/csrc/adapter/ORBextractor.h -> C ABI -> liborb_gpu.so): Frame::ExtractORB calls
// (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors) exactly as in the reference (src/Frame.cc:247-253); the
// stereo constructor runs the two extractors from two std::threads (:78-81) and then the reference's CPU
// ComputeStereoMatches on the pyramids the adapter downloaded into mvImagePyramid.  Built by tests/cpp/Makefile.
#define private public
#include "Frame.h"
#undef private

#include <cstdint>
#include <cstring>
#include <new>
#include <stdexcept>
#include <vector>

using namespace ORB_SLAM2;

namespace {
cv::Mat gray(const uint8_t* px, int w, int h) {
  cv::Mat m(h, w, CV_8U);
  std::memcpy(m.data, px, (size_t)w * h);
  return m;
}
cv::Mat calib(float fx, float fy, float cx, float cy) {
  cv::Mat K(3, 3, CV_32F);
  K.at<float>(0, 0) = fx; K.at<float>(1, 1) = fy; K.at<float>(0, 2) = cx; K.at<float>(1, 2) = cy; K.at<float>(2, 2) = 1;
  return K;
}
int export_frame(const Frame& f, OrbKeyPoint* kps, uint8_t* desc, int cap) {
  const int n = (int)f.mvKeys.size();
  if (n > cap) return -1;
  if (n) std::memcpy(kps, f.mvKeys.data(), (size_t)n * sizeof(OrbKeyPoint));
  for (int i = 0; i < n; ++i) std::memcpy(desc + (size_t)i * 32, f.mDescriptors.ptr(i), 32);
  return n;
}
}  // namespace

extern "C" {

// Frame(imGray, timeStamp, extractor, voc, K, distCoef, bf, thDepth) (src/Frame.cc:174-226), then one
// GetFeaturesInArea(x, y, r, minLevel, maxLevel) on the frame it built.  Returns the number of keypoints (or < 0).
int dropin_frame_mono(const uint8_t* img, int w, int h, int nfeatures, float scale, int nlevels, int ini_th, int min_th,
                      const float* K4, OrbKeyPoint* kps, uint8_t* desc, int cap, const float* window5,
                      int32_t* area, int* n_area, float* grid4) {
  ORBextractor ex(nfeatures, scale, nlevels, ini_th, min_th);
  ex.mbDownloadPyramid = false;
  ORBVocabulary voc;
  cv::Mat K = calib(K4[0], K4[1], K4[2], K4[3]), dist(4, 1, CV_32F);
  Frame::mbInitialComputations = true;  // image bounds and grid pitch of THIS image size (:209-224)
  Frame f(gray(img, w, h), 0.0, &ex, &voc, K, dist, 40.0f, 35.0f);
  if (ex.mLastStatus != ORB_GPU_OK) return -100 + ex.mLastStatus;
  const std::vector<size_t> v = f.GetFeaturesInArea(window5[0], window5[1], window5[2], (int)window5[3], (int)window5[4]);
  *n_area = (int)v.size();
  for (size_t i = 0; i < v.size(); ++i) area[i] = (int32_t)v[i];
  grid4[0] = Frame::mnMinX; grid4[1] = Frame::mnMinY;
  grid4[2] = Frame::mfGridElementWidthInv; grid4[3] = Frame::mfGridElementHeightInv;
  return export_frame(f, kps, desc, cap);
}

// Frame(imLeft, imRight, timeStamp, extractorLeft, extractorRight, voc, K, distCoef, bf, thDepth) (:61-117): two
// extractor threads, then ComputeStereoMatches.  u_right / depth [n] = mvuRight / mvDepth.
int dropin_frame_stereo(const uint8_t* left, const uint8_t* right, int w, int h, int nfeatures, float scale, int nlevels,
                        int ini_th, int min_th, const float* K4, float bf, OrbKeyPoint* kps, uint8_t* desc, int cap,
                        float* u_right, float* depth, OrbKeyPoint* kps_r, uint8_t* desc_r, int* n_r,
                        uint64_t* level_sums /* [2][nlevels] byte sums of mvImagePyramid, left then right */) {
  ORBextractor exl(nfeatures, scale, nlevels, ini_th, min_th), exr(nfeatures, scale, nlevels, ini_th, min_th);
  ORBVocabulary voc;
  cv::Mat K = calib(K4[0], K4[1], K4[2], K4[3]), dist(4, 1, CV_32F);
  Frame::mbInitialComputations = true;
  // The reference's stereo constructor calls ComputeStereoMatches (src/Frame.cc:90), which reads the member `mb`
  // (:496), BEFORE it assigns mb = mbf / fx (:114): the value used is whatever the object's storage held.  In the
  // reference that storage is the stack slot of the previous frame's temporary (Tracking.cc: mCurrentFrame =
  // Frame(...)), i.e. the previous frame's mb; the harness puts the same value there before constructing in place.
  alignas(Frame) static thread_local unsigned char storage[sizeof(Frame)];
  std::memset(storage, 0, sizeof(storage));
  reinterpret_cast<Frame*>(storage)->mb = bf / K4[0];
  try {
    Frame* fp = new (storage) Frame(gray(left, w, h), gray(right, w, h), 0.0, &exl, &exr, &voc, K, dist, bf, 35.0f);
    struct Guard { Frame* p; ~Guard() { p->~Frame(); } } guard{fp};
    Frame& f = *fp;
    if (exl.mLastStatus != ORB_GPU_OK || exr.mLastStatus != ORB_GPU_OK) return -100;
    const int n = export_frame(f, kps, desc, cap);
    for (int i = 0; i < n; ++i) { u_right[i] = f.mvuRight[i]; depth[i] = f.mvDepth[i]; }
    if (kps_r) {  // what the right extractor delivered, and both pyramids as ComputeStereoMatches saw them
      *n_r = (int)f.mvKeysRight.size();
      if (*n_r <= cap) {
        std::memcpy(kps_r, f.mvKeysRight.data(), (size_t)*n_r * sizeof(OrbKeyPoint));
        for (int i = 0; i < *n_r; ++i) std::memcpy(desc_r + (size_t)i * 32, f.mDescriptorsRight.ptr(i), 32);
      }
      ORBextractor* e2[2] = {&exl, &exr};
      for (int s = 0; s < 2; ++s)
        for (int l = 0; l < nlevels; ++l) {
          const cv::Mat& m = e2[s]->mvImagePyramid[l];
          uint64_t sum = 0;
          for (int y = 0; y < m.rows; ++y) for (int x = 0; x < m.cols; ++x) sum += m.ptr(y)[x] * (uint64_t)(1 + ((x + 3 * y) & 7));
          level_sums[s * nlevels + l] = sum;
        }
    }
    return n;
  } catch (const std::out_of_range&) {
    return -2;  // a correlation window left its level image (OpenCV asserts there)
  }
}

}  // extern "C"
