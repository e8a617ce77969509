// Drop-in replacement for ORB-SLAM2's include/ORBextractor.h (class ORB_SLAM2::ORBextractor,
// reference include/ORBextractor.h:45-111) on top of the C ABI in include/orb_gpu.h.
// Same constructor, operator(), getters and public mvImagePyramid, so Frame.cc / Tracking.cc
// compile against it unchanged; link with liborb_gpu.so.  Header-only; needs OpenCV core only
// for cv::Mat / cv::KeyPoint / InputArray / OutputArray (as the reference does).
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <cassert>
#include <cstdio>
#include <cstring>
#include <vector>

#include <opencv2/core/core.hpp>

#include "orb_gpu.h"

namespace ORB_SLAM2 {

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

  ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
      : nfeatures(nfeatures), scaleFactor(scaleFactor), nlevels(nlevels), iniThFAST(iniThFAST),
        minThFAST(minThFAST) {
    mvImagePyramid.resize(nlevels);
    // tables are needed before the first image (Frame.cc:69-75 reads them in the constructor);
    // a small handle provides them, the working handle is (re)created per image size.
    Recreate(640, 480);
  }
  ~ORBextractor() { orb_gpu_destroy(mHandle); }
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // Compute the ORB features and descriptors on an image (mask ignored, as in the reference).
  void operator()(cv::InputArray _image, cv::InputArray /*mask*/, std::vector<cv::KeyPoint>& _keypoints,
                  cv::OutputArray _descriptors) {
    if (_image.empty()) return;
    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);
    if (image.cols > mMaxW || image.rows > mMaxH) Recreate(std::max(image.cols, mMaxW), std::max(image.rows, mMaxH));
    const int capacity = ORB_GPU_CAPACITY(nfeatures, nlevels);
    static_assert(sizeof(cv::KeyPoint) == sizeof(OrbKeyPoint), "cv::KeyPoint layout");
    _keypoints.resize(capacity);
    mDesc.resize((size_t)capacity * 32);
    int n = 0;
    const int rc = orb_gpu_extract(mHandle, image.data, image.cols, image.rows, (size_t)image.step,
                                   reinterpret_cast<OrbKeyPoint*>(_keypoints.data()), mDesc.data(), capacity, &n);
    // The reference has no error channel (void).  ERR_CAPACITY means some level had more candidates than its key
    // arrays hold: the keypoints of the other levels are valid and are kept; any other failure reports no keypoints.
    // Either way the status stays in mLastStatus and the message in orb_gpu_last_error().
    if (rc != ORB_GPU_OK && rc != ORB_GPU_ERR_CAPACITY) n = 0;
    if (rc != ORB_GPU_OK) std::fprintf(stderr, "ORBextractor (GPU): status %d: %s\n", rc, orb_gpu_last_error());
    mLastStatus = rc;
    _keypoints.resize(n);
    if (n == 0) {
      _descriptors.release();
    } else {
      _descriptors.create(n, 32, CV_8U);
      cv::Mat d = _descriptors.getMat();
      for (int i = 0; i < n; ++i) std::memcpy(d.ptr(i), &mDesc[(size_t)i * 32], 32);
    }
    if (mbDownloadPyramid && (rc == ORB_GPU_OK || rc == ORB_GPU_ERR_CAPACITY)) {
      for (int l = 0; l < nlevels; ++l) {
        int w = 0, h = 0;
        orb_gpu_get_pyramid(mHandle, 0, l, 0, nullptr, 0, &w, &h);
        mvImagePyramid[l].create(h, w, CV_8UC1);
        orb_gpu_get_pyramid(mHandle, 0, l, 0, mvImagePyramid[l].data, (size_t)mvImagePyramid[l].step, &w, &h);
      }
    }
  }

  int inline GetLevels() { return nlevels; }
  float inline GetScaleFactor() { return (float)scaleFactor; }
  std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

  std::vector<cv::Mat> mvImagePyramid;  // read by Frame::ComputeStereoMatches (Frame.cc:473,563,575,580)
  bool mbDownloadPyramid = true;        // monocular callers can switch the D2H of the pyramid off
  int mLastStatus = ORB_GPU_OK;
  OrbGpuExtractor* Handle() { return mHandle; }
  // ---- not in the reference: which GPU this extractor computes on.  New objects take the process-wide default (0
  // unless changed); SetDevice moves an existing object (its handle is rebuilt on the new device).
  static int& DefaultDevice() { static int d = 0; return d; }
  void SetDevice(int device) { if (device != mDevice) { mDevice = device; Recreate(mMaxW, mMaxH); } }
  int GetDevice() const { return mDevice; }

 protected:
  void Recreate(int w, int h) {
    if (mHandle) orb_gpu_destroy(mHandle);
    mHandle = nullptr;
    mLastStatus = orb_gpu_create(nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST, w, h, 1, mDevice, &mHandle);
    mMaxW = w; mMaxH = h;
    mvScaleFactor.assign(nlevels, 0.f); mvInvScaleFactor.assign(nlevels, 0.f);
    mvLevelSigma2.assign(nlevels, 0.f); mvInvLevelSigma2.assign(nlevels, 0.f);
    mnFeaturesPerLevel.assign(nlevels, 0);
    if (mHandle) {
      orb_gpu_get_scale_tables(mHandle, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                               mvInvLevelSigma2.data());
      orb_gpu_get_features_per_level(mHandle, mnFeaturesPerLevel.data(), nullptr);
    }
  }

  int nfeatures;
  double scaleFactor;
  int nlevels;
  int iniThFAST;
  int minThFAST;
  std::vector<int> mnFeaturesPerLevel;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
  OrbGpuExtractor* mHandle = nullptr;
  int mDevice = DefaultDevice();
  int mMaxW = 0, mMaxH = 0;
  std::vector<uint8_t> mDesc;
};

}  // namespace ORB_SLAM2
#endif  // ORBEXTRACTOR_H
