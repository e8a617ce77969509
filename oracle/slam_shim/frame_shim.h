// TEST INFRASTRUCTURE.  Force-included when the reference's src/Frame.cc is compiled VERBATIM (oracle/Makefile,
// target `refframe`): keeps the reference's own include/Frame.h and replaces what it drags in -- MapPoint, KeyFrame
// (slam_shim.h), ORBVocabulary, ORBextractor and Converter -- by stand-ins with the members Frame.cc touches, through
// the include guards of the real headers.  Used to pin Frame::ComputeStereoMatches (src/Frame.cc:466-640),
// AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea (:230-245, :327-392) and isInFrustum (:271-325).
#pragma once
#define SLAM_SHIM_REAL_FRAME
#define ORBVOCABULARY_H
#ifndef SLAM_SHIM_GPU_EXTRACTOR  // frame_gpu_shim.h puts the product's drop-in ORBextractor.h there instead
#define ORBEXTRACTOR_H
#endif
#define CONVERTER_H
#include <climits>  // INT_MAX reaches the reference's Frame.cc through OpenCV's headers

#include "slam_shim.h"
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"

namespace ORB_SLAM2 {

class ORBVocabulary {  // Frame::ComputeBoW only forwards to it
 public:
  void transform(const std::vector<cv::Mat>&, DBoW2::BowVector&, DBoW2::FeatureVector&, int) {}
};

#ifndef SLAM_SHIM_GPU_EXTRACTOR
class ORBextractor {  // the image pyramid and the scale tables are what ComputeStereoMatches reads
 public:
  void operator()(const cv::Mat&, const cv::Mat&, std::vector<cv::KeyPoint>&, cv::Mat&) {}
  int GetLevels() { return (int)mvScaleFactor.size(); }
  float GetScaleFactor() { return mvScaleFactor.size() > 1 ? mvScaleFactor[1] : 1.0f; }
  std::vector<float> GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }
  std::vector<cv::Mat> mvImagePyramid;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
};

#endif  // SLAM_SHIM_GPU_EXTRACTOR

class Converter {
 public:
  static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& d) {
    std::vector<cv::Mat> v;
    for (int i = 0; i < d.rows; ++i) v.push_back(d.row(i));
    return v;
  }
};

}  // namespace ORB_SLAM2
