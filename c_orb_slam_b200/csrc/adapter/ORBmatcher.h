// Drop-in replacement for ORB-SLAM2's include/ORBmatcher.h (class ORB_SLAM2::ORBmatcher, reference
// include/ORBmatcher.h:34-100) on top of the C ABI in include/orb_gpu.h.  Same include guard, same constructor,
// the same eleven search methods, static DescriptorDistance and the three thresholds, so Tracking.cc,
// LocalMapping.cc and LoopClosing.cc compile against it unchanged; the bodies are in adapter/ORBmatcher.cc, which
// replaces src/ORBmatcher.cc in the reference's CMakeLists.txt.  Link with liborb_gpu.so.
//
// What stays on the host is what the reference does with pointers: projecting a MapPoint with the pose in hand
// (the reference's own cv::Mat expressions), walking the two DBoW2::FeatureVector maps for equal nodes, writing
// MapPoint* into the result vectors, MapPoint::Replace / AddObservation / KeyFrame::AddMapPoint.  Candidate
// search (GetFeaturesInArea order), every Hamming distance, best / second best, the thresholds and ratio tests,
// the in-order claims and the rotation histogram run on the GPU.
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <set>
#include <utility>
#include <vector>

#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>

#include "MapPoint.h"
#include "KeyFrame.h"
#include "Frame.h"

namespace ORB_SLAM2 {

class ORBmatcher {
 public:
  ORBmatcher(float nnratio = 0.6, bool checkOri = true);

  // Hamming distance of two 256-bit descriptors (reference src/ORBmatcher.cc:1647-1663).  One pair is host work;
  // the batched device form is orb_gpu_hamming_pairs_device.
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);

  // Tracking::SearchLocalPoints (reference src/ORBmatcher.cc:45-129)
  int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3);
  // Tracking::TrackWithMotionModel (:1328-1470)
  int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);
  // Tracking::Relocalization (:1472-1599)
  int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound,
                         const float th, const int ORBdist);
  // LoopClosing::ComputeSim3 (:290-403)
  int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                         std::vector<MapPoint*>& vpMatched, int th);
  // Tracking::TrackReferenceKeyFrame / Relocalization (:159-288), LoopClosing::ComputeSim3 (:522-655)
  int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);
  int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);
  // Tracking::MonocularInitialization (:405-520)
  int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                              std::vector<int>& vnMatches12, int windowSize = 10);
  // LocalMapping::CreateNewMapPoints (:657-823)
  int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                             std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo);
  // LoopClosing::ComputeSim3 (:1102-1326)
  int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12,
                   const cv::Mat& R12, const cv::Mat& t12, const float th);
  // LocalMapping::SearchInNeighbors (:825-975), LoopClosing::SearchAndFuse (:977-1100)
  int Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th = 3.0);
  int Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
           std::vector<MapPoint*>& vpReplacePoint);

 public:
  static const int TH_LOW;
  static const int TH_HIGH;
  static const int HISTO_LENGTH;

  // ---- not in the reference: which GPU this matcher computes on (the three SLAM threads each build their own
  // ORBmatcher objects; the default for new objects is process-wide and can be changed before they are built)
  void SetDevice(int device) { mnDevice = device; }
  int GetDevice() const { return mnDevice; }
  static void SetDefaultDevice(int device);
  static int GetDefaultDevice();
  // status of the last C-ABI call of this object (ORB_GPU_OK or a negative OrbGpuStatus); the reference's methods
  // have no error channel, so a failed call reports zero matches and leaves its outputs untouched
  int LastStatus() const { return mnLastStatus; }

 protected:
  float RadiusByViewingCos(const float& viewCos);

  float mfNNratio;
  bool mbCheckOrientation;
  int mnDevice;
  int mnLastStatus;
};

}  // namespace ORB_SLAM2

#endif  // ORBMATCHER_H
