// TEST INFRASTRUCTURE.  Stand-ins for ORB_SLAM2::Frame / KeyFrame / MapPoint with exactly the members that the
// reference's src/ORBmatcher.cc touches, so that file compiles UNMODIFIED (oracle/Makefile, target `refmatch`):
// the include guards of the reference's Frame.h / KeyFrame.h / MapPoint.h are defined here and this header is
// force-included, so `#include "Frame.h"` in include/ORBmatcher.h becomes a no-op.  The real classes drag in
// ORBextractor, ORBVocabulary, the Map, g2o types and OpenCV and cannot be built in this image.
//
// What is restated rather than verbatim: Frame / KeyFrame::GetFeaturesInArea (src/Frame.cc:327-380,
// src/KeyFrame.cc:569-608) forward to the oracle's FrameGrid; MapPoint::PredictScale returns a level the test
// chose (the adapter computes it on the host, src/MapPoint.cc:385-417).  Every GetFeaturesInArea call is
// RECORDED (target, map point last touched, x, y, r, levels): those are the windows the reference derived, and
// the tests hand them to the restated oracle / the GPU entry points as inputs.
#pragma once
#ifndef SLAM_SHIM_REAL_FRAME   // frame_shim.h keeps the reference's own Frame.h
#define FRAME_H
#endif
#define KEYFRAME_H
#define MAPPOINT_H

#include <map>
#include <set>
#include <vector>

#include "opencv2/core/core.hpp"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#include "orb_oracle.h"

using namespace std;  // leaked by the headers this one replaces; include/ORBmatcher.h relies on it

namespace ORB_SLAM2 {

class MapPoint;
class KeyFrame;
class Frame;

struct AreaCall {
  const void* target;  // Frame or KeyFrame searched
  const MapPoint* mp;  // map point whose isBad / GetWorldPos / GetDescriptor ran last
  float x, y, r;
  int min_level, max_level;
};
struct ShimLog {
  std::vector<AreaCall> calls;
  std::vector<std::pair<const MapPoint*, size_t>> map_point_reads;  // KeyFrame::GetMapPoint(idx) by the point in hand
  const MapPoint* last_mp = nullptr;
  static ShimLog& get() { static thread_local ShimLog l; return l; }
};

struct ShimGrid {  // AssignFeaturesToGrid + GetFeaturesInArea, restated in oracle/orb_oracle.cc
  orbo::FrameGrid grid;
  std::vector<orbo::KeyPoint> kps;
  void build(const std::vector<cv::KeyPoint>& k, float minX, float minY, float invW, float invH) {
    kps.resize(k.size());
    static_assert(sizeof(orbo::KeyPoint) == sizeof(cv::KeyPoint), "KeyPoint layout");
    if (!k.empty()) std::memcpy(kps.data(), k.data(), k.size() * sizeof(cv::KeyPoint));
    grid.minX = minX; grid.minY = minY; grid.invW = invW; grid.invH = invH;
    grid.build(kps.data(), (int)kps.size());
  }
  std::vector<size_t> area(const void* target, float x, float y, float r, int minLevel, int maxLevel) const {
    ShimLog& l = ShimLog::get();
    l.calls.push_back(AreaCall{target, l.last_mp, x, y, r, minLevel, maxLevel});
    std::vector<int> out;
    grid.features_in_area(kps.data(), x, y, r, minLevel, maxLevel, out);
    return std::vector<size_t>(out.begin(), out.end());
  }
};

class MapPoint {
 public:
  cv::Mat GetWorldPos() { touch(); return mWorldPos.clone(); }
  cv::Mat GetNormal() { return mNormalVector.clone(); }
  cv::Mat GetDescriptor() { touch(); return mDescriptor.clone(); }
  bool isBad() { touch(); return mbBad; }
  int Observations() { return nObs; }
  float GetMinDistanceInvariance() { return min_dist_inv; }   // 0.8f * mfMinDistance, src/MapPoint.cc:373-377
  float GetMaxDistanceInvariance() { return max_dist_inv; }   // 1.2f * mfMaxDistance, src/MapPoint.cc:379-383
  int PredictScale(const float&, KeyFrame*) { return predicted_level; }
  int PredictScale(const float&, Frame*) { return predicted_level; }
  bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) != 0; }
  int GetIndexInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) ? (int)mObservations[pKF] : -1; }
  // the map mutations of Fuse are logged, not applied (the adapter applies them on the host in the product)
  void AddObservation(KeyFrame* pKF, size_t idx) { added.push_back(std::make_pair(pKF, idx)); }
  void Replace(MapPoint* pMP) { replaced_by = pMP; }

  float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
  bool mbTrackInView = false;
  int mnTrackScaleLevel = 0;
  float mTrackViewCos = 0;

  cv::Mat mWorldPos, mNormalVector, mDescriptor;
  bool mbBad = false;
  int nObs = 1, predicted_level = 0;
  float min_dist_inv = 0, max_dist_inv = 1e9f;
  std::map<KeyFrame*, size_t> mObservations;
  std::vector<std::pair<KeyFrame*, size_t>> added;
  MapPoint* replaced_by = nullptr;

 private:
  void touch() const { ShimLog::get().last_mp = this; }
};

#ifndef SLAM_SHIM_REAL_FRAME
class Frame {
 public:
  std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                        const int maxLevel = -1) const {
    return grid.area(this, x, y, r, minLevel, maxLevel);
  }
  static float fx, fy, cx, cy, invfx, invfy;
  static float mnMinX, mnMaxX, mnMinY, mnMaxY;
  static float mfGridElementWidthInv, mfGridElementHeightInv;
  float mbf = 0, mb = 0;
  int N = 0;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
  std::vector<float> mvuRight, mvDepth;
  DBoW2::FeatureVector mFeatVec;
  cv::Mat mDescriptors;
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<bool> mvbOutlier;
  cv::Mat mTcw;
  std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  ShimGrid grid;
};

#endif  // SLAM_SHIM_REAL_FRAME

class KeyFrame {
 public:
  KeyFrame(float fx_, float fy_, float cx_, float cy_, float mbf_, float mb_, int n, int minX, int minY, int maxX,
           int maxY)
      : fx(fx_), fy(fy_), cx(cx_), cy(cy_), invfx(1.0f / fx_), invfy(1.0f / fy_), mbf(mbf_), mb(mb_), mThDepth(0),
        N(n), mnMinX(minX), mnMinY(minY), mnMaxX(maxX), mnMaxY(maxY) {}
  cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }        // src/KeyFrame.cc:109-113
  cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }             // src/KeyFrame.cc:115-119
  cv::Mat GetCameraCenter() { return Ow.clone(); }
  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
  std::set<MapPoint*> GetMapPoints() {                                               // src/KeyFrame.cc:225-238
    std::set<MapPoint*> s;
    for (MapPoint* p : mvpMapPoints) if (p && !p->isBad()) s.insert(p);
    return s;
  }
  MapPoint* GetMapPoint(const size_t& idx) {
    ShimLog& l = ShimLog::get();
    l.map_point_reads.push_back(std::make_pair(l.last_mp, idx));
    return mvpMapPoints[idx];
  }
  void AddMapPoint(MapPoint* pMP, const size_t& idx) {              // src/KeyFrame.cc:201-205
    added.push_back(std::make_pair(pMP, idx));
    mvpMapPoints[idx] = pMP;
  }
  std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const {
    return grid.area(this, x, y, r, -1, -1);
  }
  bool IsInImage(const float& x, const float& y) const {                            // src/KeyFrame.cc:610-613
    return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY);
  }
  const float fx, fy, cx, cy, invfx, invfy, mbf, mb, mThDepth;
  const int N;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
  std::vector<float> mvuRight, mvDepth;
  cv::Mat mDescriptors;
  DBoW2::FeatureVector mFeatVec;
  std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  const int mnMinX, mnMinY, mnMaxX, mnMaxY;
  float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;  // include/KeyFrame.h:132-133 (read by the drop-in matcher)
  std::vector<MapPoint*> mvpMapPoints;
  cv::Mat Tcw, Ow;
  std::vector<std::pair<MapPoint*, size_t>> added;
  ShimGrid grid;
};

}  // namespace ORB_SLAM2
