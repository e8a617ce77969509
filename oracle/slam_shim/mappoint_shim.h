// TEST INFRASTRUCTURE.  Force-included when the reference's src/MapPoint.cc is compiled VERBATIM with its own
// include/MapPoint.h (oracle/Makefile, target `refmappoint`): stand-ins for KeyFrame, Frame and Map with the members
// MapPoint.cc touches, behind the include guards of the real headers.  Used to pin
// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307, row N4).
#pragma once
#define KEYFRAME_H
#define FRAME_H
#define MAP_H
#include <climits>
#include <map>
#include <mutex>
#include <set>
#include <vector>

#include "opencv2/core/core.hpp"

using namespace std;  // leaked by the headers this one replaces; include/ORBmatcher.h relies on it

namespace ORB_SLAM2 {

class MapPoint;

class Map {
 public:
  void EraseMapPoint(MapPoint*) {}
  std::mutex mMutexPointCreation;
};

class KeyFrame {
 public:
  cv::Mat GetCameraCenter() { return Ow.clone(); }
  bool isBad() { return bad; }
  void EraseMapPointMatch(MapPoint*) {}
  void EraseMapPointMatch(const size_t&) {}
  void ReplaceMapPointMatch(const size_t&, MapPoint*) {}
  long unsigned int mnId = 0, mnFrameId = 0;
  int mnScaleLevels = 8;
  float mfLogScaleFactor = 0.18232156f;
  std::vector<float> mvScaleFactors, mvuRight;
  std::vector<cv::KeyPoint> mvKeysUn;
  cv::Mat mDescriptors, Ow;
  bool bad = false;
};

class Frame {
 public:
  cv::Mat GetCameraCenter() { return Ow.clone(); }
  long unsigned int mnId = 0;
  int mnScaleLevels = 8;
  float mfLogScaleFactor = 0.18232156f;
  std::vector<float> mvScaleFactors;
  std::vector<cv::KeyPoint> mvKeysUn;
  cv::Mat mDescriptors, Ow;
};

}  // namespace ORB_SLAM2
