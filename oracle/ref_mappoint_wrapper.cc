// TEST INFRASTRUCTURE.  C entry point around the reference's src/MapPoint.cc compiled VERBATIM with its own
// include/MapPoint.h over oracle/slam_shim/mappoint_shim.h.  Built by oracle/Makefile (target `refmappoint`) into
// oracle/_ref/libref_mappoint.so.  Pins the restated distinctive_descriptor of orb_oracle.cc (row N4).
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "MapPoint.h"
#include "ORBmatcher.h"
#include "orb_oracle.h"

using namespace ORB_SLAM2;

// The reference's src/ORBmatcher.cc needs the full Frame / KeyFrame stand-ins and is built into libref_matcher.so,
// where tests/test_ref_matcher.py shows its DescriptorDistance equal to the oracle's; this library takes the oracle's.
int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
  return orbo::descriptor_distance(a.ptr<uint8_t>(), b.ptr<uint8_t>());
}

extern "C" {

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307) for `npoints` map points: point p is observed by
// key frames offsets[p] .. offsets[p+1]-1 (one descriptor each; bad[k]: that key frame isBad()).  best[p] = index
// (within the point's observations, in std::map<KeyFrame*> order = the order given) of the descriptor it kept, -1 if
// it kept none.
void refp_distinctive_descriptors(const uint8_t* desc, const uint8_t* bad, const int32_t* offsets, int npoints,
                                  int32_t* best) {
  Map map;
  for (int p = 0; p < npoints; ++p) {
    const int n = offsets[p + 1] - offsets[p];
    // key frames in one array so that their addresses (the std::map order of mObservations) ascend with the index
    std::vector<KeyFrame> kfs((size_t)n);
    KeyFrame ref;
    const float origin[3] = {0, 0, 0};
    cv::Mat pos(3, 1, CV_32F);
    std::memcpy(pos.data, origin, sizeof(origin));
    MapPoint mp(pos, &ref, &map);
    for (int k = 0; k < n; ++k) {
      kfs[k].mDescriptors = cv::Mat(1, 32, CV_8U);
      std::memcpy(kfs[k].mDescriptors.data, desc + (size_t)(offsets[p] + k) * 32, 32);
      kfs[k].bad = bad && bad[offsets[p] + k];
      kfs[k].mvuRight.assign(1, -1.0f);
      mp.AddObservation(&kfs[k], 0);
    }
    mp.ComputeDistinctiveDescriptors();
    cv::Mat d = mp.GetDescriptor();
    best[p] = -1;
    if (!d.empty())
      for (int k = 0; k < n && best[p] < 0; ++k)
        if (!kfs[k].bad && std::memcmp(d.data, kfs[k].mDescriptors.data, 32) == 0) best[p] = k;
  }
}

}  // extern "C"
