// TEST INFRASTRUCTURE.  C entry points around the reference's src/Frame.cc compiled VERBATIM from /root/reference
// over oracle/slam_shim/frame_shim.h (the reference's own include/Frame.h; stand-ins for what it includes).
// Built by oracle/Makefile (target `refframe`) into oracle/_ref/libref_frame.so.  Pins the restated
// Frame::ComputeStereoMatches (stereo_match + stereo_refine of orb_oracle.cc) and the grid functions.
#define private public  // Frame::AssignFeaturesToGrid is private; access specifiers do not change the layout
#include "Frame.h"
#undef private

#include <chrono>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <vector>

#include "ORBmatcher.h"

using namespace ORB_SLAM2;

// ORBmatcher::DescriptorDistance and TH_HIGH / TH_LOW, which ComputeStereoMatches uses (:471, :541), come from the
// reference's src/ORBmatcher.cc compiled into this library as well (against the real Frame.h here).

namespace {
std::vector<cv::KeyPoint> kp_vec(const orbo::KeyPoint* k, int n) {
  std::vector<cv::KeyPoint> v((size_t)n);
  if (n) std::memcpy(v.data(), k, (size_t)n * sizeof(cv::KeyPoint));
  return v;
}
cv::Mat bytes_mat(const uint8_t* d, int rows, int cols) {
  cv::Mat m(rows, cols, CV_8U);
  if (rows && cols) std::memcpy(m.data, d, (size_t)rows * cols);
  return m;
}
void fill_extractor(ORBextractor& e, const uint8_t* const* levels, const int32_t* sizes, int nlevels, const float* sf,
                    const float* inv_sf) {
  for (int l = 0; l < nlevels; ++l) e.mvImagePyramid.push_back(bytes_mat(levels[l], sizes[2 * l + 1], sizes[2 * l]));
  e.mvScaleFactor.assign(sf, sf + nlevels);
  e.mvInvScaleFactor.assign(inv_sf, inv_sf + nlevels);
}
}  // namespace

static thread_local double g_last_call_seconds = 0.0;  // inside Frame::ComputeStereoMatches alone (bench.py's CPU baseline)

extern "C" {

double reff_last_call_seconds() { return g_last_call_seconds; }

// Frame::ComputeStereoMatches (src/Frame.cc:466-640) on given keypoints, descriptors and unblurred pyramids
// (sizes[2l], sizes[2l+1] = width, height of level l).  u_right / depth [nl] = mvuRight / mvDepth.
// Returns 0, or -1 when a correlation window left its level image (OpenCV would assert in rowRange / colRange).
int reff_compute_stereo_matches(const orbo::KeyPoint* kl, const uint8_t* dl, int nl, const orbo::KeyPoint* kr,
                                const uint8_t* dr, int nr, const uint8_t* const* levels_l,
                                const uint8_t* const* levels_r, const int32_t* sizes_l, const int32_t* sizes_r,
                                int nlevels, const float* sf, const float* inv_sf, float mb, float mbf, float* u_right,
                                float* depth) {
  ORBextractor el, er;
  fill_extractor(el, levels_l, sizes_l, nlevels, sf, inv_sf);
  fill_extractor(er, levels_r, sizes_r, nlevels, sf, inv_sf);
  Frame f;
  f.mpORBextractorLeft = &el;
  f.mpORBextractorRight = &er;
  f.N = nl;
  f.mvKeys = kp_vec(kl, nl);
  f.mvKeysRight = kp_vec(kr, nr);
  f.mDescriptors = bytes_mat(dl, nl, 32);
  f.mDescriptorsRight = bytes_mat(dr, nr, 32);
  f.mvScaleFactors.assign(sf, sf + nlevels);
  f.mvInvScaleFactors.assign(inv_sf, inv_sf + nlevels);
  f.mb = mb;
  f.mbf = mbf;
  const auto t0 = std::chrono::steady_clock::now();
  try {
    f.ComputeStereoMatches();
  } catch (const std::out_of_range&) {
    return -1;
  }
  g_last_call_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  for (int i = 0; i < nl; ++i) { u_right[i] = f.mvuRight[i]; depth[i] = f.mvDepth[i]; }
  return 0;
}

// Frame::AssignFeaturesToGrid + GetFeaturesInArea (src/Frame.cc:230-245, :327-392): the candidate order that the
// tie-breaks of every windowed search depend on.  Same arguments as orbo_features_in_area.
int reff_features_in_area(const orbo::KeyPoint* kps, int nk, float minX, float minY, float invW, float invH, float x,
                          float y, float r, int minLevel, int maxLevel, int32_t* out, int cap) {
  Frame::mnMinX = minX; Frame::mnMinY = minY;
  Frame::mfGridElementWidthInv = invW; Frame::mfGridElementHeightInv = invH;
  Frame f;
  f.N = nk;
  f.mvKeysUn = kp_vec(kps, nk);
  f.AssignFeaturesToGrid();
  const std::vector<size_t> v = f.GetFeaturesInArea(x, y, r, minLevel, maxLevel);
  for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int32_t)v[i];
  return (int)v.size();
}

}  // extern "C"
