// TEST INFRASTRUCTURE.  C entry points around the reference's src/ORBmatcher.cc compiled VERBATIM from
// /root/reference over oracle/slam_shim (stand-in Frame / KeyFrame / MapPoint / cv::Mat; see slam_shim.h).
// Built by oracle/Makefile (target `refmatch`) into oracle/_ref/libref_matcher.so; nothing of the reference
// is copied into this repository.  The flat-array signatures mirror oracle_capi.cc so a test can hand the same
// inputs to the restated oracle (orbo_*), to this library (refm_*) and to the GPU.
//
// For the projection searches the reference computes its own windows (pose * point, focal lengths, radius,
// level range).  Every GetFeaturesInArea call is recorded by the shim; refm_* return those windows as
// orbo::WindowQuery records, which are then the INPUT of the oracle / GPU call that must reproduce the matches.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <vector>

#include "ORBmatcher.h"

using namespace ORB_SLAM2;

float Frame::fx = 0, Frame::fy = 0, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 0, Frame::invfy = 0;
float Frame::mnMinX = 0, Frame::mnMaxX = 0, Frame::mnMinY = 0, Frame::mnMaxY = 0;
float Frame::mfGridElementWidthInv = 0, Frame::mfGridElementHeightInv = 0;

namespace {

struct Camera {  // fx, fy, cx, cy, mbf, mb, minX, minY, maxX, maxY, grid invW, invH
  float fx, fy, cx, cy, mbf, mb, minX, minY, maxX, maxY, invW, invH;
};

cv::Mat desc_mat(const uint8_t* d, int n) {
  cv::Mat m(n, 32, CV_8U);
  if (n) std::memcpy(m.data, d, (size_t)n * 32);
  return m;
}
cv::Mat float_mat(const float* v, int r, int c) {
  cv::Mat m(r, c, CV_32F);
  std::memcpy(m.data, v, (size_t)r * c * 4);
  return m;
}
std::vector<cv::KeyPoint> kp_vec(const orbo::KeyPoint* k, int n) {
  std::vector<cv::KeyPoint> v((size_t)n);
  if (n) std::memcpy(v.data(), k, (size_t)n * sizeof(cv::KeyPoint));
  return v;
}
void set_frame_statics(const Camera& c) {
  Frame::fx = c.fx; Frame::fy = c.fy; Frame::cx = c.cx; Frame::cy = c.cy;
  Frame::invfx = 1.0f / c.fx; Frame::invfy = 1.0f / c.fy;
  Frame::mnMinX = c.minX; Frame::mnMinY = c.minY; Frame::mnMaxX = c.maxX; Frame::mnMaxY = c.maxY;
  Frame::mfGridElementWidthInv = c.invW; Frame::mfGridElementHeightInv = c.invH;
}
void fill_frame(Frame& F, const Camera& c, const orbo::KeyPoint* kps, const uint8_t* desc, const float* u_right,
                int n, const float* scale_factors, int nlevels) {
  set_frame_statics(c);
  F.N = n;
  F.mbf = c.mbf; F.mb = c.mb;
  F.mvKeys = kp_vec(kps, n);
  F.mvKeysUn = F.mvKeys;
  F.mDescriptors = desc_mat(desc, n);
  F.mvuRight.assign((size_t)n, -1.0f);
  if (u_right) F.mvuRight.assign(u_right, u_right + n);
  F.mvpMapPoints.assign((size_t)n, nullptr);
  F.mvbOutlier.assign((size_t)n, false);
  if (scale_factors) F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
  F.grid.build(F.mvKeysUn, c.minX, c.minY, c.invW, c.invH);
}
std::unique_ptr<KeyFrame> make_keyframe(const Camera& c, const orbo::KeyPoint* kps, const uint8_t* desc,
                                        const float* u_right, int n, const float* scale_factors,
                                        const float* level_sigma2, int nlevels) {
  std::unique_ptr<KeyFrame> k(new KeyFrame(c.fx, c.fy, c.cx, c.cy, c.mbf, c.mb, n, (int)c.minX, (int)c.minY,
                                           (int)c.maxX, (int)c.maxY));
  k->mvKeys = kp_vec(kps, n);
  k->mvKeysUn = k->mvKeys;
  k->mDescriptors = desc_mat(desc, n);
  k->mvuRight.assign((size_t)n, -1.0f);
  if (u_right) k->mvuRight.assign(u_right, u_right + n);
  k->mvpMapPoints.assign((size_t)n, nullptr);
  if (scale_factors) k->mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
  if (level_sigma2) {
    k->mvLevelSigma2.assign(level_sigma2, level_sigma2 + nlevels);
    k->mvInvLevelSigma2.resize((size_t)nlevels);
    for (int i = 0; i < nlevels; ++i) k->mvInvLevelSigma2[i] = 1.0f / level_sigma2[i];
  }
  k->grid.build(k->mvKeysUn, c.minX, c.minY, c.invW, c.invH);
  return k;
}
void fill_featvec(DBoW2::FeatureVector& fv, const int32_t* node, const int32_t* start, const int32_t* idx, int nn) {
  for (int i = 0; i < nn; ++i)
    for (int j = start[i]; j < start[i + 1]; ++j) fv[(DBoW2::NodeId)node[i]].push_back((unsigned)idx[j]);
}
int index_of(const std::map<const MapPoint*, int>& m, const MapPoint* p) {
  if (!p) return -1;
  auto it = m.find(p);
  return it == m.end() ? -1 : it->second;
}
// pose [R | t; 0 0 0 1] from row-major 3x4
cv::Mat pose44(const float* rt) {
  cv::Mat T(4, 4, CV_32F);
  for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) T.at<float>(r, c) = rt[4 * r + c];
  T.at<float>(3, 3) = 1.0f;
  return T;
}
struct PointCloud {  // map points from flat arrays: position, normal, descriptor, observations, distances, level
  std::vector<std::unique_ptr<MapPoint>> own;
  std::map<const MapPoint*, int> index;
  MapPoint* add(int i, const float* pos, const float* normal, const uint8_t* desc, int n_obs) {
    own.emplace_back(new MapPoint());
    MapPoint* p = own.back().get();
    if (pos) p->mWorldPos = float_mat(pos, 3, 1);
    if (normal) p->mNormalVector = float_mat(normal, 3, 1);
    if (desc) p->mDescriptor = desc_mat(desc, 1);
    p->nObs = n_obs;
    index[p] = i;
    return p;
  }
};
// windows recorded during the last reference call -> orbo::WindowQuery per source point
void windows_from_log(const std::map<const MapPoint*, int>& index, orbo::WindowQuery* out) {
  for (const AreaCall& c : ShimLog::get().calls) {
    const int i = index_of(index, c.mp);
    if (i < 0) continue;
    orbo::WindowQuery& w = out[i];
    w.u = c.x; w.v = c.y; w.radius = c.r;
    w.min_level = c.min_level; w.max_level = c.max_level;
    w.valid = 1;
  }
}

}  // namespace

extern "C" {

int refm_descriptor_distance(const uint8_t* a, const uint8_t* b) {
  return ORBmatcher::DescriptorDistance(desc_mat(a, 1), desc_mat(b, 1));
}
void refm_constants(int* th_high, int* th_low, int* histo_length) {
  *th_high = ORBmatcher::TH_HIGH; *th_low = ORBmatcher::TH_LOW; *histo_length = ORBmatcher::HISTO_LENGTH;
}

// src/ORBmatcher.cc:45-129.  Same arguments as orbo_search_by_projection (th_high is the reference's constant).
int refm_search_by_projection(const orbo::ProjQuery* q, const uint8_t* qdesc, int nq, const orbo::KeyPoint* kps,
                              const uint8_t* kdesc, const float* u_right, int nk, float minX, float minY,
                              float invW, float invH, const float* scale_factors, float th, float nnratio,
                              int /*th_high*/, uint8_t* kp_blocked, int32_t* match_of_kp) {
  Camera cam{1, 1, 0, 0, 0, 0, minX, minY, 0, 0, invW, invH};
  Frame F;
  int nlevels = 1;  // the caller's table covers every level that occurs
  for (int i = 0; i < nq; ++i) nlevels = std::max(nlevels, q[i].level + 1);
  for (int i = 0; i < nk; ++i) nlevels = std::max(nlevels, kps[i].octave + 1);
  fill_frame(F, cam, kps, kdesc, u_right, nk, scale_factors, nlevels);
  MapPoint holder;  // keypoints that already hold a map point with observations (:87-89)
  holder.nObs = 1;
  for (int i = 0; i < nk; ++i) if (kp_blocked[i]) F.mvpMapPoints[i] = &holder;
  PointCloud pc;
  std::vector<MapPoint*> vp;
  for (int i = 0; i < nq; ++i) {
    MapPoint* p = pc.add(i, nullptr, nullptr, qdesc + (size_t)i * 32, q[i].has_obs ? 1 : 0);
    p->mTrackProjX = q[i].u; p->mTrackProjY = q[i].v; p->mTrackProjXR = q[i].ur;
    p->mnTrackScaleLevel = q[i].level; p->mTrackViewCos = q[i].view_cos; p->mbTrackInView = q[i].valid != 0;
    vp.push_back(p);
  }
  ORBmatcher matcher(nnratio, true);
  const int n = matcher.SearchByProjection(F, vp, th);
  for (int i = 0; i < nk; ++i) {
    match_of_kp[i] = index_of(pc.index, F.mvpMapPoints[i]);
    kp_blocked[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;
  }
  return n;
}

// src/ORBmatcher.cc:159-288
int refm_search_by_bow_frame(const int32_t* node1, const int32_t* start1, const int32_t* idx1, int nn1,
                             const int32_t* node2, const int32_t* start2, const int32_t* idx2, int nn2,
                             const orbo::KeyPoint* kps1, const uint8_t* d1, const uint8_t* good1, int n1,
                             const orbo::KeyPoint* kps2, const uint8_t* d2, int n2, float nnratio, int check_ori,
                             int32_t* match_of_2) {
  Camera cam{1, 1, 0, 0, 0, 0, 0, 0, 640, 480, 0.1f, 0.1f};
  auto kf = make_keyframe(cam, kps1, d1, nullptr, n1, nullptr, nullptr, 0);
  fill_featvec(kf->mFeatVec, node1, start1, idx1, nn1);
  PointCloud pc;
  for (int i = 0; i < n1; ++i) if (good1[i]) kf->mvpMapPoints[i] = pc.add(i, nullptr, nullptr, nullptr, 1);
  Frame F;
  fill_frame(F, cam, kps2, d2, nullptr, n2, nullptr, 0);
  fill_featvec(F.mFeatVec, node2, start2, idx2, nn2);
  std::vector<MapPoint*> matches;
  ORBmatcher matcher(nnratio, check_ori != 0);
  const int n = matcher.SearchByBoW(kf.get(), F, matches);
  for (int i = 0; i < n2; ++i) match_of_2[i] = index_of(pc.index, matches[i]);
  return n;
}

// src/ORBmatcher.cc:522-655
int refm_search_by_bow_keyframes(const int32_t* node1, const int32_t* start1, const int32_t* idx1, int nn1,
                                 const int32_t* node2, const int32_t* start2, const int32_t* idx2, int nn2,
                                 const orbo::KeyPoint* kps1, const uint8_t* d1, const uint8_t* good1, int n1,
                                 const orbo::KeyPoint* kps2, const uint8_t* d2, const uint8_t* good2, int n2,
                                 float nnratio, int check_ori, int32_t* match12) {
  Camera cam{1, 1, 0, 0, 0, 0, 0, 0, 640, 480, 0.1f, 0.1f};
  auto kf1 = make_keyframe(cam, kps1, d1, nullptr, n1, nullptr, nullptr, 0);
  auto kf2 = make_keyframe(cam, kps2, d2, nullptr, n2, nullptr, nullptr, 0);
  fill_featvec(kf1->mFeatVec, node1, start1, idx1, nn1);
  fill_featvec(kf2->mFeatVec, node2, start2, idx2, nn2);
  PointCloud pc1, pc2;
  for (int i = 0; i < n1; ++i) if (good1[i]) kf1->mvpMapPoints[i] = pc1.add(i, nullptr, nullptr, nullptr, 1);
  for (int i = 0; i < n2; ++i) if (good2[i]) kf2->mvpMapPoints[i] = pc2.add(i, nullptr, nullptr, nullptr, 1);
  std::vector<MapPoint*> matches;
  ORBmatcher matcher(nnratio, check_ori != 0);
  const int n = matcher.SearchByBoW(kf1.get(), kf2.get(), matches);
  for (int i = 0; i < n1; ++i) match12[i] = index_of(pc2.index, matches[i]);
  return n;
}

// src/ORBmatcher.cc:657-823.  The epipole (ex, ey) is what the reference derives from the two poses (:663-670):
// key frame 1 is centred at (ex, ey, 1) and key frame 2 is the identity camera with unit focal length, which
// reproduces exactly those two floats.
int refm_search_for_triangulation(const int32_t* node1, const int32_t* start1, const int32_t* idx1, int nn1,
                                  const int32_t* node2, const int32_t* start2, const int32_t* idx2, int nn2,
                                  const orbo::KeyPoint* kps1, const uint8_t* d1, const float* ur1,
                                  const uint8_t* has1, int n1, const orbo::KeyPoint* kps2, const uint8_t* d2,
                                  const float* ur2, const uint8_t* has2, int n2, const float* F12, float ex,
                                  float ey, const float* sf2, const float* sigma2_2, int only_stereo,
                                  int check_ori, int32_t* match12) {
  Camera cam{1, 1, 0, 0, 0, 0, 0, 0, 640, 480, 0.1f, 0.1f};
  auto kf1 = make_keyframe(cam, kps1, d1, ur1, n1, nullptr, nullptr, 0);
  int nlevels = 1;
  for (int i = 0; i < n2; ++i) nlevels = std::max(nlevels, kps2[i].octave + 1);
  auto kf2 = make_keyframe(cam, kps2, d2, ur2, n2, sf2, sigma2_2, nlevels);
  fill_featvec(kf1->mFeatVec, node1, start1, idx1, nn1);
  fill_featvec(kf2->mFeatVec, node2, start2, idx2, nn2);
  PointCloud pc1, pc2;
  for (int i = 0; i < n1; ++i) if (has1[i]) kf1->mvpMapPoints[i] = pc1.add(i, nullptr, nullptr, nullptr, 1);
  for (int i = 0; i < n2; ++i) if (has2[i]) kf2->mvpMapPoints[i] = pc2.add(i, nullptr, nullptr, nullptr, 1);
  const float centre[3] = {ex, ey, 1.0f};
  kf1->Ow = float_mat(centre, 3, 1);
  const float ident[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
  kf2->Tcw = pose44(ident);
  std::vector<std::pair<size_t, size_t>> pairs;
  ORBmatcher matcher(0.6f, check_ori != 0);
  const int n = matcher.SearchForTriangulation(kf1.get(), kf2.get(), float_mat(F12, 3, 3), pairs, only_stereo != 0);
  for (int i = 0; i < n1; ++i) match12[i] = -1;
  for (const auto& p : pairs) match12[p.first] = (int32_t)p.second;
  return n;
}

// src/ORBmatcher.cc:405-520.  prev_matched is updated in place as the reference does (:514-517).
int refm_search_for_initialization(const orbo::KeyPoint* kps1, const uint8_t* d1, int n1, float* prev_matched,
                                   const orbo::KeyPoint* kps2, const uint8_t* d2, int n2, float minX, float minY,
                                   float invW, float invH, int window, float nnratio, int check_ori,
                                   int32_t* match12) {
  Camera cam{1, 1, 0, 0, 0, 0, minX, minY, 0, 0, invW, invH};
  Frame F1, F2;
  fill_frame(F1, cam, kps1, d1, nullptr, n1, nullptr, 0);
  fill_frame(F2, cam, kps2, d2, nullptr, n2, nullptr, 0);
  std::vector<cv::Point2f> prev((size_t)n1);
  for (int i = 0; i < n1; ++i) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
  std::vector<int> m12;
  ORBmatcher matcher(nnratio, check_ori != 0);
  const int n = matcher.SearchForInitialization(F1, F2, prev, m12, window);
  for (int i = 0; i < n1; ++i) {
    match12[i] = m12[i];
    prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y;
  }
  return n;
}

// src/ORBmatcher.cc:1328-1470, SearchByProjection(CurrentFrame, LastFrame, th, bMono).
// cam12: fx fy cx cy mbf mb minX minY maxX maxY invW invH.  Tcw_cur / Tcw_last: row-major 3x4.
// Last frame: n_last features with octave / angle in kps_last; has_mp / outlier flags; world position, descriptor
// and observation count of each feature's map point.  cur_blocked: current keypoints that already hold a map
// point with observations.  Outputs: match_of_kp[n_cur] (last-frame feature or -1), windows[n_last] = what the
// reference searched (u, v, radius, ur, level range, angle, valid, blocks).
int refm_search_last_frame(const float* cam12, const float* Tcw_cur, const float* Tcw_last,
                           const orbo::KeyPoint* kps_cur, const uint8_t* d_cur, const float* ur_cur, int n_cur,
                           const float* scale_factors, int nlevels, const uint8_t* cur_blocked,
                           const orbo::KeyPoint* kps_last, int n_last, const uint8_t* has_mp,
                           const uint8_t* outlier, const float* world_pos, const uint8_t* mp_desc,
                           const int32_t* mp_obs, float th, int mono, int check_ori, int32_t* match_of_kp,
                           orbo::WindowQuery* windows) {
  Camera cam;
  std::memcpy(&cam, cam12, sizeof(cam));
  Frame cur, last;
  fill_frame(cur, cam, kps_cur, d_cur, ur_cur, n_cur, scale_factors, nlevels);
  cur.mTcw = pose44(Tcw_cur);
  MapPoint holder;
  holder.nObs = 1;
  for (int i = 0; i < n_cur; ++i) if (cur_blocked && cur_blocked[i]) cur.mvpMapPoints[i] = &holder;
  std::vector<uint8_t> no_desc((size_t)std::max(n_last, 1) * 32, 0);
  fill_frame(last, cam, kps_last, no_desc.data(), nullptr, n_last, scale_factors, nlevels);
  last.mTcw = pose44(Tcw_last);
  PointCloud pc;
  for (int i = 0; i < n_last; ++i) {
    if (has_mp[i]) last.mvpMapPoints[i] = pc.add(i, world_pos + 3 * i, nullptr, mp_desc + (size_t)i * 32, mp_obs[i]);
    last.mvbOutlier[i] = outlier[i] != 0;
  }
  for (int i = 0; i < n_last; ++i) {
    windows[i] = orbo::WindowQuery();
    std::memset(&windows[i], 0, sizeof(orbo::WindowQuery));
    windows[i].angle = kps_last[i].angle;
    windows[i].blocks = has_mp[i] && mp_obs[i] > 0;
  }
  ShimLog::get().calls.clear();
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int n = matcher.SearchByProjection(cur, last, th, mono != 0);
  windows_from_log(pc.index, windows);
  // ur of :1409 for the stereo gate, from the same expressions on the same matrix stand-in
  const cv::Mat Rcw = cur.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = cur.mTcw.rowRange(0, 3).col(3);
  for (int i = 0; i < n_last; ++i)
    if (windows[i].valid) {
      cv::Mat x3Dc = Rcw * last.mvpMapPoints[i]->mWorldPos + tcw;
      const float invzc = 1.0 / x3Dc.at<float>(2);
      windows[i].ur = windows[i].u - cur.mbf * invzc;
    }
  for (int i = 0; i < n_cur; ++i) match_of_kp[i] = index_of(pc.index, cur.mvpMapPoints[i]);
  return n;
}

}  // extern "C"
