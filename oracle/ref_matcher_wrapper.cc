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
#include <chrono>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <set>
#include <thread>
#include <vector>

#include "ORBmatcher.h"

using namespace ORB_SLAM2;

float Frame::fx = 0, Frame::fy = 0, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 0, Frame::invfy = 0;
float Frame::mnMinX = 0, Frame::mnMaxX = 0, Frame::mnMinY = 0, Frame::mnMaxY = 0;
float Frame::mfGridElementWidthInv = 0, Frame::mfGridElementHeightInv = 0;

namespace {

// seconds spent inside the matcher method of the last refm_* call on this thread (building the stand-in objects from
// the flat arrays is outside): what bench.py reports as the CPU baseline of a matcher call
thread_local double g_last_call_seconds = 0.0;
struct CallTimer {
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  ~CallTimer() { g_last_call_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
};
#define REFM_TIMED(expr) ([&] { CallTimer timer__; return (expr); }())

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
  k->mfGridElementWidthInv = c.invW; k->mfGridElementHeightInv = c.invH;
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

double refm_last_call_seconds() { return g_last_call_seconds; }

// Brute-force top-2 as config 5 of BASELINE.json defines it, around the matcher's own DescriptorDistance
// (src/ORBmatcher.cc:1647-1663 in the verbatim build): for each query an ascending scan of the database with the
// update rule of :216-225.  `threads` host threads split the queries.  The reference has no such loop of its own (it
// relocalises through the BoW inverted file); this is the CPU baseline bench.py times next to the GPU kernel.
int refm_bruteforce_top2(const uint8_t* q, int nq, const uint8_t* db, int ndb, int threads, int32_t* idx,
                         int32_t* best, int32_t* second) {
  auto header = [](const uint8_t* row) {
    cv::Mat m;                       // a header over the caller's row, as cv::Mat::row() gives the reference
    m.rows = 1; m.cols = 32; m.step = 32; m.data = const_cast<uint8_t*>(row);
    return m;
  };
  std::vector<cv::Mat> rows((size_t)ndb);
  for (int j = 0; j < ndb; ++j) rows[j] = header(db + (size_t)j * 32);
  CallTimer timer;
  auto work = [&](int lo, int hi) {
    for (int i = lo; i < hi; ++i) {
      const cv::Mat dq = header(q + (size_t)i * 32);
      int bestDist1 = 256, bestIdx = -1, bestDist2 = 256;
      for (int j = 0; j < ndb; ++j) {
        const int dist = ORBmatcher::DescriptorDistance(dq, rows[j]);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx = j; }
        else if (dist < bestDist2) { bestDist2 = dist; }
      }
      idx[i] = bestIdx; best[i] = bestDist1; second[i] = bestDist2;
    }
  };
  threads = std::max(1, std::min(threads, nq));
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) pool.emplace_back(work, (int)((long long)nq * t / threads), (int)((long long)nq * (t + 1) / threads));
  for (auto& t : pool) t.join();
  return 0;
}

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
  const int n = REFM_TIMED(matcher.SearchByProjection(F, vp, th));
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
  const int n = REFM_TIMED(matcher.SearchByBoW(kf.get(), F, matches));
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
  const int n = REFM_TIMED(matcher.SearchByBoW(kf1.get(), kf2.get(), matches));
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
  const int n = REFM_TIMED(matcher.SearchForTriangulation(kf1.get(), kf2.get(), float_mat(F12, 3, 3), pairs, only_stereo != 0));
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
  const int n = REFM_TIMED(matcher.SearchForInitialization(F1, F2, prev, m12, window));
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
  const int n = REFM_TIMED(matcher.SearchByProjection(cur, last, th, mono != 0));
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

// ---- projection searches over a cloud of map points ------------------------------------------------------------
// cloud arrays (n points): world_pos[3n], normal[3n] (may be null), desc[32n], flags[n] (bit 0 bad, bit 1 already
// found / already in the key frame), obs[n], dist_inv[2n] (min, max distance invariance), pred_level[n].
namespace {
struct CloudArgs {
  int n; const float* world_pos; const float* normal; const uint8_t* desc; const uint8_t* flags; const int32_t* obs;
  const float* dist_inv; const int32_t* pred_level;
};
MapPoint* add_cloud_point(PointCloud& pc, const CloudArgs& c, int i) {
  MapPoint* p = pc.add(i, c.world_pos + 3 * i, c.normal ? c.normal + 3 * i : nullptr, c.desc + (size_t)i * 32,
                       c.obs ? c.obs[i] : 1);
  p->mbBad = c.flags && (c.flags[i] & 1);
  if (c.dist_inv) { p->min_dist_inv = c.dist_inv[2 * i]; p->max_dist_inv = c.dist_inv[2 * i + 1]; }
  if (c.pred_level) p->predicted_level = c.pred_level[i];
  return p;
}
void clear_windows(orbo::WindowQuery* w, int n) { if (n) std::memset(w, 0, (size_t)n * sizeof(orbo::WindowQuery)); }
void reset_log() { ShimLog::get().calls.clear(); ShimLog::get().map_point_reads.clear(); ShimLog::get().last_mp = nullptr; }
// KeyFrame::SetPose (src/KeyFrame.cc:75-92): Ow = -Rcw^T * tcw
void set_kf_pose(KeyFrame& k, const float* rt) {
  k.Tcw = pose44(rt);
  cv::Mat Rcw = k.Tcw.rowRange(0, 3).colRange(0, 3);
  cv::Mat tcw = k.Tcw.rowRange(0, 3).col(3);
  cv::Mat Rwc = Rcw.t();
  k.Ow = -Rwc * tcw;
}
}  // namespace

extern "C" {

// src/ORBmatcher.cc:1472-1599, SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist): the cloud is
// the key frame's map points by feature (has_mp[i] == 0: none); kps_kf carries the source angles.
int refm_search_keyframe_points(const float* cam12, const float* Tcw_cur, const orbo::KeyPoint* kps_cur,
                                const uint8_t* d_cur, int n_cur, const float* scale_factors, int nlevels,
                                const uint8_t* cur_has_mp, const orbo::KeyPoint* kps_kf, const uint8_t* has_mp,
                                int n, const float* world_pos, const uint8_t* desc, const uint8_t* flags,
                                const float* dist_inv, const int32_t* pred_level, float th, int orb_dist,
                                int check_ori, int32_t* match_of_kp, orbo::WindowQuery* windows) {
  Camera cam;
  std::memcpy(&cam, cam12, sizeof(cam));
  Frame cur;
  fill_frame(cur, cam, kps_cur, d_cur, nullptr, n_cur, scale_factors, nlevels);
  cur.mTcw = pose44(Tcw_cur);
  MapPoint holder;
  for (int i = 0; i < n_cur; ++i) if (cur_has_mp && cur_has_mp[i]) cur.mvpMapPoints[i] = &holder;
  std::vector<uint8_t> zero((size_t)std::max(n, 1) * 32, 0);
  auto kf = make_keyframe(cam, kps_kf, zero.data(), nullptr, n, scale_factors, nullptr, nlevels);
  CloudArgs c{n, world_pos, nullptr, desc, flags, nullptr, dist_inv, pred_level};
  PointCloud pc;
  std::set<MapPoint*> found;
  clear_windows(windows, n);
  for (int i = 0; i < n; ++i) {
    windows[i].angle = kps_kf[i].angle;
    windows[i].blocks = 1;
    if (!has_mp[i]) continue;
    kf->mvpMapPoints[i] = add_cloud_point(pc, c, i);
    if (flags[i] & 2) found.insert(kf->mvpMapPoints[i]);
  }
  reset_log();
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int nm = REFM_TIMED(matcher.SearchByProjection(cur, kf.get(), found, th, orb_dist));
  windows_from_log(pc.index, windows);
  for (int i = 0; i < n_cur; ++i) match_of_kp[i] = index_of(pc.index, cur.mvpMapPoints[i]);
  return nm;
}

// src/ORBmatcher.cc:290-403, SearchByProjection(pKF, Scw, vpPoints, vpMatched, th).  Scw row-major 3x4 (scaled
// rotation | translation).  kf_matched[i]: vpMatched[i] is already set.  Cloud flag bit 1: the point is one of
// those already in vpMatched.  match_of_kp[n_kf] = cloud point now in vpMatched[i], or -1.
int refm_search_sim3_projection(const float* cam12, const float* Scw, const orbo::KeyPoint* kps_kf,
                                const uint8_t* d_kf, int n_kf, const float* scale_factors, int nlevels,
                                const uint8_t* kf_matched, int n, const float* world_pos, const float* normal,
                                const uint8_t* desc, const uint8_t* flags, const float* dist_inv,
                                const int32_t* pred_level, int th, int32_t* match_of_kp,
                                orbo::WindowQuery* windows) {
  Camera cam;
  std::memcpy(&cam, cam12, sizeof(cam));
  auto kf = make_keyframe(cam, kps_kf, d_kf, nullptr, n_kf, scale_factors, nullptr, nlevels);
  CloudArgs c{n, world_pos, normal, desc, flags, nullptr, dist_inv, pred_level};
  PointCloud pc;
  std::vector<MapPoint*> points, matched((size_t)n_kf, nullptr);
  MapPoint holder;
  std::vector<int> found_points;
  for (int i = 0; i < n; ++i) {
    points.push_back(add_cloud_point(pc, c, i));
    if (flags[i] & 2) found_points.push_back(i);
  }
  // already matched key-frame features hold either one of the cloud's "already found" points or a foreign point
  size_t next = 0;
  for (int i = 0; i < n_kf; ++i)
    if (kf_matched[i]) matched[i] = next < found_points.size() ? points[found_points[next++]] : &holder;
  std::vector<MapPoint*> before = matched;
  clear_windows(windows, n);
  for (int i = 0; i < n; ++i) windows[i].blocks = 1;
  reset_log();
  ORBmatcher matcher(0.75f, true);
  const int nm = REFM_TIMED(matcher.SearchByProjection(kf.get(), pose44(Scw), points, matched, th));
  windows_from_log(pc.index, windows);
  for (int i = 0; i < n; ++i)
    if (windows[i].valid) { windows[i].min_level = pred_level[i] - 1; windows[i].max_level = pred_level[i]; }  // :372-375
  for (int i = 0; i < n_kf; ++i) match_of_kp[i] = matched[i] != before[i] ? index_of(pc.index, matched[i]) : -1;
  return nm;
}

// src/ORBmatcher.cc:825-975, Fuse(pKF, vpMapPoints, th).  kf_obs[i] >= 0: key-frame feature i holds a map point
// with that many observations (-1: none).  Cloud flag bit 1: IsInKeyFrame(pKF).  match_of_query[n] = the
// key-frame feature the point was fused with (what pKF->GetMapPoint(bestIdx) was asked for), or -1;
// action[n]: 0 none, 1 AddObservation + AddMapPoint, 2 pMP->Replace(pMPinKF), 3 pMPinKF->Replace(pMP).
int refm_fuse(const float* cam12, const float* Tcw, const orbo::KeyPoint* kps_kf, const uint8_t* d_kf,
              const float* ur_kf, int n_kf, const float* scale_factors, const float* level_sigma2, int nlevels,
              const int32_t* kf_obs, int n, const float* world_pos, const float* normal, const uint8_t* desc,
              const uint8_t* flags, const int32_t* obs, const float* dist_inv, const int32_t* pred_level, float th,
              int32_t* match_of_query, int32_t* action, orbo::WindowQuery* windows) {
  Camera cam;
  std::memcpy(&cam, cam12, sizeof(cam));
  auto kf = make_keyframe(cam, kps_kf, d_kf, ur_kf, n_kf, scale_factors, level_sigma2, nlevels);
  set_kf_pose(*kf, Tcw);
  PointCloud in_kf;
  for (int i = 0; i < n_kf; ++i)
    if (kf_obs[i] >= 0) kf->mvpMapPoints[i] = in_kf.add(i, nullptr, nullptr, nullptr, kf_obs[i]);
  CloudArgs c{n, world_pos, normal, desc, flags, obs, dist_inv, pred_level};
  PointCloud pc;
  std::vector<MapPoint*> points;
  for (int i = 0; i < n; ++i) {
    points.push_back(add_cloud_point(pc, c, i));
    if (flags[i] & 2) points.back()->mObservations[kf.get()] = 0;
  }
  clear_windows(windows, n);
  reset_log();
  ORBmatcher matcher(0.6f, true);
  const int nf = REFM_TIMED(matcher.Fuse(kf.get(), points, th));
  windows_from_log(pc.index, windows);
  const cv::Mat Rcw = kf->GetRotation(), tcw = kf->GetTranslation();
  for (int i = 0; i < n; ++i) {
    match_of_query[i] = -1;
    action[i] = 0;
    if (!windows[i].valid) continue;
    windows[i].min_level = pred_level[i] - 1; windows[i].max_level = pred_level[i];  // :906-909
    cv::Mat p3Dc = Rcw * points[i]->mWorldPos + tcw;                                 // ur of :869
    const float invz = 1 / p3Dc.at<float>(2);
    windows[i].ur = windows[i].u - kf->mbf * invz;
  }
  for (const auto& r : ShimLog::get().map_point_reads) {
    const int i = index_of(pc.index, r.first);
    if (i >= 0) match_of_query[i] = (int32_t)r.second;
  }
  for (int i = 0; i < n; ++i) {
    if (!points[i]->added.empty()) action[i] = 1;
    else if (points[i]->replaced_by) action[i] = 2;
  }
  for (const auto& q : in_kf.own)
    if (q->replaced_by) { const int i = index_of(pc.index, q->replaced_by); if (i >= 0 && !action[i]) action[i] = 3; }
  return nf;
}

// src/ORBmatcher.cc:1102-1326, SearchBySim3.  Each key frame: pose (row-major 3x4), keypoints, descriptors and the
// cloud of its map points by feature (has_mp).  pre12[i] >= 0: vpMatches12[i] is already the map point of feature
// pre12[i] of key frame 2.  Out: match12[n1] (feature of key frame 2 or -1; pre-set entries stay), the windows of
// both passes.
int refm_search_by_sim3(const float* cam12, const float* T1w, const float* T2w, float s12, const float* R12,
                        const float* t12, const float* scale_factors, int nlevels,
                        const orbo::KeyPoint* kps1, const uint8_t* d1, const uint8_t* has1, int n1,
                        const float* pos1, const uint8_t* flags1, const float* dist_inv1, const int32_t* lvl1,
                        const orbo::KeyPoint* kps2, const uint8_t* d2, const uint8_t* has2, int n2,
                        const float* pos2, const uint8_t* flags2, const float* dist_inv2, const int32_t* lvl2,
                        const int32_t* pre12, float th, int32_t* match12, orbo::WindowQuery* win12,
                        orbo::WindowQuery* win21) {
  Camera cam;
  std::memcpy(&cam, cam12, sizeof(cam));
  auto kf1 = make_keyframe(cam, kps1, d1, nullptr, n1, scale_factors, nullptr, nlevels);
  auto kf2 = make_keyframe(cam, kps2, d2, nullptr, n2, scale_factors, nullptr, nlevels);
  set_kf_pose(*kf1, T1w);
  set_kf_pose(*kf2, T2w);
  // the descriptor a map point reports is that of its own feature
  CloudArgs c1{n1, pos1, nullptr, d1, flags1, nullptr, dist_inv1, lvl1}, c2{n2, pos2, nullptr, d2, flags2, nullptr, dist_inv2, lvl2};
  PointCloud pc1, pc2;
  for (int i = 0; i < n1; ++i) if (has1[i]) kf1->mvpMapPoints[i] = add_cloud_point(pc1, c1, i);
  for (int i = 0; i < n2; ++i) if (has2[i]) {
    kf2->mvpMapPoints[i] = add_cloud_point(pc2, c2, i);
    kf2->mvpMapPoints[i]->mObservations[kf2.get()] = (size_t)i;
  }
  std::vector<MapPoint*> m12((size_t)n1, nullptr);
  for (int i = 0; i < n1; ++i) if (pre12[i] >= 0 && kf2->mvpMapPoints[pre12[i]]) m12[i] = kf2->mvpMapPoints[pre12[i]];
  clear_windows(win12, n1);
  clear_windows(win21, n2);
  reset_log();
  ORBmatcher matcher(0.75f, true);
  const int nf = REFM_TIMED(matcher.SearchBySim3(kf1.get(), kf2.get(), m12, s12, float_mat(R12, 3, 3), float_mat(t12, 3, 1), th));
  windows_from_log(pc1.index, win12);
  windows_from_log(pc2.index, win21);
  for (int i = 0; i < n1; ++i) if (win12[i].valid) { win12[i].min_level = lvl1[i] - 1; win12[i].max_level = lvl1[i]; }
  for (int i = 0; i < n2; ++i) if (win21[i].valid) { win21[i].min_level = lvl2[i] - 1; win21[i].max_level = lvl2[i]; }
  for (int i = 0; i < n1; ++i) match12[i] = index_of(pc2.index, m12[i]);
  return nf;
}

// src/ORBmatcher.cc:977-1100, Fuse(pKF, Scw, vpPoints, th, vpReplacePoint).  kf_obs[i] >= 0: feature i holds a map
// point.  Cloud flag bit 1: the point is the one key-frame feature (i mod n_kf) already holds (spAlreadyFound).
// match_of_query[n] = fused key-frame feature or -1; replace[n] = feature whose point went into vpReplacePoint.
int refm_fuse_sim3(const float* cam12, const float* Scw, const orbo::KeyPoint* kps_kf, const uint8_t* d_kf, int n_kf,
                   const float* scale_factors, int nlevels, const int32_t* kf_obs, int n, const float* world_pos,
                   const float* normal, const uint8_t* desc, const uint8_t* flags, const float* dist_inv,
                   const int32_t* pred_level, float th, int32_t* match_of_query, int32_t* replace,
                   orbo::WindowQuery* windows) {
  Camera cam;
  std::memcpy(&cam, cam12, sizeof(cam));
  auto kf = make_keyframe(cam, kps_kf, d_kf, nullptr, n_kf, scale_factors, nullptr, nlevels);
  PointCloud in_kf;
  for (int i = 0; i < n_kf; ++i)
    if (kf_obs[i] >= 0) kf->mvpMapPoints[i] = in_kf.add(i, nullptr, nullptr, nullptr, kf_obs[i]);
  CloudArgs c{n, world_pos, normal, desc, flags, nullptr, dist_inv, pred_level};
  PointCloud pc;
  std::vector<MapPoint*> points, repl((size_t)n, nullptr);
  for (int i = 0; i < n; ++i) {
    points.push_back(add_cloud_point(pc, c, i));
    if ((flags[i] & 2) && n_kf) kf->mvpMapPoints[i % n_kf] = points.back();  // the key frame already sees this point
  }
  clear_windows(windows, n);
  reset_log();
  ORBmatcher matcher(0.8f, true);
  const int nf = REFM_TIMED(matcher.Fuse(kf.get(), pose44(Scw), points, th, repl));
  windows_from_log(pc.index, windows);
  for (int i = 0; i < n; ++i) {
    match_of_query[i] = -1;
    replace[i] = repl[i] ? index_of(in_kf.index, repl[i]) : -1;
    if (repl[i] && replace[i] < 0) replace[i] = -2 - index_of(pc.index, repl[i]);  // a cloud point sitting in the key frame
    if (windows[i].valid) { windows[i].min_level = pred_level[i] - 1; windows[i].max_level = pred_level[i]; }
  }
  for (const auto& r : ShimLog::get().map_point_reads) {
    const int i = index_of(pc.index, r.first);
    if (i >= 0) match_of_query[i] = (int32_t)r.second;
  }
  return nf;
}

}  // extern "C"
