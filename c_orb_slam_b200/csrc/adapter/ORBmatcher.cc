// Bodies of the drop-in ORB_SLAM2::ORBmatcher (adapter/ORBmatcher.h): replaces the reference's src/ORBmatcher.cc.
// Every method keeps the reference's signature and visible effects; what it computes itself is only what the
// reference computes from host objects before its candidate loops (poses, projections, the scale-invariance and
// viewing-angle tests, PredictScale, the walk over equal vocabulary nodes) and what it does with pointers after
// them.  The loops themselves -- GetFeaturesInArea, DescriptorDistance, best / second best, thresholds, claims,
// ComputeThreeMaxima -- are C-ABI calls into liborb_gpu.so (include/orb_gpu.h).  Line numbers cite the reference.
#include "ORBmatcher.h"

#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "orb_gpu.h"
#include "ORBmatcher_gpu.h"

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = ORB_GPU_TH_HIGH;           // src/ORBmatcher.cc:37
const int ORBmatcher::TH_LOW = ORB_GPU_TH_LOW;             // :38
const int ORBmatcher::HISTO_LENGTH = ORB_GPU_HISTO_LENGTH; // :39

namespace {

int g_default_device = 0;

static_assert(sizeof(cv::KeyPoint) == sizeof(OrbKeyPoint), "cv::KeyPoint layout");

inline const OrbKeyPoint* Keys(const std::vector<cv::KeyPoint>& v) {
  return reinterpret_cast<const OrbKeyPoint*>(v.data());
}
// descriptor rows of a cv::Mat (possibly strided) as one dense [n][32] block
void DenseDescriptors(const cv::Mat& D, int n, std::vector<uint8_t>& out) {
  out.resize((size_t)n * 32);
  for (int i = 0; i < n; ++i) std::memcpy(&out[(size_t)i * 32], D.ptr(i), 32);
}
inline OrbFrameGrid GridOf(const Frame&) {
  OrbFrameGrid g = {Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv};
  return g;
}
inline OrbFrameGrid GridOf(const KeyFrame* pKF) {  // src/KeyFrame.cc:569-608 reads these four
  OrbFrameGrid g = {(float)pKF->mnMinX, (float)pKF->mnMinY, pKF->mfGridElementWidthInv, pKF->mfGridElementHeightInv};
  return g;
}
int Report(int rc, const char* what) {
  if (rc != ORB_GPU_OK) std::fprintf(stderr, "ORBmatcher(gpu): %s failed: %d (%s)\n", what, rc, orb_gpu_last_error());
  return rc;
}

// One row per source point of a best-only projection search, filled by the caller's geometry.
struct WindowBatch {
  std::vector<OrbWindowQuery> q;
  std::vector<uint8_t> desc;
  explicit WindowBatch(size_t n) : q(n), desc(n * 32, 0) {
    if (n) std::memset(q.data(), 0, n * sizeof(OrbWindowQuery));
  }
  void Set(size_t i, float u, float v, float radius, float ur, int minLevel, int maxLevel, float angle, bool blocks,
           const cv::Mat& d) {
    OrbWindowQuery& o = q[i];
    o.u = u; o.v = v; o.radius = radius; o.ur = ur;
    o.min_level = minLevel; o.max_level = maxLevel;
    o.angle = angle; o.valid = 1; o.blocks = blocks ? 1 : 0;
    std::memcpy(&desc[i * 32], d.ptr(), 32);
  }
};

// The target of a windowed search: undistorted keypoints, descriptors, right coordinates, grid.
struct Target {
  const OrbKeyPoint* kps;
  std::vector<uint8_t> desc;
  const float* uRight;
  int n;
  OrbFrameGrid grid;
  Target(const Frame& F) : kps(Keys(F.mvKeysUn)), uRight(F.mvuRight.empty() ? NULL : F.mvuRight.data()), n(F.N), grid(GridOf(F)) {
    DenseDescriptors(F.mDescriptors, n, desc);
  }
  Target(const KeyFrame* pKF)
      : kps(Keys(pKF->mvKeysUn)), uRight(pKF->mvuRight.empty() ? NULL : pKF->mvuRight.data()), n(pKF->N), grid(GridOf(pKF)) {
    DenseDescriptors(pKF->mDescriptors, n, desc);
  }
};

// The walk of two DBoW2 feature vectors by the reference's pair of iterators (:175-264, :545-632, :686-789):
// equal node ids only, ascending; for such a node the indices of map 2 become a candidate range and `keep1(idx1)`
// decides which features of map 1 get a row (the reference's `continue`s before its inner loop).
template <class Keep1>
void BucketRows(const DBoW2::FeatureVector& fv1, const DBoW2::FeatureVector& fv2, Keep1 keep1,
                std::vector<OrbBucketQuery>& rows, std::vector<int32_t>& cand) {
  DBoW2::FeatureVector::const_iterator f1it = fv1.begin(), f2it = fv2.begin();
  const DBoW2::FeatureVector::const_iterator f1end = fv1.end(), f2end = fv2.end();
  while (f1it != f1end && f2it != f2end) {
    if (f1it->first == f2it->first) {
      const int32_t begin = (int32_t)cand.size();
      for (size_t i2 = 0; i2 < f2it->second.size(); ++i2) cand.push_back((int32_t)f2it->second[i2]);
      const int32_t end = (int32_t)cand.size();
      for (size_t i1 = 0; i1 < f1it->second.size(); ++i1) {
        const unsigned int idx1 = f1it->second[i1];
        if (!keep1(idx1)) continue;
        OrbBucketQuery r = {(int32_t)idx1, begin, end};
        rows.push_back(r);
      }
      ++f1it;
      ++f2it;
    } else if (f1it->first < f2it->first) {
      f1it = fv1.lower_bound(f2it->first);
    } else {
      f2it = fv2.lower_bound(f1it->first);
    }
  }
}

}  // namespace

void ORBmatcher::SetDefaultDevice(int device) { g_default_device = device; }
int ORBmatcher::GetDefaultDevice() { return g_default_device; }

ORBmatcher::ORBmatcher(float nnratio, bool checkOri)
    : mfNNratio(nnratio), mbCheckOrientation(checkOri), mnDevice(g_default_device), mnLastStatus(ORB_GPU_OK) {}

float ORBmatcher::RadiusByViewingCos(const float& viewCos) {  // :131-137 (the device evaluates the same rule)
  if (viewCos > 0.998) return 2.5;
  return 4.0;
}

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {  // :1647-1663
  const uint8_t* pa = a.ptr();
  const uint8_t* pb = b.ptr();
  int dist = 0;
  for (int i = 0; i < 4; ++i) {
    uint64_t x, y;
    std::memcpy(&x, pa + 8 * i, 8);
    std::memcpy(&y, pb + 8 * i, 8);
    dist += __builtin_popcountll(x ^ y);
  }
  return dist;
}

// :45-129
int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th) {
  return gpu::SearchByProjection<Frame, MapPoint>(F, vpMapPoints, th, mfNNratio, mnDevice, &mnLastStatus);
}

// :1328-1470
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
  return gpu::SearchByProjectionLast<Frame, MapPoint>(CurrentFrame, LastFrame, th, bMono, mbCheckOrientation, mnDevice,
                                                      &mnLastStatus);
}

// :1472-1599
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound,
                                   const float th, const int ORBdist) {
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat Ow = -Rcw.t() * tcw;

  const std::vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
  WindowBatch w(vpMPs.size());
  for (size_t i = 0, iend = vpMPs.size(); i < iend; i++) {
    MapPoint* pMP = vpMPs[i];
    if (!pMP) continue;
    if (pMP->isBad() || sAlreadyFound.count(pMP)) continue;
    // Project (:1497-1510)
    cv::Mat x3Dw = pMP->GetWorldPos();
    cv::Mat x3Dc = Rcw * x3Dw + tcw;
    const float xc = x3Dc.at<float>(0);
    const float yc = x3Dc.at<float>(1);
    const float invzc = 1.0 / x3Dc.at<float>(2);
    const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
    const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
    if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
    if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
    // Depth must be inside the scale pyramid of the image (:1513-1521)
    cv::Mat PO = x3Dw - Ow;
    float dist3D = cv::norm(PO);
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    int nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame);
    const float radius = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
    w.Set(i, u, v, radius, 0.f, nPredictedLevel - 1, nPredictedLevel + 1, pKF->mvKeysUn[i].angle, true,
          pMP->GetDescriptor());
  }
  Target t(CurrentFrame);
  std::vector<uint8_t> blocked(t.n, 0);
  for (int i = 0; i < t.n; ++i) blocked[i] = CurrentFrame.mvpMapPoints[i] != NULL;  // :1541
  std::vector<int32_t> match(t.n, -1);
  int nmatches = 0;
  mnLastStatus = Report(orb_gpu_search_windowed(w.q.data(), w.desc.data(), (int)w.q.size(), t.kps, t.desc.data(), NULL,
                                                t.n, t.grid, ORBdist, /*gate=*/0, NULL, 0, mbCheckOrientation ? 1 : 0,
                                                mnDevice, blocked.data(), match.data(), NULL, &nmatches),
                        "SearchByProjection(Frame, KeyFrame)");
  if (mnLastStatus != ORB_GPU_OK) return 0;
  for (int i = 0; i < t.n; ++i) {
    if (match[i] >= 0) CurrentFrame.mvpMapPoints[i] = vpMPs[match[i]];              // :1557
    else if (match[i] == -2) CurrentFrame.mvpMapPoints[i] = static_cast<MapPoint*>(NULL);  // :1591
  }
  return nmatches;
}

// :290-403
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                                   std::vector<MapPoint*>& vpMatched, int th) {
  const float& fx = pKF->fx;
  const float& fy = pKF->fy;
  const float& cx = pKF->cx;
  const float& cy = pKF->cy;
  // Decompose Scw (:299-303)
  cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
  const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
  cv::Mat Rcw = sRcw / scw;
  cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
  cv::Mat Ow = -Rcw.t() * tcw;
  std::set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
  spAlreadyFound.erase(static_cast<MapPoint*>(NULL));

  WindowBatch w(vpPoints.size());
  for (int iMP = 0, iendMP = vpPoints.size(); iMP < iendMP; iMP++) {
    MapPoint* pMP = vpPoints[iMP];
    if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc = Rcw * p3Dw + tcw;
    if (p3Dc.at<float>(2) < 0.0) continue;
    const float invz = 1 / p3Dc.at<float>(2);
    const float x = p3Dc.at<float>(0) * invz;
    const float y = p3Dc.at<float>(1) * invz;
    const float u = fx * x + cx;
    const float v = fy * y + cy;
    if (!pKF->IsInImage(u, v)) continue;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    cv::Mat PO = p3Dw - Ow;
    const float dist = cv::norm(PO);
    if (dist < minDistance || dist > maxDistance) continue;
    cv::Mat Pn = pMP->GetNormal();
    if (PO.dot(Pn) < 0.5 * dist) continue;
    int nPredictedLevel = pMP->PredictScale(dist, pKF);
    const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
    w.Set(iMP, u, v, radius, 0.f, nPredictedLevel - 1, nPredictedLevel, 0.f, true, pMP->GetDescriptor());  // :378-381
  }
  Target t(pKF);
  std::vector<uint8_t> blocked(t.n, 0);
  for (int i = 0; i < t.n; ++i) blocked[i] = vpMatched[i] != NULL;  // :375
  std::vector<int32_t> match(t.n, -1);
  int nmatches = 0;
  mnLastStatus = Report(orb_gpu_search_windowed(w.q.data(), w.desc.data(), (int)w.q.size(), t.kps, t.desc.data(), NULL,
                                                t.n, t.grid, TH_LOW, 0, NULL, 0, /*check_orientation=*/0, mnDevice,
                                                blocked.data(), match.data(), NULL, &nmatches),
                        "SearchByProjection(KeyFrame, Scw)");
  if (mnLastStatus != ORB_GPU_OK) return 0;
  for (int i = 0; i < t.n; ++i)
    if (match[i] >= 0) vpMatched[i] = vpPoints[match[i]];  // :396
  return nmatches;
}

// :159-288
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches) {
  const std::vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
  vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
  std::vector<OrbBucketQuery> rows;
  std::vector<int32_t> cand;
  BucketRows(pKF->mFeatVec, F.mFeatVec,
             [&](unsigned int idx) { MapPoint* p = vpMapPointsKF[idx]; return p && !p->isBad(); },  // :191-197
             rows, cand);
  const int n1 = (int)vpMapPointsKF.size(), n2 = F.N;
  std::vector<uint8_t> d1, d2;
  DenseDescriptors(pKF->mDescriptors, n1, d1);
  DenseDescriptors(F.mDescriptors, n2, d2);
  std::vector<int32_t> m12(n1, -1), m21(n2, -1);
  int nmatches = 0;
  mnLastStatus = Report(orb_gpu_search_by_bow(rows.data(), (int)rows.size(), cand.data(), (int)cand.size(),
                                              Keys(pKF->mvKeysUn), d1.data(), n1, Keys(F.mvKeys), d2.data(), NULL, n2,
                                              mfNNratio, TH_LOW, mbCheckOrientation ? 1 : 0, mnDevice, m12.data(),
                                              m21.data(), &nmatches),
                        "SearchByBoW(KeyFrame, Frame)");
  if (mnLastStatus != ORB_GPU_OK) return 0;
  for (int i = 0; i < n2; ++i)
    if (m21[i] >= 0) vpMapPointMatches[i] = vpMapPointsKF[m21[i]];  // :232
  return nmatches;
}

// :522-655
int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12) {
  const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
  const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
  vpMatches12 = std::vector<MapPoint*>(vpMapPoints1.size(), static_cast<MapPoint*>(NULL));
  const int n1 = (int)vpMapPoints1.size(), n2 = (int)vpMapPoints2.size();
  std::vector<uint8_t> skip2(n2, 0);
  for (int i = 0; i < n2; ++i) skip2[i] = !vpMapPoints2[i] || vpMapPoints2[i]->isBad();  // :574-580
  std::vector<OrbBucketQuery> rows;
  std::vector<int32_t> cand;
  BucketRows(pKF1->mFeatVec, pKF2->mFeatVec,
             [&](unsigned int idx) { MapPoint* p = vpMapPoints1[idx]; return p && !p->isBad(); },  // :558-562
             rows, cand);
  std::vector<uint8_t> d1, d2;
  DenseDescriptors(pKF1->mDescriptors, n1, d1);
  DenseDescriptors(pKF2->mDescriptors, n2, d2);
  std::vector<int32_t> m12(n1, -1), m21(n2, -1);
  int nmatches = 0;
  mnLastStatus = Report(orb_gpu_search_by_bow(rows.data(), (int)rows.size(), cand.data(), (int)cand.size(),
                                              Keys(pKF1->mvKeysUn), d1.data(), n1, Keys(pKF2->mvKeysUn), d2.data(),
                                              skip2.data(), n2, mfNNratio, TH_LOW - 1 /* strict '<', :598 */,
                                              mbCheckOrientation ? 1 : 0, mnDevice, m12.data(), m21.data(), &nmatches),
                        "SearchByBoW(KeyFrame, KeyFrame)");
  if (mnLastStatus != ORB_GPU_OK) return 0;
  for (int i = 0; i < n1; ++i)
    if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];  // :602
  return nmatches;
}

// :405-520
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                        std::vector<int>& vnMatches12, int windowSize) {
  const int n1 = (int)F1.mvKeysUn.size(), n2 = (int)F2.mvKeysUn.size();
  vnMatches12 = std::vector<int>(n1, -1);
  std::vector<float> prev((size_t)n1 * 2);
  for (int i = 0; i < n1; ++i) {
    prev[2 * i] = vbPrevMatched[i].x;
    prev[2 * i + 1] = vbPrevMatched[i].y;
  }
  std::vector<uint8_t> d1, d2;
  DenseDescriptors(F1.mDescriptors, n1, d1);
  DenseDescriptors(F2.mDescriptors, n2, d2);
  std::vector<int32_t> m12(n1, -1);
  int nmatches = 0;
  mnLastStatus = Report(orb_gpu_search_for_initialization(Keys(F1.mvKeysUn), d1.data(), n1, prev.data(),
                                                          Keys(F2.mvKeysUn), d2.data(), n2, GridOf(F2), windowSize,
                                                          mfNNratio, mbCheckOrientation ? 1 : 0, mnDevice, m12.data(),
                                                          &nmatches),
                        "SearchForInitialization");
  if (mnLastStatus != ORB_GPU_OK) return 0;
  for (int i = 0; i < n1; ++i) vnMatches12[i] = m12[i];
  // Update prev matched (:514-517)
  for (size_t i1 = 0, iend1 = vnMatches12.size(); i1 < iend1; i1++)
    if (vnMatches12[i1] >= 0) vbPrevMatched[i1] = F2.mvKeysUn[vnMatches12[i1]].pt;
  return nmatches;
}

// :657-823
int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                       std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo) {
  // Compute epipole in second image (:663-670)
  cv::Mat Cw = pKF1->GetCameraCenter();
  cv::Mat R2w = pKF2->GetRotation();
  cv::Mat t2w = pKF2->GetTranslation();
  cv::Mat C2 = R2w * Cw + t2w;
  const float invz = 1.0f / C2.at<float>(2);
  const float ex = pKF2->fx * C2.at<float>(0) * invz + pKF2->cx;
  const float ey = pKF2->fy * C2.at<float>(1) * invz + pKF2->cy;

  const int n1 = pKF1->N, n2 = pKF2->N;
  std::vector<uint8_t> skip2(n2, 0);
  for (int i = 0; i < n2; ++i) skip2[i] = pKF2->GetMapPoint(i) != NULL;  // :722-726
  std::vector<OrbBucketQuery> rows;
  std::vector<int32_t> cand;
  BucketRows(pKF1->mFeatVec, pKF2->mFeatVec,
             [&](unsigned int idx1) {
               if (pKF1->GetMapPoint(idx1)) return false;                    // :699-703
               if (bOnlyStereo && !(pKF1->mvuRight[idx1] >= 0)) return false;  // :705-709
               return true;
             },
             rows, cand);
  std::vector<uint8_t> d1, d2;
  DenseDescriptors(pKF1->mDescriptors, n1, d1);
  DenseDescriptors(pKF2->mDescriptors, n2, d2);
  float F[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) F[3 * r + c] = F12.at<float>(r, c);
  std::vector<int32_t> m12(n1, -1);
  int nmatches = 0;
  mnLastStatus = Report(
      orb_gpu_search_for_triangulation(rows.data(), (int)rows.size(), cand.data(), (int)cand.size(), Keys(pKF1->mvKeysUn),
                                       d1.data(), pKF1->mvuRight.empty() ? NULL : pKF1->mvuRight.data(), n1,
                                       Keys(pKF2->mvKeysUn), d2.data(),
                                       pKF2->mvuRight.empty() ? NULL : pKF2->mvuRight.data(), skip2.data(), n2, F, ex, ey,
                                       pKF2->mvScaleFactors.data(), pKF2->mvLevelSigma2.data(),
                                       (int)pKF2->mvScaleFactors.size(), bOnlyStereo ? 1 : 0, mbCheckOrientation ? 1 : 0,
                                       mnDevice, m12.data(), &nmatches),
      "SearchForTriangulation");
  vMatchedPairs.clear();
  if (mnLastStatus != ORB_GPU_OK) return 0;
  vMatchedPairs.reserve(nmatches);
  for (size_t i = 0, iend = m12.size(); i < iend; i++) {  // :815-820
    if (m12[i] < 0) continue;
    vMatchedPairs.push_back(std::make_pair(i, (size_t)m12[i]));
  }
  return nmatches;
}

// :825-975
int ORBmatcher::Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th) {
  cv::Mat Rcw = pKF->GetRotation();
  cv::Mat tcw = pKF->GetTranslation();
  const float& fx = pKF->fx;
  const float& fy = pKF->fy;
  const float& cx = pKF->cx;
  const float& cy = pKF->cy;
  const float& bf = pKF->mbf;
  cv::Mat Ow = pKF->GetCameraCenter();

  const int nMPs = vpMapPoints.size();
  WindowBatch w(nMPs);
  for (int i = 0; i < nMPs; i++) {
    MapPoint* pMP = vpMapPoints[i];
    if (!pMP) continue;
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc = Rcw * p3Dw + tcw;
    if (p3Dc.at<float>(2) < 0.0f) continue;
    const float invz = 1 / p3Dc.at<float>(2);
    const float x = p3Dc.at<float>(0) * invz;
    const float y = p3Dc.at<float>(1) * invz;
    const float u = fx * x + cx;
    const float v = fy * y + cy;
    if (!pKF->IsInImage(u, v)) continue;
    const float ur = u - bf * invz;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    cv::Mat PO = p3Dw - Ow;
    const float dist3D = cv::norm(PO);
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    cv::Mat Pn = pMP->GetNormal();
    if (PO.dot(Pn) < 0.5 * dist3D) continue;
    int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
    const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
    w.Set(i, u, v, radius, ur, nPredictedLevel - 1, nPredictedLevel, 0.f, false, pMP->GetDescriptor());  // :909-912
  }
  Target t(pKF);
  std::vector<uint8_t> blocked(t.n, 0);
  std::vector<int32_t> ofKp(t.n, -1), ofQuery(nMPs, -1);
  int n = 0;
  // chi-square gate 5.99 / 7.8 on the reprojection error (:914-938), best <= TH_LOW (:952), nothing is claimed
  mnLastStatus = Report(orb_gpu_search_windowed(w.q.data(), w.desc.data(), nMPs, t.kps, t.desc.data(), t.uRight, t.n,
                                                t.grid, TH_LOW, /*gate=*/2, pKF->mvInvLevelSigma2.data(),
                                                (int)pKF->mvInvLevelSigma2.size(), 0, mnDevice, blocked.data(),
                                                ofKp.data(), ofQuery.data(), &n),
                        "Fuse(KeyFrame, MapPoints)");
  if (mnLastStatus != ORB_GPU_OK) return 0;
  // If there is already a MapPoint replace otherwise add new measurement (:951-971), in the reference's order.  A
  // Replace() of an earlier iteration can have made a later point bad or put it into the key frame; the reference
  // tests that at the top of its iteration (:849), so it is tested again here before the point acts.
  int nFused = 0;
  for (int i = 0; i < nMPs; i++) {
    const int bestIdx = ofQuery[i];
    if (bestIdx < 0) continue;
    MapPoint* pMP = vpMapPoints[i];
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
    MapPoint* pMPinKF = pKF->GetMapPoint(bestIdx);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) {
        if (pMPinKF->Observations() > pMP->Observations())
          pMP->Replace(pMPinKF);
        else
          pMPinKF->Replace(pMP);
      }
    } else {
      pMP->AddObservation(pKF, bestIdx);
      pKF->AddMapPoint(pMP, bestIdx);
    }
    nFused++;
  }
  return nFused;
}

// :977-1100
int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
                     std::vector<MapPoint*>& vpReplacePoint) {
  const float& fx = pKF->fx;
  const float& fy = pKF->fy;
  const float& cx = pKF->cx;
  const float& cy = pKF->cy;
  cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
  const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
  cv::Mat Rcw = sRcw / scw;
  cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
  cv::Mat Ow = -Rcw.t() * tcw;
  const std::set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();

  const int nPoints = vpPoints.size();
  WindowBatch w(nPoints);
  for (int iMP = 0; iMP < nPoints; iMP++) {
    MapPoint* pMP = vpPoints[iMP];
    if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc = Rcw * p3Dw + tcw;
    if (p3Dc.at<float>(2) < 0.0f) continue;
    const float invz = 1.0 / p3Dc.at<float>(2);
    const float x = p3Dc.at<float>(0) * invz;
    const float y = p3Dc.at<float>(1) * invz;
    const float u = fx * x + cx;
    const float v = fy * y + cy;
    if (!pKF->IsInImage(u, v)) continue;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    cv::Mat PO = p3Dw - Ow;
    const float dist3D = cv::norm(PO);
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    cv::Mat Pn = pMP->GetNormal();
    if (PO.dot(Pn) < 0.5 * dist3D) continue;
    const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
    const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
    w.Set(iMP, u, v, radius, 0.f, nPredictedLevel - 1, nPredictedLevel, 0.f, false, pMP->GetDescriptor());  // :1067-1068
  }
  Target t(pKF);
  std::vector<uint8_t> blocked(t.n, 0);
  std::vector<int32_t> ofKp(t.n, -1), ofQuery(nPoints, -1);
  int n = 0;
  mnLastStatus = Report(orb_gpu_search_windowed(w.q.data(), w.desc.data(), nPoints, t.kps, t.desc.data(), NULL, t.n,
                                                t.grid, TH_LOW, 0, NULL, 0, 0, mnDevice, blocked.data(), ofKp.data(),
                                                ofQuery.data(), &n),
                        "Fuse(KeyFrame, Scw)");
  if (mnLastStatus != ORB_GPU_OK) return 0;
  int nFused = 0;
  for (int iMP = 0; iMP < nPoints; iMP++) {  // :1081-1096
    const int bestIdx = ofQuery[iMP];
    if (bestIdx < 0) continue;
    MapPoint* pMP = vpPoints[iMP];
    if (pMP->isBad()) continue;  // as at the top of the reference's iteration (:1005); nothing in this call sets it
    MapPoint* pMPinKF = pKF->GetMapPoint(bestIdx);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) vpReplacePoint[iMP] = pMPinKF;
    } else {
      pMP->AddObservation(pKF, bestIdx);
      pKF->AddMapPoint(pMP, bestIdx);
    }
    nFused++;
  }
  return nFused;
}

// :1102-1326
int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12,
                             const cv::Mat& R12, const cv::Mat& t12, const float th) {
  const float& fx = pKF1->fx;
  const float& fy = pKF1->fy;
  const float& cx = pKF1->cx;
  const float& cy = pKF1->cy;
  cv::Mat R1w = pKF1->GetRotation();
  cv::Mat t1w = pKF1->GetTranslation();
  cv::Mat R2w = pKF2->GetRotation();
  cv::Mat t2w = pKF2->GetTranslation();
  cv::Mat sR12 = s12 * R12;
  cv::Mat sR21 = (1.0 / s12) * R12.t();
  cv::Mat t21 = -sR21 * t12;

  const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
  const int N1 = vpMapPoints1.size();
  const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
  const int N2 = vpMapPoints2.size();
  std::vector<bool> vbAlreadyMatched1(N1, false);
  std::vector<bool> vbAlreadyMatched2(N2, false);
  for (int i = 0; i < N1; i++) {  // :1132-1142
    MapPoint* pMP = vpMatches12[i];
    if (pMP) {
      vbAlreadyMatched1[i] = true;
      int idx2 = pMP->GetIndexInKeyFrame(pKF2);
      if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
    }
  }

  // Transform from KF1 to KF2 and search (:1148-1225)
  WindowBatch w12(N1);
  for (int i1 = 0; i1 < N1; i1++) {
    MapPoint* pMP = vpMapPoints1[i1];
    if (!pMP || vbAlreadyMatched1[i1]) continue;
    if (pMP->isBad()) continue;
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc1 = R1w * p3Dw + t1w;
    cv::Mat p3Dc2 = sR21 * p3Dc1 + t21;
    if (p3Dc2.at<float>(2) < 0.0) continue;
    const float invz = 1.0 / p3Dc2.at<float>(2);
    const float x = p3Dc2.at<float>(0) * invz;
    const float y = p3Dc2.at<float>(1) * invz;
    const float u = fx * x + cx;
    const float v = fy * y + cy;
    if (!pKF2->IsInImage(u, v)) continue;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    const float dist3D = cv::norm(p3Dc2);
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    const int nPredictedLevel = pMP->PredictScale(dist3D, pKF2);
    const float radius = th * pKF2->mvScaleFactors[nPredictedLevel];
    w12.Set(i1, u, v, radius, 0.f, nPredictedLevel - 1, nPredictedLevel, 0.f, false, pMP->GetDescriptor());
  }
  // Transform from KF2 to KF1 and search (:1228-1305)
  WindowBatch w21(N2);
  for (int i2 = 0; i2 < N2; i2++) {
    MapPoint* pMP = vpMapPoints2[i2];
    if (!pMP || vbAlreadyMatched2[i2]) continue;
    if (pMP->isBad()) continue;
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc2 = R2w * p3Dw + t2w;
    cv::Mat p3Dc1 = sR12 * p3Dc2 + t12;
    if (p3Dc1.at<float>(2) < 0.0) continue;
    const float invz = 1.0 / p3Dc1.at<float>(2);
    const float x = p3Dc1.at<float>(0) * invz;
    const float y = p3Dc1.at<float>(1) * invz;
    const float u = fx * x + cx;
    const float v = fy * y + cy;
    if (!pKF1->IsInImage(u, v)) continue;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    const float dist3D = cv::norm(p3Dc1);
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    const int nPredictedLevel = pMP->PredictScale(dist3D, pKF1);
    const float radius = th * pKF1->mvScaleFactors[nPredictedLevel];
    w21.Set(i2, u, v, radius, 0.f, nPredictedLevel - 1, nPredictedLevel, 0.f, false, pMP->GetDescriptor());
  }

  Target T1(pKF1), T2(pKF2);
  std::vector<int32_t> vnMatch1(N1, -1), vnMatch2(N2, -1);
  {
    std::vector<uint8_t> blocked(T2.n, 0);
    std::vector<int32_t> ofKp(T2.n, -1);
    int n = 0;
    mnLastStatus = Report(orb_gpu_search_windowed(w12.q.data(), w12.desc.data(), N1, T2.kps, T2.desc.data(), NULL, T2.n,
                                                  T2.grid, TH_HIGH, 0, NULL, 0, 0, mnDevice, blocked.data(), ofKp.data(),
                                                  vnMatch1.data(), &n),
                          "SearchBySim3 (1 -> 2)");
    if (mnLastStatus != ORB_GPU_OK) return 0;
  }
  {
    std::vector<uint8_t> blocked(T1.n, 0);
    std::vector<int32_t> ofKp(T1.n, -1);
    int n = 0;
    mnLastStatus = Report(orb_gpu_search_windowed(w21.q.data(), w21.desc.data(), N2, T1.kps, T1.desc.data(), NULL, T1.n,
                                                  T1.grid, TH_HIGH, 0, NULL, 0, 0, mnDevice, blocked.data(), ofKp.data(),
                                                  vnMatch2.data(), &n),
                          "SearchBySim3 (2 -> 1)");
    if (mnLastStatus != ORB_GPU_OK) return 0;
  }
  // Check agreement (:1307-1323)
  int nFound = 0;
  for (int i1 = 0; i1 < N1; i1++) {
    int idx2 = vnMatch1[i1];
    if (idx2 >= 0) {
      int idx1 = vnMatch2[idx2];
      if (idx1 == i1) {
        vpMatches12[i1] = vpMapPoints2[idx2];
        nFound++;
      }
    }
  }
  return nFound;
}

}  // namespace ORB_SLAM2
