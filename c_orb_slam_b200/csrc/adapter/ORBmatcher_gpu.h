// Building blocks of the drop-in ORB_SLAM2::ORBmatcher (adapter/ORBmatcher.h, adapter/ORBmatcher.cc): the two
// Frame-to-Frame projection searches and the stereo helpers as templates over the reference's Frame / MapPoint
// types, so this header does not pull in Frame.h (DBoW2, Eigen, ...) and can also be instantiated on its own
// (tests/cpp/dropin_matcher.cc does, with the reference's real Frame).  Pointer bookkeeping
// (F.mvpMapPoints[i] = pMP) stays on the host exactly as in the reference; candidate search, Hamming distances,
// top-2, ratio test and the ordered claims run on the GPU.  `device` = CUDA ordinal; `status` (optional) receives
// the OrbGpuStatus of the C-ABI call.
#ifndef ORBMATCHER_GPU_H
#define ORBMATCHER_GPU_H

#include <cstring>
#include <vector>

#include "orb_gpu.h"

namespace ORB_SLAM2 {
namespace gpu {

// ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, const float th)
// (src/ORBmatcher.cc:45-129).  Returns nmatches and assigns F.mvpMapPoints like the reference.
template <class FrameT, class MapPointT>
int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, float th, float nnratio,
                       int device = 0, int* status = nullptr) {
  const int nq = (int)vpMapPoints.size(), nk = F.N;
  std::vector<OrbProjQuery> q(nq);
  std::vector<uint8_t> qdesc((size_t)nq * 32), blocked(nk, 0);
  for (int i = 0; i < nq; ++i) {
    MapPointT* p = vpMapPoints[i];
    OrbProjQuery& o = q[i];
    o.valid = p->mbTrackInView && !p->isBad();
    o.u = p->mTrackProjX; o.v = p->mTrackProjY; o.ur = p->mTrackProjXR;
    o.level = p->mnTrackScaleLevel; o.view_cos = p->mTrackViewCos;
    o.has_obs = p->Observations() > 0;
    o.pad_[0] = o.pad_[1] = 0;
    if (o.valid) {
      const cv::Mat d = p->GetDescriptor();
      std::memcpy(&qdesc[(size_t)i * 32], d.ptr(), 32);
    }
  }
  for (int i = 0; i < nk; ++i)
    blocked[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;  // :87-89
  static_assert(sizeof(cv::KeyPoint) == sizeof(OrbKeyPoint), "cv::KeyPoint layout");
  std::vector<uint8_t> kdesc((size_t)nk * 32);
  for (int i = 0; i < nk; ++i) std::memcpy(&kdesc[(size_t)i * 32], F.mDescriptors.ptr(i), 32);
  OrbFrameGrid g = {FrameT::mnMinX, FrameT::mnMinY, FrameT::mfGridElementWidthInv, FrameT::mfGridElementHeightInv};
  std::vector<int32_t> match(nk, -1);
  int nmatches = 0;
  const int rc = orb_gpu_search_by_projection(
      q.data(), qdesc.data(), nq, reinterpret_cast<const OrbKeyPoint*>(F.mvKeysUn.data()), kdesc.data(),
      F.mvuRight.empty() ? nullptr : F.mvuRight.data(), nk, g, F.mvScaleFactors.data(),
      (int)F.mvScaleFactors.size(), th, nnratio, ORB_GPU_TH_HIGH, device, blocked.data(), match.data(), &nmatches);
  if (status) *status = rc;
  if (rc != ORB_GPU_OK) return 0;
  for (int i = 0; i < nk; ++i)
    if (match[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[match[i]];  // :123
  return nmatches;
}

// ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono)
// (src/ORBmatcher.cc:1328-1470).  The projection of every last-frame map point is computed on the host with
// the reference's own expressions (:1339-1390); the candidate loop, the claims and the rotation histogram
// (:1392-1467) run in orb_gpu_search_windowed.
template <class FrameT, class MapPointT>
int SearchByProjectionLast(FrameT& CurrentFrame, const FrameT& LastFrame, float th, bool bMono,
                           bool checkOrientation, int device = 0, int* status = nullptr) {
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat twc = -Rcw.t() * tcw;
  const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat tlc = Rlw * twc + tlw;
  const bool bForward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  const bool bBackward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  const int nq = LastFrame.N, nk = CurrentFrame.N;
  std::vector<OrbWindowQuery> q(nq);
  std::vector<uint8_t> qdesc((size_t)nq * 32), blocked(nk, 0);
  for (int i = 0; i < nq; ++i) {
    OrbWindowQuery& o = q[i];
    std::memset(&o, 0, sizeof(o));
    MapPointT* pMP = LastFrame.mvpMapPoints[i];
    if (!pMP || LastFrame.mvbOutlier[i]) continue;
    cv::Mat x3Dw = pMP->GetWorldPos();
    cv::Mat x3Dc = Rcw * x3Dw + tcw;
    const float xc = x3Dc.at<float>(0), yc = x3Dc.at<float>(1);
    const float invzc = 1.0 / x3Dc.at<float>(2);
    if (invzc < 0) continue;
    const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
    const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
    if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
    if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
    const int nLastOctave = LastFrame.mvKeys[i].octave;
    o.u = u; o.v = v;
    o.radius = th * CurrentFrame.mvScaleFactors[nLastOctave];
    o.ur = u - CurrentFrame.mbf * invzc;
    if (bForward) { o.min_level = nLastOctave; o.max_level = -1; }            // :1386
    else if (bBackward) { o.min_level = 0; o.max_level = nLastOctave; }       // :1388
    else { o.min_level = nLastOctave - 1; o.max_level = nLastOctave + 1; }    // :1390
    o.angle = LastFrame.mvKeysUn[i].angle;
    o.valid = 1;
    o.blocks = pMP->Observations() > 0;                                       // :1403-1405 seen by later points
    std::memcpy(&qdesc[(size_t)i * 32], pMP->GetDescriptor().ptr(), 32);
  }
  for (int i = 0; i < nk; ++i)
    blocked[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;
  std::vector<uint8_t> kdesc((size_t)nk * 32);
  for (int i = 0; i < nk; ++i) std::memcpy(&kdesc[(size_t)i * 32], CurrentFrame.mDescriptors.ptr(i), 32);
  OrbFrameGrid g = {FrameT::mnMinX, FrameT::mnMinY, FrameT::mfGridElementWidthInv, FrameT::mfGridElementHeightInv};
  std::vector<int32_t> match(nk, -1);
  int nmatches = 0;
  const int rc = orb_gpu_search_windowed(
      q.data(), qdesc.data(), nq, reinterpret_cast<const OrbKeyPoint*>(CurrentFrame.mvKeysUn.data()), kdesc.data(),
      CurrentFrame.mvuRight.empty() ? nullptr : CurrentFrame.mvuRight.data(), nk, g, ORB_GPU_TH_HIGH,
      /*gate=*/1, nullptr, 0, checkOrientation ? 1 : 0, device, blocked.data(), match.data(), nullptr, &nmatches);
  if (status) *status = rc;
  if (rc != ORB_GPU_OK) return 0;
  for (int i = 0; i < nk; ++i) {
    if (match[i] >= 0) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[match[i]];
    else if (match[i] == -2) CurrentFrame.mvpMapPoints[i] = static_cast<MapPointT*>(NULL);  // :1461
  }
  return nmatches;
}

// Hamming stage of Frame::ComputeStereoMatches (src/Frame.cc:466-549): fills bestIdxR / bestDist
// per left keypoint; the caller continues with the sub-pixel refinement (:551-640) unchanged.
template <class FrameT>
int StereoHammingMatches(const FrameT& F, int nRows, float minD, float maxD, std::vector<int32_t>& bestIdxR,
                         std::vector<int32_t>& bestDist, int device = 0) {
  const int nl = F.N, nr = (int)F.mvKeysRight.size();
  bestIdxR.assign(nl, -1);
  bestDist.assign(nl, ORB_GPU_TH_HIGH);
  std::vector<uint8_t> dl((size_t)nl * 32), dr((size_t)nr * 32);
  for (int i = 0; i < nl; ++i) std::memcpy(&dl[(size_t)i * 32], F.mDescriptors.ptr(i), 32);
  for (int i = 0; i < nr; ++i) std::memcpy(&dr[(size_t)i * 32], F.mDescriptorsRight.ptr(i), 32);
  return orb_gpu_stereo_match(reinterpret_cast<const OrbKeyPoint*>(F.mvKeys.data()), dl.data(), nl,
                              reinterpret_cast<const OrbKeyPoint*>(F.mvKeysRight.data()), dr.data(), nr, nRows,
                              F.mvScaleFactors.data(), (int)F.mvScaleFactors.size(), minD, maxD, ORB_GPU_TH_HIGH,
                              device, bestIdxR.data(), bestDist.data());
}

// Frame::ComputeStereoMatches (src/Frame.cc:466-640) whole -- Hamming stage, sub-pixel refinement on the
// pyramids both extractors kept on the device, median filter.  With this the extractors can run with
// mbDownloadPyramid = false: no image returns to the host for stereo.
template <class FrameT>
void ComputeStereoMatches(FrameT& F) {
  const int nl = F.N, nr = (int)F.mvKeysRight.size();
  F.mvuRight.assign(nl, -1.0f);
  F.mvDepth.assign(nl, -1.0f);
  std::vector<uint8_t> dl((size_t)nl * 32), dr((size_t)nr * 32);
  for (int i = 0; i < nl; ++i) std::memcpy(&dl[(size_t)i * 32], F.mDescriptors.ptr(i), 32);
  for (int i = 0; i < nr; ++i) std::memcpy(&dr[(size_t)i * 32], F.mDescriptorsRight.ptr(i), 32);
  orb_gpu_compute_stereo_matches(F.mpORBextractorLeft->Handle(), 0, F.mpORBextractorRight->Handle(), 0,
                                 reinterpret_cast<const OrbKeyPoint*>(F.mvKeys.data()), dl.data(), nl,
                                 reinterpret_cast<const OrbKeyPoint*>(F.mvKeysRight.data()), dr.data(), nr, F.mbf,
                                 F.mb, F.mvuRight.data(), F.mvDepth.data());
}

}  // namespace gpu
}  // namespace ORB_SLAM2
#endif  // ORBMATCHER_GPU_H
