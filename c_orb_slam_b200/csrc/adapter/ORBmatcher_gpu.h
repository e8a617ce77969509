// Adapters that let the reference's ORBmatcher call sites use the C ABI (include/orb_gpu.h).
// They are templates over the reference's Frame / MapPoint types so this header does not pull in
// Frame.h (DBoW2, Eigen, ...): inside ORB-SLAM2, src/ORBmatcher.cc instantiates them with its own
// classes.  Pointer bookkeeping (F.mvpMapPoints[i] = pMP) stays on the host
// exactly as in the reference; candidate search, Hamming distances, top-2, ratio test and the
// ordered claims run on the GPU.
#ifndef ORBMATCHER_GPU_H
#define ORBMATCHER_GPU_H

#include <vector>

#include "orb_gpu.h"

namespace ORB_SLAM2 {
namespace gpu {

// ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, const float th)
// (src/ORBmatcher.cc:45-129).  Returns nmatches and assigns F.mvpMapPoints like the reference.
template <class FrameT, class MapPointT>
int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, float th, float nnratio,
                       int device = 0) {
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
  if (rc != ORB_GPU_OK) return 0;
  for (int i = 0; i < nk; ++i)
    if (match[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[match[i]];  // :123
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

}  // namespace gpu
}  // namespace ORB_SLAM2
#endif  // ORBMATCHER_GPU_H
