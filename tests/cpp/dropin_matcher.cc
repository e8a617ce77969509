// TEST INFRASTRUCTURE.  The product's matcher adapters (c_orb_slam_b200/csrc/adapter/ORBmatcher_gpu.h, templates over
// the reference's Frame / MapPoint types) instantiated with the reference's REAL Frame (include/Frame.h, src/Frame.cc
// verbatim) and run side by side with the reference's own ORBmatcher (src/ORBmatcher.cc verbatim, same library) on
// copies of the same Frame objects.  Both leave their result in Frame::mvpMapPoints; the harness reports them as
// indices.  Built by tests/cpp/Makefile.
#define private public
#include "Frame.h"
#undef private

#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <vector>

#include "ORBmatcher.h"
#include "adapter/ORBmatcher_gpu.h"

using namespace ORB_SLAM2;

namespace {
cv::Mat gray(const uint8_t* px, int w, int h) {
  cv::Mat m(h, w, CV_8U);
  std::memcpy(m.data, px, (size_t)w * h);
  return m;
}
cv::Mat fmat(const float* v, int r, int c) {
  cv::Mat m(r, c, CV_32F);
  std::memcpy(m.data, v, (size_t)r * c * 4);
  return m;
}
cv::Mat pose44(const float* rt) {
  cv::Mat T(4, 4, CV_32F);
  for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) T.at<float>(r, c) = rt[4 * r + c];
  T.at<float>(3, 3) = 1.0f;
  return T;
}
cv::Mat calib(const float* K4) {
  cv::Mat K(3, 3, CV_32F);
  K.at<float>(0, 0) = K4[0]; K.at<float>(1, 1) = K4[1]; K.at<float>(0, 2) = K4[2]; K.at<float>(1, 2) = K4[3];
  K.at<float>(2, 2) = 1;
  return K;
}
struct Points {
  std::vector<std::unique_ptr<MapPoint>> own;
  std::map<const MapPoint*, int> index;
  MapPoint* add(int i, const uint8_t* desc, int obs) {
    own.emplace_back(new MapPoint());
    MapPoint* p = own.back().get();
    p->mDescriptor = cv::Mat(1, 32, CV_8U);
    std::memcpy(p->mDescriptor.data, desc, 32);
    p->nObs = obs;
    index[p] = i;
    return p;
  }
  int of(const MapPoint* p) const {
    if (!p) return -1;
    auto it = index.find(p);
    return it == index.end() ? -2 : it->second;  // -2: the pre-existing holder point
  }
};
// a frame through the reference's monocular constructor on the drop-in extractor
std::unique_ptr<Frame> make_frame(ORBextractor& ex, ORBVocabulary& voc, const uint8_t* img, int w, int h, const float* K4,
                                  float bf) {
  cv::Mat K = calib(K4), dist(4, 1, CV_32F);
  Frame::mbInitialComputations = true;
  return std::unique_ptr<Frame>(new Frame(gray(img, w, h), 0.0, &ex, &voc, K, dist, bf, 35.0f));
}
}  // namespace

extern "C" {

// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) (src/ORBmatcher.cc:45-129): reference vs adapter on
// copies of one Frame.  proj: OrbProjQuery rows (the MapPoint tracking fields), u_right / holder: per keypoint.
// match_ref / match_gpu [N]: query index now in mvpMapPoints[i], -1 none, -2 the pre-existing point.
int dropin_search_by_projection(const uint8_t* img, int w, int h, int nfeatures, const float* K4, const OrbProjQuery* proj,
                                const uint8_t* qdesc, int nq, const float* u_right, const uint8_t* holder, float th,
                                float nnratio, int32_t* match_ref, int32_t* match_gpu, int* n_ref, int* n_gpu) {
  ORBextractor ex(nfeatures, 1.2f, 8, 20, 7);
  ex.mbDownloadPyramid = false;
  ORBVocabulary voc;
  std::unique_ptr<Frame> f = make_frame(ex, voc, img, w, h, K4, 40.0f);
  const int N = f->N;
  MapPoint held;
  held.nObs = 1;
  for (int i = 0; i < N; ++i) {
    if (u_right) f->mvuRight[i] = u_right[i];
    if (holder && holder[i]) f->mvpMapPoints[i] = &held;
  }
  Points pts;
  std::vector<MapPoint*> vp;
  for (int i = 0; i < nq; ++i) {
    MapPoint* p = pts.add(i, qdesc + (size_t)i * 32, proj[i].has_obs ? 1 : 0);
    p->mTrackProjX = proj[i].u; p->mTrackProjY = proj[i].v; p->mTrackProjXR = proj[i].ur;
    p->mnTrackScaleLevel = proj[i].level; p->mTrackViewCos = proj[i].view_cos; p->mbTrackInView = proj[i].valid != 0;
    vp.push_back(p);
  }
  Frame fr(*f), fg(*f);
  ORBmatcher matcher(nnratio, true);
  *n_ref = matcher.SearchByProjection(fr, vp, th);
  *n_gpu = gpu::SearchByProjection(fg, vp, th, nnratio);
  for (int i = 0; i < N; ++i) { match_ref[i] = pts.of(fr.mvpMapPoints[i]); match_gpu[i] = pts.of(fg.mvpMapPoints[i]); }
  return N;
}

// ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (:1328-1470): reference vs adapter.  The current
// frame comes from the image; the last frame is given by its keypoints (octave, angle), map points (world position,
// descriptor, observations), outlier flags and pose.
int dropin_search_last_frame(const uint8_t* img, int w, int h, int nfeatures, const float* K4, float bf,
                             const float* Tcw_cur, const float* Tcw_last, const float* u_right, const uint8_t* holder,
                             const OrbKeyPoint* kps_last, int n_last, const uint8_t* has_mp, const uint8_t* outlier,
                             const float* world_pos, const uint8_t* mp_desc, const int32_t* mp_obs, float th, int mono,
                             int check_ori, int32_t* match_ref, int32_t* match_gpu, int* n_ref, int* n_gpu) {
  ORBextractor ex(nfeatures, 1.2f, 8, 20, 7);
  ex.mbDownloadPyramid = false;
  ORBVocabulary voc;
  std::unique_ptr<Frame> cur = make_frame(ex, voc, img, w, h, K4, bf);
  const int N = cur->N;
  cur->SetPose(pose44(Tcw_cur));
  MapPoint held;
  held.nObs = 1;
  for (int i = 0; i < N; ++i) {
    if (u_right) cur->mvuRight[i] = u_right[i];
    if (holder && holder[i]) cur->mvpMapPoints[i] = &held;
  }
  Frame last;
  last.N = n_last;
  last.mvKeys.resize((size_t)n_last);
  if (n_last) std::memcpy(last.mvKeys.data(), kps_last, (size_t)n_last * sizeof(OrbKeyPoint));
  last.mvKeysUn = last.mvKeys;
  last.mvpMapPoints.assign((size_t)n_last, nullptr);
  last.mvbOutlier.assign((size_t)n_last, false);
  last.SetPose(pose44(Tcw_last));
  Points pts;
  for (int i = 0; i < n_last; ++i) {
    if (has_mp[i]) {
      MapPoint* p = pts.add(i, mp_desc + (size_t)i * 32, mp_obs[i]);
      p->mWorldPos = fmat(world_pos + 3 * i, 3, 1);
      last.mvpMapPoints[i] = p;
    }
    last.mvbOutlier[i] = outlier[i] != 0;
  }
  Frame fr(*cur), fg(*cur);
  ORBmatcher matcher(0.9f, check_ori != 0);
  *n_ref = matcher.SearchByProjection(fr, last, th, mono != 0);
  *n_gpu = gpu::SearchByProjectionLast<Frame, MapPoint>(fg, last, th, mono != 0, check_ori != 0);
  for (int i = 0; i < N; ++i) { match_ref[i] = pts.of(fr.mvpMapPoints[i]); match_gpu[i] = pts.of(fg.mvpMapPoints[i]); }
  return N;
}

// Frame::ComputeStereoMatches: the reference's CPU function (run by its stereo constructor) vs the adapter's
// gpu::ComputeStereoMatches (whole function on the device, pyramids never leave it) and gpu::StereoHammingMatches.
int dropin_stereo_adapter(const uint8_t* left, const uint8_t* right, int w, int h, int nfeatures, const float* K4, float bf,
                          float* ur_ref, float* dp_ref, float* ur_gpu, float* dp_gpu, int32_t* best_idx, int32_t* best_dist) {
  ORBextractor exl(nfeatures, 1.2f, 8, 20, 7), exr(nfeatures, 1.2f, 8, 20, 7);
  ORBVocabulary voc;
  cv::Mat K = calib(K4), dist(4, 1, CV_32F);
  Frame::mbInitialComputations = true;
  alignas(Frame) static thread_local unsigned char storage[sizeof(Frame)];  // see dropin_frame.cc: mb is read early
  std::memset(storage, 0, sizeof(storage));
  reinterpret_cast<Frame*>(storage)->mb = bf / K4[0];
  Frame* fp = new (storage) Frame(gray(left, w, h), gray(right, w, h), 0.0, &exl, &exr, &voc, K, dist, bf, 35.0f);
  struct Guard { Frame* p; ~Guard() { p->~Frame(); } } guard{fp};
  Frame& f = *fp;
  const int n = f.N;
  for (int i = 0; i < n; ++i) { ur_ref[i] = f.mvuRight[i]; dp_ref[i] = f.mvDepth[i]; }
  gpu::ComputeStereoMatches(f);
  for (int i = 0; i < n; ++i) { ur_gpu[i] = f.mvuRight[i]; dp_gpu[i] = f.mvDepth[i]; }
  std::vector<int32_t> bi, bd;
  const int nrows = exl.mvImagePyramid[0].rows;
  if (gpu::StereoHammingMatches(f, nrows, 0.0f, f.mbf / f.mb, bi, bd) != ORB_GPU_OK) return -1;
  for (int i = 0; i < n; ++i) { best_idx[i] = bi[i]; best_dist[i] = bd[i]; }
  return n;
}

}  // extern "C"
