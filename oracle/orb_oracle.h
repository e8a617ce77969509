// TEST INFRASTRUCTURE (CPU oracle) -- not part of the shipped product path.
//
// CPU restatement of the reference's ORB front-end (ORB-SLAM2, /root/reference):
//   extractor  src/ORBextractor.cc:72-147, 410-853, 1034-1132; include/ORBextractor.h:32-111
//   matcher    src/ORBmatcher.cc:37-137, 1601-1663 (+ variant bodies); src/Frame.cc:230-245,
//              327-392, 466-549
// Pixel arithmetic (resize/blur/FAST/fastAtan2) comes from cvprim.h, pinned to cv2 4.13.0
// because the reference neither vendors nor pins OpenCV ("parity unpinned by the reference
// itself"; pinned here against cv2 outputs, see tests/golden/ and tests/test_oracle_vs_cv2.py).
// The quadtree's equal-size tie-break follows the canonical creation-sequence rule of
// SURVEY.md Appendix C (the reference's own order depends on heap addresses, :684).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// use this code.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

#include "cvprim.h"

namespace orbo {

struct KeyPoint {  // layout == cv::KeyPoint (28 bytes)
  float x, y, size, angle, response;
  int32_t octave, class_id;
};

struct Image {
  int w = 0, h = 0;
  std::vector<uint8_t> px;  // w*h, dense
  const uint8_t* row(int y) const { return px.data() + (size_t)y * w; }
  uint8_t* row(int y) { return px.data() + (size_t)y * w; }
};

struct LevelCandidate { float x, y, response; };  // coords relative to the 16-px border

class Extractor {
 public:
  Extractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);

  // ORBextractor::operator() (src/ORBextractor.cc:1043-1105); mask is ignored there too.
  void extract(const uint8_t* gray, int w, int h, size_t stride, std::vector<KeyPoint>& kps,
               std::vector<uint8_t>& desc);

  // tables of the constructor (:410-470)
  int nfeatures, nlevels, iniTh, minTh;
  double scaleFactorD;
  std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
  std::vector<int> quota;
  int umax[16];

  // stage outputs of the last extract() call (for stage-by-stage parity tests)
  std::vector<Image> pyramid;                          // mvImagePyramid ROIs (no 19-px pad)
  std::vector<Image> blurred;                          // per-level GaussianBlur (empty if no kps)
  std::vector<std::vector<LevelCandidate>> candidates;  // vToDistributeKeys per level
  std::vector<std::vector<KeyPoint>> level_kps;         // after distribute + orientation

  // per-stage wall clock of the last call, seconds: pyramid, fast, quadtree, orient, blur, desc
  double stage_s[6] = {0, 0, 0, 0, 0, 0};

  void compute_pyramid(const uint8_t* gray, int w, int h, size_t stride);
  void detect_level(int level, std::vector<LevelCandidate>& out) const;
  // DistributeOctTree (:539-763).  Returns indices into `keys` in output (list) order.
  static std::vector<int> distribute_quadtree(const std::vector<LevelCandidate>& keys, int minX,
                                              int maxX, int minY, int maxY, int N);
  float ic_angle(const Image& img, int cx, int cy) const;
  static void brief_descriptor(const Image& blurred, const KeyPoint& kp, uint8_t* out32);
};

// ------------------------------- matcher ------------------------------------------------
// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663), SWAR popcount as there.
int descriptor_distance(const uint8_t* a, const uint8_t* b);

struct BruteResult { int32_t idx; int32_t best, second; };
// M9 (defined in SURVEY 8(a)): ascending scan with the update rule of ORBmatcher.cc:216-225,
// best=second=256 initially, idx=-1 if the database is empty.
void match_bruteforce(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, BruteResult* out);

struct FrameGrid {  // Frame::AssignFeaturesToGrid / GetFeaturesInArea (src/Frame.cc:230-245,327-392)
  float minX, minY, invW, invH;  // mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv
  static const int COLS = 64, ROWS = 48;  // include/Frame.h:37-38
  std::vector<std::vector<int>> cells;    // [ix*ROWS+iy]
  void build(const KeyPoint* kps, int n);
  void features_in_area(const KeyPoint* kps, float x, float y, float r, int minLevel,
                        int maxLevel, std::vector<int>& out) const;
};

struct ProjQuery {   // the MapPoint fields SearchByProjection reads (ORBmatcher.cc:52-66)
  float u, v, ur;    // mTrackProjX, mTrackProjY, mTrackProjXR
  int32_t level;     // mnTrackScaleLevel
  float view_cos;    // mTrackViewCos
  uint8_t valid;     // mbTrackInView && !isBad()
  uint8_t has_obs;   // Observations() > 0 (what a claim by this point means for later points)
};
// M2: ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) (:45-129).
// kp_blocked[i] in/out: keypoint already holds a MapPoint with Observations()>0 (:87-89).
// match_of_kp[i] out: index of the query assigned to keypoint i in this call, else -1.
int search_by_projection(const ProjQuery* q, const uint8_t* qdesc, int nq, const KeyPoint* kps,
                         const uint8_t* kdesc, const float* u_right, int nk,
                         const FrameGrid& grid, const float* scale_factors, float th,
                         float nnratio, int th_high, uint8_t* kp_blocked, int32_t* match_of_kp);

// The per-source-point quantities the best-only projection searches compute before their candidate loop
// (M3 ORBmatcher.cc:1355-1390, M4 :1497-1527, M5 :337-362): projection, radius, level window.
struct WindowQuery {
  float u, v;          // projected position in the target frame
  float radius;        // th * mvScaleFactors[level]
  float ur;            // u - mbf*invzc (stereo gate of M3, :1407-1413)
  int32_t min_level, max_level;  // GetFeaturesInArea(u, v, radius, minLevel, maxLevel) arguments
  float angle;         // angle of the source keypoint (rotation histogram, :1432-1441)
  uint8_t valid;       // the point passed every geometric test of the caller
  uint8_t blocks;      // its claim hides the keypoint from later points (M3: Observations()>0; M4/M5: always)
  uint8_t pad_[2];
};
// Rotation histogram + ComputeThreeMaxima (ORBmatcher.cc:1601-1642): returns the three kept bins (or -1).
void compute_three_maxima(const int* histo_sizes, int L, int& ind1, int& ind2, int& ind3);
// M3/M4/M5: best-only search in a window, accept best <= th_accept, in-order claims, optional rotation
// consistency (mbCheckOrientation).  kp_blocked in/out; match_of_kp[i] = last query assigned to keypoint i or -1
// (-2 when any of its assignments fell in a pruned histogram bin: reset to NULL at :1448-1467).  Returns nmatches.
// gate: 0 none; 1 stereo radius gate of M2/M3 (u_right[i] > 0 and |ur - u_right[i]| > radius -> skip);
// 2 chi-square gate of ORBmatcher::Fuse (:904-931): e2 * inv_level_sigma2[octave] > 5.99 (mono) / 7.8 (stereo,
// u_right[i] >= 0).  match_of_query[i] (optional) = keypoint chosen by query i or -1 (SearchBySim3 :1190-1227
// needs the per-source result: its two passes claim nothing and are cross-checked on the host).
int search_windowed(const WindowQuery* q, const uint8_t* qdesc, int nq, const KeyPoint* kps,
                    const uint8_t* kdesc, const float* u_right, int nk, const FrameGrid& grid,
                    int th_accept, int gate, const float* inv_level_sigma2, bool check_orientation,
                    uint8_t* kp_blocked, int32_t* match_of_kp, int32_t* match_of_query);

// DBoW2::FeatureVector as flat arrays: node ids ascending (std::map order), start[n_nodes + 1] into idx[]
// (the feature indices of a node in insertion order, Thirdparty/DBoW2/DBoW2/FeatureVector.cpp:28-41).
struct FeatVec {
  const int32_t* node;
  const int32_t* start;
  const int32_t* idx;
  int n_nodes;
};

// M6: ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) (src/ORBmatcher.cc:159-288).  good_mp1[i]: key-frame
// feature i holds a map point that is not bad (:194-200).  match_of_2[n2] = key-frame feature whose map
// point the frame feature received, or -1.  Returns nmatches.
int search_by_bow_frame(const FeatVec& fv1, const FeatVec& fv2, const KeyPoint* kps1, const uint8_t* d1,
                        const uint8_t* good_mp1, int n1, const KeyPoint* kps2, const uint8_t* d2, int n2,
                        float nnratio, bool check_orientation, int32_t* match_of_2);

// M6: ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...) (:522-655): strict best < TH_LOW (:598).
// match12[n1] = feature of key frame 2 or -1.
int search_by_bow_keyframes(const FeatVec& fv1, const FeatVec& fv2, const KeyPoint* kps1, const uint8_t* d1,
                            const uint8_t* good_mp1, int n1, const KeyPoint* kps2, const uint8_t* d2,
                            const uint8_t* good_mp2, int n2, float nnratio, bool check_orientation,
                            int32_t* match12);

// M6: ORBmatcher::SearchForTriangulation (:657-823) with CheckDistEpipolarLine (:138-157).  has_mp*: the
// feature already has a map point (:699-703, :722-726); F12 row-major 3x3; (ex, ey) the epipole (:663-670).
int search_for_triangulation(const FeatVec& fv1, const FeatVec& fv2, const KeyPoint* kps1, const uint8_t* d1,
                             const float* ur1, const uint8_t* has_mp1, int n1, const KeyPoint* kps2,
                             const uint8_t* d2, const float* ur2, const uint8_t* has_mp2, int n2,
                             const float* F12, float ex, float ey, const float* scale_factors2,
                             const float* level_sigma2_2, bool only_stereo, bool check_orientation,
                             int32_t* match12);

// M7: ORBmatcher::SearchForInitialization (:405-520).  prev_matched[2*i] = (x, y) of vbPrevMatched[i]
// (not updated here; the caller applies :514-517).  match12[n1] = index in frame 2 or -1.
int search_for_initialization(const KeyPoint* kps1, const uint8_t* d1, int n1, const float* prev_matched,
                              const KeyPoint* kps2, const uint8_t* d2, int n2, const FrameGrid& grid2,
                              int window, float nnratio, bool check_orientation, int32_t* match12);

// M10: Hamming stage of Frame::ComputeStereoMatches (src/Frame.cc:466-549).
// best_idx_r[i] = right index or -1; best_dist[i] = distance (TH_HIGH if none).
void stereo_match(const KeyPoint* kl, const uint8_t* dl, int nl, const KeyPoint* kr,
                  const uint8_t* dr, int nr, int n_rows, const float* scale_factors,
                  float min_d, float max_d, int th_high, int32_t* best_idx_r,
                  int32_t* best_dist);

// N1: sub-pixel refinement and outlier filter of Frame::ComputeStereoMatches (src/Frame.cc:551-640) on the
// unblurred pyramids of both extractors (mvImagePyramid).  pyr_l / pyr_r: one Image per level.  Inputs are the
// Hamming stage's best_idx_r / best_dist.  u_right / depth [nl] as mvuRight / mvDepth (-1 = none).  Returns the
// number of keypoints whose 11x11 windows left the level image (the reference would throw in cv::Mat::rowRange /
// colRange there; they are skipped).
int stereo_refine(const KeyPoint* kl, int nl, const KeyPoint* kr, const int32_t* best_idx_r,
                  const int32_t* best_dist, const Image* pyr_l, const Image* pyr_r,
                  const float* scale_factors, const float* inv_scale_factors, float min_d, float max_d,
                  float bf, float* u_right, float* depth);

// N3: DBoW2 vocabulary tree as flat arrays (node 0 = root; children of node i in Node::children order).
struct VocTree {
  int n_nodes;
  const int32_t* child_begin;  // [n_nodes + 1]
  const int32_t* children;
  const uint8_t* desc;         // [n_nodes][32]
  const int32_t* word_id;      // [n_nodes]
  const double* weight;        // [n_nodes]
  int depth;                   // m_L
};
// TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1213-1259)
void voc_transform_feature(const VocTree& t, const uint8_t* feature, int levelsup, uint32_t* word_id,
                           double* weight, uint32_t* nid);
// TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1127-1192) for TF_IDF
// weighting and L1 normalisation (what ORBvoc.txt declares).  bow_id / bow_val: the BowVector in key order,
// returns its size; feature_node[i] = node of feature i or 0xffffffff when the word is stopped (w <= 0).
int voc_transform(const VocTree& t, const uint8_t* features, int n, int levelsup, uint32_t* bow_id,
                  double* bow_val, uint32_t* feature_node);

// N4: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307) for the n descriptors of one map point.
// Returns BestIdx (:288-301), -1 when n == 0.
int distinctive_descriptor(const uint8_t* desc, int n);

}  // namespace orbo
