// TEST INFRASTRUCTURE (CPU oracle) -- flat C entry points over orb_oracle.h / cvprim.h for
// ctypes (tests/, __graft_entry__.smoke(), bench.py cpu_baseline / --impl reference only).
#include <cstring>
#include <thread>
#include <vector>

#include "orb_oracle.h"

using namespace orbo;

extern "C" {

void orbo_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t ss, uint8_t* dst, int dw,
                           int dh, size_t ds) {
  resize_linear_u8(src, sw, sh, ss, dst, dw, dh, ds);
}
void orbo_resize_axis_table(int S, int D, int* ofs, int16_t* a0, int16_t* a1) {
  resize_axis_table(S, D, ofs, a0, a1);
}
void orbo_gaussian_blur7_u8(const uint8_t* src, int w, int h, size_t ss, uint8_t* dst, size_t ds) {
  gaussian_blur7_u8(src, w, h, ss, dst, ds);
}
void orbo_copy_make_border_u8(const uint8_t* src, int w, int h, size_t ss, uint8_t* dst, size_t ds,
                              int t, int b, int l, int r) {
  copy_make_border_reflect101_u8(src, w, h, ss, dst, ds, t, b, l, r);
}
// returns the number of keypoints found (may exceed cap; only cap are written), xys = x,y,score
int orbo_fast9_nms(const uint8_t* img, int w, int h, size_t stride, int threshold, int32_t* xys,
                   int cap) {
  std::vector<FastPoint> v;
  fast9_nms(img, w, h, stride, threshold, v);
  for (int i = 0; i < (int)v.size() && i < cap; ++i) {
    xys[3 * i] = v[i].x; xys[3 * i + 1] = v[i].y; xys[3 * i + 2] = v[i].score;
  }
  return (int)v.size();
}
float orbo_fast_atan2(float y, float x) { return fast_atan2_deg(y, x); }
void orbo_fast_atan2_n(const float* y, const float* x, float* out, int n) {
  for (int i = 0; i < n; ++i) out[i] = fast_atan2_deg(y[i], x[i]);
}
int orbo_cv_round(double v) { return cv_round(v); }

struct OrboExtractor {
  Extractor ex;
  std::vector<KeyPoint> kps;
  std::vector<uint8_t> desc;
  OrboExtractor(int n, float s, int l, int a, int b) : ex(n, s, l, a, b) {}
};

OrboExtractor* orbo_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh,
                                     int minTh) {
  return new OrboExtractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void orbo_extractor_destroy(OrboExtractor* e) { delete e; }
// runs the extractor; returns the keypoint count (results are held by the handle)
int orbo_extractor_run(OrboExtractor* e, const uint8_t* gray, int w, int h, size_t stride) {
  e->ex.extract(gray, w, h, stride, e->kps, e->desc);
  return (int)e->kps.size();
}
void orbo_extractor_results(OrboExtractor* e, KeyPoint* kps, uint8_t* desc) {
  if (!e->kps.empty()) {
    std::memcpy(kps, e->kps.data(), e->kps.size() * sizeof(KeyPoint));
    std::memcpy(desc, e->desc.data(), e->desc.size());
  }
}
void orbo_extractor_tables(OrboExtractor* e, float* scale, float* inv_scale, float* sigma2,
                           float* inv_sigma2, int* quota, int* umax16) {
  const Extractor& x = e->ex;
  for (int i = 0; i < x.nlevels; ++i) {
    scale[i] = x.scale[i]; inv_scale[i] = x.inv_scale[i];
    sigma2[i] = x.sigma2[i]; inv_sigma2[i] = x.inv_sigma2[i]; quota[i] = x.quota[i];
  }
  for (int i = 0; i < 16; ++i) umax16[i] = x.umax[i];
}
void orbo_extractor_level_size(OrboExtractor* e, int level, int* w, int* h) {
  *w = e->ex.pyramid[level].w; *h = e->ex.pyramid[level].h;
}
void orbo_extractor_level_image(OrboExtractor* e, int level, int blurred, uint8_t* dst) {
  const Image& im = blurred ? e->ex.blurred[level] : e->ex.pyramid[level];
  if (!im.px.empty()) std::memcpy(dst, im.px.data(), im.px.size());
}
int orbo_extractor_level_has_blur(OrboExtractor* e, int level) {
  return !e->ex.blurred[level].px.empty();
}
int orbo_extractor_num_candidates(OrboExtractor* e, int level) {
  return (int)e->ex.candidates[level].size();
}
void orbo_extractor_candidates(OrboExtractor* e, int level, float* xyr) {
  const auto& c = e->ex.candidates[level];
  for (size_t i = 0; i < c.size(); ++i) {
    xyr[3 * i] = c[i].x; xyr[3 * i + 1] = c[i].y; xyr[3 * i + 2] = c[i].response;
  }
}
int orbo_extractor_num_level_kps(OrboExtractor* e, int level) {
  return (int)e->ex.level_kps[level].size();
}
void orbo_extractor_level_kps(OrboExtractor* e, int level, KeyPoint* out) {
  const auto& c = e->ex.level_kps[level];
  if (!c.empty()) std::memcpy(out, c.data(), c.size() * sizeof(KeyPoint));
}
void orbo_extractor_stage_seconds(OrboExtractor* e, double* out6) {
  for (int i = 0; i < 6; ++i) out6[i] = e->ex.stage_s[i];
}
// standalone quadtree on a candidate list (x,y,response triples); returns count, writes indices
int orbo_distribute_quadtree(const float* xyr, int n, int minX, int maxX, int minY, int maxY, int N,
                             int32_t* out_idx, int cap) {
  std::vector<LevelCandidate> keys(n);
  for (int i = 0; i < n; ++i) keys[i] = {xyr[3 * i], xyr[3 * i + 1], xyr[3 * i + 2]};
  std::vector<int> sel = Extractor::distribute_quadtree(keys, minX, maxX, minY, maxY, N);
  for (int i = 0; i < (int)sel.size() && i < cap; ++i) out_idx[i] = sel[i];
  return (int)sel.size();
}

// Batched extraction on `threads` host threads (frames are independent; the reference already
// runs two extractors in parallel, src/Frame.cc:78-81).  counts[f] = keypoints of frame f;
// kps/desc (optional) receive up to cap per frame.  Returns total keypoints.
long orbo_extract_batch(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh,
                        const uint8_t* frames, int nframes, int w, int h, size_t stride,
                        size_t frame_stride, int threads, int* counts, KeyPoint* kps,
                        uint8_t* desc, int cap) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  std::vector<long> totals(threads, 0);
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      Extractor ex(nfeatures, scaleFactor, nlevels, iniTh, minTh);
      std::vector<KeyPoint> k;
      std::vector<uint8_t> d;
      for (int f = t; f < nframes; f += threads) {
        ex.extract(frames + (size_t)f * frame_stride, w, h, stride, k, d);
        if (counts) counts[f] = (int)k.size();
        totals[t] += (long)k.size();
        const int n = std::min<int>((int)k.size(), cap);
        if (kps && n) std::memcpy(kps + (size_t)f * cap, k.data(), (size_t)n * sizeof(KeyPoint));
        if (desc && n) std::memcpy(desc + (size_t)f * cap * 32, d.data(), (size_t)n * 32);
      }
    });
  for (auto& th : pool) th.join();
  long total = 0;
  for (long v : totals) total += v;
  return total;
}

int orbo_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

// out: idx[nq], best[nq], second[nq]; `threads` host threads split the queries.
void orbo_match_bruteforce(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int32_t* idx,
                           int32_t* best, int32_t* second, int threads) {
  std::vector<BruteResult> r(nq);
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      const int lo = (int)((int64_t)nq * t / threads), hi = (int)((int64_t)nq * (t + 1) / threads);
      if (hi > lo) match_bruteforce(q + (size_t)lo * 32, hi - lo, db, ndb, r.data() + lo);
    });
  for (auto& th : pool) th.join();
  for (int i = 0; i < nq; ++i) { idx[i] = r[i].idx; best[i] = r[i].best; second[i] = r[i].second; }
}

int orbo_search_by_projection(const ProjQuery* q, const uint8_t* qdesc, int nq, const KeyPoint* kps,
                              const uint8_t* kdesc, const float* u_right, int nk, float minX,
                              float minY, float invW, float invH, const float* scale_factors,
                              float th, float nnratio, int th_high, uint8_t* kp_blocked,
                              int32_t* match_of_kp) {
  FrameGrid g;
  g.minX = minX; g.minY = minY; g.invW = invW; g.invH = invH;
  g.build(kps, nk);
  return search_by_projection(q, qdesc, nq, kps, kdesc, u_right, nk, g, scale_factors, th, nnratio,
                              th_high, kp_blocked, match_of_kp);
}

int orbo_search_windowed(const WindowQuery* q, const uint8_t* qdesc, int nq, const KeyPoint* kps,
                         const uint8_t* kdesc, const float* u_right, int nk, float minX, float minY,
                         float invW, float invH, int th_accept, int gate, const float* inv_level_sigma2,
                         int check_orientation, uint8_t* kp_blocked, int32_t* match_of_kp,
                         int32_t* match_of_query) {
  FrameGrid g;
  g.minX = minX; g.minY = minY; g.invW = invW; g.invH = invH;
  g.build(kps, nk);
  return search_windowed(q, qdesc, nq, kps, kdesc, u_right, nk, g, th_accept, gate, inv_level_sigma2,
                         check_orientation != 0, kp_blocked, match_of_kp, match_of_query);
}

static FeatVec mk_fv(const int32_t* node, const int32_t* start, const int32_t* idx, int n) {
  FeatVec f; f.node = node; f.start = start; f.idx = idx; f.n_nodes = n; return f;
}

int orbo_search_by_bow_frame(const int32_t* node1, const int32_t* start1, const int32_t* idx1, int nn1,
                             const int32_t* node2, const int32_t* start2, const int32_t* idx2, int nn2,
                             const KeyPoint* kps1, const uint8_t* d1, const uint8_t* good1, int n1,
                             const KeyPoint* kps2, const uint8_t* d2, int n2, float nnratio, int check_ori,
                             int32_t* match_of_2) {
  return search_by_bow_frame(mk_fv(node1, start1, idx1, nn1), mk_fv(node2, start2, idx2, nn2), kps1, d1, good1, n1,
                             kps2, d2, n2, nnratio, check_ori != 0, match_of_2);
}

int orbo_search_by_bow_keyframes(const int32_t* node1, const int32_t* start1, const int32_t* idx1, int nn1,
                                 const int32_t* node2, const int32_t* start2, const int32_t* idx2, int nn2,
                                 const KeyPoint* kps1, const uint8_t* d1, const uint8_t* good1, int n1,
                                 const KeyPoint* kps2, const uint8_t* d2, const uint8_t* good2, int n2,
                                 float nnratio, int check_ori, int32_t* match12) {
  return search_by_bow_keyframes(mk_fv(node1, start1, idx1, nn1), mk_fv(node2, start2, idx2, nn2), kps1, d1, good1,
                                 n1, kps2, d2, good2, n2, nnratio, check_ori != 0, match12);
}

int orbo_search_for_triangulation(const int32_t* node1, const int32_t* start1, const int32_t* idx1, int nn1,
                                  const int32_t* node2, const int32_t* start2, const int32_t* idx2, int nn2,
                                  const KeyPoint* kps1, const uint8_t* d1, const float* ur1, const uint8_t* has1,
                                  int n1, const KeyPoint* kps2, const uint8_t* d2, const float* ur2,
                                  const uint8_t* has2, int n2, const float* F12, float ex, float ey,
                                  const float* sf2, const float* sigma2_2, int only_stereo, int check_ori,
                                  int32_t* match12) {
  return search_for_triangulation(mk_fv(node1, start1, idx1, nn1), mk_fv(node2, start2, idx2, nn2), kps1, d1, ur1,
                                  has1, n1, kps2, d2, ur2, has2, n2, F12, ex, ey, sf2, sigma2_2, only_stereo != 0,
                                  check_ori != 0, match12);
}

int orbo_search_for_initialization(const KeyPoint* kps1, const uint8_t* d1, int n1, const float* prev_matched,
                                   const KeyPoint* kps2, const uint8_t* d2, int n2, float minX, float minY,
                                   float invW, float invH, int window, float nnratio, int check_ori,
                                   int32_t* match12) {
  FrameGrid g;
  g.minX = minX; g.minY = minY; g.invW = invW; g.invH = invH;
  g.build(kps2, n2);
  return search_for_initialization(kps1, d1, n1, prev_matched, kps2, d2, n2, g, window, nnratio, check_ori != 0,
                                   match12);
}

// candidate list of one GetFeaturesInArea query, in the reference's order
int orbo_features_in_area(const KeyPoint* kps, int nk, float minX, float minY, float invW,
                          float invH, float x, float y, float r, int minLevel, int maxLevel,
                          int32_t* out, int cap) {
  FrameGrid g;
  g.minX = minX; g.minY = minY; g.invW = invW; g.invH = invH;
  g.build(kps, nk);
  std::vector<int> v;
  g.features_in_area(kps, x, y, r, minLevel, maxLevel, v);
  for (int i = 0; i < (int)v.size() && i < cap; ++i) out[i] = v[i];
  return (int)v.size();
}

void orbo_stereo_match(const KeyPoint* kl, const uint8_t* dl, int nl, const KeyPoint* kr,
                       const uint8_t* dr, int nr, int n_rows, const float* scale_factors,
                       float min_d, float max_d, int th_high, int32_t* best_idx_r,
                       int32_t* best_dist) {
  stereo_match(kl, dl, nl, kr, dr, nr, n_rows, scale_factors, min_d, max_d, th_high, best_idx_r,
               best_dist);
}

// levels_*: nlevels dense images (w, h in sizes[2*l], sizes[2*l+1])
int orbo_stereo_refine(const KeyPoint* kl, int nl, const KeyPoint* kr, const int32_t* best_idx_r,
                       const int32_t* best_dist, const uint8_t* const* levels_l, const uint8_t* const* levels_r,
                       const int32_t* sizes_l, const int32_t* sizes_r, int nlevels, const float* sf,
                       const float* inv_sf, float min_d, float max_d, float bf, float* u_right, float* depth) {
  std::vector<Image> pl((size_t)nlevels), pr((size_t)nlevels);
  for (int l = 0; l < nlevels; ++l) {
    pl[l].w = sizes_l[2 * l]; pl[l].h = sizes_l[2 * l + 1];
    pl[l].px.assign(levels_l[l], levels_l[l] + (size_t)pl[l].w * pl[l].h);
    pr[l].w = sizes_r[2 * l]; pr[l].h = sizes_r[2 * l + 1];
    pr[l].px.assign(levels_r[l], levels_r[l] + (size_t)pr[l].w * pr[l].h);
  }
  return stereo_refine(kl, nl, kr, best_idx_r, best_dist, pl.data(), pr.data(), sf, inv_sf, min_d, max_d, bf,
                       u_right, depth);
}

void orbo_distinctive_descriptors(const uint8_t* desc, const int32_t* offsets, int npoints, int32_t* best) {
  for (int p = 0; p < npoints; ++p)
    best[p] = distinctive_descriptor(desc + (size_t)offsets[p] * 32, offsets[p + 1] - offsets[p]);
}

int orbo_voc_transform(int n_nodes, const int32_t* child_begin, const int32_t* children, const uint8_t* node_desc,
                       const int32_t* word_id, const double* weight, int depth, const uint8_t* features, int n,
                       int levelsup, uint32_t* f_word, double* f_weight, uint32_t* f_node, uint32_t* bow_id,
                       double* bow_val, uint32_t* feature_node) {
  VocTree t;
  t.n_nodes = n_nodes; t.child_begin = child_begin; t.children = children; t.desc = node_desc;
  t.word_id = word_id; t.weight = weight; t.depth = depth;
  for (int i = 0; i < n; ++i)
    voc_transform_feature(t, features + (size_t)i * 32, levelsup, f_word + i, f_weight + i, f_node + i);
  return voc_transform(t, features, n, levelsup, bow_id, bow_val, feature_node);
}

}  // extern "C"
