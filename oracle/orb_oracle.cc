// TEST INFRASTRUCTURE (CPU oracle) -- see orb_oracle.h.  Build with -ffp-contract=off.
#include "orb_oracle.h"

#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <map>
#include <cstring>

namespace orbo {

static const int kPatchSize = 31;      // ORBextractor.cc:72
static const int kHalfPatch = 15;      // :73
static const int kEdgeThreshold = 19;  // :74

static const int8_t kPattern[256 * 4] = {
#include "../c_orb_slam_b200/csrc/brief_pattern.inc"
};

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---------------------------------------------------------------------------------------------
// constructor tables -- ORBextractor.cc:410-470.  `scaleFactor` is a double member initialised
// from a float argument (ORBextractor.h:98), so products below mix float and double as there.
// ---------------------------------------------------------------------------------------------
Extractor::Extractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniTh_, int minTh_)
    : nfeatures(nfeatures_), nlevels(nlevels_), iniTh(iniTh_), minTh(minTh_),
      scaleFactorD((double)scaleFactor_) {
  scale.assign(nlevels, 1.0f);
  sigma2.assign(nlevels, 1.0f);
  for (int i = 1; i < nlevels; ++i) {
    scale[i] = (float)((double)scale[i - 1] * scaleFactorD);  // :421
    sigma2[i] = scale[i] * scale[i];                          // :422
  }
  inv_scale.resize(nlevels);
  inv_sigma2.resize(nlevels);
  for (int i = 0; i < nlevels; ++i) {
    inv_scale[i] = 1.0f / scale[i];
    inv_sigma2[i] = 1.0f / sigma2[i];
  }
  quota.assign(nlevels, 0);
  const float factor = (float)(1.0 / scaleFactorD);  // 1.0f / double -> double -> float (:434)
  float desired = (float)nfeatures * (1 - factor) /
                  (1 - (float)std::pow((double)factor, (double)nlevels));  // :435
  int sum = 0;
  for (int l = 0; l < nlevels - 1; ++l) {
    quota[l] = cv_round(desired);
    sum += quota[l];
    desired *= factor;
  }
  quota[nlevels - 1] = std::max(nfeatures - sum, 0);

  // circular patch row extents (:454-469)
  const int vmax = cv_floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
  const int vmin = cv_ceil(kHalfPatch * std::sqrt(2.f) / 2);
  const double hp2 = kHalfPatch * kHalfPatch;
  for (int v = 0; v < 16; ++v) umax[v] = 0;
  for (int v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
  for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
    while (umax[v0] == umax[v0 + 1]) ++v0;
    umax[v] = v0;
    ++v0;
  }
}

// ---------------------------------------------------------------------------------------------
// ComputePyramid -- :1107-1132.  The 19-px REFLECT_101 pad of the reference buffer is never
// read by the extractor (FAST starts at 16, the orientation/BRIEF discs stay inside the image),
// so only the ROI is kept.
// ---------------------------------------------------------------------------------------------
void Extractor::compute_pyramid(const uint8_t* gray, int w, int h, size_t stride) {
  pyramid.assign(nlevels, Image());
  for (int l = 0; l < nlevels; ++l) {
    Image& im = pyramid[l];
    im.w = cv_round((float)w * inv_scale[l]);  // sized from the ORIGINAL image (:1112)
    im.h = cv_round((float)h * inv_scale[l]);
    im.px.resize((size_t)im.w * im.h);
    if (l == 0) {
      for (int y = 0; y < h; ++y) std::memcpy(im.row(y), gray + (size_t)y * stride, w);
    } else {
      const Image& p = pyramid[l - 1];
      resize_linear_u8(p.px.data(), p.w, p.h, p.w, im.px.data(), im.w, im.h, im.w);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// per-cell FAST with threshold retry -- ComputeKeyPointsOctTree cell loop, :765-829
// ---------------------------------------------------------------------------------------------
void Extractor::detect_level(int level, std::vector<LevelCandidate>& out) const {
  out.clear();
  const Image& im = pyramid[level];
  const int minBX = kEdgeThreshold - 3, minBY = minBX;
  const int maxBX = im.w - kEdgeThreshold + 3, maxBY = im.h - kEdgeThreshold + 3;
  const float W = 30;
  const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
  const int nCols = (int)(width / W), nRows = (int)(height / W);
  if (nCols <= 0 || nRows <= 0) return;
  const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
  std::vector<FastPoint> cell;
  for (int i = 0; i < nRows; ++i) {
    const float iniY = (float)(minBY + i * hCell);
    float maxY = iniY + hCell + 6;
    if (iniY >= maxBY - 3) continue;  // :794
    if (maxY > maxBY) maxY = (float)maxBY;
    for (int j = 0; j < nCols; ++j) {
      const float iniX = (float)(minBX + j * wCell);
      float maxX = iniX + wCell + 6;
      if (iniX >= maxBX - 6) continue;  // :803
      if (maxX > maxBX) maxX = (float)maxBX;
      const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
      const uint8_t* roi = im.row(y0) + x0;
      fast9_nms(roi, cw, ch, im.w, iniTh, cell);
      if (cell.empty()) fast9_nms(roi, cw, ch, im.w, minTh, cell);  // :812-816
      for (const FastPoint& p : cell)
        out.push_back({(float)p.x + (float)(j * wCell), (float)p.y + (float)(i * hCell),
                       (float)p.score});
    }
  }
}

// ---------------------------------------------------------------------------------------------
// DistributeOctTree -- :539-763 with DivideNode :481-537, as the ordered algorithm of
// SURVEY.md Appendix C.  Nodes live in a vector; the vector index is the creation sequence
// number, which is the canonical equal-size tie-break.
// ---------------------------------------------------------------------------------------------
namespace {
struct QNode {
  int ulx, uly, brx, bry;  // UL and BR corners (UR=(brx,uly), BL=(ulx,bry))
  std::vector<int> keys;   // indices into the candidate array, stable order
  int prev = -1, next = -1;
  bool no_more = false;
  bool alive = true;
};
struct QList {
  std::vector<QNode> nodes;
  int head = -1, tail = -1, size = 0;
  int push_front(QNode&& n) {
    int id = (int)nodes.size();
    nodes.push_back(std::move(n));
    nodes[id].prev = -1;
    nodes[id].next = head;
    if (head >= 0) nodes[head].prev = id; else tail = id;
    head = id;
    ++size;
    return id;
  }
  int push_back(QNode&& n) {
    int id = (int)nodes.size();
    nodes.push_back(std::move(n));
    nodes[id].next = -1;
    nodes[id].prev = tail;
    if (tail >= 0) nodes[tail].next = id; else head = id;
    tail = id;
    ++size;
    return id;
  }
  int erase(int id) {  // returns the following node
    QNode& n = nodes[id];
    int nx = n.next;
    if (n.prev >= 0) nodes[n.prev].next = n.next; else head = n.next;
    if (n.next >= 0) nodes[n.next].prev = n.prev; else tail = n.prev;
    n.alive = false;
    n.keys.clear();
    n.keys.shrink_to_fit();
    --size;
    return nx;
  }
};

// DivideNode (:481-537): children c[0..3] = n1 (x<mid,y<mid), n2 (x>=mid,y<mid), n3, n4.
void divide(const QNode& p, const std::vector<LevelCandidate>& keys, QNode c[4]) {
  const int halfX = (int)std::ceil((float)(p.brx - p.ulx) / 2);
  const int halfY = (int)std::ceil((float)(p.bry - p.uly) / 2);
  const int mx = p.ulx + halfX, my = p.uly + halfY;
  c[0].ulx = p.ulx; c[0].uly = p.uly; c[0].brx = mx;    c[0].bry = my;
  c[1].ulx = mx;    c[1].uly = p.uly; c[1].brx = p.brx; c[1].bry = my;
  c[2].ulx = p.ulx; c[2].uly = my;    c[2].brx = mx;    c[2].bry = p.bry;
  c[3].ulx = mx;    c[3].uly = my;    c[3].brx = p.brx; c[3].bry = p.bry;
  for (int k : p.keys) {
    const LevelCandidate& kp = keys[k];
    const bool left = kp.x < (float)mx, top = kp.y < (float)my;  // int->float compare (:515-525)
    c[left ? (top ? 0 : 2) : (top ? 1 : 3)].keys.push_back(k);
  }
  for (int i = 0; i < 4; ++i) c[i].no_more = (c[i].keys.size() == 1);
}
}  // namespace

std::vector<int> Extractor::distribute_quadtree(const std::vector<LevelCandidate>& keys, int minX,
                                                int maxX, int minY, int maxY, int N) {
  std::vector<int> result;
  const int nIni = (int)std::round((float)(maxX - minX) / (float)(maxY - minY));  // :543
  if (nIni <= 0) return result;  // portrait shapes divide by zero in the reference; rejected
  const float hX = (float)(maxX - minX) / nIni;
  QList L;
  L.nodes.reserve(keys.size() * 2 + 16);
  std::vector<int> roots(nIni);
  for (int i = 0; i < nIni; ++i) {
    QNode n;
    n.ulx = (int)(hX * (float)i);
    n.brx = (int)(hX * (float)(i + 1));
    n.uly = 0;
    n.bry = maxY - minY;
    roots[i] = L.push_back(std::move(n));
  }
  for (int k = 0; k < (int)keys.size(); ++k) L.nodes[roots[(int)(keys[k].x / hX)]].keys.push_back(k);
  for (int it = L.head; it >= 0;) {  // :572-585
    QNode& n = L.nodes[it];
    if (n.keys.size() == 1) { n.no_more = true; it = n.next; }
    else if (n.keys.empty()) it = L.erase(it);
    else it = n.next;
  }

  std::vector<std::pair<int, int>> expandable;  // (nKeys, node id == creation sequence)
  auto split_into_front = [&](int id, int* nToExpand) {
    QNode c[4];
    divide(L.nodes[id], keys, c);
    for (int i = 0; i < 4; ++i) {
      if (c[i].keys.empty()) continue;
      const int nk = (int)c[i].keys.size();
      const int cid = L.push_front(std::move(c[i]));
      if (nk > 1) {
        if (nToExpand) ++*nToExpand;
        expandable.push_back({nk, cid});
      }
    }
  };

  bool finish = false;
  while (!finish) {
    const int prevSize = L.size;
    int nToExpand = 0;
    expandable.clear();
    for (int it = L.head; it >= 0;) {  // :594-665
      if (L.nodes[it].no_more) { it = L.nodes[it].next; continue; }
      split_into_front(it, &nToExpand);
      it = L.erase(it);
    }
    if (L.size >= N || L.size == prevSize) {
      finish = true;
    } else if (L.size + nToExpand * 3 > N) {  // careful phase, :673-737
      while (!finish) {
        const int prev2 = L.size;
        std::vector<std::pair<int, int>> prev = expandable;
        expandable.clear();
        std::sort(prev.begin(), prev.end());  // (nKeys, seq) ascending; walked from the back
        for (int j = (int)prev.size() - 1; j >= 0; --j) {
          split_into_front(prev[j].second, nullptr);
          L.erase(prev[j].second);
          if (L.size >= N) break;
        }
        if (L.size >= N || L.size == prev2) finish = true;
      }
    }
  }

  result.reserve(L.size);
  for (int it = L.head; it >= 0; it = L.nodes[it].next) {  // :741-760
    const std::vector<int>& ks = L.nodes[it].keys;
    int best = ks[0];
    float maxR = keys[best].response;
    for (size_t k = 1; k < ks.size(); ++k)
      if (keys[ks[k]].response > maxR) { best = ks[k]; maxR = keys[best].response; }
    result.push_back(best);
  }
  return result;
}

// ---------------------------------------------------------------------------------------------
// IC_Angle -- :77-104
// ---------------------------------------------------------------------------------------------
float Extractor::ic_angle(const Image& img, int cx, int cy) const {
  int m01 = 0, m10 = 0;
  const uint8_t* c = img.row(cy) + cx;
  const int step = img.w;
  for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
  for (int v = 1; v <= kHalfPatch; ++v) {
    int vsum = 0;
    const int d = umax[v];
    for (int u = -d; u <= d; ++u) {
      const int vp = c[u + v * step], vm = c[u - v * step];
      vsum += vp - vm;
      m10 += u * (vp + vm);
    }
    m01 += v * vsum;
  }
  return fast_atan2_deg((float)m01, (float)m10);
}

// ---------------------------------------------------------------------------------------------
// computeOrbDescriptor -- :106-147
// ---------------------------------------------------------------------------------------------
void Extractor::brief_descriptor(const Image& img, const KeyPoint& kp, uint8_t* out32) {
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  const float angle = kp.angle * factorPI;
  const float a = std::cos(angle), b = std::sin(angle);  // float overloads, as with `using namespace std`
  const uint8_t* center = img.row(cv_round(kp.y)) + cv_round(kp.x);
  const int step = img.w;
  auto tap = [&](int px, int py) -> int {
    const int ry = cv_round((float)px * b + (float)py * a);
    const int rx = cv_round((float)px * a - (float)py * b);
    return center[ry * step + rx];
  };
  for (int i = 0; i < 32; ++i) {
    int val = 0;
    for (int j = 0; j < 8; ++j) {
      const int8_t* t = &kPattern[(i * 8 + j) * 4];
      val |= (tap(t[0], t[1]) < tap(t[2], t[3])) << j;
    }
    out32[i] = (uint8_t)val;
  }
}

// ---------------------------------------------------------------------------------------------
// operator() -- :1043-1105 and ComputeKeyPointsOctTree :765-853
// ---------------------------------------------------------------------------------------------
void Extractor::extract(const uint8_t* gray, int w, int h, size_t stride, std::vector<KeyPoint>& kps,
                        std::vector<uint8_t>& desc) {
  kps.clear();
  desc.clear();
  for (double& s : stage_s) s = 0;
  if (!gray || w <= 0 || h <= 0) return;  // :1046
  double t0 = now_s();
  compute_pyramid(gray, w, h, stride);
  double t1 = now_s();
  stage_s[0] = t1 - t0;

  candidates.assign(nlevels, {});
  level_kps.assign(nlevels, {});
  blurred.assign(nlevels, Image());
  const int minB = kEdgeThreshold - 3;
  for (int l = 0; l < nlevels; ++l) {
    t0 = now_s();
    detect_level(l, candidates[l]);
    t1 = now_s();
    stage_s[1] += t1 - t0;
    const Image& im = pyramid[l];
    std::vector<int> sel = distribute_quadtree(candidates[l], minB, im.w - kEdgeThreshold + 3, minB,
                                               im.h - kEdgeThreshold + 3, quota[l]);
    const int scaledPatch = (int)((float)kPatchSize * scale[l]);  // const int = int*float (:837)
    for (int k : sel) {
      KeyPoint kp;
      kp.x = candidates[l][k].x + (float)minB;
      kp.y = candidates[l][k].y + (float)minB;
      kp.size = (float)scaledPatch;
      kp.angle = -1.f;
      kp.response = candidates[l][k].response;
      kp.octave = l;
      kp.class_id = -1;
      level_kps[l].push_back(kp);
    }
    stage_s[2] += now_s() - t1;
  }
  t0 = now_s();
  for (int l = 0; l < nlevels; ++l)
    for (KeyPoint& kp : level_kps[l]) kp.angle = ic_angle(pyramid[l], cv_round(kp.x), cv_round(kp.y));
  stage_s[3] = now_s() - t0;

  size_t total = 0;
  for (int l = 0; l < nlevels; ++l) total += level_kps[l].size();
  kps.reserve(total);
  desc.assign(total * 32, 0);
  size_t off = 0;
  for (int l = 0; l < nlevels; ++l) {
    std::vector<KeyPoint>& lk = level_kps[l];
    if (lk.empty()) continue;
    t0 = now_s();
    Image& bl = blurred[l];
    bl.w = pyramid[l].w;
    bl.h = pyramid[l].h;
    bl.px.resize(pyramid[l].px.size());
    gaussian_blur7_u8(pyramid[l].px.data(), bl.w, bl.h, bl.w, bl.px.data(), bl.w);
    t1 = now_s();
    stage_s[4] += t1 - t0;
    for (size_t i = 0; i < lk.size(); ++i) brief_descriptor(bl, lk[i], &desc[(off + i) * 32]);
    stage_s[5] += now_s() - t1;
    off += lk.size();
    for (const KeyPoint& kp : lk) {
      KeyPoint o = kp;
      if (l != 0) { o.x *= scale[l]; o.y *= scale[l]; }  // :1095-1101
      kps.push_back(o);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// matcher
// ---------------------------------------------------------------------------------------------
int descriptor_distance(const uint8_t* a, const uint8_t* b) {
  int dist = 0;
  for (int i = 0; i < 8; ++i) {
    uint32_t x, y;
    std::memcpy(&x, a + 4 * i, 4);
    std::memcpy(&y, b + 4 * i, 4);
    uint32_t v = x ^ y;
    v = v - ((v >> 1) & 0x55555555u);
    v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
    dist += (int)((((v + (v >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
  }
  return dist;
}

void match_bruteforce(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, BruteResult* out) {
  for (int i = 0; i < nq; ++i) {
    int best = 256, second = 256, idx = -1;
    const uint64_t* a = (const uint64_t*)(q + (size_t)i * 32);
    for (int64_t j = 0; j < ndb; ++j) {
      const uint64_t* b = (const uint64_t*)(db + (size_t)j * 32);
      const int d = __builtin_popcountll(a[0] ^ b[0]) + __builtin_popcountll(a[1] ^ b[1]) +
                    __builtin_popcountll(a[2] ^ b[2]) + __builtin_popcountll(a[3] ^ b[3]);
      if (d < best) { second = best; best = d; idx = (int)j; }
      else if (d < second) second = d;
    }
    out[i] = {idx, best, second};
  }
}

void FrameGrid::build(const KeyPoint* kps, int n) {
  cells.assign(COLS * ROWS, {});
  for (int i = 0; i < n; ++i) {  // AssignFeaturesToGrid + PosInGrid (Frame.cc:237-244,382-392)
    const int px = (int)std::round((kps[i].x - minX) * invW);
    const int py = (int)std::round((kps[i].y - minY) * invH);
    if (px < 0 || px >= COLS || py < 0 || py >= ROWS) continue;
    cells[px * ROWS + py].push_back(i);
  }
}

void FrameGrid::features_in_area(const KeyPoint* kps, float x, float y, float r, int minLevel,
                                 int maxLevel, std::vector<int>& out) const {
  out.clear();
  const int c0 = std::max(0, (int)std::floor((x - minX - r) * invW));
  if (c0 >= COLS) return;
  const int c1 = std::min(COLS - 1, (int)std::ceil((x - minX + r) * invW));
  if (c1 < 0) return;
  const int r0 = std::max(0, (int)std::floor((y - minY - r) * invH));
  if (r0 >= ROWS) return;
  const int r1 = std::min(ROWS - 1, (int)std::ceil((y - minY + r) * invH));
  if (r1 < 0) return;
  const bool check = (minLevel > 0) || (maxLevel >= 0);
  for (int ix = c0; ix <= c1; ++ix)
    for (int iy = r0; iy <= r1; ++iy)
      for (int id : cells[ix * ROWS + iy]) {
        const KeyPoint& kp = kps[id];
        if (check) {
          if (kp.octave < minLevel) continue;
          if (maxLevel >= 0 && kp.octave > maxLevel) continue;
        }
        if (std::fabs(kp.x - x) < r && std::fabs(kp.y - y) < r) out.push_back(id);
      }
}

int search_by_projection(const ProjQuery* q, const uint8_t* qdesc, int nq, const KeyPoint* kps,
                         const uint8_t* kdesc, const float* u_right, int nk,
                         const FrameGrid& grid, const float* sf, float th, float nnratio,
                         int th_high, uint8_t* blocked, int32_t* match_of_kp) {
  int nmatches = 0;
  const bool bFactor = th != 1.0;
  for (int i = 0; i < nk; ++i) match_of_kp[i] = -1;
  std::vector<int> cand;
  for (int i = 0; i < nq; ++i) {
    const ProjQuery& p = q[i];
    if (!p.valid) continue;
    float r = p.view_cos > 0.998 ? 2.5f : 4.0f;  // RadiusByViewingCos (:131-137)
    if (bFactor) r *= th;
    grid.features_in_area(kps, p.u, p.v, r * sf[p.level], p.level - 1, p.level, cand);
    if (cand.empty()) continue;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int idx : cand) {
      if (blocked[idx]) continue;  // :87-89
      if (u_right && u_right[idx] > 0) {
        const float er = std::fabs(p.ur - u_right[idx]);
        if (er > r * sf[p.level]) continue;
      }
      const int dist = descriptor_distance(qdesc + (size_t)i * 32, kdesc + (size_t)idx * 32);
      if (dist < bestDist) {
        bestDist2 = bestDist; bestDist = dist;
        bestLevel2 = bestLevel; bestLevel = kps[idx].octave;
        bestIdx = idx;
      } else if (dist < bestDist2) {
        bestLevel2 = kps[idx].octave;
        bestDist2 = dist;
      }
    }
    if (bestDist <= th_high) {
      if (bestLevel == bestLevel2 && (float)bestDist > nnratio * (float)bestDist2) continue;
      match_of_kp[bestIdx] = i;  // F.mvpMapPoints[bestIdx]=pMP (:123)
      blocked[bestIdx] = p.has_obs;
      ++nmatches;
    }
  }
  return nmatches;
}

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1601-1642)
void compute_three_maxima(const int* hs, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; ++i) {
    const int s = hs[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

// Candidate loop, acceptance, claim and rotation consistency of ORBmatcher::SearchByProjection(Cur, Last)
// (src/ORBmatcher.cc:1385-1467); the (Cur, KF) form (:1527-1597) and the (KF, Scw) form (:363-400) are
// the same loop with other gates / thresholds.
int search_windowed(const WindowQuery* q, const uint8_t* qdesc, int nq, const KeyPoint* kps,
                    const uint8_t* kdesc, const float* u_right, int nk, const FrameGrid& grid,
                    int th_accept, int gate, const float* inv_level_sigma2, bool check_orientation,
                    uint8_t* blocked, int32_t* match_of_kp, int32_t* match_of_query) {
  const int HISTO_LENGTH = 30;
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];
  const float factor = 1.0f / HISTO_LENGTH;
  for (int i = 0; i < nk; ++i) match_of_kp[i] = -1;
  if (match_of_query) for (int i = 0; i < nq; ++i) match_of_query[i] = -1;
  std::vector<int> cand;
  for (int i = 0; i < nq; ++i) {
    const WindowQuery& p = q[i];
    if (!p.valid) continue;
    grid.features_in_area(kps, p.u, p.v, p.radius, p.min_level, p.max_level, cand);
    if (cand.empty()) continue;
    int bestDist = 256, bestIdx2 = -1;
    for (int i2 : cand) {
      if (blocked[i2]) continue;  // :1403-1405 / :1541-1542 / :375
      if (gate == 1 && u_right && u_right[i2] > 0) {
        const float er = std::fabs(p.ur - u_right[i2]);
        if (er > p.radius) continue;
      }
      if (gate == 2) {  // Fuse :904-931
        const float ex = p.u - kps[i2].x, ey = p.v - kps[i2].y;
        if (u_right && u_right[i2] >= 0) {
          const float er = p.ur - u_right[i2];
          const float e2 = ex * ex + ey * ey + er * er;
          if (e2 * inv_level_sigma2[kps[i2].octave] > 7.8) continue;
        } else {
          const float e2 = ex * ex + ey * ey;
          if (e2 * inv_level_sigma2[kps[i2].octave] > 5.99) continue;
        }
      }
      const int dist = descriptor_distance(qdesc + (size_t)i * 32, kdesc + (size_t)i2 * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
    }
    if (bestDist <= th_accept) {
      if (match_of_query) match_of_query[i] = bestIdx2;
      match_of_kp[bestIdx2] = i;
      blocked[bestIdx2] = p.blocks;
      ++nmatches;
      if (check_orientation) {
        float rot = p.angle - kps[bestIdx2].angle;
        if (rot < 0.0) rot += 360.0f;
        int bin = (int)std::round(rot * factor);
        if (bin == HISTO_LENGTH) bin = 0;
        rotHist[bin].push_back(bestIdx2);
      }
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, sizes[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; ++i) sizes[i] = (int)rotHist[i].size();
    compute_three_maxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; ++i)
      if (i != ind1 && i != ind2 && i != ind3)
        for (int idx : rotHist[i]) { match_of_kp[idx] = -2; blocked[idx] = 0; --nmatches; }  // NULL after :1461
    if (match_of_query)
      for (int i = 0; i < nq; ++i)
        if (match_of_query[i] >= 0 && match_of_kp[match_of_query[i]] == -2) match_of_query[i] = -1;
  }
  return nmatches;
}

namespace {
const int kThLow = 50;  // ORBmatcher::TH_LOW (src/ORBmatcher.cc:37)
const int kHistoLength = 30;  // ORBmatcher::HISTO_LENGTH (:39)

inline int rot_bin(float a1, float a2) {  // :240-247 and its copies
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * (1.0f / kHistoLength));
  if (bin == kHistoLength) bin = 0;
  return bin;
}

// the two-iterator walk over equal node ids (:178-266); lower_bound on a std::map == advance to first >= key
template <class Fn>
void for_equal_nodes(const FeatVec& a, const FeatVec& b, Fn fn) {
  int ia = 0, ib = 0;
  while (ia < a.n_nodes && ib < b.n_nodes) {
    if (a.node[ia] == b.node[ib]) { fn(ia, ib); ++ia; ++ib; }
    else if (a.node[ia] < b.node[ib]) { while (ia < a.n_nodes && a.node[ia] < b.node[ib]) ++ia; }
    else { while (ib < b.n_nodes && b.node[ib] < a.node[ia]) ++ib; }
  }
}

// prune every bin but the three maxima; calls drop(element) for the pruned entries, returns how many
template <class Fn>
int prune_rotation(std::vector<int>* rotHist, Fn drop) {
  int ind1 = -1, ind2 = -1, ind3 = -1, sizes[kHistoLength], n = 0;
  for (int i = 0; i < kHistoLength; ++i) sizes[i] = (int)rotHist[i].size();
  compute_three_maxima(sizes, kHistoLength, ind1, ind2, ind3);
  for (int i = 0; i < kHistoLength; ++i) {
    if (i == ind1 || i == ind2 || i == ind3) continue;
    for (int e : rotHist[i]) n += drop(e);
  }
  return n;
}
}  // namespace

int search_by_bow_frame(const FeatVec& fv1, const FeatVec& fv2, const KeyPoint* kps1, const uint8_t* d1,
                        const uint8_t* good_mp1, int n1, const KeyPoint* kps2, const uint8_t* d2, int n2,
                        float nnratio, bool check_orientation, int32_t* match_of_2) {
  (void)n1;
  for (int i = 0; i < n2; ++i) match_of_2[i] = -1;
  int nmatches = 0;
  std::vector<int> rotHist[kHistoLength];
  for_equal_nodes(fv1, fv2, [&](int a, int b) {
    for (int iKF = fv1.start[a]; iKF < fv1.start[a + 1]; ++iKF) {
      const int realIdxKF = fv1.idx[iKF];
      if (!good_mp1[realIdxKF]) continue;  // :194-200
      int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
      for (int iF = fv2.start[b]; iF < fv2.start[b + 1]; ++iF) {
        const int realIdxF = fv2.idx[iF];
        if (match_of_2[realIdxF] >= 0) continue;  // :209
        const int dist = descriptor_distance(d1 + (size_t)realIdxKF * 32, d2 + (size_t)realIdxF * 32);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
        else if (dist < bestDist2) { bestDist2 = dist; }
      }
      if (bestDist1 <= kThLow && (float)bestDist1 < nnratio * (float)bestDist2) {  // :228-230
        match_of_2[bestIdxF] = realIdxKF;
        if (check_orientation) rotHist[rot_bin(kps1[realIdxKF].angle, kps2[bestIdxF].angle)].push_back(bestIdxF);
        ++nmatches;
      }
    }
  });
  if (check_orientation)
    nmatches -= prune_rotation(rotHist, [&](int idxF) { match_of_2[idxF] = -1; return 1; });  // :268-285
  return nmatches;
}

int search_by_bow_keyframes(const FeatVec& fv1, const FeatVec& fv2, const KeyPoint* kps1, const uint8_t* d1,
                            const uint8_t* good_mp1, int n1, const KeyPoint* kps2, const uint8_t* d2,
                            const uint8_t* good_mp2, int n2, float nnratio, bool check_orientation,
                            int32_t* match12) {
  for (int i = 0; i < n1; ++i) match12[i] = -1;
  std::vector<uint8_t> matched2((size_t)n2, 0);
  int nmatches = 0;
  std::vector<int> rotHist[kHistoLength];
  for_equal_nodes(fv1, fv2, [&](int a, int b) {
    for (int i1 = fv1.start[a]; i1 < fv1.start[a + 1]; ++i1) {
      const int idx1 = fv1.idx[i1];
      if (!good_mp1[idx1]) continue;  // :559-563
      int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
      for (int i2 = fv2.start[b]; i2 < fv2.start[b + 1]; ++i2) {
        const int idx2 = fv2.idx[i2];
        if (matched2[idx2] || !good_mp2[idx2]) continue;  // :576-580
        const int dist = descriptor_distance(d1 + (size_t)idx1 * 32, d2 + (size_t)idx2 * 32);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
        else if (dist < bestDist2) { bestDist2 = dist; }
      }
      if (bestDist1 < kThLow && (float)bestDist1 < nnratio * (float)bestDist2) {  // :598-600 (strict)
        match12[idx1] = bestIdx2;
        matched2[bestIdx2] = 1;
        if (check_orientation) rotHist[rot_bin(kps1[idx1].angle, kps2[bestIdx2].angle)].push_back(idx1);
        ++nmatches;
      }
    }
  });
  if (check_orientation)
    nmatches -= prune_rotation(rotHist, [&](int idx1) { match12[idx1] = -1; return 1; });  // :634-652
  return nmatches;
}

static bool check_dist_epipolar_line(const KeyPoint& kp1, const KeyPoint& kp2, const float* F12,
                                     const float* level_sigma2_2) {  // :138-157
  const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
  const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
  const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
  const float num = a * kp2.x + b * kp2.y + c;
  const float den = a * a + b * b;
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * level_sigma2_2[kp2.octave];
}

int search_for_triangulation(const FeatVec& fv1, const FeatVec& fv2, const KeyPoint* kps1, const uint8_t* d1,
                             const float* ur1, const uint8_t* has_mp1, int n1, const KeyPoint* kps2,
                             const uint8_t* d2, const float* ur2, const uint8_t* has_mp2, int n2,
                             const float* F12, float ex, float ey, const float* scale_factors2,
                             const float* level_sigma2_2, bool only_stereo, bool check_orientation,
                             int32_t* match12) {
  for (int i = 0; i < n1; ++i) match12[i] = -1;
  std::vector<uint8_t> matched2((size_t)n2, 0);  // never set by the reference (:677), kept for the gate at :724
  int nmatches = 0;
  std::vector<int> rotHist[kHistoLength];
  for_equal_nodes(fv1, fv2, [&](int a, int b) {
    for (int i1 = fv1.start[a]; i1 < fv1.start[a + 1]; ++i1) {
      const int idx1 = fv1.idx[i1];
      if (has_mp1[idx1]) continue;  // :699-703
      const bool bStereo1 = ur1 && ur1[idx1] >= 0;
      if (only_stereo && !bStereo1) continue;
      const KeyPoint& kp1 = kps1[idx1];
      int bestDist = kThLow, bestIdx2 = -1;
      for (int i2 = fv2.start[b]; i2 < fv2.start[b + 1]; ++i2) {
        const int idx2 = fv2.idx[i2];
        if (matched2[idx2] || has_mp2[idx2]) continue;
        const bool bStereo2 = ur2 && ur2[idx2] >= 0;
        if (only_stereo && !bStereo2) continue;
        const int dist = descriptor_distance(d1 + (size_t)idx1 * 32, d2 + (size_t)idx2 * 32);
        if (dist > kThLow || dist > bestDist) continue;  // :737 (an equal distance replaces the best)
        const KeyPoint& kp2 = kps2[idx2];
        if (!bStereo1 && !bStereo2) {
          const float distex = ex - kp2.x, distey = ey - kp2.y;
          if (distex * distex + distey * distey < 100 * scale_factors2[kp2.octave]) continue;  // :742-748
        }
        if (check_dist_epipolar_line(kp1, kp2, F12, level_sigma2_2)) { bestIdx2 = idx2; bestDist = dist; }
      }
      if (bestIdx2 >= 0) {
        match12[idx1] = bestIdx2;
        ++nmatches;
        if (check_orientation) rotHist[rot_bin(kp1.angle, kps2[bestIdx2].angle)].push_back(idx1);
      }
    }
  });
  if (check_orientation)
    nmatches -= prune_rotation(rotHist, [&](int idx1) { match12[idx1] = -1; return 1; });  // :790-808
  return nmatches;
}

int search_for_initialization(const KeyPoint* kps1, const uint8_t* d1, int n1, const float* prev_matched,
                              const KeyPoint* kps2, const uint8_t* d2, int n2, const FrameGrid& grid2,
                              int window, float nnratio, bool check_orientation, int32_t* match12) {
  int nmatches = 0;
  for (int i = 0; i < n1; ++i) match12[i] = -1;
  std::vector<int> rotHist[kHistoLength];
  std::vector<int> vMatchedDistance((size_t)n2, INT_MAX), vnMatches21((size_t)n2, -1);
  std::vector<int> cand;
  for (int i1 = 0; i1 < n1; ++i1) {
    const int level1 = kps1[i1].octave;
    if (level1 > 0) continue;  // :421-423
    grid2.features_in_area(kps2, prev_matched[2 * i1], prev_matched[2 * i1 + 1], (float)window, level1, level1, cand);
    if (cand.empty()) continue;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : cand) {
      const int dist = descriptor_distance(d1 + (size_t)i1 * 32, d2 + (size_t)i2 * 32);
      if (vMatchedDistance[i2] <= dist) continue;  // :444-445
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) { bestDist2 = dist; }
    }
    if (bestDist <= kThLow && bestDist < (float)bestDist2 * nnratio) {  // :459-461
      if (vnMatches21[bestIdx2] >= 0) { match12[vnMatches21[bestIdx2]] = -1; --nmatches; }  // steal-back :463-467
      match12[i1] = bestIdx2;
      vnMatches21[bestIdx2] = i1;
      vMatchedDistance[bestIdx2] = bestDist;
      ++nmatches;
      if (check_orientation) rotHist[rot_bin(kps1[i1].angle, kps2[bestIdx2].angle)].push_back(i1);
    }
  }
  if (check_orientation)
    nmatches -= prune_rotation(rotHist, [&](int idx1) {  // :486-509: entries stolen earlier are already -1
      if (match12[idx1] >= 0) { match12[idx1] = -1; return 1; }
      return 0;
    });
  return nmatches;
}

void stereo_match(const KeyPoint* kl, const uint8_t* dl, int nl, const KeyPoint* kr,
                  const uint8_t* dr, int nr, int n_rows, const float* sf, float min_d,
                  float max_d, int th_high, int32_t* best_idx_r, int32_t* best_dist) {
  std::vector<std::vector<int>> rows(n_rows);
  for (int i = 0; i < nr; ++i) {  // Frame.cc:483-493
    const float y = kr[i].y, r = 2.0f * sf[kr[i].octave];
    const int maxr = (int)std::ceil(y + r), minr = (int)std::floor(y - r);
    for (int yi = minr; yi <= maxr; ++yi)
      if (yi >= 0 && yi < n_rows) rows[yi].push_back(i);  // (the reference does not bounds-check)
  }
  for (int i = 0; i < nl; ++i) {
    best_idx_r[i] = -1;
    best_dist[i] = th_high;
    const int level = kl[i].octave;
    const float vL = kl[i].y, uL = kl[i].x;
    const int row = (int)vL;
    if (row < 0 || row >= n_rows) continue;
    const std::vector<int>& cand = rows[row];
    if (cand.empty()) continue;
    const float minU = uL - max_d, maxU = uL - min_d;
    if (maxU < 0) continue;
    int bd = th_high, bi = -1;
    for (int ir : cand) {
      if (kr[ir].octave < level - 1 || kr[ir].octave > level + 1) continue;
      const float uR = kr[ir].x;
      if (uR >= minU && uR <= maxU) {
        const int d = descriptor_distance(dl + (size_t)i * 32, dr + (size_t)ir * 32);
        if (d < bd) { bd = d; bi = ir; }
      }
    }
    best_idx_r[i] = bi;
    best_dist[i] = bd;
  }
}

int stereo_refine(const KeyPoint* kl, int nl, const KeyPoint* kr, const int32_t* best_idx_r,
                  const int32_t* best_dist, const Image* pyr_l, const Image* pyr_r,
                  const float* sf, const float* inv_sf, float minD, float maxD, float mbf,
                  float* u_right, float* depth) {
  const int thOrbDist = (100 + 50) / 2;  // (TH_HIGH+TH_LOW)/2, Frame.cc:471
  int out_of_image = 0;
  std::vector<std::pair<int, int>> vDistIdx;
  for (int iL = 0; iL < nl; ++iL) { u_right[iL] = -1.0f; depth[iL] = -1.0f; }
  for (int iL = 0; iL < nl; ++iL) {
    if (best_idx_r[iL] < 0 || !(best_dist[iL] < thOrbDist)) continue;  // :552
    const KeyPoint& kpL = kl[iL];
    const float uL = kpL.x;
    const float uR0 = kr[best_idx_r[iL]].x;
    const float scaleFactor = inv_sf[kpL.octave];
    const float scaleduL = std::round(kpL.x * scaleFactor);
    const float scaledvL = std::round(kpL.y * scaleFactor);
    const float scaleduR0 = std::round(uR0 * scaleFactor);
    const int w = 5, L = 5;
    const Image& imL = pyr_l[kpL.octave];
    const Image& imR = pyr_r[kpL.octave];
    const int r0 = (int)(scaledvL - w), r1 = (int)(scaledvL + w + 1);
    const int c0 = (int)(scaleduL - w), c1 = (int)(scaleduL + w + 1);
    if (r0 < 0 || r1 > imL.h || c0 < 0 || c1 > imL.w) { ++out_of_image; continue; }  // cv::Mat::rowRange asserts
    const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
    if (iniu < 0 || endu >= imR.w) continue;  // :575-578
    if ((int)(scaleduR0 - L - w) < 0 || r1 > imR.h) { ++out_of_image; continue; }
    float IL[11][11];
    const float cl = (float)imL.row(r0 + w)[c0 + w];
    for (int r = 0; r < 11; ++r)
      for (int c = 0; c < 11; ++c) IL[r][c] = (float)imL.row(r0 + r)[c0 + c] - cl;  // :564-566
    int bestDist = INT_MAX, bestincR = 0;
    float vDists[2 * 5 + 1];
    for (int incR = -L; incR <= +L; ++incR) {
      const int cr0 = (int)(scaleduR0 + incR - w);
      const float cr = (float)imR.row(r0 + w)[cr0 + w];
      double acc = 0;  // cv::norm(IL, IR, NORM_L1) on CV_32F accumulates in double; the terms are integers
      for (int r = 0; r < 11; ++r)
        for (int c = 0; c < 11; ++c) acc += std::fabs(IL[r][c] - ((float)imR.row(r0 + r)[cr0 + c] - cr));
      const float dist = (float)acc;
      if (dist < bestDist) { bestDist = (int)dist; bestincR = incR; }
      vDists[L + incR] = dist;
    }
    if (bestincR == -L || bestincR == L) continue;
    const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
    const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
    if (deltaR < -1 || deltaR > 1) continue;
    float bestuR = sf[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
    float disparity = (uL - bestuR);
    if (disparity >= minD && disparity < maxD) {
      if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
      depth[iL] = mbf / disparity;
      u_right[iL] = bestuR;
      vDistIdx.push_back(std::pair<int, int>(bestDist, iL));
    }
  }
  if (vDistIdx.empty()) return out_of_image;  // (the reference indexes vDistIdx[0] regardless, :627)
  std::sort(vDistIdx.begin(), vDistIdx.end());
  const float median = vDistIdx[vDistIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;
  for (int i = (int)vDistIdx.size() - 1; i >= 0; --i) {
    if (vDistIdx[i].first < thDist) break;
    u_right[vDistIdx[i].second] = -1;
    depth[vDistIdx[i].second] = -1;
  }
  return out_of_image;
}

void voc_transform_feature(const VocTree& t, const uint8_t* feature, int levelsup, uint32_t* word_id,
                           double* weight, uint32_t* nid) {
  const int nid_level = t.depth - levelsup;
  if (nid_level <= 0 && nid != NULL) *nid = 0;  // root
  int final_id = 0, current_level = 0;
  do {
    ++current_level;
    const int32_t* nodes = t.children + t.child_begin[final_id];
    const int count = t.child_begin[final_id + 1] - t.child_begin[final_id];
    final_id = nodes[0];
    double best_d = descriptor_distance(feature, t.desc + (size_t)final_id * 32);  // FORB::distance
    for (int c = 1; c < count; ++c) {
      const int id = nodes[c];
      const double d = descriptor_distance(feature, t.desc + (size_t)id * 32);
      if (d < best_d) { best_d = d; final_id = id; }
    }
    if (nid != NULL && current_level == nid_level) *nid = (uint32_t)final_id;
  } while (t.child_begin[final_id + 1] > t.child_begin[final_id]);  // !isLeaf()
  *word_id = (uint32_t)t.word_id[final_id];
  *weight = t.weight[final_id];
}

int voc_transform(const VocTree& t, const uint8_t* features, int n, int levelsup, uint32_t* bow_id,
                  double* bow_val, uint32_t* feature_node) {
  std::map<uint32_t, double> v;  // BowVector
  for (int i = 0; i < n; ++i) {
    uint32_t id = 0, nid = 0;
    double w = 0;
    voc_transform_feature(t, features + (size_t)i * 32, levelsup, &id, &w, &nid);
    feature_node[i] = 0xffffffffu;
    if (w > 0) {  // not stopped
      std::map<uint32_t, double>::iterator it = v.find(id);  // BowVector::addWeight (BowVector.cpp:34-46)
      if (it != v.end()) it->second += w; else v[id] = w;
      feature_node[i] = nid;  // FeatureVector::addFeature(nid, i_feature)
    }
  }
  double norm = 0.0;  // BowVector::normalize(L1) (BowVector.cpp:62-84)
  for (auto& kv : v) norm += std::fabs(kv.second);
  if (norm > 0.0)
    for (auto& kv : v) kv.second /= norm;
  int k = 0;
  for (auto& kv : v) { bow_id[k] = kv.first; bow_val[k] = kv.second; ++k; }
  return k;
}

int distinctive_descriptor(const uint8_t* desc, int n) {
  if (n <= 0) return -1;
  const size_t N = (size_t)n;
  std::vector<float> Distances(N * N);  // float Distances[N][N] (:275)
  for (size_t i = 0; i < N; i++) {
    Distances[i * N + i] = 0;
    for (size_t j = i + 1; j < N; j++) {
      const int distij = descriptor_distance(desc + i * 32, desc + j * 32);
      Distances[i * N + j] = distij;
      Distances[j * N + i] = distij;
    }
  }
  int BestMedian = INT_MAX, BestIdx = 0;
  for (size_t i = 0; i < N; i++) {
    std::vector<int> vDists(Distances.begin() + i * N, Distances.begin() + (i + 1) * N);
    std::sort(vDists.begin(), vDists.end());
    const int median = vDists[0.5 * (N - 1)];
    if (median < BestMedian) { BestMedian = median; BestIdx = (int)i; }
  }
  return BestIdx;
}

}  // namespace orbo
