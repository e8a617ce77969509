// TEST INFRASTRUCTURE (CPU oracle) -- see cvprim.h.  Arithmetic pinned to cv2 4.13.0.
#include "cvprim.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace orbo {

int cv_round(double v) { return (int)std::lrint(v); }
int cv_floor(double v) {
  int i = (int)v;
  return i - (i > v);
}
int cv_ceil(double v) {
  int i = (int)v;
  return i + (i < v);
}

static inline int16_t sat_short_rint(float v) {
  long r = std::lrintf(v);
  return (int16_t)std::min<long>(32767, std::max<long>(-32768, r));
}

// ---------------------------------------------------------------------------------------------
// resize, INTER_LINEAR, 8UC1.  Model: SURVEY.md Appendix A.1 (call site ORBextractor.cc:1120).
// ---------------------------------------------------------------------------------------------
void resize_axis_table(int S, int D, int* ofs, int16_t* a0, int16_t* a1) {
  const double inv_scale = (double)D / (double)S;
  const double scale = 1.0 / inv_scale;
  for (int d = 0; d < D; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = cv_floor(f);
    f -= (float)s;
    if (s < 0) { f = 0.f; s = 0; }
    if (s >= S - 1) { f = 0.f; s = S - 1; }
    ofs[d] = s;
    a0[d] = sat_short_rint((1.f - f) * 2048.f);
    a1[d] = sat_short_rint(f * 2048.f);
  }
}

void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride,
                      uint8_t* dst, int dw, int dh, size_t dstride) {
  std::vector<int> xo(dw), yo(dh);
  std::vector<int16_t> ax0(dw), ax1(dw), ay0(dh), ay1(dh);
  resize_axis_table(sw, dw, xo.data(), ax0.data(), ax1.data());
  resize_axis_table(sh, dh, yo.data(), ay0.data(), ay1.data());
  std::vector<int> r0(dw), r1(dw);
  int cached0 = -1, cached1 = -1;
  auto hrow = [&](int sy, std::vector<int>& out) {
    const uint8_t* s = src + (size_t)sy * sstride;
    for (int d = 0; d < dw; ++d) {
      int x0 = xo[d], x1 = std::min(x0 + 1, sw - 1);
      out[d] = (int)s[x0] * ax0[d] + (int)s[x1] * ax1[d];
    }
  };
  for (int y = 0; y < dh; ++y) {
    int sy0 = yo[y], sy1 = std::min(sy0 + 1, sh - 1);
    if (cached1 == sy0) { std::swap(r0, r1); std::swap(cached0, cached1); }
    if (cached0 != sy0) { hrow(sy0, r0); cached0 = sy0; }
    if (cached1 != sy1) {
      if (sy1 == sy0) { r1 = r0; } else { hrow(sy1, r1); }
      cached1 = sy1;
    }
    const int b0 = ay0[y], b1 = ay1[y];
    uint8_t* o = dst + (size_t)y * dstride;
    for (int d = 0; d < dw; ++d)
      o[d] = (uint8_t)((((b0 * (r0[d] >> 4)) >> 16) + ((b1 * (r1[d] >> 4)) >> 16) + 2) >> 2);
  }
}

// ---------------------------------------------------------------------------------------------
// borders
// ---------------------------------------------------------------------------------------------
static inline int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) {
    if (p < 0) p = -p;
    else p = 2 * (len - 1) - p;
  }
  return p;
}

void copy_make_border_reflect101_u8(const uint8_t* src, int w, int h, size_t sstride,
                                    uint8_t* dst, size_t dstride, int top, int bottom,
                                    int left, int right) {
  const int W = w + left + right, H = h + top + bottom;
  for (int y = 0; y < H; ++y) {
    const uint8_t* s = src + (size_t)reflect101(y - top, h) * sstride;
    uint8_t* d = dst + (size_t)y * dstride;
    for (int x = 0; x < W; ++x) d[x] = s[reflect101(x - left, w)];
  }
}

// ---------------------------------------------------------------------------------------------
// GaussianBlur 7x7 sigma 2, 8UC1, REFLECT_101.  Model: SURVEY.md Appendix A.2 (:1086).
// ---------------------------------------------------------------------------------------------
void gaussian_blur7_u8(const uint8_t* src, int w, int h, size_t sstride,
                       uint8_t* dst, size_t dstride) {
  static const int k[7] = {18, 34, 48, 56, 48, 34, 18};
  std::vector<uint32_t> hbuf((size_t)w * h);
  std::vector<int> xi(w + 6);
  for (int x = -3; x < w + 3; ++x) xi[x + 3] = reflect101(x, w);
  for (int y = 0; y < h; ++y) {
    const uint8_t* s = src + (size_t)y * sstride;
    uint32_t* hr = &hbuf[(size_t)y * w];
    for (int x = 0; x < w; ++x) {
      uint32_t a = 0;
      for (int i = 0; i < 7; ++i) a += (uint32_t)k[i] * s[xi[x + i]];
      hr[x] = a;
    }
  }
  for (int y = 0; y < h; ++y) {
    const uint32_t* rows[7];
    for (int j = 0; j < 7; ++j) rows[j] = &hbuf[(size_t)reflect101(y + j - 3, h) * w];
    uint8_t* d = dst + (size_t)y * dstride;
    for (int x = 0; x < w; ++x) {
      uint32_t v = 0;
      for (int j = 0; j < 7; ++j) v += (uint32_t)k[j] * rows[j][x];
      d[x] = (uint8_t)((v + 32768u) >> 16);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// FAST-9/16 with 3x3 non-max suppression.  Model: SURVEY.md Appendix A.3 (:809, :814).
// ---------------------------------------------------------------------------------------------
static const int kCircle[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},  {3, 0},  {3, -1},
                                   {2, -2}, {1, -3},  {0, -3},  {-1, -3}, {-2, -2}, {-3, -1},
                                   {-3, 0}, {-3, 1},  {-2, 2},  {-1, 3}};

int fast9_best(const uint8_t* p, size_t stride) {
  int d[25];
  const int v = p[0];
  for (int k = 0; k < 16; ++k)
    d[k] = v - (int)p[(ptrdiff_t)kCircle[k][1] * (ptrdiff_t)stride + kCircle[k][0]];
  for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
  int best = -1000;
  for (int k = 0; k < 16; ++k) {
    int mn = d[k], mx = d[k];
    for (int j = 1; j < 9; ++j) {
      mn = std::min(mn, d[k + j]);
      mx = std::max(mx, d[k + j]);
    }
    best = std::max(best, std::max(mn, -mx));
  }
  return best;
}

void fast9_nms(const uint8_t* img, int w, int h, size_t stride, int threshold,
               std::vector<FastPoint>& out) {
  out.clear();
  if (w < 7 || h < 7) return;
  // score map (0 = not a corner at this threshold), then strict 8-neighbour maximum.
  static thread_local std::vector<int> sc;
  sc.assign((size_t)w * h, 0);
  const ptrdiff_t st = (ptrdiff_t)stride;
  for (int y = 3; y < h - 3; ++y) {
    const uint8_t* row = img + (size_t)y * stride;
    for (int x = 3; x < w - 3; ++x) {
      // Any 9 contiguous circle pixels contain one of each opposite pair (0,8), (4,12), (2,10),
      // (6,14): reject unless all four pairs have a member brighter (or all darker) than +-t.
      const uint8_t* p = row + x;
      const int v = p[0], lo = v - threshold, hi = v + threshold;
      const int c0 = p[3 * st], c8 = p[-3 * st], c4 = p[3], c12 = p[-3];
      const bool dark = (c0 < lo || c8 < lo) && (c4 < lo || c12 < lo);
      const bool bright = (c0 > hi || c8 > hi) && (c4 > hi || c12 > hi);
      if (!dark && !bright) continue;
      const int c2 = p[2 * st + 2], c10 = p[-2 * st - 2], c6 = p[-2 * st + 2], c14 = p[2 * st - 2];
      const bool dark2 = dark && (c2 < lo || c10 < lo) && (c6 < lo || c14 < lo);
      const bool bright2 = bright && (c2 > hi || c10 > hi) && (c6 > hi || c14 > hi);
      if (!dark2 && !bright2) continue;
      int b = fast9_best(p, stride);
      if (b > threshold) sc[(size_t)y * w + x] = b - 1;
    }
  }
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) {
      const int s = sc[(size_t)y * w + x];
      if (s <= 0) continue;  // a score-0 "corner" (threshold 0) is never a strict maximum
      bool keep = true;
      for (int dy = -1; dy <= 1 && keep; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          if (!dx && !dy) continue;
          if (sc[(size_t)(y + dy) * w + (x + dx)] >= s) { keep = false; break; }
        }
      if (keep) out.push_back({x, y, s});
    }
}

// ---------------------------------------------------------------------------------------------
// fastAtan2.  Model: SURVEY.md Appendix A.4 (:103).  Every operation rounds to float32
// (build with -ffp-contract=off); this form is bit-identical to cv2.fastAtan2 (probe:
// 0 mismatches on 2e5 random int-valued inputs).
// ---------------------------------------------------------------------------------------------
float fast_atan2_deg(float y, float x) {
  const float s = (float)(180.0 / 3.14159265358979323846);
  const float p1 = 0.9997878412794807f * s;
  const float p3 = -0.3258083974640975f * s;
  const float p5 = 0.1555786518463281f * s;
  const float p7 = -0.04432655554792128f * s;
  const float eps = (float)DBL_EPSILON;
  const float ax = std::fabs(x), ay = std::fabs(y);
  float a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + eps);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + eps);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

}  // namespace orbo
