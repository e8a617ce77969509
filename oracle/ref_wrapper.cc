// TEST INFRASTRUCTURE.  Flat C entry points around the reference's ORB_SLAM2::ORBextractor
// (include/ORBextractor.h, src/ORBextractor.cc -- compiled unmodified from /root/reference by
// oracle/Makefile into oracle/_ref/libref_orb.so).
#include <thread>
#include <vector>

#include "ORBextractor.h"

struct RefHandle {
  ORB_SLAM2::ORBextractor ex;
  std::vector<cv::KeyPoint> kps;
  cv::Mat desc;
  RefHandle(int n, float s, int l, int a, int b) : ex(n, s, l, a, b) {}
};

extern "C" void ref_arena_reset();

static int run(RefHandle* h, const uint8_t* gray, int w, int ht, size_t stride) {
  ref_arena_reset();
  cv::Mat img(ht, w, CV_8UC1, (void*)gray, stride);
  h->kps.clear();
  h->desc.release();
  h->ex(img, cv::Mat(), h->kps, h->desc);
  return (int)h->kps.size();
}

extern "C" {
RefHandle* ref_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
  return new RefHandle(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void ref_extractor_destroy(RefHandle* h) { delete h; }
int ref_extractor_run(RefHandle* h, const uint8_t* gray, int w, int ht, size_t stride) {
  return run(h, gray, w, ht, stride);
}
void ref_extractor_results(RefHandle* h, cv::KeyPoint* kps, uint8_t* desc) {
  for (size_t i = 0; i < h->kps.size(); ++i) {
    kps[i] = h->kps[i];
    std::memcpy(desc + i * 32, h->desc.ptr((int)i), 32);
  }
}
void ref_extractor_tables(RefHandle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2) {
  std::vector<float> a = h->ex.GetScaleFactors(), b = h->ex.GetInverseScaleFactors(),
                     c = h->ex.GetScaleSigmaSquares(), d = h->ex.GetInverseScaleSigmaSquares();
  for (size_t i = 0; i < a.size(); ++i) { scale[i] = a[i]; inv_scale[i] = b[i]; sigma2[i] = c[i]; inv_sigma2[i] = d[i]; }
}
void ref_extractor_level_size(RefHandle* h, int level, int* w, int* ht) {
  *w = h->ex.mvImagePyramid[level].cols; *ht = h->ex.mvImagePyramid[level].rows;
}
void ref_extractor_level_image(RefHandle* h, int level, uint8_t* dst) {
  const cv::Mat& m = h->ex.mvImagePyramid[level];
  for (int y = 0; y < m.rows; ++y) std::memcpy(dst + (size_t)y * m.cols, m.ptr(y), m.cols);
}
long ref_extract_batch(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh,
                       const uint8_t* frames, int nframes, int w, int ht, size_t stride,
                       size_t frame_stride, int threads, int* counts) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  std::vector<long> totals(threads, 0);
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t]() {
      RefHandle h(nfeatures, scaleFactor, nlevels, iniTh, minTh);
      for (int f = t; f < nframes; f += threads) {
        int n = run(&h, frames + (size_t)f * frame_stride, w, ht, stride);
        if (counts) counts[f] = n;
        totals[t] += n;
      }
    });
  for (auto& th : pool) th.join();
  long total = 0;
  for (long v : totals) total += v;
  return total;
}
}

#ifdef REF_SEQ_ARENA
// Second build of the same unmodified reference (oracle/_ref/libref_orb_seq.so): std::list
// nodes of ExtractorNode come from a per-thread bump arena that never reuses memory within one
// extractor call, so node ADDRESS order == node CREATION order.  The reference's
// (nKeys, pointer) sort (ORBextractor.cc:684) then becomes exactly the canonical
// (nKeys, creation sequence) rule of SURVEY.md Appendix C, which lets the restated oracle be
// checked against the reference's literal control flow without touching its source.
#include <cstdlib>
#include <new>
#include <sys/mman.h>

namespace {
const size_t kNodeBytes = sizeof(std::_List_node<ORB_SLAM2::ExtractorNode>);
const size_t kArenaBytes = size_t(1) << 30;  // virtual reservation, touched lazily
struct Arena {
  char* base = nullptr;
  size_t used = 0;
  void* take(size_t n) {
    if (!base) {
      base = (char*)mmap(nullptr, kArenaBytes, PROT_READ | PROT_WRITE,
                         MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
      if (base == MAP_FAILED) std::abort();
    }
    n = (n + 15) & ~size_t(15);
    if (used + n > kArenaBytes) std::abort();
    void* p = base + used;
    used += n;
    return p;
  }
  bool owns(void* p) const { return base && (char*)p >= base && (char*)p < base + kArenaBytes; }
};
thread_local Arena g_arena;
}  // namespace

void* operator new(size_t n) {
  if (n == kNodeBytes) return g_arena.take(n);
  void* p = std::malloc(n ? n : 1);
  if (!p) throw std::bad_alloc();
  return p;
}
void operator delete(void* p) noexcept {
  if (!p || g_arena.owns(p)) return;
  std::free(p);
}
void operator delete(void* p, size_t) noexcept { operator delete(p); }
extern "C" void ref_arena_reset() { g_arena.used = 0; }
#else
extern "C" void ref_arena_reset() {}
#endif
