// C ABI of the B200-native ORB front-end: handle management, host-side tables and the
// extraction pipeline.  See include/orb_gpu.h for the contract and the reference citations.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <map>
#include <vector>

#include "common.cuh"

namespace orbgpu {

static std::atomic<long long> g_launches{0};
static thread_local std::string g_last_error;

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int use_cached_pool(int device) {
  static std::atomic<unsigned long long> done{0};
  ORB_CUDA_TRY(cudaSetDevice(device));
  if (device >= 0 && device < 64 && !(done.load() >> device & 1ull)) {
    cudaMemPool_t pool;
    ORB_CUDA_TRY(cudaDeviceGetDefaultMemPool(&pool, device));
    unsigned long long keep = ~0ull;  // never trim: the scratch of the matcher calls is a few MB
    ORB_CUDA_TRY(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    done.fetch_or(1ull << device);
  }
  return ORB_GPU_OK;
}
void set_last_error(const char* what, cudaError_t e) {
  g_last_error = std::string(what) + ": " + cudaGetErrorString(e);
}
void set_last_message(const char* msg) { g_last_error = msg; }

static inline int cv_round(double v) { return (int)std::lrint(v); }  // round-half-even, like cvRound
static inline int cv_floor(double v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil(double v) { int i = (int)v; return i + (i < v); }
static inline int pow2ceil(long long v) { int p = 1; while (p < v) p <<= 1; return p; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// cv::resize INTER_LINEAR coefficient table of one axis (OpenCV 4.13 8U path): 11-bit weights.
static void resize_axis_table(int S, int D, int* ofs, short* a0, short* a1) {
  const double scale = 1.0 / ((double)D / (double)S);
  for (int d = 0; d < D; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = cv_floor(f);
    f -= (float)s;
    if (s < 0) { f = 0.f; s = 0; }
    if (s >= S - 1) { f = 0.f; s = S - 1; }
    ofs[d] = s;
    a0[d] = (short)std::lrintf((1.f - f) * 2048.f);
    a1[d] = (short)std::lrintf(f * 2048.f);
  }
}

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  int ensure(size_t need) {
    if (need <= bytes) return ORB_GPU_OK;
    if (p) cudaFree(p);
    p = nullptr; bytes = 0;
    ORB_CUDA_TRY(cudaMalloc(&p, need));
    bytes = need;
    return ORB_GPU_OK;
  }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
};

}  // namespace orbgpu

using namespace orbgpu;

struct OrbGpuExtractor {
  int device = 0;
  cudaStream_t stream = nullptr;
  // constructor arguments and tables (reference src/ORBextractor.cc:410-470)
  int nfeatures = 0, nlevels = 0, ini_th = 0, min_th = 0;
  float scale_factor_f = 0;
  double scale_factor = 0;  // the reference keeps a double member (include/ORBextractor.h:98)
  std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
  std::vector<int> quota;
  int umax[16];
  int max_w = 0, max_h = 0, max_batch = 0;
  // plan of the current geometry
  Plan plan;
  int plan_w = 0, plan_h = 0, plan_batch = 0;
  std::vector<size_t> lvl_off;  // byte offset of each level inside pyr/blur (per batch capacity)
  DevBuf pyr, blur, tables, cell_tab, cell_kp, cell_cnt, keys, pay, sel, sel_cnt, cand_cnt, status;
  DevBuf out_kps, out_desc, out_n, stage;
  Scratch sc;
  int last_nframes = 0;
  bool have_result = false;
  // level 0 of the last call (handle-owned copy, or the caller's device frames)
  const uint8_t* last_l0 = nullptr;
  int last_l0_pitch = 0;
  long long last_l0_stride = 0;
  // optional per-stage CUDA-event timing (bench/roofline)
  bool timing = false;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // host-pointer path: frames go through in chunks, H2D of chunk i+1 and D2H of chunk i-1 overlap
  // the kernels of chunk i (one copy stream per direction, events per chunk)
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaStream_t s_alt[7] = {};  // extra compute streams: chunk c runs on stream c % n
  std::vector<cudaEvent_t> ev_in, ev_done;
  // single-chunk host calls (one frame, a stereo pair, a small batch) are launch-bound: the kernel sequence
  // is captured into a CUDA graph on the second call with the same geometry and buffers and replayed after
  std::vector<long long> graph_key, graph_seen;
  cudaGraphExec_t graph_exec = nullptr;
  int graph_nlaunch = 0;
  int plan_generation = 0;
  // one- to four-frame host calls go through pinned staging blocks (one async copy each way, exact-size unpack)
  void* pin_in = nullptr; size_t pin_in_bytes = 0;
  void* pin_out = nullptr; size_t pin_out_bytes = 0;
  // blur depends only on the pyramid: it runs on a side stream next to FAST + quadtree and joins before the
  // descriptors (one side stream and two events per compute stream, created on first use)
  struct Fork { cudaStream_t side = nullptr; cudaEvent_t forked = nullptr, joined = nullptr; };
  std::map<cudaStream_t, Fork> forks;
};

static bool fork_enabled() {
  static const bool on = [] { const char* e = getenv("ORB_GPU_NO_FORK"); return !(e && atoi(e)); }();
  return on;
}
static int fork_point() {
  static const int at = [] { const char* e = getenv("ORB_GPU_FORK_AT"); return e ? atoi(e) : 1; }();
  return at;
}
// Device-resident batches of at least 2 * 64 frames are split into this many sub-batches on separate streams, so
// the latency-bound stages of one (quadtree) overlap the issue-bound stages of another (ORB_GPU_DEVICE_SPLIT).
static int device_split() {
  static const int n = [] { const char* e = getenv("ORB_GPU_DEVICE_SPLIT"); return e ? atoi(e) : 1; }();
  return std::max(1, std::min(8, n));
}
static int fork_for(OrbGpuExtractor* h, cudaStream_t s, OrbGpuExtractor::Fork** out) {
  auto it = h->forks.find(s);
  if (it == h->forks.end()) {
    OrbGpuExtractor::Fork f;
    ORB_CUDA_TRY(cudaStreamCreateWithFlags(&f.side, cudaStreamNonBlocking));
    ORB_CUDA_TRY(cudaEventCreateWithFlags(&f.forked, cudaEventDisableTiming));
    ORB_CUDA_TRY(cudaEventCreateWithFlags(&f.joined, cudaEventDisableTiming));
    it = h->forks.emplace(s, f).first;
  }
  *out = &it->second;
  return ORB_GPU_OK;
}

// Frames per chunk of the host-pointer path.  Chunk c runs on compute stream c % 4, so the kernels of up to four
// chunks interleave on the SMs: one chunk's launch gaps and partial last waves are filled by its neighbours.
// Small chunks (3/32 of the batch, 16..64 frames) keep the first upload short and the overlap fine-grained;
// measured on 512 TUM frames: 4 x 128 frames on one stream 4.09 ms, 16 x 32 on four streams 3.57 ms; with the
// kernels of the end of round 2 chunks of 48 beat 32 (3.32 vs 3.45 ms: fewer, fuller launches) and 96 loses (3.58).
// ORB_GPU_HOST_CHUNK / ORB_GPU_HOST_STREAMS override.
static int host_chunk(int nb) {
  static const int forced = [] { const char* e = getenv("ORB_GPU_HOST_CHUNK"); return e ? atoi(e) : 0; }();
  if (forced > 0) return forced;
  return std::max(16, std::min(64, (3 * nb + 31) / 32));
}
static int host_streams() {
  static const int n = [] { const char* e = getenv("ORB_GPU_HOST_STREAMS"); return e ? atoi(e) : 4; }();
  return std::max(1, std::min(8, n));
}

static bool graphs_enabled() {
  static const bool on = [] { const char* e = getenv("ORB_GPU_NO_GRAPH"); return !(e && atoi(e)); }();
  return on;
}

// Chunk boundaries of one host-pointer batch: [0, b1, b2, ..., nb].  The first upload is latency the kernels cannot
// hide and the last chunk's kernels and download are latency the uploads cannot hide, so the default schedule ramps
// up at the head and down at the tail.  ORB_GPU_HOST_SCHEDULE="a,b,c/y,z" gives the chunk sizes explicitly for
// experiments: head a, b, then c repeated, then the tail y, z.
static std::vector<int> schedule_from(const std::vector<int>& head, const std::vector<int>& tail, int nb) {
  std::vector<int> b(1, 0);
  int tail_sum = 0;
  for (int t : tail) tail_sum += t;
  for (size_t i = 0; b.back() < nb - tail_sum; ++i)
    b.push_back(std::min(nb - tail_sum, b.back() + std::max(1, head[std::min(i, head.size() - 1)])));
  for (int t : tail)
    if (b.back() < nb) b.push_back(std::min(nb, b.back() + std::max(1, t)));
  if (b.back() < nb) b.push_back(nb);
  return b;
}
static std::vector<int> host_schedule(int nb) {
  static const std::pair<std::vector<int>, std::vector<int>> forced = [] {
    std::pair<std::vector<int>, std::vector<int>> v;
    const char* e = getenv("ORB_GPU_HOST_SCHEDULE");
    bool in_tail = false;
    while (e && *e) {
      (in_tail ? v.second : v.first).push_back(atoi(e));
      const char* c = strpbrk(e, ",/");
      if (!c) break;
      if (*c == '/') in_tail = true;
      e = c + 1;
    }
    return v;
  }();
  if (!forced.first.empty()) {
    int tail_sum = 0;
    for (int t : forced.second) tail_sum += t;
    return schedule_from(forced.first, tail_sum < nb ? forced.second : std::vector<int>(), nb);
  }
  // default: chunks of c frames with a third-size first chunk (the kernels start after a third of the upload time) and
  // a half-size last one (half the tail behind the last upload); measured on 512 TUM frames (head, body / tail):
  // 16,48/24 3.32 ms; 16,48/16 3.35; 16,64/16 3.33; 24,48/24 3.33; 16,32/16 3.45; 8,16,32/16,8 3.47; 16,96/16 3.58
  const int c = host_chunk(nb);
  if (nb >= 4 * c && c >= 16) return schedule_from({std::max(8, c / 3), c}, {c / 2}, nb);
  std::vector<int> b(1, 0);
  while (b.back() < nb) b.push_back(std::min(nb, b.back() + c));
  return b;
}

// Plan / scratch of the frames [f0, ...) of a batch: every per-frame array advanced by f0 frames.
static Plan plan_at(const Plan& p, int f0) {
  Plan q = p;
  for (int l = 0; l < p.nlevels; ++l) {
    q.lv[l].img += (size_t)f0 * p.lv[l].frame_stride;
    q.lv[l].blur += (size_t)f0 * p.lv[l].bframe_stride;
  }
  return q;
}
static Scratch scratch_at(const Scratch& sc, const Plan& p, int f0) {
  Scratch q = sc;
  q.cell_kp += (size_t)f0 * p.cells_per_frame * p.cell_cap;
  q.cell_cnt += (size_t)f0 * p.cells_per_frame;
  q.keys += (size_t)f0 * p.keys_per_frame;
  q.pay += (size_t)f0 * p.keys_per_frame;
  q.sel += (size_t)f0 * p.slots_per_frame;
  q.sel_cnt += (size_t)f0 * p.nlevels;
  q.cand_cnt += (size_t)f0 * p.nlevels;
  return q;
}

static int build_tables(OrbGpuExtractor* h) {
  const int L = h->nlevels;
  h->scale.assign(L, 1.0f);
  h->sigma2.assign(L, 1.0f);
  for (int i = 1; i < L; ++i) {
    h->scale[i] = (float)((double)h->scale[i - 1] * h->scale_factor);  // float * double member
    h->sigma2[i] = h->scale[i] * h->scale[i];
  }
  h->inv_scale.resize(L);
  h->inv_sigma2.resize(L);
  for (int i = 0; i < L; ++i) {
    h->inv_scale[i] = 1.0f / h->scale[i];
    h->inv_sigma2[i] = 1.0f / h->sigma2[i];
  }
  h->quota.assign(L, 0);
  const float factor = (float)(1.0 / h->scale_factor);
  float desired = (float)h->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)L));
  int sum = 0;
  for (int l = 0; l < L - 1; ++l) {
    h->quota[l] = cv_round(desired);
    sum += h->quota[l];
    desired *= factor;
  }
  h->quota[L - 1] = std::max(h->nfeatures - sum, 0);
  // rows of the circular patch (:454-469)
  const int vmax = cv_floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
  const int vmin = cv_ceil(kHalfPatch * std::sqrt(2.f) / 2);
  const double hp2 = kHalfPatch * kHalfPatch;
  for (int v = 0; v < 16; ++v) h->umax[v] = 0;
  for (int v = 0; v <= vmax; ++v) h->umax[v] = cv_round(std::sqrt(hp2 - v * v));
  for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
    while (h->umax[v0] == h->umax[v0 + 1]) ++v0;
    h->umax[v] = v0;
    ++v0;
  }
  return ORB_GPU_OK;
}

// Builds the plan for (w, h, batch): level geometry, resize tables, scratch layout; grows the
// device buffers when needed.  Level 0 of the plan points at the handle's own level-0 buffer;
// the device-input entry point overrides it with the caller's frames.
static int ensure_plan(OrbGpuExtractor* h, int w, int ht, int batch) {
  if (h->plan_w == w && h->plan_h == ht && batch <= h->plan_batch) return ORB_GPU_OK;
  const int L = h->nlevels;
  Plan p;
  std::memset(&p, 0, sizeof(p));
  p.nlevels = L; p.width = w; p.height = ht;
  p.ini_th = h->ini_th; p.min_th = h->min_th;
  for (int v = 0; v < 16; ++v) p.umax[v] = h->umax[v];
  size_t table_bytes = 0, pyr_bytes = 0;
  int cells = 0, slots = 0, cell_cap = 1;
  long long keys = 0;
  h->lvl_off.assign(L, 0);
  std::vector<size_t> tab_off(L, 0);
  for (int l = 0; l < L; ++l) {
    LevelGeom& g = p.lv[l];
    g.w = cv_round((float)w * h->inv_scale[l]);
    g.h = cv_round((float)ht * h->inv_scale[l]);
    g.pitch = (int)align_up((size_t)g.w, 64);
    g.frame_stride = (long long)g.pitch * g.h;
    g.bpitch = g.pitch; g.bframe_stride = g.frame_stride;
    g.max_bx = g.w - kBorder; g.max_by = g.h - kBorder;
    g.box_w = g.max_bx - kBorder; g.box_h = g.max_by - kBorder;
    if (g.box_w < 30 || g.box_h < 30) return ORB_GPU_ERR_ARG;  // nCols/nRows would be 0 (:781-782)
    const float width = (float)g.box_w, height = (float)g.box_h;
    g.ncols = (int)(width / 30.f); g.nrows = (int)(height / 30.f);
    g.wcell = (int)std::ceil(width / g.ncols); g.hcell = (int)std::ceil(height / g.nrows);
    if (g.wcell + 6 > 250 || g.hcell + 6 > 250) return ORB_GPU_ERR_ARG;
    g.cell_base = cells;
    cells += g.ncols * g.nrows;
    cell_cap = std::max(cell_cap, ((g.wcell + 1) / 2) * ((g.hcell + 1) / 2));
    g.n_ini = (int)std::round(width / height);  // :543
    if (g.n_ini < 1 || g.n_ini > 255) return ORB_GPU_ERR_ARG;  // portrait shapes: nIni == 0
    g.hx = width / (float)g.n_ini;
    g.quota = h->quota[l];
    // candidate capacity of the level: one per 16 pixels (the synthetic textured frames have one per ~150), at most
    // 2^18 -- a 4096 x 4096 level 0 with one candidate per 64 pixels still fits (keys and payload: 12 bytes each)
    g.key_cap = pow2ceil(std::min<long long>(262144, std::max<long long>(1024, (long long)g.w * g.h / 16)));
    g.key_base = keys;
    keys += g.key_cap;
    g.slot_base = slots;
    slots += g.quota + 3;
    g.scale = h->scale[l];
    g.patch_size = (float)(int)(31.f * h->scale[l]);  // const int scaledPatchSize = 31*scale (:837)
    h->lvl_off[l] = pyr_bytes;
    pyr_bytes += align_up((size_t)g.frame_stride * batch, 256);
    tab_off[l] = table_bytes;
    table_bytes += align_up((size_t)g.w * 8 + (size_t)g.h * 8 + (l > 0 ? (size_t)p.lv[l - 1].h * 2 : 0), 16);
    if (quadtree_smem_bytes(p, l) > 227 * 1024) return ORB_GPU_ERR_CAPACITY;
  }
  p.cells_per_frame = cells;
  p.cell_cap = cell_cap;
  p.keys_per_frame = keys;
  p.slots_per_frame = slots;

  cudaStreamSynchronize(h->stream);
  int rc;
  if ((rc = h->pyr.ensure(pyr_bytes)) || (rc = h->blur.ensure(pyr_bytes)) ||
      (rc = h->tables.ensure(table_bytes)) || (rc = h->cell_tab.ensure((size_t)cells * sizeof(uint2))) ||
      (rc = h->cell_kp.ensure((size_t)batch * cells * cell_cap * 4)) ||
      (rc = h->cell_cnt.ensure((size_t)batch * cells * 4)) ||
      (rc = h->keys.ensure((size_t)batch * keys * 8)) || (rc = h->pay.ensure((size_t)batch * keys * 4)) ||
      (rc = h->sel.ensure((size_t)batch * slots * 4)) || (rc = h->sel_cnt.ensure((size_t)batch * L * 4)) ||
      (rc = h->cand_cnt.ensure((size_t)batch * L * 4)) || (rc = h->status.ensure(16)))
    return rc;

  // resize tables: per level [xofs int][yofs int][ax0,ax1 short][ay0,ay1 short]
  std::vector<uint8_t> tab(table_bytes, 0);
  for (int l = 0; l < L; ++l) {
    LevelGeom& g = p.lv[l];
    g.img = (uint8_t*)h->pyr.p + h->lvl_off[l];
    g.blur = (uint8_t*)h->blur.p + h->lvl_off[l];
    uint8_t* base = tab.data() + tab_off[l];
    int* xofs = (int*)base; int* yofs = xofs + g.w;
    short* ax0 = (short*)(yofs + g.h); short* ax1 = ax0 + g.w;
    short* ay0 = ax1 + g.w; short* ay1 = ay0 + g.h;
    if (l > 0) {
      resize_axis_table(p.lv[l - 1].w, g.w, xofs, ax0, ax1);
      resize_axis_table(p.lv[l - 1].h, g.h, yofs, ay0, ay1);
      // inverse of yofs for the tensor-core resize: the destination row whose upper source row is s, or -1 (for a
      // scale factor > 1 every source row is the upper tap of at most one destination row)
      short* yinv = ay1 + g.h;
      for (int sy = 0; sy < p.lv[l - 1].h; ++sy) yinv[sy] = -1;
      for (int dy = 0; dy < g.h; ++dy) yinv[yofs[dy]] = (short)dy;
    }
    uint8_t* dbase = (uint8_t*)h->tables.p + tab_off[l];
    g.xofs = (const int*)dbase; g.yofs = g.xofs + g.w;
    g.ax0 = (const short*)(g.yofs + g.h); g.ax1 = g.ax0 + g.w;
    g.ay0 = g.ax1 + g.w; g.ay1 = g.ay0 + g.h;
    g.yinv = l > 0 ? g.ay1 + g.h : nullptr;
  }
  // FAST cell ROIs (:789-806): the skip rules differ for rows (maxBorderY-3) and columns (maxBorderX-6)
  std::vector<uint2> ctab(cells);
  for (int l = 0; l < L; ++l) {
    const LevelGeom& g = p.lv[l];
    for (int i = 0; i < g.nrows; ++i)
      for (int j = 0; j < g.ncols; ++j) {
        const int ini_y = kBorder + i * g.hcell, ini_x = kBorder + j * g.wcell;
        const int max_y = std::min(ini_y + g.hcell + 6, g.max_by), max_x = std::min(ini_x + g.wcell + 6, g.max_bx);
        int cw = max_x - ini_x, ch = max_y - ini_y;
        if (ini_y >= g.max_by - 3 || ini_x >= g.max_bx - 6 || cw < 7 || ch < 7) cw = ch = 0;
        ctab[g.cell_base + i * g.ncols + j] =
            make_uint2((unsigned)ini_x | ((unsigned)ini_y << 16), (unsigned)cw | ((unsigned)ch << 8) | ((unsigned)l << 16));
      }
  }
  ORB_CUDA_TRY(cudaMemcpyAsync(h->cell_tab.p, ctab.data(), ctab.size() * sizeof(uint2), cudaMemcpyHostToDevice, h->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(h->tables.p, tab.data(), table_bytes, cudaMemcpyHostToDevice, h->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(h->stream));  // `tab` dies at scope exit
  h->sc.cell_tab = (const uint2*)h->cell_tab.p;
  h->sc.cell_kp = (uint32_t*)h->cell_kp.p; h->sc.cell_cnt = (uint32_t*)h->cell_cnt.p;
  h->sc.keys = (uint64_t*)h->keys.p; h->sc.pay = (uint32_t*)h->pay.p;
  h->sc.sel = (uint32_t*)h->sel.p; h->sc.sel_cnt = (int*)h->sel_cnt.p;
  h->sc.cand_cnt = (int*)h->cand_cnt.p; h->sc.status = (int*)h->status.p;
  h->plan = p;
  ++h->plan_generation;
  h->plan_w = w; h->plan_h = ht; h->plan_batch = batch;
  h->have_result = false;
  return ORB_GPU_OK;
}

// Runs all stages for `nframes` frames whose level-0 images are described by plan.lv[0].
static int run_pipeline(OrbGpuExtractor* h, const Plan& p, const Scratch& sc, int nframes, OrbKeyPoint* d_kps,
                        uint8_t* d_desc, int capacity, int* d_n_out, cudaStream_t s, bool reset_status = true) {
  int rc;
  if (reset_status) ORB_CUDA_TRY(cudaMemsetAsync(sc.status, 0, sizeof(int), s));
  const bool tm = h->timing;
  if (tm) ORB_CUDA_TRY(cudaEventRecord(h->ev[0], s));
  for (int l = 1; l < p.nlevels; ++l)
    if ((rc = launch_resize_level(p, l, nframes, s))) return rc;
  if (tm) ORB_CUDA_TRY(cudaEventRecord(h->ev[1], s));
  // per-stage timing needs the stages one after another on `s`; otherwise blur (which depends only on the pyramid)
  // forks off onto the side stream, before FAST by default (ORB_GPU_FORK_AT=2: after FAST, next to the quadtree).
  // Measured on 512 TUM frames: 2.614 ms (before FAST), 2.629 (after), 2.648 (no fork) -- the stages are bound by
  // instruction issue, so running them side by side buys ~1 %
  OrbGpuExtractor::Fork* fk = nullptr;
  if (!tm && fork_enabled() && (rc = fork_for(h, s, &fk))) return rc;
  auto fork_blur = [&]() -> int {
    ORB_CUDA_TRY(cudaEventRecord(fk->forked, s));
    ORB_CUDA_TRY(cudaStreamWaitEvent(fk->side, fk->forked, 0));
    const int r = launch_blur(p, nframes, fk->side);
    if (r) return r;
    ORB_CUDA_TRY(cudaEventRecord(fk->joined, fk->side));
    return ORB_GPU_OK;
  };
  if (fk && fork_point() == 1 && (rc = fork_blur())) return rc;
  if ((rc = launch_fast_cells(p, sc, nframes, s))) return rc;
  if (tm) ORB_CUDA_TRY(cudaEventRecord(h->ev[2], s));
  if (fk && fork_point() != 1 && (rc = fork_blur())) return rc;
  if ((rc = launch_quadtree(p, sc, nframes, s))) return rc;
  if (tm) ORB_CUDA_TRY(cudaEventRecord(h->ev[3], s));
  if (fk) ORB_CUDA_TRY(cudaStreamWaitEvent(s, fk->joined, 0));
  else if ((rc = launch_blur(p, nframes, s))) return rc;
  if (tm) ORB_CUDA_TRY(cudaEventRecord(h->ev[4], s));
  if ((rc = launch_orient_desc(p, sc, nframes, d_kps, d_desc, capacity, d_n_out, s))) return rc;
  if (tm) ORB_CUDA_TRY(cudaEventRecord(h->ev[5], s));
  return ORB_GPU_OK;
}

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

const char* orb_gpu_version(void) { return "orb-b200 0.1 (sm_100a)"; }
const char* orb_gpu_last_error(void) { return g_last_error.c_str(); }
int64_t orb_gpu_launch_count(void) { return (int64_t)g_launches.load(); }

int orb_gpu_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                   int max_width, int max_height, int max_batch, int device, OrbGpuExtractor** out) {
  if (!out) return ORB_GPU_ERR_ARG;
  *out = nullptr;
  if (nfeatures < 1 || nfeatures > 60000 || nlevels < 1 || nlevels > kMaxLevels || !(scaleFactor > 1.0f) || scaleFactor > 2.0f ||
      iniThFAST < 1 || iniThFAST > 254 || minThFAST < 1 || minThFAST > iniThFAST || max_width < 1 ||
      max_height < 1 || max_width > kMaxDim || max_height > kMaxDim || max_batch < 1)
    return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(device));
  OrbGpuExtractor* h = new OrbGpuExtractor();
  h->device = device;
  h->nfeatures = nfeatures; h->nlevels = nlevels; h->ini_th = iniThFAST; h->min_th = minThFAST;
  h->scale_factor_f = scaleFactor; h->scale_factor = (double)scaleFactor;
  h->max_w = max_width; h->max_h = max_height; h->max_batch = max_batch;
  build_tables(h);
  cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { set_last_error("cudaStreamCreate", e); delete h; return ORB_GPU_ERR_CUDA; }
  int rc = upload_constants();
  if (!rc && (cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking) != cudaSuccess ||
              cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking) != cudaSuccess ||
              [&] { for (cudaStream_t& a : h->s_alt)
                      if (cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking) != cudaSuccess) return true;
                    return false; }())) {
    set_last_error("copy streams", cudaGetLastError());
    rc = ORB_GPU_ERR_CUDA;
  }
  if (rc) {
    if (h->s_h2d) cudaStreamDestroy(h->s_h2d);
    if (h->s_d2h) cudaStreamDestroy(h->s_d2h);
    for (cudaStream_t a : h->s_alt) if (a) cudaStreamDestroy(a);
    cudaStreamDestroy(h->stream); delete h; return rc;
  }
  *out = h;
  return ORB_GPU_OK;
}

int orb_gpu_destroy(OrbGpuExtractor* h) {
  if (!h) return ORB_GPU_OK;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  cudaStreamSynchronize(h->s_h2d); cudaStreamSynchronize(h->s_d2h);
  for (cudaStream_t a : h->s_alt) if (a) cudaStreamSynchronize(a);
  if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
  if (h->pin_in) cudaFreeHost(h->pin_in);
  if (h->pin_out) cudaFreeHost(h->pin_out);
  for (auto& kv : h->forks) {
    cudaStreamSynchronize(kv.second.side);
    cudaStreamDestroy(kv.second.side);
    cudaEventDestroy(kv.second.forked);
    cudaEventDestroy(kv.second.joined);
  }
  for (cudaEvent_t e : h->ev_in) cudaEventDestroy(e);
  for (cudaEvent_t e : h->ev_done) cudaEventDestroy(e);
  cudaStreamDestroy(h->s_h2d); cudaStreamDestroy(h->s_d2h);
  for (cudaStream_t a : h->s_alt) if (a) cudaStreamDestroy(a);
  DevBuf* bufs[] = {&h->pyr, &h->blur, &h->tables, &h->cell_tab, &h->cell_kp, &h->cell_cnt, &h->keys, &h->pay, &h->sel,
                    &h->sel_cnt, &h->cand_cnt, &h->status, &h->out_kps, &h->out_desc, &h->out_n, &h->stage};
  for (DevBuf* b : bufs) b->release();
  for (int i = 0; i < 6; ++i) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  cudaStreamDestroy(h->stream);
  delete h;
  return ORB_GPU_OK;
}

// Per-stage device times of the LAST extract call on this handle: pyramid, fast, quadtree, blur,
// orient+descriptor (milliseconds, CUDA events on the stream the kernels ran on).
int orb_gpu_set_stage_timing(OrbGpuExtractor* h, int enable) {
  if (!h) return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(h->device));
  if (enable && !h->ev[0])
    for (int i = 0; i < 6; ++i) ORB_CUDA_TRY(cudaEventCreate(&h->ev[i]));
  h->timing = enable != 0;
  return ORB_GPU_OK;
}
int orb_gpu_get_stage_ms(OrbGpuExtractor* h, float* ms5) {
  if (!h || !ms5) return ORB_GPU_ERR_ARG;
  if (!h->timing || !h->have_result) return ORB_GPU_ERR_STATE;
  ORB_CUDA_TRY(cudaSetDevice(h->device));
  ORB_CUDA_TRY(cudaEventSynchronize(h->ev[5]));
  for (int i = 0; i < 5; ++i) ORB_CUDA_TRY(cudaEventElapsedTime(&ms5[i], h->ev[i], h->ev[i + 1]));
  return ORB_GPU_OK;
}

int orb_gpu_get_levels(const OrbGpuExtractor* h) { return h ? h->nlevels : ORB_GPU_ERR_ARG; }
float orb_gpu_get_scale_factor(const OrbGpuExtractor* h) { return h ? (float)h->scale_factor : 0.f; }
int orb_gpu_get_scale_tables(const OrbGpuExtractor* h, float* scale, float* inv_scale, float* sigma2,
                             float* inv_sigma2) {
  if (!h) return ORB_GPU_ERR_ARG;
  for (int i = 0; i < h->nlevels; ++i) {
    if (scale) scale[i] = h->scale[i];
    if (inv_scale) inv_scale[i] = h->inv_scale[i];
    if (sigma2) sigma2[i] = h->sigma2[i];
    if (inv_sigma2) inv_sigma2[i] = h->inv_sigma2[i];
  }
  return ORB_GPU_OK;
}
int orb_gpu_get_features_per_level(const OrbGpuExtractor* h, int* quota, int* umax16) {
  if (!h) return ORB_GPU_ERR_ARG;
  if (quota) for (int i = 0; i < h->nlevels; ++i) quota[i] = h->quota[i];
  if (umax16) for (int i = 0; i < 16; ++i) umax16[i] = h->umax[i];
  return ORB_GPU_OK;
}

static int plan_batch_for(const OrbGpuExtractor* h, int w, int ht, int nb) {
  return std::max(nb, (h->plan_w == w && h->plan_h == ht) ? h->plan_batch : 0);
}

int orb_gpu_extract_batch_device(OrbGpuExtractor* h, const uint8_t* d_frames, int nframes, int width,
                                 int height, size_t stride, size_t frame_stride, OrbKeyPoint* d_kps,
                                 uint8_t* d_desc, int capacity, int* d_n_out, int* d_status,
                                 void* stream) {
  if (!h || !d_frames || !d_kps || !d_desc || !d_n_out || nframes < 1 || capacity < 1)
    return ORB_GPU_ERR_ARG;
  if (width < 1 || height < 1 || width > h->max_w || height > h->max_h || nframes > h->max_batch ||
      stride < (size_t)width || frame_stride < stride * (size_t)height)
    return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(h->device));
  int rc = ensure_plan(h, width, height, plan_batch_for(h, width, height, nframes));
  if (rc) return rc;
  cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
  Plan p = h->plan;
  if ((stride & 3) == 0 && (frame_stride & 3) == 0 && (reinterpret_cast<uintptr_t>(d_frames) & 3) == 0) {
    p.lv[0].img = const_cast<uint8_t*>(d_frames);  // level 0 aliases the caller's frames
    p.lv[0].pitch = (int)stride;
    p.lv[0].frame_stride = (long long)frame_stride;
  } else {
    // rows not 4-byte aligned (e.g. 1241-px KITTI rows): repack into the handle's level-0 buffer
    if ((rc = launch_repack_rows(d_frames, stride, frame_stride, width, height, p.lv[0], nframes, s))) return rc;
  }
  const int nsplit = (h->timing || nframes < 128) ? 1 : std::min(device_split(), nframes / 64);
  if (nsplit <= 1) {
    if ((rc = run_pipeline(h, p, h->sc, nframes, d_kps, d_desc, capacity, d_n_out, s))) return rc;
  } else {
    ORB_CUDA_TRY(cudaMemsetAsync(h->sc.status, 0, sizeof(int), s));
    while ((int)h->ev_in.size() < nsplit) {
      cudaEvent_t a = nullptr, b = nullptr;
      ORB_CUDA_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
      ORB_CUDA_TRY(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
      h->ev_in.push_back(a); h->ev_done.push_back(b);
    }
    ORB_CUDA_TRY(cudaEventRecord(h->ev_in[0], s));  // everything queued on `s` so far (repack, earlier calls)
    for (int k = 0; k < nsplit; ++k) {
      const int c0 = (int)((long long)nframes * k / nsplit), cn = (int)((long long)nframes * (k + 1) / nsplit) - c0;
      cudaStream_t sk = k ? h->s_alt[k - 1] : s;
      if (k) ORB_CUDA_TRY(cudaStreamWaitEvent(sk, h->ev_in[0], 0));
      Plan pc = plan_at(p, c0);
      if ((rc = run_pipeline(h, pc, scratch_at(h->sc, p, c0), cn, d_kps + (size_t)c0 * capacity,
                             d_desc + (size_t)c0 * capacity * 32, capacity, d_n_out + c0, sk, false)))
        return rc;
      if (k) {
        ORB_CUDA_TRY(cudaEventRecord(h->ev_done[k], sk));
        ORB_CUDA_TRY(cudaStreamWaitEvent(s, h->ev_done[k], 0));
      }
    }
  }
  if (d_status)
    ORB_CUDA_TRY(cudaMemcpyAsync(d_status, h->sc.status, sizeof(int), cudaMemcpyDeviceToDevice, s));
  h->last_nframes = nframes;
  h->have_result = true;
  h->last_l0 = p.lv[0].img; h->last_l0_pitch = p.lv[0].pitch; h->last_l0_stride = p.lv[0].frame_stride;
  return ORB_GPU_OK;
}

int orb_gpu_extract_batch(OrbGpuExtractor* h, const uint8_t* frames, int nframes, int width, int height,
                          size_t stride, size_t frame_stride, OrbKeyPoint* kps, uint8_t* desc,
                          int capacity, int* n_out) {
  if (!h || !n_out || nframes < 0 || capacity < 0) return ORB_GPU_ERR_ARG;
  for (int f = 0; f < nframes; ++f) n_out[f] = 0;
  if (nframes == 0 || width == 0 || height == 0 || !frames) return ORB_GPU_OK;  // empty image (:1046)
  if (!kps || !desc || capacity < 1) return ORB_GPU_ERR_ARG;
  if (width < 0 || height < 0 || width > h->max_w || height > h->max_h || stride < (size_t)width)
    return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(h->device));
  cudaStream_t s = h->stream;
  int status_all = ORB_GPU_OK;
  for (int f0 = 0; f0 < nframes; f0 += h->max_batch) {
    const int nb = std::min(h->max_batch, nframes - f0);
    int rc = ensure_plan(h, width, height, plan_batch_for(h, width, height, nb));
    if (rc) return rc;
    if ((rc = h->out_kps.ensure((size_t)h->plan_batch * capacity * sizeof(OrbKeyPoint))) ||
        (rc = h->out_desc.ensure((size_t)h->plan_batch * capacity * 32)) ||
        (rc = h->out_n.ensure((size_t)h->plan_batch * sizeof(int))))
      return rc;
    const Plan& p = h->plan;
    const LevelGeom& l0 = p.lv[0];
    const std::vector<int> bounds = host_schedule(nb);
    const int nchunks = (int)bounds.size() - 1;
    while ((int)h->ev_in.size() < nchunks) {
      cudaEvent_t a = nullptr, b = nullptr;
      ORB_CUDA_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
      ORB_CUDA_TRY(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
      h->ev_in.push_back(a); h->ev_done.push_back(b);
    }
    ORB_CUDA_TRY(cudaMemsetAsync(h->sc.status, 0, sizeof(int), s));
    // earlier work on the compute stream (a previous call's kernels) must not race the new uploads
    ORB_CUDA_TRY(cudaEventRecord(h->ev_done[0], s));
    ORB_CUDA_TRY(cudaStreamWaitEvent(h->s_h2d, h->ev_done[0], 0));
    // Uploads are plain 1-D copies of the caller's layout into a staging buffer (a 2-D copy with a row pitch
    // change runs far below PCIe speed); level 0 then aliases the staging buffer when its rows are 4-byte
    // aligned, else the frames are repacked on the device.
    const size_t frame_bytes = stride * (size_t)height;
    if ((rc = h->stage.ensure((size_t)(nb - 1) * frame_stride + frame_bytes + 16))) return rc;
    uint8_t* stage = (uint8_t*)h->stage.p;
    const bool alias = (stride & 3) == 0 && (frame_stride & 3) == 0;
    if (nchunks == 1 && !h->timing && graphs_enabled()) {
      // one chunk: everything in order on the compute stream, kernels replayed from a graph
      OrbKeyPoint* dk = (OrbKeyPoint*)h->out_kps.p;
      uint8_t* dd = (uint8_t*)h->out_desc.p;
      int* dn = (int*)h->out_n.p;
      const size_t in_bytes = (size_t)(nb - 1) * frame_stride + frame_bytes;
      const size_t kp_bytes = (size_t)nb * capacity * sizeof(OrbKeyPoint), de_bytes = (size_t)nb * capacity * 32;
      const bool pinned = nb <= 4;  // latency path: the caller's (pageable) arrays are copied through pinned staging
      if (pinned) {
        const size_t out_bytes = kp_bytes + de_bytes + (size_t)nb * sizeof(int) + 64;
        if (h->pin_in_bytes < in_bytes) {
          if (h->pin_in) cudaFreeHost(h->pin_in);
          h->pin_in = nullptr; h->pin_in_bytes = 0;
          ORB_CUDA_TRY(cudaHostAlloc(&h->pin_in, in_bytes, cudaHostAllocDefault));
          h->pin_in_bytes = in_bytes;
        }
        if (h->pin_out_bytes < out_bytes) {
          if (h->pin_out) cudaFreeHost(h->pin_out);
          h->pin_out = nullptr; h->pin_out_bytes = 0;
          ORB_CUDA_TRY(cudaHostAlloc(&h->pin_out, out_bytes, cudaHostAllocDefault));
          h->pin_out_bytes = out_bytes;
        }
        std::memcpy(h->pin_in, frames + (size_t)f0 * frame_stride, in_bytes);
        ORB_CUDA_TRY(cudaMemcpyAsync(stage, h->pin_in, in_bytes, cudaMemcpyHostToDevice, s));
      } else {
        ORB_CUDA_TRY(cudaMemcpyAsync(stage, frames + (size_t)f0 * frame_stride, in_bytes, cudaMemcpyHostToDevice, s));
      }
      Plan pc = p;
      if (alias) { pc.lv[0].img = stage; pc.lv[0].pitch = (int)stride; pc.lv[0].frame_stride = (long long)frame_stride; }
      auto enqueue = [&]() -> int {
        if (!alias) {
          const int rr = launch_repack_rows(stage, stride, frame_stride, width, height, pc.lv[0], nb, s);
          if (rr) return rr;
        }
        return run_pipeline(h, pc, h->sc, nb, dk, dd, capacity, dn, s, false);
      };
      const std::vector<long long> key = {nb, width, height, (long long)stride, (long long)frame_stride, capacity,
                                          h->plan_generation, (long long)(uintptr_t)stage, (long long)(uintptr_t)dk,
                                          (long long)(uintptr_t)dd, (long long)(uintptr_t)dn};
      if (h->graph_exec && key == h->graph_key) {
        ORB_CUDA_TRY(cudaGraphLaunch(h->graph_exec, s));
        count_launch(h->graph_nlaunch);
      } else if (key == h->graph_seen) {  // second call with these buffers: capture, then replay from now on
        if (h->graph_exec) { cudaGraphExecDestroy(h->graph_exec); h->graph_exec = nullptr; }
        cudaGraph_t g = nullptr;
        const long long before = g_launches.load();
        ORB_CUDA_TRY(cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed));
        rc = enqueue();
        cudaError_t e = cudaStreamEndCapture(s, &g);
        if (rc) { if (g) cudaGraphDestroy(g); return rc; }
        if (e != cudaSuccess) { set_last_error("cudaStreamEndCapture", e); return ORB_GPU_ERR_CUDA; }
        e = cudaGraphInstantiate(&h->graph_exec, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) { h->graph_exec = nullptr; set_last_error("cudaGraphInstantiate", e); return ORB_GPU_ERR_CUDA; }
        h->graph_nlaunch = (int)(g_launches.load() - before);
        h->graph_key = key;
        ORB_CUDA_TRY(cudaGraphLaunch(h->graph_exec, s));
      } else {
        h->graph_seen = key;
        if ((rc = enqueue())) return rc;
      }
      if (pinned) {
        uint8_t* po = (uint8_t*)h->pin_out;
        int* pn = (int*)(po + kp_bytes + de_bytes);
        int* pstat = pn + nb;
        ORB_CUDA_TRY(cudaMemcpyAsync(po, dk, kp_bytes, cudaMemcpyDeviceToHost, s));
        ORB_CUDA_TRY(cudaMemcpyAsync(po + kp_bytes, dd, de_bytes, cudaMemcpyDeviceToHost, s));
        ORB_CUDA_TRY(cudaMemcpyAsync(pn, dn, (size_t)nb * sizeof(int), cudaMemcpyDeviceToHost, s));
        ORB_CUDA_TRY(cudaMemcpyAsync(pstat, h->sc.status, sizeof(int), cudaMemcpyDeviceToHost, s));
        ORB_CUDA_TRY(cudaStreamSynchronize(s));
        for (int f = 0; f < nb; ++f) {  // exactly n_out keypoints and descriptor rows per frame
          const int n = std::min(pn[f], capacity);
          n_out[f0 + f] = pn[f];
          std::memcpy(kps + (size_t)(f0 + f) * capacity, po + (size_t)f * capacity * sizeof(OrbKeyPoint), (size_t)n * sizeof(OrbKeyPoint));
          std::memcpy(desc + (size_t)(f0 + f) * capacity * 32, po + kp_bytes + (size_t)f * capacity * 32, (size_t)n * 32);
        }
        if (*pstat) status_all = *pstat;
        h->last_nframes = nb;
        h->have_result = true;
        if (alias) { h->last_l0 = stage; h->last_l0_pitch = (int)stride; h->last_l0_stride = (long long)frame_stride; }
        else { h->last_l0 = l0.img; h->last_l0_pitch = l0.pitch; h->last_l0_stride = l0.frame_stride; }
        continue;
      }
      ORB_CUDA_TRY(cudaMemcpyAsync(n_out + f0, dn, (size_t)nb * sizeof(int), cudaMemcpyDeviceToHost, s));
      ORB_CUDA_TRY(cudaMemcpyAsync(kps + (size_t)f0 * capacity, dk, (size_t)nb * capacity * sizeof(OrbKeyPoint),
                                   cudaMemcpyDeviceToHost, s));
      ORB_CUDA_TRY(cudaMemcpyAsync(desc + (size_t)f0 * capacity * 32, dd, (size_t)nb * capacity * 32,
                                   cudaMemcpyDeviceToHost, s));
    } else {
    static const bool timeline = [] { const char* e = getenv("ORB_GPU_TIMELINE"); return e && atoi(e); }();
    std::vector<cudaEvent_t> tl;  // debug: t0, then per chunk (h2d done, compute start, compute end, d2h end)
    if (timeline) {
      tl.resize(1 + 4 * (size_t)nchunks);
      for (cudaEvent_t& e : tl) cudaEventCreate(&e);
      cudaEventRecord(tl[0], h->s_h2d);
    }
    for (int c = 0; c < nchunks; ++c) {  // all queued on the H2D stream, one event per chunk
      const int c0 = bounds[c], cn = bounds[c + 1] - c0;
      ORB_CUDA_TRY(cudaMemcpyAsync(stage + (size_t)c0 * frame_stride, frames + (size_t)(f0 + c0) * frame_stride,
                                   (size_t)(cn - 1) * frame_stride + frame_bytes, cudaMemcpyHostToDevice, h->s_h2d));
      ORB_CUDA_TRY(cudaEventRecord(h->ev_in[c], h->s_h2d));
      if (timeline) cudaEventRecord(tl[1 + 4 * c], h->s_h2d);
    }
    for (int c = 0; c < nchunks; ++c) {  // kernels of chunk c, then its results leave on the D2H stream
      const int c0 = bounds[c], cn = bounds[c + 1] - c0;
      OrbKeyPoint* dk = (OrbKeyPoint*)h->out_kps.p + (size_t)c0 * capacity;
      uint8_t* dd = (uint8_t*)h->out_desc.p + (size_t)c0 * capacity * 32;
      int* dn = (int*)h->out_n.p + c0;
      const int si = c % host_streams();
      cudaStream_t sc = si ? h->s_alt[si - 1] : s;
      // ev_in[c] was recorded on the upload stream, which itself waited for the status reset and for earlier work on
      // the compute stream: the extra streams need no further ordering (and must not wait for chunk 0's kernels)
      ORB_CUDA_TRY(cudaStreamWaitEvent(sc, h->ev_in[c], 0));
      if (timeline) cudaEventRecord(tl[2 + 4 * c], sc);
      Plan pc = plan_at(p, c0);
      if (alias) {
        pc.lv[0].img = stage + (size_t)c0 * frame_stride;
        pc.lv[0].pitch = (int)stride;
        pc.lv[0].frame_stride = (long long)frame_stride;
      } else {
        if ((rc = launch_repack_rows(stage + (size_t)c0 * frame_stride, stride, frame_stride, width, height, pc.lv[0],
                                     cn, sc)))
          return rc;
      }
      if ((rc = run_pipeline(h, pc, scratch_at(h->sc, p, c0), cn, dk, dd, capacity, dn, sc, false))) return rc;
      ORB_CUDA_TRY(cudaEventRecord(h->ev_done[c], sc));
      if (timeline) cudaEventRecord(tl[3 + 4 * c], sc);
      ORB_CUDA_TRY(cudaStreamWaitEvent(h->s_d2h, h->ev_done[c], 0));
      ORB_CUDA_TRY(cudaMemcpyAsync(n_out + f0 + c0, dn, (size_t)cn * sizeof(int), cudaMemcpyDeviceToHost, h->s_d2h));
      ORB_CUDA_TRY(cudaMemcpyAsync(kps + (size_t)(f0 + c0) * capacity, dk, (size_t)cn * capacity * sizeof(OrbKeyPoint),
                                   cudaMemcpyDeviceToHost, h->s_d2h));
      ORB_CUDA_TRY(cudaMemcpyAsync(desc + (size_t)(f0 + c0) * capacity * 32, dd, (size_t)cn * capacity * 32,
                                   cudaMemcpyDeviceToHost, h->s_d2h));
      if (timeline) cudaEventRecord(tl[4 + 4 * c], h->s_d2h);
    }
    if (timeline) {
      cudaDeviceSynchronize();
      for (int c = 0; c < nchunks; ++c) {
        float a = 0, b = 0, d = 0, e = 0;
        cudaEventElapsedTime(&a, tl[0], tl[1 + 4 * c]); cudaEventElapsedTime(&b, tl[0], tl[2 + 4 * c]);
        cudaEventElapsedTime(&d, tl[0], tl[3 + 4 * c]); cudaEventElapsedTime(&e, tl[0], tl[4 + 4 * c]);
        fprintf(stderr, "[orb_gpu timeline] chunk %d (%d frames): h2d done %.3f  compute %.3f..%.3f  d2h done %.3f ms\n", c,
                bounds[c + 1] - bounds[c], a, b, d, e);
      }
      for (cudaEvent_t& e : tl) cudaEventDestroy(e);
    }
    }
    int dev_status = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(&dev_status, h->sc.status, sizeof(int), cudaMemcpyDeviceToHost, (nchunks == 1 && !h->timing && graphs_enabled()) ? s : h->s_d2h));
    ORB_CUDA_TRY(cudaStreamSynchronize(s));
    for (cudaStream_t a : h->s_alt) ORB_CUDA_TRY(cudaStreamSynchronize(a));
    ORB_CUDA_TRY(cudaStreamSynchronize(h->s_d2h));
    if (dev_status) status_all = dev_status;
    h->last_nframes = nb;
    h->have_result = true;
    if (alias) { h->last_l0 = stage; h->last_l0_pitch = (int)stride; h->last_l0_stride = (long long)frame_stride; }
    else { h->last_l0 = l0.img; h->last_l0_pitch = l0.pitch; h->last_l0_stride = l0.frame_stride; }
  }
  return status_all;
}

int orb_gpu_extract(OrbGpuExtractor* h, const uint8_t* gray, int width, int height, size_t stride,
                    OrbKeyPoint* kps, uint8_t* desc, int capacity, int* n_out) {
  return orb_gpu_extract_batch(h, gray, 1, width, height, stride, stride * (size_t)height, kps, desc,
                               capacity, n_out);
}

int orb_gpu_get_pyramid(OrbGpuExtractor* h, int frame, int level, int blurred, uint8_t* dst,
                        size_t dst_stride, int* width, int* height) {
  if (!h || level < 0 || level >= h->nlevels) return ORB_GPU_ERR_ARG;
  if (!h->have_result) return ORB_GPU_ERR_STATE;
  if (frame < 0 || frame >= h->last_nframes) return ORB_GPU_ERR_ARG;
  const LevelGeom& g = h->plan.lv[level];
  if (width) *width = g.w;
  if (height) *height = g.h;
  if (!dst) return ORB_GPU_OK;
  if (dst_stride < (size_t)g.w) return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(h->device));
  const uint8_t* src;
  size_t spitch;
  if (blurred) { src = g.blur + (size_t)frame * g.bframe_stride; spitch = (size_t)g.bpitch; }
  else if (level == 0) { src = h->last_l0 + (size_t)frame * h->last_l0_stride; spitch = (size_t)h->last_l0_pitch; }
  else { src = g.img + (size_t)frame * g.frame_stride; spitch = (size_t)g.pitch; }
  ORB_CUDA_TRY(cudaDeviceSynchronize());
  ORB_CUDA_TRY(cudaMemcpy2D(dst, dst_stride, src, spitch, g.w, g.h, cudaMemcpyDeviceToHost));
  return ORB_GPU_OK;
}

int orb_gpu_get_candidates(OrbGpuExtractor* h, int frame, int level, float* xyr, int cap, int* n) {
  if (!h || !n || level < 0 || level >= h->nlevels) return ORB_GPU_ERR_ARG;
  if (!h->have_result) return ORB_GPU_ERR_STATE;
  if (frame < 0 || frame >= h->last_nframes) return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(h->device));
  ORB_CUDA_TRY(cudaDeviceSynchronize());
  int cnt = 0;
  ORB_CUDA_TRY(cudaMemcpy(&cnt, h->sc.cand_cnt + frame * h->nlevels + level, sizeof(int), cudaMemcpyDeviceToHost));
  *n = cnt;
  const int m = std::min(cnt, cap);
  if (m <= 0 || !xyr) return ORB_GPU_OK;
  float* d = nullptr;
  ORB_CUDA_TRY(cudaMalloc(&d, (size_t)m * 3 * sizeof(float)));
  int rc = launch_export_candidates(h->plan, h->sc, frame, level, d, m, h->stream);
  if (!rc) {
    cudaError_t e = cudaMemcpyAsync(xyr, d, (size_t)m * 3 * sizeof(float), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) { set_last_error("get_candidates copy", e); rc = ORB_GPU_ERR_CUDA; }
  }
  cudaFree(d);
  return rc;
}

// ---- Frame::ComputeStereoMatches on the pyramids of two extractors (stereo_kernels.cu) ------------
static int fill_stereo_side(const OrbGpuExtractor* h, StereoSide& s) {
  if (!h->have_result) return ORB_GPU_ERR_STATE;
  for (int l = 0; l < kMaxLevels; ++l) {
    const bool on = l < h->nlevels;
    const LevelGeom& g = h->plan.lv[on ? l : 0];
    s.img[l] = !on ? nullptr : (l == 0 ? h->last_l0 : g.img);
    s.pitch[l] = !on ? 0 : (l == 0 ? h->last_l0_pitch : g.pitch);
    s.frame_stride[l] = !on ? 0 : (l == 0 ? h->last_l0_stride : g.frame_stride);
    s.w[l] = on ? g.w : 0;
    s.h[l] = on ? g.h : 0;
  }
  return ORB_GPU_OK;
}

static int stereo_params(const OrbGpuExtractor* left, const OrbGpuExtractor* right, float bf, float b,
                         StereoParams& P) {
  if (left->nlevels != right->nlevels || left->plan.width != right->plan.width ||
      left->plan.height != right->plan.height || left->scale_factor_f != right->scale_factor_f || !(b > 0))
    return ORB_GPU_ERR_ARG;
  for (int l = 0; l < kMaxLevels; ++l) {
    P.scale[l] = l < left->nlevels ? left->scale[l] : 0.f;
    P.inv_scale[l] = l < left->nlevels ? left->inv_scale[l] : 0.f;
  }
  P.min_d = 0.f;       // Frame.cc:497
  P.max_d = bf / b;    // :498  (mbf/mb)
  P.bf = bf;
  P.n_rows = left->plan.height;   // mvImagePyramid[0].rows (:473)
  return ORB_GPU_OK;
}

int orb_gpu_stereo_matches_device(OrbGpuExtractor* left, int frame_l0, int frame_l_step, OrbGpuExtractor* right,
                                  int frame_r0, int frame_r_step, int npairs, const OrbKeyPoint* d_kps_l,
                                  const uint8_t* d_desc_l, const int* d_n_l, int capacity_l,
                                  const OrbKeyPoint* d_kps_r, const uint8_t* d_desc_r, const int* d_n_r,
                                  int capacity_r, float bf, float b, float* d_u_right, float* d_depth,
                                  int* d_sad, void* stream) {
  if (!left || !right || npairs < 1 || !d_kps_l || !d_desc_l || !d_n_l || !d_kps_r || !d_desc_r || !d_n_r ||
      capacity_l < 1 || capacity_r < 1 || capacity_r > 65535 || !d_u_right || !d_depth || !d_sad)
    return ORB_GPU_ERR_ARG;
  const int last_l = frame_l0 + (npairs - 1) * frame_l_step, last_r = frame_r0 + (npairs - 1) * frame_r_step;
  if (frame_l0 < 0 || frame_r0 < 0 || last_l < 0 || last_r < 0) return ORB_GPU_ERR_ARG;
  StereoSide L, R;
  StereoParams P;
  int rc;
  if ((rc = fill_stereo_side(left, L)) || (rc = fill_stereo_side(right, R)) ||
      (rc = stereo_params(left, right, bf, b, P)))
    return rc;
  if (std::max(frame_l0, last_l) >= left->last_nframes || std::max(frame_r0, last_r) >= right->last_nframes)
    return ORB_GPU_ERR_ARG;
  L.img_frame0 = L.kp_frame0 = frame_l0; L.img_frame_step = L.kp_frame_step = frame_l_step;
  R.img_frame0 = R.kp_frame0 = frame_r0; R.img_frame_step = R.kp_frame_step = frame_r_step;
  L.kps = d_kps_l; L.desc = d_desc_l; L.n = d_n_l; L.capacity = capacity_l;
  R.kps = d_kps_r; R.desc = d_desc_r; R.n = d_n_r; R.capacity = capacity_r;
  ORB_CUDA_TRY(cudaSetDevice(left->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : left->stream;
  return launch_stereo_pairs(L, R, P, npairs, capacity_l, d_u_right, d_depth, d_sad, s);
}

int orb_gpu_compute_stereo_matches(OrbGpuExtractor* left, int frame_l, OrbGpuExtractor* right, int frame_r,
                                   const OrbKeyPoint* kps_l, const uint8_t* desc_l, int nl,
                                   const OrbKeyPoint* kps_r, const uint8_t* desc_r, int nr, float bf, float b,
                                   float* u_right, float* depth) {
  if (!left || !right || nl < 0 || nr < 0 || nr > 65535 || (nl > 0 && (!kps_l || !desc_l || !u_right || !depth)) ||
      (nr > 0 && (!kps_r || !desc_r)))
    return ORB_GPU_ERR_ARG;
  StereoSide L, R;
  StereoParams P;
  int rc;
  if ((rc = fill_stereo_side(left, L)) || (rc = fill_stereo_side(right, R)) ||
      (rc = stereo_params(left, right, bf, b, P)))
    return rc;
  if (frame_l < 0 || frame_l >= left->last_nframes || frame_r < 0 || frame_r >= right->last_nframes)
    return ORB_GPU_ERR_ARG;
  for (int i = 0; i < nl; ++i) { u_right[i] = -1.0f; depth[i] = -1.0f; }
  if (nl == 0 || nr == 0) return ORB_GPU_OK;
  { const int rc_pool = use_cached_pool(left->device); if (rc_pool) return rc_pool; }
  // both extractors may still be running on their own streams
  ORB_CUDA_TRY(cudaStreamSynchronize(left->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(right->stream));
  TmpDevW kl, dl, kr, dr, nn, ur, dp, sad;
  if ((rc = kl.alloc((size_t)nl * sizeof(OrbKeyPoint))) || (rc = dl.alloc((size_t)nl * 32)) ||
      (rc = kr.alloc((size_t)nr * sizeof(OrbKeyPoint))) || (rc = dr.alloc((size_t)nr * 32)) ||
      (rc = nn.alloc(8)) || (rc = ur.alloc((size_t)nl * 4)) || (rc = dp.alloc((size_t)nl * 4)) ||
      (rc = sad.alloc((size_t)nl * 4)))
    return rc;
  ORB_CUDA_TRY(cudaStreamSynchronize(0));  // the pool hands the buffers out in legacy-stream order
  cudaStream_t s = left->stream;
  const int counts[2] = {nl, nr};
  ORB_CUDA_TRY(cudaMemcpyAsync(kl.p, kps_l, (size_t)nl * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice, s));
  ORB_CUDA_TRY(cudaMemcpyAsync(dl.p, desc_l, (size_t)nl * 32, cudaMemcpyHostToDevice, s));
  ORB_CUDA_TRY(cudaMemcpyAsync(kr.p, kps_r, (size_t)nr * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice, s));
  ORB_CUDA_TRY(cudaMemcpyAsync(dr.p, desc_r, (size_t)nr * 32, cudaMemcpyHostToDevice, s));
  ORB_CUDA_TRY(cudaMemcpyAsync(nn.p, counts, 8, cudaMemcpyHostToDevice, s));
  L.img_frame0 = frame_l; L.img_frame_step = 0; L.kp_frame0 = 0; L.kp_frame_step = 0;
  R.img_frame0 = frame_r; R.img_frame_step = 0; R.kp_frame0 = 0; R.kp_frame_step = 0;
  L.kps = (const OrbKeyPoint*)kl.p; L.desc = (const uint8_t*)dl.p; L.n = (const int*)nn.p; L.capacity = nl;
  R.kps = (const OrbKeyPoint*)kr.p; R.desc = (const uint8_t*)dr.p; R.n = (const int*)nn.p + 1; R.capacity = nr;
  if ((rc = launch_stereo_pairs(L, R, P, 1, nl, (float*)ur.p, (float*)dp.p, (int*)sad.p, s))) return rc;
  ORB_CUDA_TRY(cudaMemcpyAsync(u_right, ur.p, (size_t)nl * 4, cudaMemcpyDeviceToHost, s));
  ORB_CUDA_TRY(cudaMemcpyAsync(depth, dp.p, (size_t)nl * 4, cudaMemcpyDeviceToHost, s));
  ORB_CUDA_TRY(cudaStreamSynchronize(s));
  return ORB_GPU_OK;
}

}  // extern "C"
