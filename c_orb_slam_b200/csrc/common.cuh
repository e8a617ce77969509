// Shared declarations of the B200-native ORB front-end (device plan, error plumbing).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/orb_gpu.h"

namespace orbgpu {

constexpr int kMaxLevels = ORB_GPU_MAX_LEVELS;
constexpr int kEdge = 19;          // EDGE_THRESHOLD, reference src/ORBextractor.cc:74
constexpr int kBorder = kEdge - 3; // FAST region starts 16 px inside the image (:773-776)
constexpr int kHalfPatch = 15;     // HALF_PATCH_SIZE (:73)
constexpr int kCodeDepth = 12;     // quadtree path digits kept per key (boxes up to 4096 px)
constexpr int kMaxDim = 4096;

// Geometry of one pyramid level for one (width,height); built on the host (plan.cc logic in
// orb_gpu.cu), passed to kernels by value inside Plan.
struct LevelGeom {
  int w, h;                 // level image size, cvRound((float)W * invScale[l]) (:1112)
  int pitch;                // bytes per row of img/blur
  long long frame_stride;   // bytes between consecutive frames of this level
  uint8_t* img;             // [B][h][pitch] (level 0 may alias caller memory)
  uint8_t* blur;            // [B][h][bpitch] (always handle-owned)
  int bpitch;               // bytes per row of blur
  long long bframe_stride;  // bytes between frames of blur
  // resize tables (level >= 1): source taps and 11-bit weights per destination column/row
  const int* xofs; const short* ax0; const short* ax1;
  const int* yofs; const short* ay0; const short* ay1;
  const short* yinv;        // [h of level l-1]: destination row with yofs == source row, or -1
  // FAST cell grid (:778-806)
  int ncols, nrows, wcell, hcell;
  int max_bx, max_by;       // maxBorderX/Y = w-16, h-16
  int cell_base;            // first cell of this level inside a frame's cell arrays
  // quadtree (:539-563)
  int n_ini;                // number of root nodes
  float hx;                 // root width, (float)(W')/nIni
  int box_w, box_h;         // W' = maxBorderX-minBorderX, H'
  int quota;                // mnFeaturesPerLevel[l]
  int key_cap;              // capacity (power of two) of the candidate/key arrays of this level
  long long key_base;       // offset of this level inside a frame's key scratch (elements)
  int slot_base;            // first output slot of this level inside a frame (quota+3 slots)
  float scale;              // mvScaleFactor[l]
  float patch_size;         // (float)(int)(31 * scale)  (:837)
};

struct Plan {
  int nlevels;
  int width, height;
  int cells_per_frame;      // sum over levels of nrows*ncols
  int cell_cap;             // slots per cell (upper bound of NMS survivors in one cell)
  long long keys_per_frame; // sum of key_cap
  int slots_per_frame;      // sum of quota+3
  int ini_th, min_th;
  int umax[16];
  LevelGeom lv[kMaxLevels];
};

// Per-batch device scratch shared by the extraction kernels.
struct Scratch {
  const uint2* cell_tab;  // [cells_per_frame] FAST cell ROIs: x | y<<16, cw | ch<<8 | level<<16 (cw == 0: skipped)
  uint32_t* cell_kp;   // [B][cells_per_frame][cell_cap]  lx | ly<<8 | score<<16
  uint32_t* cell_cnt;  // [B][cells_per_frame]            all | hi<<16
  uint64_t* keys;      // [B][keys_per_frame]             quadtree sort keys (global fallback)
  uint32_t* pay;       // [B][keys_per_frame]             x | y<<12 | response<<24 by arrival
  uint32_t* sel;       // [B][slots_per_frame]            selected keypoints, list order
  int* sel_cnt;        // [B][nlevels]
  int* cand_cnt;       // [B][nlevels]                    candidates per level (diagnostics/tests)
  int* status;         // [1] sticky device-side status (0 or ORB_GPU_ERR_CAPACITY)
};

// One side of a batch of stereo pairs (stereo_kernels.cu): the unblurred pyramid an extractor left in HBM and
// the keypoints / descriptors of its frames.  Pair i reads image frame img_frame0 + i*img_frame_step and the
// keypoint arrays of frame kp_frame0 + i*kp_frame_step ([frame][capacity] layout, n[frame] valid entries).
struct StereoSide {
  const uint8_t* img[kMaxLevels];
  long long frame_stride[kMaxLevels];
  int pitch[kMaxLevels], w[kMaxLevels], h[kMaxLevels];
  int img_frame0, img_frame_step, kp_frame0, kp_frame_step;
  const OrbKeyPoint* kps;
  const uint8_t* desc;
  const int* n;
  int capacity;
};
struct StereoParams {
  float scale[kMaxLevels], inv_scale[kMaxLevels];  // mvScaleFactors / mvInvScaleFactors of the left extractor
  float min_d, max_d, bf;                          // Frame.cc:496-498; mbf
  int n_rows;                                      // mvImagePyramid[0].rows (:473)
};

// Peer-memory exchange of the database-sharded brute-force matcher (exchange_kernels.cu): every rank's gather buffer
// [2][world][3][nq_max] and flag words, mapped into this process (CUDA IPC or same-process pointers).
constexpr int kExMaxWorld = 16;
struct ExPeers {
  int32_t* buf[kExMaxWorld];    // rank r's gather buffer
  unsigned* flag[kExMaxWorld];  // rank r's flags: [kExMaxWorld] arrival counters of the stand-alone exchange kernel,
                                // then [world][qb_max] sequence numbers per source rank and query block (fused form)
};
// What the tcgen05 brute-force kernel needs to finish a step itself: the block that completes a query block stores the
// merged triples into every rank's gather buffer, releases the call's sequence number for that query block to every
// rank, waits for the same from every rank and merges the `world` slots.
struct FusedExchange {
  ExPeers peers;
  int rank, world, nq_max, qb_max, parity;
  unsigned seq;
};

void count_launch(int n = 1);
void set_last_error(const char* what, cudaError_t e);
void set_last_message(const char* msg);  // what orb_gpu_last_error() returns, for errors that are not CUDA's

#define ORB_CUDA_TRY(expr)                                  \
  do {                                                      \
    cudaError_t e__ = (expr);                               \
    if (e__ != cudaSuccess) {                               \
      ::orbgpu::set_last_error(#expr, e__);                 \
      return ORB_GPU_ERR_CUDA;                              \
    }                                                       \
  } while (0)

// Scratch buffer of one host-pointer call: stream-ordered allocation from the device's default memory pool on
// the legacy default stream (the stream those calls launch on).  use_cached_pool() raises the pool's release
// threshold once per device, so after the first call of a given size the allocation is a free-list lookup
// instead of a cudaMalloc / cudaFree pair (the matcher calls are ~100 us of work behind ~15 buffers).
int use_cached_pool(int device);
struct TmpDevW {
  void* p = nullptr;
  ~TmpDevW() { if (p) cudaFreeAsync(p, 0); }
  int alloc(size_t bytes) { ORB_CUDA_TRY(cudaMallocAsync(&p, bytes ? bytes : 16, 0)); return ORB_GPU_OK; }
  template <class T> T* as() { return (T*)p; }
};

// ---- launchers implemented in extract_kernels.cu -------------------------------------------
int launch_repack_rows(const uint8_t* d_frames, size_t stride, size_t frame_stride, int width, int height,
                       const LevelGeom& l0, int nframes, cudaStream_t s);
int launch_resize_level(const Plan& p, int level, int nframes, cudaStream_t s);
int launch_fast_cells(const Plan& p, const Scratch& sc, int nframes, cudaStream_t s);
int launch_quadtree(const Plan& p, const Scratch& sc, int nframes, cudaStream_t s);
int launch_blur(const Plan& p, int nframes, cudaStream_t s);
// blur_tc_kernels.cu: the levels in level_mask on the tensor cores (TMA in, tcgen05.mma, TMA out)
bool blur_tc_level_ok(const LevelGeom& g, int nframes);
int launch_blur_tc(const Plan& p, int nframes, unsigned level_mask, cudaStream_t s);
// resize_tc_kernels.cu: level l from l-1 with the horizontal pass on the tensor cores; false / not launched when the
// level pair does not qualify (TMA alignment, frame gaps, scale factor above ~1.85)
bool resize_tc_level_ok(const LevelGeom& src, const LevelGeom& dst, int nframes);
int launch_resize_tc(const LevelGeom& src, const LevelGeom& dst, int nframes, cudaStream_t s);
int resize_variant();  // 0: byte-SIMD (default), 1: tensor-core form for batches (orb_gpu_set_resize_variant)
// cached 2-D u8 tensor map (CUtensorMap, 128 bytes) of [rows][w] with `pitch` bytes per row and a box_w x box_h box
int tensor_map_2d_u8(void* out_map, const void* base, int w, long long rows, int pitch, int box_w, int box_h, bool swizzle128);
bool tma_available();
int launch_orient_desc(const Plan& p, const Scratch& sc, int nframes, OrbKeyPoint* d_kps,
                       uint8_t* d_desc, int capacity, int* d_n_out, cudaStream_t s);
int launch_export_candidates(const Plan& p, const Scratch& sc, int frame, int level, float* d_xyr,
                             int cap, cudaStream_t s);
int launch_stereo_pairs(const StereoSide& L, const StereoSide& R, const StereoParams& P, int npairs,
                        int max_left, float* d_u_right, float* d_depth, int* d_sad, cudaStream_t s);
int upload_constants();
// ---- match_tc5_kernels.cu: brute force on tcgen05 (256 queries per block, slices of whole 256-row tiles)
int launch_bruteforce_tc5(dim3 grid, cudaStream_t s, const uint4* q, int nq, const uint4* db, long long ndb,
                          long long slice_rows, int32_t idx_base, int32_t* o_idx, int32_t* o_best,
                          int32_t* o_second, unsigned* g_thr, unsigned* g_ticket, int32_t* f_idx, int32_t* f_best,
                          int32_t* f_second, const FusedExchange* fx = nullptr);
// brute force over this rank's rows (match_kernels.cu).  With fx, the tcgen05 form and a grid of at most one block per
// SM the exchange across the ranks happens inside the same launch and *fused = 1; else *fused = 0 and the caller
// exchanges d_idx / d_best / d_second afterwards.
int bruteforce_device_impl(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb, int32_t idx_base, int32_t* d_idx,
                           int32_t* d_best, int32_t* d_second, cudaStream_t s, const FusedExchange* fx, int* fused);
size_t quadtree_smem_bytes(const Plan& p, int level);

}  // namespace orbgpu
