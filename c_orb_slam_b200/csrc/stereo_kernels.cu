// Frame::ComputeStereoMatches on the device (reference src/Frame.cc:466-640): Hamming stage (:476-549),
// sub-pixel refinement by 11x11 SAD over 11 shifts with a parabola fit (:551-620) and the median outlier
// filter (:623-639).  It reads the unblurred pyramids the two extractors left in HBM (mvImagePyramid,
// include/ORBextractor.h:85), so no image ever returns to the host for stereo.
//
// Layout: one warp per left keypoint; grid.y = stereo pair.  The warp first scans the right keypoints of its
// pair for the best Hamming distance inside the row band, then, when that distance is below
// (TH_HIGH+TH_LOW)/2, stages the left 11x11 window and the right 11x21 strip in shared memory and computes the
// 11 SAD values as integers (the reference's float sums are integer-valued and < 2^24, so they are exact).
// A second kernel (one block per pair) finds the median SAD by a bitwise search and removes the outliers.
#include <climits>

#include "common.cuh"

namespace orbgpu {

constexpr int kWin = 5;         // w, Frame.cc:561
constexpr int kShift = 5;       // L, :568
constexpr int kWinSide = 2 * kWin + 1;               // 11
constexpr int kStripW = kWinSide + 2 * kShift;       // 21
constexpr int kStereoWarps = 8;

__device__ __forceinline__ int hamming256(const uint4* a, const uint4* b) {
  const uint4 a0 = a[0], a1 = a[1], b0 = b[0], b1 = b[1];
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(kStereoWarps * 32)
stereo_pairs_kernel(StereoSide L, StereoSide R, StereoParams P, float* __restrict__ u_right,
                    float* __restrict__ depth, int* __restrict__ sad /*[npairs][cap_l]*/) {
  __shared__ int s_l[kStereoWarps][kWinSide * kWinSide];
  __shared__ int s_r[kStereoWarps][kWinSide * kStripW];
  __shared__ int s_sad[kStereoWarps][2 * kShift + 1];
  const int pair = blockIdx.y, wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int iL = blockIdx.x * kStereoWarps + wib;
  const int fl = L.kp_frame0 + pair * L.kp_frame_step, fr = R.kp_frame0 + pair * R.kp_frame_step;
  const int nl = min(L.n[fl], L.capacity), nr = min(R.n[fr], R.capacity);
  if (iL >= nl) return;
  const OrbKeyPoint* kl = L.kps + (size_t)fl * L.capacity;
  const OrbKeyPoint* kr = R.kps + (size_t)fr * R.capacity;
  const uint4* dl = (const uint4*)(L.desc + (size_t)fl * L.capacity * 32);
  const uint4* dr = (const uint4*)(R.desc + (size_t)fr * R.capacity * 32);
  const size_t o = (size_t)pair * L.capacity + iL;
  const OrbKeyPoint kpL = kl[iL];

  // ---- Hamming stage (:507-549) ----
  const int row = (int)kpL.y;
  const float min_u = __fsub_rn(kpL.x, P.max_d), max_u = __fsub_rn(kpL.x, P.min_d);
  unsigned best = 0xffffffffu;
  if (row >= 0 && row < P.n_rows && !(max_u < 0)) {
    for (int i = lane; i < nr; i += 32) {
      const OrbKeyPoint kpR = kr[i];
      const float r = __fmul_rn(2.0f, P.scale[kpR.octave]);   // :487
      const int maxr = (int)ceilf(__fadd_rn(kpR.y, r)), minr = (int)floorf(__fsub_rn(kpR.y, r));
      if (row < minr || row > maxr) continue;
      if (kpR.octave < kpL.octave - 1 || kpR.octave > kpL.octave + 1) continue;   // :533
      if (!(kpR.x >= min_u && kpR.x <= max_u)) continue;                          // :538
      const unsigned d = (unsigned)hamming256(dl + 2 * (size_t)iL, dr + 2 * (size_t)i);
      best = min(best, (d << 16) | (unsigned)i);
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
  const int best_dist = best == 0xffffffffu ? INT_MAX : (int)(best >> 16);
  float out_u = -1.0f, out_d = -1.0f;
  int out_sad = -1;
  // bestDist starts at TH_HIGH with a strict '<' (:522,:543); refined only below thOrbDist (:552)
  const bool refine = best_dist < ORB_GPU_TH_HIGH && best_dist < (ORB_GPU_TH_HIGH + ORB_GPU_TH_LOW) / 2;
  if (refine) {  // warp-uniform
    const int level = kpL.octave;
    const float uR0 = kr[best & 0xffffu].x;
    const float inv = P.inv_scale[level];
    const int su = (int)roundf(__fmul_rn(kpL.x, inv)), sv = (int)roundf(__fmul_rn(kpL.y, inv));
    const float sur_f = roundf(__fmul_rn(uR0, inv));
    const int sur = (int)sur_f;
    const int wl = L.w[level], hl = L.h[level], wr = R.w[level], hr = R.h[level];
    // the reference would throw in cv::Mat::rowRange/colRange when a window leaves the level image; skipped here
    const bool inside = sv - kWin >= 0 && sv + kWin < hl && su - kWin >= 0 && su + kWin < wl && sv + kWin < hr;
    const float iniu = __fadd_rn(sur_f, (float)(kShift - kWin)), endu = __fadd_rn(sur_f, (float)(kShift + kWin + 1));
    const bool in_r = !(iniu < 0 || endu >= (float)wr) && sur - kShift - kWin >= 0;   // :575-578
    if (inside && in_r) {
      const uint8_t* pl = L.img[level] + (size_t)(L.img_frame0 + pair * L.img_frame_step) * L.frame_stride[level] +
                          (size_t)(sv - kWin) * L.pitch[level] + (su - kWin);
      const uint8_t* pr = R.img[level] + (size_t)(R.img_frame0 + pair * R.img_frame_step) * R.frame_stride[level] +
                          (size_t)(sv - kWin) * R.pitch[level] + (sur - kShift - kWin);
      for (int i = lane; i < kWinSide * kWinSide; i += 32)
        s_l[wib][i] = pl[(size_t)(i / kWinSide) * L.pitch[level] + (i % kWinSide)];
      for (int i = lane; i < kWinSide * kStripW; i += 32)
        s_r[wib][i] = pr[(size_t)(i / kStripW) * R.pitch[level] + (i % kStripW)];
      if (lane < 2 * kShift + 1) s_sad[wib][lane] = 0;
      __syncwarp();
      const int cl = s_l[wib][kWin * kWinSide + kWin];
      for (int i = lane; i < (2 * kShift + 1) * kWinSide; i += 32) {   // (shift, row) pairs
        const int sh = i / kWinSide, r = i % kWinSide;                 // sh = incR + L
        const int cr = s_r[wib][kWin * kStripW + sh + kWin];           // centre of the shifted right window
        int acc = 0;
#pragma unroll
        for (int c = 0; c < kWinSide; ++c)
          acc += abs((s_l[wib][r * kWinSide + c] - cl) - (s_r[wib][r * kStripW + sh + c] - cr));
        atomicAdd(&s_sad[wib][sh], acc);
      }
      __syncwarp();
      if (lane == 0) {
        int bestDist = INT_MAX, bestinc = 0;
        for (int sh = 0; sh <= 2 * kShift; ++sh)
          if (s_sad[wib][sh] < bestDist) { bestDist = s_sad[wib][sh]; bestinc = sh - kShift; }
        if (bestinc != -kShift && bestinc != kShift) {   // :596-597
          const float dist1 = (float)s_sad[wib][kShift + bestinc - 1], dist2 = (float)s_sad[wib][kShift + bestinc],
                      dist3 = (float)s_sad[wib][kShift + bestinc + 1];
          const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2)));
          const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3), den);   // :604
          if (!(deltaR < -1 || deltaR > 1)) {
            float bestuR = __fmul_rn(P.scale[level], __fadd_rn(__fadd_rn(sur_f, (float)bestinc), deltaR));   // :610
            float disparity = __fsub_rn(kpL.x, bestuR);
            if (disparity >= P.min_d && disparity < P.max_d) {
              if (disparity <= 0) {   // :616-620 (0.01 is a double literal)
                disparity = 0.01f;
                bestuR = (float)((double)kpL.x - 0.01);
              }
              out_d = __fdiv_rn(P.bf, disparity);
              out_u = bestuR;
              out_sad = bestDist;
            }
          }
        }
      }
    }
  }
  if (lane == 0) { u_right[o] = out_u; depth[o] = out_d; sad[o] = out_sad; }
}

// median of the SAD values of the accepted matches and removal of those >= 1.5 * 1.4 * median (:623-639).
// vDistIdx[size/2].first of the sorted pairs = the (size/2)-th smallest distance.
__global__ void __launch_bounds__(1024)
stereo_median_kernel(const int* __restrict__ n_left, int kp_frame0, int kp_frame_step, int capacity,
                     float* __restrict__ u_right,
                     float* __restrict__ depth, const int* __restrict__ sad) {
  __shared__ int s_cnt;
  const int pair = blockIdx.x, t = threadIdx.x, T = blockDim.x;
  const int n = min(n_left[kp_frame0 + pair * kp_frame_step], capacity);
  const size_t base = (size_t)pair * capacity;
  if (t == 0) s_cnt = 0;
  __syncthreads();
  int local = 0;
  for (int i = t; i < n; i += T) local += sad[base + i] >= 0;
  if (local) atomicAdd(&s_cnt, local);
  __syncthreads();
  const int total = s_cnt;
  if (total == 0) return;
  const int k = total / 2;
  int lo = 0;  // largest v with #{x < v} <= k  ==  k-th smallest (0-based)
  for (int bit = 16; bit >= 0; --bit) {
    const int cand = lo | (1 << bit);
    __syncthreads();
    if (t == 0) s_cnt = 0;
    __syncthreads();
    local = 0;
    for (int i = t; i < n; i += T) {
      const int v = sad[base + i];
      local += v >= 0 && v < cand;
    }
    if (local) atomicAdd(&s_cnt, local);
    __syncthreads();
    if (s_cnt <= k) lo = cand;
  }
  const float th = __fmul_rn(1.5f * 1.4f, (float)lo);
  for (int i = t; i < n; i += T) {
    const int v = sad[base + i];
    if (v >= 0 && !((float)v < th)) { u_right[base + i] = -1.0f; depth[base + i] = -1.0f; }
  }
}

int launch_stereo_pairs(const StereoSide& L, const StereoSide& R, const StereoParams& P, int npairs,
                        int max_left, float* d_u_right, float* d_depth, int* d_sad, cudaStream_t s) {
  if (npairs <= 0 || max_left <= 0) return ORB_GPU_OK;
  dim3 grid((max_left + kStereoWarps - 1) / kStereoWarps, npairs);
  stereo_pairs_kernel<<<grid, kStereoWarps * 32, 0, s>>>(L, R, P, d_u_right, d_depth, d_sad);
  stereo_median_kernel<<<npairs, 1024, 0, s>>>(L.n, L.kp_frame0, L.kp_frame_step, L.capacity, d_u_right, d_depth,
                                                d_sad);
  count_launch(2);
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

}  // namespace orbgpu
