// Windowed (guided) matching kernels: SearchByProjection-style search and the stereo row-band
// search.  Reference: ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th)
// src/ORBmatcher.cc:45-129; Frame::AssignFeaturesToGrid / GetFeaturesInArea / PosInGrid
// src/Frame.cc:230-245, 327-392; Frame::ComputeStereoMatches Hamming stage src/Frame.cc:466-549.
//
// The reference walks the map points in order and lets a point claim a keypoint, which hides it
// from later points (:87-89, :123).  Here the per-query candidate lists (in GetFeaturesInArea
// order) and all Hamming distances are computed in parallel, and the claims are found as the
// fixpoint of "query i sees the claims of queries j < i": each round re-evaluates every query
// against the previous round's claims; query i is final after at most i+1 rounds and the loop
// ends when a round changes nothing, at which point the state equals the sequential result.
#include <algorithm>
#include <climits>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace orbgpu {

constexpr int kGridCols = 64, kGridRows = 48;  // FRAME_GRID_COLS/ROWS, include/Frame.h:37-38
constexpr int kCells = kGridCols * kGridRows;

__device__ __forceinline__ int hamming_rows(const uint4* a, const uint4* b) {
  const uint4 a0 = a[0], a1 = a[1], b0 = b[0], b1 = b[1];
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// ---- grid: keypoint indices sorted by (cell, index); cell = ix*48 + iy (mGrid[ix][iy]) --------
__global__ void __launch_bounds__(1024)
grid_build_kernel(const OrbKeyPoint* __restrict__ kps, int nk, int nkpad, OrbFrameGrid g,
                  uint32_t* __restrict__ sorted /*[nkpad]*/, int* __restrict__ cell_start /*[kCells+1]*/) {
  extern __shared__ uint32_t s_key[];
  for (int i = threadIdx.x; i < nkpad; i += blockDim.x) {
    uint32_t key = 0xffffffffu;
    if (i < nk) {
      const int px = (int)roundf(__fmul_rn(__fsub_rn(kps[i].x, g.min_x), g.inv_w));  // PosInGrid
      const int py = (int)roundf(__fmul_rn(__fsub_rn(kps[i].y, g.min_y), g.inv_h));
      if (px >= 0 && px < kGridCols && py >= 0 && py < kGridRows)
        key = ((uint32_t)(px * kGridRows + py) << 16) | (uint32_t)i;
    }
    s_key[i] = key;
  }
  __syncthreads();
  for (int k = 2; k <= nkpad; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (nkpad >> 1); t += blockDim.x) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int l = i | j;
        const uint32_t x = s_key[i], y = s_key[l];
        if ((x > y) == ((i & k) == 0)) { s_key[i] = y; s_key[l] = x; }
      }
      __syncthreads();
    }
  for (int i = threadIdx.x; i < nkpad; i += blockDim.x) sorted[i] = s_key[i];
  for (int c = threadIdx.x; c <= kCells; c += blockDim.x) {
    const uint32_t target = (uint32_t)c << 16;
    int lo = 0, hi = nkpad;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (s_key[mid] < target) lo = mid + 1; else hi = mid;
    }
    cell_start[c] = lo;
  }
}

// ---- the kTopK best candidates of a query, sorted by (distance, position): one warp ------------------------------
constexpr int kTopK = 8;
__device__ __forceinline__ void topk_of_query(const uint32_t* entries /*may have been written by this warp*/, int beg,
                                              int end, uint2* row /*[kTopK]: (dist << 20 | position, entry)*/, int lane) {
  unsigned last = 0;
  for (int t = 0; t < kTopK; ++t) {
    unsigned best = 0xffffffffu;
    for (int e = beg + lane; e < end; e += 32) {
      const unsigned key = (((entries[e] >> 16) & 0x1ffu) << 20) | (unsigned)(e - beg);
      if ((t == 0 || key > last) && key < best) best = key;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) row[t] = make_uint2(best, best == 0xffffffffu ? 0u : entries[beg + (best & 0xfffffu)]);
    if (best == 0xffffffffu) {
      if (lane == 0) for (int u = t + 1; u < kTopK; ++u) row[u] = make_uint2(0xffffffffu, 0u);
      break;
    }
    last = best;
  }
}

// exclusive scan of in[0..n) into out[0..n], out[n] = total, by one whole block
__device__ __forceinline__ int block_scan_counts(const int* in, int n, int* out, int* s_part /*[33]*/) {
  const int T = blockDim.x, t = threadIdx.x;
  const int chunk = (n + T - 1) / T, lo = min(t * chunk, n), hi = min(lo + chunk, n);
  int sum = 0;
  for (int k = lo; k < hi; ++k) sum += in[k];
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((t & 31) >= o) incl += v;
  }
  if ((t & 31) == 31) s_part[t >> 5] = incl;
  __syncthreads();
  if (t < 32) {
    const int nw = T >> 5;
    const int w = t < nw ? s_part[t] : 0;
    int wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, wi, o);
      if (t >= o) wi += v;
    }
    if (t < nw) s_part[t] = wi - w;
    if (t == nw - 1) s_part[32] = wi;
  }
  __syncthreads();
  int run = s_part[t >> 5] + incl - sum;
  for (int k = lo; k < hi; ++k) { out[k] = run; run += in[k]; }
  const int total = s_part[32];
  if (t == 0) out[n] = total;
  return total;
}

// ---- candidates of every query, in GetFeaturesInArea order -----------------------------------
// FILL == false: count only.  Entry = kp index | dist << 16 | octave << 25.
template <bool FILL>
__global__ void __launch_bounds__(256)
window_candidates_kernel(const OrbWindowQuery* __restrict__ q, const uint4* __restrict__ qdesc, int nq,
                         const OrbKeyPoint* __restrict__ kps, const uint4* __restrict__ kdesc,
                         const float* __restrict__ u_right, OrbFrameGrid g, int gate,
                         const float* __restrict__ inv_sigma2,
                         const uint32_t* __restrict__ sorted, const int* __restrict__ cell_start,
                         int* __restrict__ counts, const int* __restrict__ offsets,
                         uint32_t* __restrict__ entries, int ent_cap, int* __restrict__ overflow,
                         uint2* __restrict__ topk = nullptr /*FILL: also the kTopK best of every query*/,
                         int* __restrict__ ticket = nullptr /*!FILL: the last block to finish scans the counts ...*/,
                         int* __restrict__ scan_out = nullptr /*... into these offsets [nq + 1] ...*/,
                         int* __restrict__ tail = nullptr /*... and resets (nmatches, total, overflow)*/) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const bool active = warp < nq;
  if (FILL) {
    if (!active) return;
    if (offsets[nq] > ent_cap) {  // the entry pool is too small: flag it, the host grows the pool and reruns
      if (warp == 0 && lane == 0) *overflow = 1;
      return;
    }
  }
  OrbWindowQuery p = OrbWindowQuery();
  if (active) p = q[warp];
  int n = 0;
  if (active && p.valid) {
    const float rad = p.radius;
    const int min_level = p.min_level, max_level = p.max_level;
    const bool check = min_level > 0 || max_level >= 0;  // bCheckLevels (Frame.cc:348)
    const float bx = __fsub_rn(p.u, g.min_x), by = __fsub_rn(p.v, g.min_y);
    const int c0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(bx, rad), g.inv_w)));
    const int c1 = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(bx, rad), g.inv_w)));
    const int r0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(by, rad), g.inv_h)));
    const int r1 = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(by, rad), g.inv_h)));
    if (c0 < kGridCols && c1 >= 0 && r0 < kGridRows && r1 >= 0) {
      uint32_t* dst = FILL ? entries + offsets[warp] : nullptr;
      for (int ix = c0; ix <= c1; ++ix) {
        // cells (ix, r0..r1) are contiguous in the sorted array
        const int beg = cell_start[ix * kGridRows + r0], end = cell_start[ix * kGridRows + r1 + 1];
        for (int e0 = beg; e0 < end; e0 += 32) {
          const int e = e0 + lane;
          bool ok = false;
          int idx = 0, oct = 0;
          if (e < end) {
            idx = (int)(sorted[e] & 0xffffu);
            const OrbKeyPoint kp = kps[idx];
            oct = kp.octave;
            ok = fabsf(__fsub_rn(kp.x, p.u)) < rad && fabsf(__fsub_rn(kp.y, p.v)) < rad;
            if (check && (oct < min_level || (max_level >= 0 && oct > max_level))) ok = false;
            if (ok && gate == 1 && u_right) {
              const float ur = u_right[idx];
              if (ur > 0 && fabsf(__fsub_rn(p.ur, ur)) > rad) ok = false;  // stereo gate (:91-96, :1407-1413)
            }
            if (ok && gate == 2) {  // chi-square gate of Fuse (:904-931)
              const float ex = __fsub_rn(p.u, kp.x), ey = __fsub_rn(p.v, kp.y);
              float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
              double lim = 5.99;
              if (u_right && u_right[idx] >= 0) {
                const float er = __fsub_rn(p.ur, u_right[idx]);
                e2 = __fadd_rn(e2, __fmul_rn(er, er));
                lim = 7.8;
              }
              if ((double)__fmul_rn(e2, inv_sigma2[oct]) > lim) ok = false;
            }
          }
          const unsigned m = __ballot_sync(0xffffffffu, ok);
          if (FILL && ok) {
            const int d = hamming_rows(qdesc + 2 * (size_t)warp, kdesc + 2 * (size_t)idx);
            dst[n + __popc(m & ((1u << lane) - 1))] = (uint32_t)idx | ((uint32_t)d << 16) | ((uint32_t)oct << 25);
          }
          n += __popc(m);
        }
      }
    }
  }
  if (FILL) {
    if (topk) {
      __syncwarp();
      topk_of_query(entries, offsets[warp], offsets[warp] + n, topk + (size_t)warp * kTopK, lane);
    }
    return;
  }
  if (active && lane == 0) counts[warp] = n;
  if (ticket) {  // every thread of every block gets here
    __shared__ int s_last, s_part[33];
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (s_last) {
      __threadfence();
      const int total = block_scan_counts(counts, nq, scan_out, s_part);
      if (threadIdx.x == 0) {
        if (tail) { tail[0] = 0; tail[1] = total; tail[2] = 0; }
        *ticket = 0;  // ready for the next call
      }
    }
  }
}

__global__ void __launch_bounds__(1024) exclusive_scan_kernel(const int* in, int n, int* out /*[n+1]*/,
                                                              int* tail = nullptr /*[3]: nmatches, total, overflow*/) {
  __shared__ int s_part[1024];
  const int T = blockDim.x, t = threadIdx.x;
  const int chunk = (n + T - 1) / T, lo = min(t * chunk, n), hi = min(lo + chunk, n);
  int sum = 0;
  for (int k = lo; k < hi; ++k) sum += in[k];
  s_part[t] = sum;
  __syncthreads();
  if (t == 0) {
    int run = 0;
    for (int k = 0; k < T; ++k) { const int v = s_part[k]; s_part[k] = run; run += v; }
    out[n] = run;
    if (tail) { tail[0] = 0; tail[1] = run; tail[2] = 0; }
  }
  __syncthreads();
  int run = s_part[t];
  for (int k = lo; k < hi; ++k) { out[k] = run; run += in[k]; }
}

// ---- the kTopK best candidates of every query (stand-alone form for the searches over vocabulary nodes) ---------
// The ordered resolution below re-evaluates every query once per round; with its candidates pre-sorted a
// round costs one thread and a handful of loads per query instead of a warp scan of the whole list.  A query
// whose kTopK best are all taken falls back to the full scan (rare).  One warp per query, any number of blocks.
__global__ void __launch_bounds__(256)
topk_entries_kernel(const int* __restrict__ offsets, const uint32_t* __restrict__ entries, int nq,
                    uint2* __restrict__ topk /*[nq][kTopK]*/, int ent_cap = INT_MAX) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= nq) return;
  if (offsets[nq] > ent_cap) return;
  topk_of_query(entries, offsets[warp], offsets[warp + 1], topk + (size_t)warp * kTopK, lane);
}

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1601-1642) on the 30 bin sizes
__device__ inline void three_maxima(const int* hist, int* keep) {
  int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
  for (int i = 0; i < ORB_GPU_HISTO_LENGTH; ++i) {
    const int sz = hist[i];
    if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
    else if (sz > max3) { max3 = sz; ind3 = i; }
  }
  if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
  else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
  keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
}

// ---- ordered claim resolution as a fixpoint ---------------------------------------------------
// TOP2: stable top-2 with the same-level ratio rule of :102-121 (M2); else best-only, accept <= th (M3/M4/M5).
// check_ori: rotation histogram over the accepted matches, ComputeThreeMaxima (:1601-1642), losers un-matched.
template <bool TOP2>
__global__ void __launch_bounds__(1024)
window_resolve_kernel(const OrbWindowQuery* __restrict__ q, int nq, const OrbKeyPoint* __restrict__ kps, int nk,
                      const int* __restrict__ offsets, const uint32_t* __restrict__ entries,
                      const uint2* __restrict__ topk, float nnratio,
                      int ratio_rule, int th_accept, int check_ori, uint8_t* blocked_g /*in/out [nk]*/,
                      int* claim_a, int* claim_b, int* minclaim_g /*[nk]*/,
                      int32_t* __restrict__ match_of_kp, int32_t* __restrict__ match_of_query,
                      int* __restrict__ nmatches, int use_smem = 0, int ent_cap = INT_MAX,
                      uint8_t* blocked_out = nullptr /*use_smem only: final flags go here instead of blocked_g*/) {
  extern __shared__ __align__(16) int s_dyn[];
  __shared__ int s_changed, s_count;
  __shared__ int s_hist[ORB_GPU_HISTO_LENGTH], s_keep[3];
  const int T = blockDim.x, t = threadIdx.x, warp = t >> 5, lane = t & 31, nwarps = T >> 5;
  if (offsets[nq] > ent_cap) return;  // entry pool overflow (flagged by the fill kernel): nothing to resolve yet
  // The fixpoint rounds touch only the claims, the per-keypoint minimum and the blocked flags: they live in shared
  // memory when they fit (use_smem: 8 nq + 5 nk + nq bytes), which turns a round from a few L2 round trips into
  // a few hundred cycles.
  int* prev = claim_a;
  int* cur = claim_b;
  int* minclaim = minclaim_g;
  uint8_t* blocked = blocked_g;
  uint8_t* qblocks = nullptr;
  if (use_smem) {
    prev = s_dyn; cur = prev + nq; minclaim = cur + nq;
    blocked = reinterpret_cast<uint8_t*>(minclaim + nk);
    qblocks = blocked + ((nk + 3) & ~3);
    for (int k = t; k < nk; k += T) blocked[k] = blocked_g[k];
    for (int i = t; i < nq; i += T) qblocks[i] = q[i].blocks;
  }
  for (int i = t; i < nq; i += T) prev[i] = -1;
  __syncthreads();
  for (int round = 0; round <= nq; ++round) {
    for (int k = t; k < nk; k += T) minclaim[k] = INT_MAX;
    if (t == 0) s_changed = 0;
    __syncthreads();
    for (int i = t; i < nq; i += T)
      if (prev[i] >= 0 && (qblocks ? qblocks[i] : q[i].blocks)) atomicMin(&minclaim[prev[i]], i);
    __syncthreads();
    for (int i = t; i < nq; i += T) {
      const int beg = offsets[i], end = offsets[i + 1];
      unsigned k1 = 0xffffffffu, k2 = 0xffffffffu;  // two smallest (dist << 20 | position) still free
      const int nent = end - beg, ntop = min(nent, kTopK);
      for (int u = 0; u < ntop; ++u) {
        const uint2 c = topk[(size_t)i * kTopK + u];
        const int idx = (int)(c.y & 0xffffu);
        if (blocked[idx] || minclaim[idx] < i) continue;  // keypoint already taken (:87-89)
        if (k1 == 0xffffffffu) { k1 = c.x; if (!TOP2) break; }
        else { k2 = c.x; break; }
      }
      if (nent > kTopK && (k1 == 0xffffffffu || (TOP2 && k2 == 0xffffffffu))) {  // the best kTopK did not settle it
        k1 = k2 = 0xffffffffu;
        for (int e = beg; e < end; ++e) {
          const uint32_t en = entries[e];
          const int idx = (int)(en & 0xffffu);
          if (blocked[idx] || minclaim[idx] < i) continue;
          const unsigned key = (((en >> 16) & 0x1ffu) << 20) | (unsigned)(e - beg);
          if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
        }
      }
      {
        int result = -1;
        if (k1 != 0xffffffffu) {
          const int best_dist = (int)(k1 >> 20);
          const uint32_t e1 = entries[beg + (k1 & 0xfffffu)];
          bool ok = best_dist <= th_accept;
          if (TOP2 && ok) {
            int best_dist2 = 256, level2 = -1;
            if (k2 != 0xffffffffu) {
              best_dist2 = (int)(k2 >> 20);
              level2 = (int)(entries[beg + (k2 & 0xfffffu)] >> 25);
            }
            const int level1 = (int)(e1 >> 25);
            if (ratio_rule == 0) {  // SearchByProjection (:115-116): only between candidates of one level
              if (level1 == level2 && (float)best_dist > __fmul_rn(nnratio, (float)best_dist2)) ok = false;
            } else {                // SearchByBoW (:230, :600): strict, against 256 when there is no second
              ok = (float)best_dist < __fmul_rn(nnratio, (float)best_dist2);
            }
          }
          if (ok) result = (int)(e1 & 0xffffu);
        }
        cur[i] = result;
        if (result != prev[i]) s_changed = 1;
      }
    }
    __syncthreads();
    int* tmp = prev; prev = cur; cur = tmp;
    const int changed = s_changed;
    __syncthreads();
    if (!changed) break;
  }
  // final state: last claimer per keypoint, blocked flags, match count
  for (int k = t; k < nk; k += T) { match_of_kp[k] = -1; }
  if (t == 0) s_count = 0;
  if (t < ORB_GPU_HISTO_LENGTH) s_hist[t] = 0;
  __syncthreads();
  int local = 0;
  for (int i = t; i < nq; i += T)
    if (prev[i] >= 0) {
      atomicMax(&match_of_kp[prev[i]], i);
      ++local;
    }
  if (local) atomicAdd(&s_count, local);
  __syncthreads();
  // a keypoint keeps the flag of its last claimer (claims by points with blocks == 0 can be overwritten)
  for (int i = t; i < nq; i += T)
    if (prev[i] >= 0 && match_of_kp[prev[i]] == i) blocked[prev[i]] = q[i].blocks;
  if (check_ori) {
    // rotation histogram (:1432-1441): bin = round(rot / 30), rot in [0, 360)
    const float factor = __fdiv_rn(1.0f, (float)ORB_GPU_HISTO_LENGTH);
    for (int i = t; i < nq; i += T)
      if (prev[i] >= 0) {
        float rot = __fsub_rn(q[i].angle, kps[prev[i]].angle);
        if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
        int bin = (int)roundf(__fmul_rn(rot, factor));
        if (bin == ORB_GPU_HISTO_LENGTH) bin = 0;
        cur[i] = bin;
        atomicAdd(&s_hist[bin], 1);
      }
    __syncthreads();
    if (t == 0) three_maxima(s_hist, s_keep);
    __syncthreads();
    local = 0;
    for (int i = t; i < nq; i += T)
      if (prev[i] >= 0) {
        const int bin = cur[i];
        if (bin != s_keep[0] && bin != s_keep[1] && bin != s_keep[2]) {  // :1456-1464
          match_of_kp[prev[i]] = -2;  // reset to NULL by the rotation check
          blocked[prev[i]] = 0;
          ++local;
        }
      }
    if (local) atomicSub(&s_count, local);
  }
  __syncthreads();
  for (int i = t; i < nq; i += T) {  // per-query result (-1: none, or un-matched by the rotation check)
    const int k = prev[i];
    match_of_query[i] = (k >= 0 && match_of_kp[k] != -2) ? k : -1;
  }
  if (use_smem) {
    uint8_t* dst = blocked_out ? blocked_out : blocked_g;
    for (int k = t; k < nk; k += T) dst[k] = blocked[k];
  }
  if (t == 0) *nmatches = s_count;
}

// ---- stereo row-band search -------------------------------------------------------------------
__global__ void __launch_bounds__(256)
stereo_match_kernel(const OrbKeyPoint* __restrict__ kl, const uint4* __restrict__ dl, int nl,
                    const OrbKeyPoint* __restrict__ kr, const uint4* __restrict__ dr, int nr, int n_rows,
                    const float* __restrict__ sf, float min_d, float max_d, int th_high,
                    int32_t* __restrict__ best_idx_r, int32_t* __restrict__ best_dist) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= nl) return;
  const OrbKeyPoint L = kl[warp];
  const int row = (int)L.y;  // vRowIndices[vL] (Frame.cc:511)
  const float min_u = __fsub_rn(L.x, max_d), max_u = __fsub_rn(L.x, min_d);
  unsigned best = 0xffffffffu;
  if (row >= 0 && row < n_rows && !(max_u < 0)) {
    for (int i = lane; i < nr; i += 32) {
      const OrbKeyPoint R = kr[i];
      const float r = __fmul_rn(2.0f, sf[R.octave]);
      const int maxr = (int)ceilf(__fadd_rn(R.y, r)), minr = (int)floorf(__fsub_rn(R.y, r));
      if (row < minr || row > maxr) continue;
      if (R.octave < L.octave - 1 || R.octave > L.octave + 1) continue;
      if (!(R.x >= min_u && R.x <= max_u)) continue;
      const unsigned d = (unsigned)hamming_rows(dl + 2 * (size_t)warp, dr + 2 * (size_t)i);
      best = min(best, (d << 16) | (unsigned)i);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0) {
    const int d = best == 0xffffffffu ? 1 << 20 : (int)(best >> 16);
    const bool ok = d < th_high;  // bestDist starts at TH_HIGH, strict '<' (:522,:543)
    best_idx_r[warp] = ok ? (int)(best & 0xffffu) : -1;
    best_dist[warp] = ok ? d : th_high;
  }
}

// ---- bucket searches (SearchByBoW x2, SearchForTriangulation): equal vocabulary nodes ----------
// A query = one feature of frame 1 with the index list of the same node in frame 2 (cand[begin..end), in the
// node's insertion order).  Entries use the layout of the windowed search (idx | dist << 16, level 0), so the
// ordered claims are resolved by window_resolve_kernel.
__global__ void __launch_bounds__(256)
bucket_entries_kernel(const OrbBucketQuery* __restrict__ q, int nq, const int32_t* __restrict__ cand,
                      const uint4* __restrict__ d1, const uint4* __restrict__ d2,
                      const int* __restrict__ offsets, uint32_t* __restrict__ entries) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= nq) return;
  const OrbBucketQuery Q = q[warp];
  const int base = offsets[warp];
  for (int e = Q.begin + lane; e < Q.end; e += 32) {
    const int idx2 = cand[e];
    const unsigned d = (unsigned)hamming_rows(d1 + 2 * (size_t)Q.index, d2 + 2 * (size_t)idx2);
    entries[base + (e - Q.begin)] = (unsigned)idx2 | (d << 16);
  }
}

struct TriangulationParams {
  float F12[9];
  float ex, ey;
  float scale2[kMaxLevels], sigma2_2[kMaxLevels];
  int only_stereo, check_ori, th_low;
};

// SearchForTriangulation (src/ORBmatcher.cc:657-823): no claims (vbMatched2 is never set, :677), so every
// query is independent: among the candidates with dist <= TH_LOW that pass the epipole and epipolar-line gates
// (:742-755, :138-157) the smallest distance wins and, on ties, the LAST one in node order (:737 skips only
// dist > bestDist).  One block: warp per query, then the rotation histogram (:790-808).
__global__ void __launch_bounds__(1024)
triangulation_kernel(const OrbBucketQuery* __restrict__ q, int nq, const int32_t* __restrict__ cand,
                     const OrbKeyPoint* __restrict__ k1, const uint4* __restrict__ d1,
                     const float* __restrict__ ur1, const OrbKeyPoint* __restrict__ k2,
                     const uint4* __restrict__ d2, const float* __restrict__ ur2,
                     const uint8_t* __restrict__ has_mp2, TriangulationParams P, int* __restrict__ bins,
                     int32_t* __restrict__ match_of_query, int* __restrict__ nmatches) {
  __shared__ int s_hist[ORB_GPU_HISTO_LENGTH], s_keep[3], s_count;
  const int T = blockDim.x, t = threadIdx.x, warp = t >> 5, lane = t & 31, nwarps = T >> 5;
  if (t < ORB_GPU_HISTO_LENGTH) s_hist[t] = 0;
  if (t == 0) s_count = 0;
  __syncthreads();
  for (int i = warp; i < nq; i += nwarps) {
    const OrbBucketQuery Q = q[i];
    const OrbKeyPoint kp1 = k1[Q.index];
    const bool stereo1 = ur1 && ur1[Q.index] >= 0;
    // epipolar line l = x1' F12 (:141-143), one rounding per operation, left to right
    const float a = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.F12[0]), __fmul_rn(kp1.y, P.F12[3])), P.F12[6]);
    const float b = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.F12[1]), __fmul_rn(kp1.y, P.F12[4])), P.F12[7]);
    const float c = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.F12[2]), __fmul_rn(kp1.y, P.F12[5])), P.F12[8]);
    const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
    unsigned best = 0xffffffffu;  // dist << 20 | (0xfffff - position): smallest distance, then last position
    for (int e = Q.begin + lane; e < Q.end; e += 32) {
      const int idx2 = cand[e];
      if (has_mp2[idx2]) continue;  // :722-726
      const bool stereo2 = ur2 && ur2[idx2] >= 0;
      if (P.only_stereo && !stereo2) continue;
      const int dist = hamming_rows(d1 + 2 * (size_t)Q.index, d2 + 2 * (size_t)idx2);
      if (dist > P.th_low) continue;
      const OrbKeyPoint kp2 = k2[idx2];
      if (!stereo1 && !stereo2) {
        const float dx = __fsub_rn(P.ex, kp2.x), dy = __fsub_rn(P.ey, kp2.y);
        if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, P.scale2[kp2.octave])) continue;
      }
      const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, kp2.x), __fmul_rn(b, kp2.y)), c);
      if (den == 0) continue;
      const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
      if (!((double)dsqr < 3.84 * (double)P.sigma2_2[kp2.octave])) continue;
      best = min(best, ((unsigned)dist << 20) | (0xfffffu - (unsigned)(e - Q.begin)));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) {
      int res = -1;
      if (best != 0xffffffffu) {
        res = cand[Q.begin + (int)(0xfffffu - (best & 0xfffffu))];
        atomicAdd(&s_count, 1);
        if (P.check_ori) {
          float rot = __fsub_rn(kp1.angle, k2[res].angle);
          if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
          int bin = (int)roundf(__fmul_rn(rot, __fdiv_rn(1.0f, (float)ORB_GPU_HISTO_LENGTH)));
          if (bin == ORB_GPU_HISTO_LENGTH) bin = 0;
          bins[i] = bin;
          atomicAdd(&s_hist[bin], 1);
        }
      }
      match_of_query[i] = res;
    }
  }
  __syncthreads();
  if (P.check_ori) {
    if (t == 0) three_maxima(s_hist, s_keep);
    __syncthreads();
    int local = 0;
    for (int i = t; i < nq; i += T)
      if (match_of_query[i] >= 0) {
        const int bin = bins[i];
        if (bin != s_keep[0] && bin != s_keep[1] && bin != s_keep[2]) { match_of_query[i] = -1; ++local; }
      }
    if (local) atomicSub(&s_count, local);
    __syncthreads();
  }
  if (t == 0) *nmatches = s_count;
}

// SearchForInitialization (src/ORBmatcher.cc:405-520).  A source keypoint (level 0 only) looks in a window of
// frame 2 around its previously matched position; a target already taken at distance <= dist is invisible
// (:444-445), otherwise the newcomer steals it (:463-471).  vMatchedDistance[i2] only ever decreases, so the
// value query i sees equals min{dist_j : j < i accepted on i2}; the sequential result is the fixpoint of
// "query i evaluates against the accepted choices of all j < i", reached by rounds as in window_resolve_kernel.
// Accepted choices are kept as per-target linked lists (head[], next[]).
__global__ void __launch_bounds__(1024)
init_resolve_kernel(const OrbKeyPoint* __restrict__ kps1, int nq, const OrbKeyPoint* __restrict__ kps2, int nk,
                    const int* __restrict__ offsets, const uint32_t* __restrict__ entries,
                    const uint2* __restrict__ topk, float nnratio,
                    int th_low, int check_ori, int* __restrict__ claim_a, int* __restrict__ claim_b,
                    int* __restrict__ dist_a, int* __restrict__ dist_b, int* __restrict__ next,
                    int* __restrict__ head /*[nk]*/, int32_t* __restrict__ match12, int* __restrict__ nmatches) {
  __shared__ int s_changed, s_count;
  __shared__ int s_hist[ORB_GPU_HISTO_LENGTH], s_keep[3];
  const int T = blockDim.x, t = threadIdx.x, warp = t >> 5, lane = t & 31, nwarps = T >> 5;
  int *prev = claim_a, *cur = claim_b, *pdist = dist_a, *cdist = dist_b;
  for (int i = t; i < nq; i += T) prev[i] = -1;
  __syncthreads();
  for (int round = 0; round <= nq; ++round) {
    for (int k = t; k < nk; k += T) head[k] = -1;
    if (t == 0) s_changed = 0;
    __syncthreads();
    for (int i = t; i < nq; i += T)
      if (prev[i] >= 0) next[i] = atomicExch(&head[prev[i]], i);
    __syncthreads();
    for (int i = t; i < nq; i += T) {
      const int beg = offsets[i], end = offsets[i + 1];
      unsigned k1 = 0xffffffffu, k2 = 0xffffffffu;  // two smallest (dist << 20 | position) still open to query i
      const int nent = end - beg, ntop = min(nent, kTopK);
      for (int u = 0; u < ntop; ++u) {
        const uint2 c = topk[(size_t)i * kTopK + u];
        const int idx = (int)(c.y & 0xffffu), dist = (int)(c.x >> 20);
        int taken = INT_MAX;  // vMatchedDistance[idx] as query i sees it
        for (int j = head[idx]; j >= 0; j = next[j])
          if (j < i) taken = min(taken, pdist[j]);
        if (taken <= dist) continue;  // :444-445
        if (k1 == 0xffffffffu) k1 = c.x; else { k2 = c.x; break; }
      }
      if (nent > kTopK && k2 == 0xffffffffu) {  // the best kTopK did not settle it: full scan
        k1 = k2 = 0xffffffffu;
        for (int e = beg; e < end; ++e) {
          const uint32_t en = entries[e];
          const int idx = (int)(en & 0xffffu), dist = (int)((en >> 16) & 0x1ffu);
          int taken = INT_MAX;
          for (int j = head[idx]; j >= 0; j = next[j])
            if (j < i) taken = min(taken, pdist[j]);
          if (taken <= dist) continue;
          const unsigned key = ((unsigned)dist << 20) | (unsigned)(e - beg);
          if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
        }
      }
      {
        int result = -1, rdist = 0;
        if (k1 != 0xffffffffu) {
          const int best = (int)(k1 >> 20);
          const float second = k2 != 0xffffffffu ? (float)(int)(k2 >> 20) : (float)INT_MAX;
          if (best <= th_low && (float)best < __fmul_rn(second, nnratio)) {  // :459-461
            result = (int)(entries[beg + (k1 & 0xfffffu)] & 0xffffu);
            rdist = best;
          }
        }
        cur[i] = result;
        cdist[i] = rdist;
        if (result != prev[i]) s_changed = 1;
      }
    }
    __syncthreads();
    int* tmp = prev; prev = cur; cur = tmp;
    tmp = pdist; pdist = cdist; cdist = tmp;
    const int changed = s_changed;
    __syncthreads();
    if (!changed) break;
  }
  // vnMatches21 = last accepted query per target; everyone else on that target was robbed (:463-467)
  for (int k = t; k < nk; k += T) head[k] = -1;
  if (t == 0) s_count = 0;
  if (t < ORB_GPU_HISTO_LENGTH) s_hist[t] = 0;
  __syncthreads();
  for (int i = t; i < nq; i += T)
    if (prev[i] >= 0) atomicMax(&head[prev[i]], i);
  __syncthreads();
  const float factor = __fdiv_rn(1.0f, (float)ORB_GPU_HISTO_LENGTH);
  int local = 0;
  for (int i = t; i < nq; i += T) {
    const int k = prev[i];
    const bool keeps = k >= 0 && head[k] == i;
    match12[i] = keeps ? k : -1;
    local += keeps;
    if (k >= 0 && check_ori) {  // the histogram also counts matches that were robbed later (:472-481)
      float rot = __fsub_rn(kps1[i].angle, kps2[k].angle);
      if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
      int bin = (int)roundf(__fmul_rn(rot, factor));
      if (bin == ORB_GPU_HISTO_LENGTH) bin = 0;
      cur[i] = bin;
      atomicAdd(&s_hist[bin], 1);
    }
  }
  if (local) atomicAdd(&s_count, local);
  __syncthreads();
  if (check_ori) {
    if (t == 0) three_maxima(s_hist, s_keep);
    __syncthreads();
    local = 0;
    for (int i = t; i < nq; i += T)
      if (match12[i] >= 0) {
        const int bin = cur[i];
        if (bin != s_keep[0] && bin != s_keep[1] && bin != s_keep[2]) { match12[i] = -1; ++local; }
      }
    if (local) atomicSub(&s_count, local);
    __syncthreads();
  }
  if (t == 0) *nmatches = s_count;
}

}  // namespace orbgpu

using namespace orbgpu;

extern "C" {

}  // extern "C"

// ---- a frame's matcher-side data resident in device memory ------------------------------------------------------
// The reference keeps mvKeysUn / mDescriptors / mvuRight and the 64 x 48 grid in the Frame (src/Frame.cc:230-253) and
// every search reads them there.  OrbGpuFrame is that state on the device: uploaded (or copied from the extractor's
// device outputs) once, grid sorted once, then any number of searches -- each of which moves only its queries in (one
// copy from a pinned staging block) and its results out (one copy), with no allocation and one synchronisation.
namespace {

struct GrowDev {   // device buffer that only grows
  void* p = nullptr; size_t bytes = 0;
  int ensure(size_t need) {
    if (need <= bytes) return ORB_GPU_OK;
    if (p) cudaFree(p);
    p = nullptr; bytes = 0;
    need += need / 2 + 256;
    ORB_CUDA_TRY(cudaMalloc(&p, need));
    bytes = need;
    return ORB_GPU_OK;
  }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
  template <class T> T* as() const { return (T*)p; }
};
struct GrowPinned {  // pinned host staging block that only grows
  void* p = nullptr; size_t bytes = 0;
  int ensure(size_t need) {
    if (need <= bytes) return ORB_GPU_OK;
    if (p) cudaFreeHost(p);
    p = nullptr; bytes = 0;
    need += need / 2 + 256;
    ORB_CUDA_TRY(cudaHostAlloc(&p, need, cudaHostAllocDefault));
    bytes = need;
    return ORB_GPU_OK;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
};
inline size_t up16(size_t v) { return (v + 15) & ~(size_t)15; }

}  // namespace

struct OrbGpuFrame {
  int device = 0;
  cudaStream_t stream = nullptr;
  // target: keypoints (undistorted), descriptors, right coordinates, grid
  int n = 0;
  bool has_ur = false, grid_built = false;
  OrbFrameGrid grid{};
  GrowDev kps, desc, ur, sorted, cell;
  // workspace of a search
  GrowDev in, out, cnt, off, ent, top, ca, cb, minc, flag;
  GrowPinned hin, hout;
};

namespace {

int frame_build_grid(OrbGpuFrame* f, const OrbKeyPoint* d_kps) {
  int nkpad = 2;
  while (nkpad < f->n) nkpad <<= 1;
  int rc;
  if ((rc = f->sorted.ensure((size_t)nkpad * 4)) || (rc = f->cell.ensure((size_t)(kCells + 1) * 4))) return rc;
  const size_t smem = (size_t)nkpad * 4;
  if (smem > 48 * 1024)
    ORB_CUDA_TRY(cudaFuncSetAttribute(grid_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  grid_build_kernel<<<1, 1024, smem, f->stream>>>(d_kps, f->n, nkpad, f->grid, f->sorted.as<uint32_t>(), f->cell.as<int>());
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  f->grid_built = true;
  return ORB_GPU_OK;
}

// One windowed search against the frame's resident target.  `extra_in` bytes at the head of the staging block are the
// caller's (a transient target for the host-pointer entry points); queries, descriptors, flags follow.
int frame_search(OrbGpuFrame* f, const OrbWindowQuery* q, const uint8_t* qdesc, int nq, bool top2, float nnratio,
                 int th_accept, int gate, const float* inv_sigma2, int nlevels, int check_ori, uint8_t* kp_blocked,
                 int32_t* match_of_kp, int32_t* match_of_query, int* nmatches, const OrbKeyPoint* host_kps = nullptr,
                 const uint8_t* host_desc = nullptr, const float* host_ur = nullptr) {
  const int nk = f->n;
  *nmatches = 0;
  for (int i = 0; i < nk; ++i) match_of_kp[i] = -1;
  if (match_of_query) for (int i = 0; i < nq; ++i) match_of_query[i] = -1;
  if (nq == 0 || nk == 0) return ORB_GPU_OK;
  if (nk > 32768) return ORB_GPU_ERR_CAPACITY;
  ORB_CUDA_TRY(cudaSetDevice(f->device));
  cudaStream_t s = f->stream;
  int rc;
  // ---- input block: [transient target: kps | desc | ur] [q | qdesc | blocked | sigma]
  const bool transient = host_kps != nullptr;
  const size_t o_kps = 0, o_desc = o_kps + (transient ? up16((size_t)nk * sizeof(OrbKeyPoint)) : 0);
  const size_t o_ur = o_desc + (transient ? (size_t)nk * 32 : 0);
  const size_t o_q = o_ur + (transient && host_ur ? up16((size_t)nk * 4) : 0);
  const size_t o_qd = o_q + (size_t)nq * sizeof(OrbWindowQuery), o_blk = o_qd + (size_t)nq * 32;
  const size_t o_sig = o_blk + up16((size_t)nk), in_bytes = o_sig + kMaxLevels * 4;
  // ---- output block: [match_of_kp | match_of_query | blocked | nmatches, total, overflow]
  const size_t r_mq = (size_t)nk * 4, r_blk = r_mq + (size_t)nq * 4, r_tail = r_blk + up16((size_t)nk), out_bytes = r_tail + 16;
  if ((rc = f->hin.ensure(in_bytes)) || (rc = f->in.ensure(in_bytes)) || (rc = f->hout.ensure(out_bytes)) ||
      (rc = f->out.ensure(out_bytes)) || (rc = f->cnt.ensure((size_t)nq * 4)) || (rc = f->off.ensure((size_t)(nq + 1) * 4)) ||
      (rc = f->top.ensure((size_t)nq * kTopK * sizeof(uint2))) || (rc = f->ca.ensure((size_t)nq * 4)) ||
      (rc = f->cb.ensure((size_t)nq * 4)) || (rc = f->minc.ensure((size_t)nk * 4)) ||
      (rc = f->ent.ensure((size_t)std::max(nq * 48, 1 << 16) * 4)))
    return rc;
  uint8_t* hin = (uint8_t*)f->hin.p;
  if (transient) {
    std::memcpy(hin + o_kps, host_kps, (size_t)nk * sizeof(OrbKeyPoint));
    std::memcpy(hin + o_desc, host_desc, (size_t)nk * 32);
    if (host_ur) std::memcpy(hin + o_ur, host_ur, (size_t)nk * 4);
  }
  std::memcpy(hin + o_q, q, (size_t)nq * sizeof(OrbWindowQuery));
  std::memcpy(hin + o_qd, qdesc, (size_t)nq * 32);
  std::memcpy(hin + o_blk, kp_blocked, (size_t)nk);
  if (gate == 2) std::memcpy(hin + o_sig, inv_sigma2, (size_t)nlevels * 4);
  ORB_CUDA_TRY(cudaMemcpyAsync(f->in.p, hin, in_bytes, cudaMemcpyHostToDevice, s));
  uint8_t* din = (uint8_t*)f->in.p;
  const OrbKeyPoint* d_kps = transient ? (const OrbKeyPoint*)(din + o_kps) : f->kps.as<OrbKeyPoint>();
  const uint4* d_desc = transient ? (const uint4*)(din + o_desc) : f->desc.as<uint4>();
  const float* d_ur = transient ? (host_ur ? (const float*)(din + o_ur) : nullptr) : (f->has_ur ? f->ur.as<float>() : nullptr);
  if (!gate) d_ur = nullptr;
  if (transient || !f->grid_built) { if ((rc = frame_build_grid(f, d_kps))) return rc; }
  const OrbWindowQuery* d_q = (const OrbWindowQuery*)(din + o_q);
  const uint4* d_qd = (const uint4*)(din + o_qd);
  uint8_t* d_blk = din + o_blk;
  const float* d_sig = (const float*)(din + o_sig);
  uint8_t* dout = (uint8_t*)f->out.p;
  int32_t* d_match = (int32_t*)dout;
  int32_t* d_mq = (int32_t*)(dout + r_mq);
  int* d_tail = (int*)(dout + r_tail);  // nmatches, total, overflow
  const int blocks = (nq * 32 + 255) / 256;
  // shared-memory state of the resolve kernel: prev, cur [nq], minclaim [nk], blocked [nk], query flags [nq]
  const size_t rs_smem = (size_t)nq * 8 + (size_t)nk * 4 + up16((size_t)nk) + up16((size_t)nq);
  const int use_smem = rs_smem <= 200 * 1024;
  if (use_smem && rs_smem > 48 * 1024) {
    if (top2) ORB_CUDA_TRY(cudaFuncSetAttribute(window_resolve_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_smem));
    else ORB_CUDA_TRY(cudaFuncSetAttribute(window_resolve_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_smem));
  }
  // counts, then (in the last block to finish) their scan into offsets and the reset of the result tail
  window_candidates_kernel<false><<<blocks, 256, 0, s>>>(d_q, d_qd, nq, d_kps, d_desc, d_ur, f->grid, gate, d_sig,
      f->sorted.as<uint32_t>(), f->cell.as<int>(), f->cnt.as<int>(), nullptr, nullptr, 0, nullptr, nullptr,
      f->flag.as<int>(), f->off.as<int>(), d_tail);
  count_launch(1);
  uint8_t* d_blk_out = use_smem ? dout + r_blk : nullptr;
  for (int attempt = 0; attempt < 2; ++attempt) {
    const int ent_cap = (int)std::min<size_t>(f->ent.bytes / 4, (size_t)INT_MAX);
    window_candidates_kernel<true><<<blocks, 256, 0, s>>>(d_q, d_qd, nq, d_kps, d_desc, d_ur, f->grid, gate, d_sig,
        f->sorted.as<uint32_t>(), f->cell.as<int>(), nullptr, f->off.as<int>(), f->ent.as<uint32_t>(), ent_cap, d_tail + 2,
        f->top.as<uint2>());  // entries and the kTopK best of every query in one pass
    if (top2)
      window_resolve_kernel<true><<<1, 1024, use_smem ? rs_smem : 0, s>>>(d_q, nq, d_kps, nk, f->off.as<int>(),
          f->ent.as<uint32_t>(), f->top.as<uint2>(), nnratio, 0, th_accept, check_ori, d_blk, f->ca.as<int>(), f->cb.as<int>(),
          f->minc.as<int>(), d_match, d_mq, d_tail, use_smem, ent_cap, d_blk_out);
    else
      window_resolve_kernel<false><<<1, 1024, use_smem ? rs_smem : 0, s>>>(d_q, nq, d_kps, nk, f->off.as<int>(),
          f->ent.as<uint32_t>(), f->top.as<uint2>(), nnratio, 0, th_accept, check_ori, d_blk, f->ca.as<int>(), f->cb.as<int>(),
          f->minc.as<int>(), d_match, d_mq, d_tail, use_smem, ent_cap, d_blk_out);
    count_launch(2);
    ORB_CUDA_TRY(cudaGetLastError());
    // the blocked flags travel back inside the output block (the resolve kernel writes them there when its state is
    // in shared memory)
    if (!d_blk_out) ORB_CUDA_TRY(cudaMemcpyAsync(dout + r_blk, d_blk, (size_t)nk, cudaMemcpyDeviceToDevice, s));
    ORB_CUDA_TRY(cudaMemcpyAsync(f->hout.p, dout, out_bytes, cudaMemcpyDeviceToHost, s));
    ORB_CUDA_TRY(cudaStreamSynchronize(s));
    const int* tail = (const int*)((const uint8_t*)f->hout.p + r_tail);
    if (!tail[2]) break;
    if (attempt == 1) return ORB_GPU_ERR_CAPACITY;
    if ((rc = f->ent.ensure((size_t)tail[1] * 4))) return rc;  // grow the entry pool to the counted total and rerun
    ORB_CUDA_TRY(cudaMemsetAsync(d_tail + 2, 0, 4, s));
  }
  const uint8_t* hout = (const uint8_t*)f->hout.p;
  std::memcpy(match_of_kp, hout, (size_t)nk * 4);
  if (match_of_query) std::memcpy(match_of_query, hout + r_mq, (size_t)nq * 4);
  std::memcpy(kp_blocked, hout + r_blk, (size_t)nk);
  *nmatches = *(const int*)(hout + r_tail);
  return ORB_GPU_OK;
}

// the host-pointer entry points keep one scratch frame per thread and device (transient target, no residency)
OrbGpuFrame* scratch_frame(int device) {
  struct Holder {
    std::vector<OrbGpuFrame*> v;
    ~Holder() { for (OrbGpuFrame* f : v) orb_gpu_frame_destroy(f); }
  };
  static thread_local Holder h;
  for (OrbGpuFrame* f : h.v) if (f->device == device) return f;
  OrbGpuFrame* f = nullptr;
  if (orb_gpu_frame_create(device, &f) != ORB_GPU_OK) return nullptr;
  h.v.push_back(f);
  return f;
}

// window of every map point as the reference computes it on the host (:63-69, RadiusByViewingCos :131-137)
void proj_to_windows(const OrbProjQuery* q, int nq, const float* scale_factors, int nlevels, float th,
                     std::vector<OrbWindowQuery>& wq) {
  wq.resize((size_t)nq);
  for (int i = 0; i < nq; ++i) {
    OrbWindowQuery& w = wq[i];
    w = OrbWindowQuery();
    const OrbProjQuery& p = q[i];
    if (!p.valid || p.level < 0 || p.level >= nlevels) continue;
    float r = (double)p.view_cos > 0.998 ? 2.5f : 4.0f;
    if (th != 1.0f) r *= th;
    w.u = p.u; w.v = p.v; w.ur = p.ur;
    w.radius = r * scale_factors[p.level];
    w.min_level = p.level - 1; w.max_level = p.level;
    w.valid = 1; w.blocks = p.has_obs;
  }
}

}  // namespace

extern "C" {

int orb_gpu_frame_create(int device, OrbGpuFrame** out) {
  if (!out) return ORB_GPU_ERR_ARG;
  *out = nullptr;
  ORB_CUDA_TRY(cudaSetDevice(device));
  OrbGpuFrame* f = new OrbGpuFrame();
  f->device = device;
  const cudaError_t e = cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { set_last_error("cudaStreamCreate (frame)", e); delete f; return ORB_GPU_ERR_CUDA; }
  if (f->flag.ensure(16) != ORB_GPU_OK || cudaMemset(f->flag.p, 0, 16) != cudaSuccess) {  // scan ticket of the count kernel
    orb_gpu_frame_destroy(f);
    return ORB_GPU_ERR_CUDA;
  }
  *out = f;
  return ORB_GPU_OK;
}

int orb_gpu_frame_destroy(OrbGpuFrame* f) {
  if (!f) return ORB_GPU_OK;
  cudaSetDevice(f->device);
  cudaStreamSynchronize(f->stream);
  for (GrowDev* b : {&f->kps, &f->desc, &f->ur, &f->sorted, &f->cell, &f->in, &f->out, &f->cnt, &f->off, &f->ent, &f->top,
                     &f->ca, &f->cb, &f->minc, &f->flag})
    b->release();
  f->hin.release(); f->hout.release();
  cudaStreamDestroy(f->stream);
  delete f;
  return ORB_GPU_OK;
}

int orb_gpu_frame_size(const OrbGpuFrame* f) { return f ? f->n : 0; }

int orb_gpu_frame_upload(OrbGpuFrame* f, const OrbKeyPoint* kps, const uint8_t* desc, const float* u_right, int n,
                         OrbFrameGrid grid) {
  if (!f || n < 0 || (n > 0 && (!kps || !desc))) return ORB_GPU_ERR_ARG;
  if (n > 32768) return ORB_GPU_ERR_CAPACITY;
  ORB_CUDA_TRY(cudaSetDevice(f->device));
  f->n = n; f->grid = grid; f->has_ur = u_right != nullptr; f->grid_built = false;
  if (n == 0) return ORB_GPU_OK;
  int rc;
  const size_t bk = (size_t)n * sizeof(OrbKeyPoint), bd = (size_t)n * 32, bu = (size_t)n * 4;
  if ((rc = f->kps.ensure(bk)) || (rc = f->desc.ensure(bd)) || (rc = f->ur.ensure(bu)) ||
      (rc = f->hin.ensure(up16(bk) + bd + bu)))
    return rc;
  uint8_t* h = (uint8_t*)f->hin.p;
  std::memcpy(h, kps, bk);
  std::memcpy(h + up16(bk), desc, bd);
  if (u_right) std::memcpy(h + up16(bk) + bd, u_right, bu);
  ORB_CUDA_TRY(cudaMemcpyAsync(f->kps.p, h, bk, cudaMemcpyHostToDevice, f->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(f->desc.p, h + up16(bk), bd, cudaMemcpyHostToDevice, f->stream));
  if (u_right) ORB_CUDA_TRY(cudaMemcpyAsync(f->ur.p, h + up16(bk) + bd, bu, cudaMemcpyHostToDevice, f->stream));
  if ((rc = frame_build_grid(f, f->kps.as<OrbKeyPoint>()))) return rc;
  ORB_CUDA_TRY(cudaStreamSynchronize(f->stream));  // the staging block is reused by the next call
  return ORB_GPU_OK;
}

int orb_gpu_frame_from_device(OrbGpuFrame* f, const OrbKeyPoint* d_kps, const uint8_t* d_desc, const float* d_u_right,
                              int n, OrbFrameGrid grid, void* stream) {
  if (!f || n < 0 || (n > 0 && (!d_kps || !d_desc))) return ORB_GPU_ERR_ARG;
  if (n > 32768) return ORB_GPU_ERR_CAPACITY;
  ORB_CUDA_TRY(cudaSetDevice(f->device));
  f->n = n; f->grid = grid; f->has_ur = d_u_right != nullptr; f->grid_built = false;
  if (n == 0) return ORB_GPU_OK;
  int rc;
  if ((rc = f->kps.ensure((size_t)n * sizeof(OrbKeyPoint))) || (rc = f->desc.ensure((size_t)n * 32)) ||
      (rc = f->ur.ensure((size_t)n * 4)))
    return rc;
  // the producer's stream (the extractor's) must have finished writing: order the copies after it
  if (stream) ORB_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(f->kps.p, d_kps, (size_t)n * sizeof(OrbKeyPoint), cudaMemcpyDeviceToDevice, f->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(f->desc.p, d_desc, (size_t)n * 32, cudaMemcpyDeviceToDevice, f->stream));
  if (d_u_right) ORB_CUDA_TRY(cudaMemcpyAsync(f->ur.p, d_u_right, (size_t)n * 4, cudaMemcpyDeviceToDevice, f->stream));
  return frame_build_grid(f, f->kps.as<OrbKeyPoint>());
}

int orb_gpu_frame_search_by_projection(OrbGpuFrame* f, const OrbProjQuery* q, const uint8_t* qdesc, int nq,
                                       const float* scale_factors, int nlevels, float th, float nnratio, int th_high,
                                       uint8_t* kp_blocked, int32_t* match_of_kp, int* nmatches) {
  if (!f || nq < 0 || !nmatches || nlevels < 1 || nlevels > kMaxLevels || !scale_factors ||
      (f->n > 0 && (!kp_blocked || !match_of_kp)) || (nq > 0 && (!q || !qdesc)))
    return ORB_GPU_ERR_ARG;
  std::vector<OrbWindowQuery> wq;
  proj_to_windows(q, nq, scale_factors, nlevels, th, wq);
  return frame_search(f, wq.data(), qdesc, nq, true, nnratio, th_high, 1, nullptr, 0, 0, kp_blocked, match_of_kp, nullptr,
                      nmatches);
}

int orb_gpu_frame_search_windowed(OrbGpuFrame* f, const OrbWindowQuery* q, const uint8_t* qdesc, int nq, int th_accept,
                                  int gate, const float* inv_level_sigma2, int nlevels, int check_orientation,
                                  uint8_t* kp_blocked, int32_t* match_of_kp, int32_t* match_of_query, int* nmatches) {
  if (!f || nq < 0 || !nmatches || gate < 0 || gate > 2 ||
      (gate == 2 && (!inv_level_sigma2 || nlevels < 1 || nlevels > kMaxLevels)) ||
      (f->n > 0 && (!kp_blocked || !match_of_kp)) || (nq > 0 && (!q || !qdesc)))
    return ORB_GPU_ERR_ARG;
  return frame_search(f, q, qdesc, nq, false, 0.f, th_accept, gate, inv_level_sigma2, nlevels, check_orientation,
                      kp_blocked, match_of_kp, match_of_query, nmatches);
}

// shared body of the host-pointer windowed searches: a transient target on this thread's scratch frame
static int run_windowed(const OrbWindowQuery* q, const uint8_t* qdesc, int nq, const OrbKeyPoint* kps,
                        const uint8_t* kdesc, const float* u_right, int nk, OrbFrameGrid grid, bool top2,
                        float nnratio, int th_accept, int gate, const float* inv_sigma2, int nlevels,
                        int check_ori, int device, uint8_t* kp_blocked, int32_t* match_of_kp,
                        int32_t* match_of_query, int* nmatches) {
  *nmatches = 0;
  for (int i = 0; i < nk; ++i) match_of_kp[i] = -1;
  if (match_of_query) for (int i = 0; i < nq; ++i) match_of_query[i] = -1;
  if (nq == 0 || nk == 0) return ORB_GPU_OK;
  if (nk > 32768) return ORB_GPU_ERR_CAPACITY;
  OrbGpuFrame* f = scratch_frame(device);
  if (!f) return ORB_GPU_ERR_CUDA;
  f->n = nk; f->grid = grid; f->has_ur = u_right != nullptr; f->grid_built = false;
  return frame_search(f, q, qdesc, nq, top2, nnratio, th_accept, gate, inv_sigma2, nlevels, check_ori, kp_blocked,
                      match_of_kp, match_of_query, nmatches, kps, kdesc, u_right);
}

int orb_gpu_search_by_projection(const OrbProjQuery* q, const uint8_t* qdesc, int nq,
                                 const OrbKeyPoint* kps, const uint8_t* kdesc, const float* u_right,
                                 int nk, OrbFrameGrid grid, const float* scale_factors, int nlevels,
                                 float th, float nnratio, int th_high, int device, uint8_t* kp_blocked,
                                 int32_t* match_of_kp, int* nmatches) {
  if (nq < 0 || nk < 0 || !nmatches || nlevels < 1 || nlevels > kMaxLevels || !scale_factors ||
      (nk > 0 && (!kps || !kdesc || !kp_blocked || !match_of_kp)) || (nq > 0 && (!q || !qdesc)))
    return ORB_GPU_ERR_ARG;
  std::vector<OrbWindowQuery> wq;
  proj_to_windows(q, nq, scale_factors, nlevels, th, wq);
  return run_windowed(wq.data(), qdesc, nq, kps, kdesc, u_right, nk, grid, true, nnratio, th_high, 1, nullptr, 0, 0,
                      device, kp_blocked, match_of_kp, nullptr, nmatches);
}

int orb_gpu_search_windowed(const OrbWindowQuery* q, const uint8_t* qdesc, int nq, const OrbKeyPoint* kps,
                            const uint8_t* kdesc, const float* u_right, int nk, OrbFrameGrid grid,
                            int th_accept, int gate, const float* inv_level_sigma2, int nlevels,
                            int check_orientation, int device, uint8_t* kp_blocked, int32_t* match_of_kp,
                            int32_t* match_of_query, int* nmatches) {
  if (nq < 0 || nk < 0 || !nmatches || gate < 0 || gate > 2 ||
      (gate == 2 && (!inv_level_sigma2 || nlevels < 1 || nlevels > kMaxLevels)) ||
      (nk > 0 && (!kps || !kdesc || !kp_blocked || !match_of_kp)) || (nq > 0 && (!q || !qdesc)))
    return ORB_GPU_ERR_ARG;
  return run_windowed(q, qdesc, nq, kps, kdesc, gate ? u_right : nullptr, nk, grid, false, 0.f, th_accept, gate,
                      inv_level_sigma2, nlevels, check_orientation, device, kp_blocked, match_of_kp,
                      match_of_query, nmatches);
}

// Candidate lists of GetFeaturesInArea for a batch of windows, in the reference's order (N2; used by tests and by
// callers that want the lists themselves).
int orb_gpu_features_in_area(const OrbWindowQuery* q, int nq, const OrbKeyPoint* kps, int nk, OrbFrameGrid grid,
                             int device, int32_t* offsets, int32_t* indices, int capacity) {
  if (nq < 0 || nk < 0 || !offsets || capacity < 0 || (capacity > 0 && !indices) || (nq > 0 && !q) || (nk > 0 && !kps))
    return ORB_GPU_ERR_ARG;
  for (int i = 0; i <= nq; ++i) offsets[i] = 0;
  if (nq == 0 || nk == 0) return ORB_GPU_OK;
  if (nk > 32768) return ORB_GPU_ERR_CAPACITY;
  { const int rc_pool = use_cached_pool(device); if (rc_pool) return rc_pool; }
  int nkpad = 2;
  while (nkpad < nk) nkpad <<= 1;
  TmpDevW dq, dqd, dk, dkd, dsorted, dcell, dcnt, doff, dent;
  int rc;
  if ((rc = dq.alloc((size_t)nq * sizeof(OrbWindowQuery))) || (rc = dqd.alloc((size_t)nq * 32)) ||
      (rc = dk.alloc((size_t)nk * sizeof(OrbKeyPoint))) || (rc = dkd.alloc((size_t)nk * 32)) ||
      (rc = dsorted.alloc((size_t)nkpad * 4)) || (rc = dcell.alloc((size_t)(kCells + 1) * 4)) ||
      (rc = dcnt.alloc((size_t)nq * 4)) || (rc = doff.alloc((size_t)(nq + 1) * 4)))
    return rc;
  ORB_CUDA_TRY(cudaMemcpy(dq.p, q, (size_t)nq * sizeof(OrbWindowQuery), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dk.p, kps, (size_t)nk * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemset(dqd.p, 0, (size_t)nq * 32));
  ORB_CUDA_TRY(cudaMemset(dkd.p, 0, (size_t)nk * 32));
  const size_t smem = (size_t)nkpad * 4;
  if (smem > 48 * 1024)
    ORB_CUDA_TRY(cudaFuncSetAttribute(grid_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  grid_build_kernel<<<1, 1024, smem>>>(dk.as<OrbKeyPoint>(), nk, nkpad, grid, dsorted.as<uint32_t>(), dcell.as<int>());
  const int blocks = (nq * 32 + 255) / 256;
  window_candidates_kernel<false><<<blocks, 256>>>(dq.as<OrbWindowQuery>(), dqd.as<uint4>(), nq, dk.as<OrbKeyPoint>(),
      dkd.as<uint4>(), nullptr, grid, 0, nullptr, dsorted.as<uint32_t>(), dcell.as<int>(), dcnt.as<int>(), nullptr,
      nullptr, 0, nullptr);
  exclusive_scan_kernel<<<1, 1024>>>(dcnt.as<int>(), nq, doff.as<int>());
  ORB_CUDA_TRY(cudaMemcpy(offsets, doff.p, (size_t)(nq + 1) * 4, cudaMemcpyDeviceToHost));
  const int total = offsets[nq];
  count_launch(3);
  if (total > capacity) return ORB_GPU_ERR_CAPACITY;  // offsets are valid: the caller can size the buffer
  if (total == 0) return ORB_GPU_OK;
  if ((rc = dent.alloc((size_t)total * 4))) return rc;
  window_candidates_kernel<true><<<blocks, 256>>>(dq.as<OrbWindowQuery>(), dqd.as<uint4>(), nq, dk.as<OrbKeyPoint>(),
      dkd.as<uint4>(), nullptr, grid, 0, nullptr, dsorted.as<uint32_t>(), dcell.as<int>(), nullptr, doff.as<int>(),
      dent.as<uint32_t>(), INT_MAX, nullptr);
  count_launch(1);
  ORB_CUDA_TRY(cudaGetLastError());
  std::vector<uint32_t> ent((size_t)total);
  ORB_CUDA_TRY(cudaMemcpy(ent.data(), dent.p, (size_t)total * 4, cudaMemcpyDeviceToHost));
  for (int i = 0; i < total; ++i) indices[i] = (int32_t)(ent[i] & 0xffffu);
  return ORB_GPU_OK;
}

static int check_bucket_queries(const OrbBucketQuery* q, int nq, int ncand2, int n1, const int32_t* cand2,
                                int n2, std::vector<int>& offsets) {
  offsets.assign((size_t)nq + 1, 0);
  for (int i = 0; i < nq; ++i) {
    if (q[i].index < 0 || q[i].index >= n1 || q[i].begin < 0 || q[i].end < q[i].begin || q[i].end > ncand2)
      return ORB_GPU_ERR_ARG;
    offsets[i + 1] = offsets[i] + (q[i].end - q[i].begin);
  }
  for (int i = 0; i < ncand2; ++i)
    if (cand2[i] < 0 || cand2[i] >= n2) return ORB_GPU_ERR_ARG;
  return ORB_GPU_OK;
}

int orb_gpu_search_by_bow(const OrbBucketQuery* q, int nq, const int32_t* cand2, int ncand2,
                          const OrbKeyPoint* kps1, const uint8_t* desc1, int n1, const OrbKeyPoint* kps2,
                          const uint8_t* desc2, const uint8_t* skip2, int n2, float nnratio, int th_accept,
                          int check_orientation, int device, int32_t* match12, int32_t* match21,
                          int* nmatches) {
  if (nq < 0 || n1 < 0 || n2 < 0 || ncand2 < 0 || !nmatches || (n1 > 0 && !match12) || (n2 > 0 && !match21) ||
      (nq > 0 && (!q || !kps1 || !desc1)) || (ncand2 > 0 && (!cand2 || !kps2 || !desc2)))
    return ORB_GPU_ERR_ARG;
  *nmatches = 0;
  for (int i = 0; i < n1; ++i) match12[i] = -1;
  for (int i = 0; i < n2; ++i) match21[i] = -1;
  if (nq == 0 || n2 == 0 || ncand2 == 0) return ORB_GPU_OK;
  if (n2 > 32768) return ORB_GPU_ERR_CAPACITY;
  std::vector<int> offsets;
  int rc;
  if ((rc = check_bucket_queries(q, nq, ncand2, n1, cand2, n2, offsets))) return rc;
  std::vector<OrbWindowQuery> wq((size_t)nq);
  for (int i = 0; i < nq; ++i) {
    wq[i] = OrbWindowQuery();
    wq[i].angle = kps1[q[i].index].angle;
    wq[i].valid = 1;
    wq[i].blocks = 1;
  }
  { const int rc_pool = use_cached_pool(device); if (rc_pool) return rc_pool; }
  TmpDevW dq, dwq, dc, d1, dk2, d2, doff, dent, dblk, dca, dcb, dmin, dmatch, dmq, dn;
  const int total = offsets[nq];
  if ((rc = dq.alloc((size_t)nq * sizeof(OrbBucketQuery))) || (rc = dwq.alloc((size_t)nq * sizeof(OrbWindowQuery))) ||
      (rc = dc.alloc((size_t)ncand2 * 4)) || (rc = d1.alloc((size_t)n1 * 32)) ||
      (rc = dk2.alloc((size_t)n2 * sizeof(OrbKeyPoint))) || (rc = d2.alloc((size_t)n2 * 32)) ||
      (rc = doff.alloc((size_t)(nq + 1) * 4)) || (rc = dent.alloc((size_t)total * 4)) || (rc = dblk.alloc((size_t)n2)) ||
      (rc = dca.alloc((size_t)nq * 4)) || (rc = dcb.alloc((size_t)nq * 4)) || (rc = dmin.alloc((size_t)n2 * 4)) ||
      (rc = dmatch.alloc((size_t)n2 * 4)) || (rc = dmq.alloc((size_t)nq * 4)) || (rc = dn.alloc(4)))
    return rc;
  ORB_CUDA_TRY(cudaMemcpy(dq.p, q, (size_t)nq * sizeof(OrbBucketQuery), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dwq.p, wq.data(), (size_t)nq * sizeof(OrbWindowQuery), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dc.p, cand2, (size_t)ncand2 * 4, cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(d1.p, desc1, (size_t)n1 * 32, cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dk2.p, kps2, (size_t)n2 * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(d2.p, desc2, (size_t)n2 * 32, cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(doff.p, offsets.data(), (size_t)(nq + 1) * 4, cudaMemcpyHostToDevice));
  if (skip2) ORB_CUDA_TRY(cudaMemcpy(dblk.p, skip2, (size_t)n2, cudaMemcpyHostToDevice));
  else ORB_CUDA_TRY(cudaMemset(dblk.p, 0, (size_t)n2));
  bucket_entries_kernel<<<(nq * 32 + 255) / 256, 256>>>(dq.as<OrbBucketQuery>(), nq, dc.as<int32_t>(), d1.as<uint4>(),
                                                        d2.as<uint4>(), doff.as<int>(), dent.as<uint32_t>());
  TmpDevW dtop;
  if ((rc = dtop.alloc((size_t)nq * kTopK * sizeof(uint2)))) return rc;
  topk_entries_kernel<<<(nq * 32 + 255) / 256, 256>>>(doff.as<int>(), dent.as<uint32_t>(), nq, dtop.as<uint2>());
  // claims, per-keypoint minima and flags of the fixpoint rounds in shared memory when they fit
  const size_t rs_smem = (size_t)nq * 8 + (size_t)n2 * 4 + (((size_t)n2 + 15) & ~(size_t)15) + (((size_t)nq + 15) & ~(size_t)15);
  const int use_smem = rs_smem <= 200 * 1024;
  if (use_smem && rs_smem > 48 * 1024)
    ORB_CUDA_TRY(cudaFuncSetAttribute(window_resolve_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_smem));
  window_resolve_kernel<true><<<1, 1024, use_smem ? rs_smem : 0>>>(dwq.as<OrbWindowQuery>(), nq, dk2.as<OrbKeyPoint>(), n2, doff.as<int>(),
      dent.as<uint32_t>(), dtop.as<uint2>(), nnratio, 1, th_accept, check_orientation, dblk.as<uint8_t>(), dca.as<int>(), dcb.as<int>(),
      dmin.as<int>(), dmatch.as<int32_t>(), dmq.as<int32_t>(), dn.as<int>(), use_smem);
  count_launch(3);
  ORB_CUDA_TRY(cudaGetLastError());
  std::vector<int32_t> mq((size_t)nq);
  ORB_CUDA_TRY(cudaMemcpy(mq.data(), dmq.p, (size_t)nq * 4, cudaMemcpyDeviceToHost));
  ORB_CUDA_TRY(cudaMemcpy(nmatches, dn.p, 4, cudaMemcpyDeviceToHost));
  for (int i = 0; i < nq; ++i)
    if (mq[i] >= 0) { match12[q[i].index] = mq[i]; match21[mq[i]] = q[i].index; }
  return ORB_GPU_OK;
}

int orb_gpu_search_for_triangulation(const OrbBucketQuery* q, int nq, const int32_t* cand2, int ncand2,
                                     const OrbKeyPoint* kps1, const uint8_t* desc1, const float* u_right1, int n1,
                                     const OrbKeyPoint* kps2, const uint8_t* desc2, const float* u_right2,
                                     const uint8_t* skip2, int n2, const float* F12, float ex, float ey,
                                     const float* scale_factors2, const float* level_sigma2_2, int nlevels,
                                     int only_stereo, int check_orientation, int device, int32_t* match12,
                                     int* nmatches) {
  if (nq < 0 || n1 < 0 || n2 < 0 || ncand2 < 0 || !nmatches || (n1 > 0 && !match12) || !F12 || !scale_factors2 ||
      !level_sigma2_2 || nlevels < 1 || nlevels > kMaxLevels || (nq > 0 && (!q || !kps1 || !desc1)) ||
      (ncand2 > 0 && (!cand2 || !kps2 || !desc2)))
    return ORB_GPU_ERR_ARG;
  *nmatches = 0;
  for (int i = 0; i < n1; ++i) match12[i] = -1;
  if (nq == 0 || n2 == 0 || ncand2 == 0) return ORB_GPU_OK;
  std::vector<int> offsets;
  int rc;
  if ((rc = check_bucket_queries(q, nq, ncand2, n1, cand2, n2, offsets))) return rc;
  TriangulationParams P;
  for (int i = 0; i < 9; ++i) P.F12[i] = F12[i];
  P.ex = ex; P.ey = ey;
  for (int i = 0; i < kMaxLevels; ++i) {
    P.scale2[i] = i < nlevels ? scale_factors2[i] : 0.f;
    P.sigma2_2[i] = i < nlevels ? level_sigma2_2[i] : 0.f;
  }
  P.only_stereo = only_stereo; P.check_ori = check_orientation; P.th_low = ORB_GPU_TH_LOW;
  { const int rc_pool = use_cached_pool(device); if (rc_pool) return rc_pool; }
  TmpDevW dq, dc, dk1, d1, du1, dk2, d2, du2, dskip, dbins, dmq, dn;
  if ((rc = dq.alloc((size_t)nq * sizeof(OrbBucketQuery))) || (rc = dc.alloc((size_t)ncand2 * 4)) ||
      (rc = dk1.alloc((size_t)n1 * sizeof(OrbKeyPoint))) || (rc = d1.alloc((size_t)n1 * 32)) ||
      (rc = du1.alloc((size_t)n1 * 4)) || (rc = dk2.alloc((size_t)n2 * sizeof(OrbKeyPoint))) ||
      (rc = d2.alloc((size_t)n2 * 32)) || (rc = du2.alloc((size_t)n2 * 4)) || (rc = dskip.alloc((size_t)n2)) ||
      (rc = dbins.alloc((size_t)nq * 4)) || (rc = dmq.alloc((size_t)nq * 4)) || (rc = dn.alloc(4)))
    return rc;
  ORB_CUDA_TRY(cudaMemcpy(dq.p, q, (size_t)nq * sizeof(OrbBucketQuery), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dc.p, cand2, (size_t)ncand2 * 4, cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dk1.p, kps1, (size_t)n1 * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(d1.p, desc1, (size_t)n1 * 32, cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dk2.p, kps2, (size_t)n2 * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(d2.p, desc2, (size_t)n2 * 32, cudaMemcpyHostToDevice));
  if (u_right1) ORB_CUDA_TRY(cudaMemcpy(du1.p, u_right1, (size_t)n1 * 4, cudaMemcpyHostToDevice));
  if (u_right2) ORB_CUDA_TRY(cudaMemcpy(du2.p, u_right2, (size_t)n2 * 4, cudaMemcpyHostToDevice));
  if (skip2) ORB_CUDA_TRY(cudaMemcpy(dskip.p, skip2, (size_t)n2, cudaMemcpyHostToDevice));
  else ORB_CUDA_TRY(cudaMemset(dskip.p, 0, (size_t)n2));
  triangulation_kernel<<<1, 1024>>>(dq.as<OrbBucketQuery>(), nq, dc.as<int32_t>(), dk1.as<OrbKeyPoint>(), d1.as<uint4>(),
      u_right1 ? du1.as<float>() : nullptr, dk2.as<OrbKeyPoint>(), d2.as<uint4>(), u_right2 ? du2.as<float>() : nullptr,
      dskip.as<uint8_t>(), P, dbins.as<int>(), dmq.as<int32_t>(), dn.as<int>());
  count_launch(1);
  ORB_CUDA_TRY(cudaGetLastError());
  std::vector<int32_t> mq((size_t)nq);
  ORB_CUDA_TRY(cudaMemcpy(mq.data(), dmq.p, (size_t)nq * 4, cudaMemcpyDeviceToHost));
  ORB_CUDA_TRY(cudaMemcpy(nmatches, dn.p, 4, cudaMemcpyDeviceToHost));
  for (int i = 0; i < nq; ++i)
    if (mq[i] >= 0) match12[q[i].index] = mq[i];
  return ORB_GPU_OK;
}

int orb_gpu_search_for_initialization(const OrbKeyPoint* kps1, const uint8_t* desc1, int n1,
                                      const float* prev_matched, const OrbKeyPoint* kps2, const uint8_t* desc2,
                                      int n2, OrbFrameGrid grid, int window, float nnratio, int check_orientation,
                                      int device, int32_t* match12, int* nmatches) {
  if (n1 < 0 || n2 < 0 || !nmatches || (n1 > 0 && (!kps1 || !desc1 || !prev_matched || !match12)) ||
      (n2 > 0 && (!kps2 || !desc2)))
    return ORB_GPU_ERR_ARG;
  *nmatches = 0;
  for (int i = 0; i < n1; ++i) match12[i] = -1;
  if (n1 == 0 || n2 == 0) return ORB_GPU_OK;
  if (n2 > 32768) return ORB_GPU_ERR_CAPACITY;
  const int nq = n1, nk = n2;
  std::vector<OrbWindowQuery> wq((size_t)nq);
  for (int i = 0; i < nq; ++i) {
    OrbWindowQuery& w = wq[i];
    w = OrbWindowQuery();
    if (kps1[i].octave > 0) continue;  // :421-423
    w.u = prev_matched[2 * i]; w.v = prev_matched[2 * i + 1];
    w.radius = (float)window;
    w.min_level = kps1[i].octave; w.max_level = kps1[i].octave;  // GetFeaturesInArea(x, y, windowSize, level1, level1)
    w.valid = 1;
  }
  { const int rc_pool = use_cached_pool(device); if (rc_pool) return rc_pool; }
  int nkpad = 2;
  while (nkpad < nk) nkpad <<= 1;
  TmpDevW dq, dqd, dk1, dk, dkd, dsorted, dcell, dcnt, doff, dent, dca, dcb, dda, ddb, dnext, dhead, dmatch, dn;
  int rc;
  if ((rc = dq.alloc((size_t)nq * sizeof(OrbWindowQuery))) || (rc = dqd.alloc((size_t)nq * 32)) ||
      (rc = dk1.alloc((size_t)nq * sizeof(OrbKeyPoint))) || (rc = dk.alloc((size_t)nk * sizeof(OrbKeyPoint))) ||
      (rc = dkd.alloc((size_t)nk * 32)) || (rc = dsorted.alloc((size_t)nkpad * 4)) ||
      (rc = dcell.alloc((size_t)(kCells + 1) * 4)) || (rc = dcnt.alloc((size_t)nq * 4)) ||
      (rc = doff.alloc((size_t)(nq + 1) * 4)) || (rc = dca.alloc((size_t)nq * 4)) || (rc = dcb.alloc((size_t)nq * 4)) ||
      (rc = dda.alloc((size_t)nq * 4)) || (rc = ddb.alloc((size_t)nq * 4)) || (rc = dnext.alloc((size_t)nq * 4)) ||
      (rc = dhead.alloc((size_t)nk * 4)) || (rc = dmatch.alloc((size_t)nq * 4)) || (rc = dn.alloc(4)))
    return rc;
  ORB_CUDA_TRY(cudaMemcpy(dq.p, wq.data(), (size_t)nq * sizeof(OrbWindowQuery), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dqd.p, desc1, (size_t)nq * 32, cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dk1.p, kps1, (size_t)nq * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dk.p, kps2, (size_t)nk * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(dkd.p, desc2, (size_t)nk * 32, cudaMemcpyHostToDevice));
  const size_t smem = (size_t)nkpad * 4;
  if (smem > 48 * 1024)
    ORB_CUDA_TRY(cudaFuncSetAttribute(grid_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  grid_build_kernel<<<1, 1024, smem>>>(dk.as<OrbKeyPoint>(), nk, nkpad, grid, dsorted.as<uint32_t>(), dcell.as<int>());
  const int blocks = (nq * 32 + 255) / 256;
  window_candidates_kernel<false><<<blocks, 256>>>(dq.as<OrbWindowQuery>(), dqd.as<uint4>(), nq, dk.as<OrbKeyPoint>(),
      dkd.as<uint4>(), nullptr, grid, 0, nullptr, dsorted.as<uint32_t>(), dcell.as<int>(), dcnt.as<int>(), nullptr,
      nullptr, 0, nullptr);
  exclusive_scan_kernel<<<1, 1024>>>(dcnt.as<int>(), nq, doff.as<int>());
  int total = 0;
  ORB_CUDA_TRY(cudaMemcpy(&total, doff.as<int>() + nq, 4, cudaMemcpyDeviceToHost));
  if ((rc = dent.alloc((size_t)total * 4))) return rc;
  window_candidates_kernel<true><<<blocks, 256>>>(dq.as<OrbWindowQuery>(), dqd.as<uint4>(), nq, dk.as<OrbKeyPoint>(),
      dkd.as<uint4>(), nullptr, grid, 0, nullptr, dsorted.as<uint32_t>(), dcell.as<int>(), nullptr, doff.as<int>(),
      dent.as<uint32_t>(), INT_MAX, nullptr);
  TmpDevW dtop;
  if ((rc = dtop.alloc((size_t)nq * kTopK * sizeof(uint2)))) return rc;
  topk_entries_kernel<<<blocks, 256>>>(doff.as<int>(), dent.as<uint32_t>(), nq, dtop.as<uint2>());
  init_resolve_kernel<<<1, 1024>>>(dk1.as<OrbKeyPoint>(), nq, dk.as<OrbKeyPoint>(), nk, doff.as<int>(),
      dent.as<uint32_t>(), dtop.as<uint2>(), nnratio, ORB_GPU_TH_LOW, check_orientation, dca.as<int>(), dcb.as<int>(), dda.as<int>(),
      ddb.as<int>(), dnext.as<int>(), dhead.as<int>(), dmatch.as<int32_t>(), dn.as<int>());
  count_launch(5);
  ORB_CUDA_TRY(cudaGetLastError());
  ORB_CUDA_TRY(cudaMemcpy(match12, dmatch.p, (size_t)nq * 4, cudaMemcpyDeviceToHost));
  ORB_CUDA_TRY(cudaMemcpy(nmatches, dn.p, 4, cudaMemcpyDeviceToHost));
  return ORB_GPU_OK;
}

int orb_gpu_stereo_match(const OrbKeyPoint* kl, const uint8_t* dl, int nl, const OrbKeyPoint* kr,
                         const uint8_t* dr, int nr, int n_rows, const float* scale_factors, int nlevels,
                         float min_d, float max_d, int th_high, int device, int32_t* best_idx_r,
                         int32_t* best_dist) {
  if (nl < 0 || nr < 0 || nlevels < 1 || nlevels > kMaxLevels || !scale_factors ||
      (nl > 0 && (!kl || !dl || !best_idx_r || !best_dist)) || (nr > 0 && (!kr || !dr)) || nr > 65535)
    return ORB_GPU_ERR_ARG;
  if (nl == 0) return ORB_GPU_OK;
  { const int rc_pool = use_cached_pool(device); if (rc_pool) return rc_pool; }
  TmpDevW a, b, c, d, e, f, g;
  int rc;
  if ((rc = a.alloc((size_t)nl * sizeof(OrbKeyPoint))) || (rc = b.alloc((size_t)nl * 32)) ||
      (rc = c.alloc((size_t)nr * sizeof(OrbKeyPoint))) || (rc = d.alloc((size_t)nr * 32)) ||
      (rc = e.alloc((size_t)nlevels * 4)) || (rc = f.alloc((size_t)nl * 4)) || (rc = g.alloc((size_t)nl * 4)))
    return rc;
  ORB_CUDA_TRY(cudaMemcpy(a.p, kl, (size_t)nl * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(b.p, dl, (size_t)nl * 32, cudaMemcpyHostToDevice));
  if (nr) {
    ORB_CUDA_TRY(cudaMemcpy(c.p, kr, (size_t)nr * sizeof(OrbKeyPoint), cudaMemcpyHostToDevice));
    ORB_CUDA_TRY(cudaMemcpy(d.p, dr, (size_t)nr * 32, cudaMemcpyHostToDevice));
  }
  ORB_CUDA_TRY(cudaMemcpy(e.p, scale_factors, (size_t)nlevels * 4, cudaMemcpyHostToDevice));
  stereo_match_kernel<<<(nl * 32 + 255) / 256, 256>>>(a.as<OrbKeyPoint>(), b.as<uint4>(), nl, c.as<OrbKeyPoint>(),
      d.as<uint4>(), nr, n_rows, e.as<float>(), min_d, max_d, th_high, f.as<int32_t>(), g.as<int32_t>());
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  ORB_CUDA_TRY(cudaMemcpy(best_idx_r, f.p, (size_t)nl * 4, cudaMemcpyDeviceToHost));
  ORB_CUDA_TRY(cudaMemcpy(best_dist, g.p, (size_t)nl * 4, cudaMemcpyDeviceToHost));
  return ORB_GPU_OK;
}

}  // extern "C"
