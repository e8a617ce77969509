// Hamming matching kernels of the B200-native ORB front-end (sm_100a).
//
// Reference: ORBmatcher::DescriptorDistance src/ORBmatcher.cc:1647-1663 (256-bit XOR+popcount),
// top-2 update rule :216-225, SearchByProjection :45-129 with Frame::GetFeaturesInArea
// src/Frame.cc:327-380, Frame::ComputeStereoMatches Hamming stage src/Frame.cc:466-549.
// Everything is integer work on the LOP3/POPC pipes; results are bit-exact with the CPU oracle.
#include <algorithm>
#include <vector>

#include <cstdlib>
#include <cstdio>
#include <mutex>

#include <atomic>

#include "common.cuh"

namespace orbgpu {

__device__ __forceinline__ int hamming256(const uint32_t q[8], const uint4 lo, const uint4 hi) {
  return __popc(q[0] ^ lo.x) + __popc(q[1] ^ lo.y) + __popc(q[2] ^ lo.z) + __popc(q[3] ^ lo.w) +
         __popc(q[4] ^ hi.x) + __popc(q[5] ^ hi.y) + __popc(q[6] ^ hi.z) + __popc(q[7] ^ hi.w);
}

// Carry-save (Harley-Seal) form: 4 POPC + LOP3 adders instead of 8 POPC.
__device__ __forceinline__ int hamming256_csa(const uint32_t q[8], const uint4 lo, const uint4 hi) {
  const uint32_t w0 = q[0] ^ lo.x, w1 = q[1] ^ lo.y, w2 = q[2] ^ lo.z, w3 = q[3] ^ lo.w;
  const uint32_t w4 = q[4] ^ hi.x, w5 = q[5] ^ hi.y, w6 = q[6] ^ hi.z, w7 = q[7] ^ hi.w;
  const uint32_t sa = w0 ^ w1 ^ w2, ca = (w0 & w1) | (w2 & (w0 | w1));
  const uint32_t sb = w3 ^ w4 ^ w5, cb = (w3 & w4) | (w5 & (w3 | w4));
  const uint32_t sc = sa ^ sb ^ w6, cc = (sa & sb) | (w6 & (sa | sb));
  const uint32_t ones = sc ^ w7, cd = sc & w7;
  const uint32_t se = ca ^ cb ^ cc, ce = (ca & cb) | (cc & (ca | cb));
  const uint32_t twos = se ^ cd, cf = se & cd;
  const uint32_t fours = ce ^ cf, eights = ce & cf;
  return __popc(ones) + 2 * __popc(twos) + 4 * __popc(fours) + 8 * __popc(eights);
}

// ---------------------------------------------------------------------------------------------
// DescriptorDistance for n independent pairs
// ---------------------------------------------------------------------------------------------
__global__ void hamming_pairs_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int n,
                                     int32_t* __restrict__ dist) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 a0 = a[2 * i], a1 = a[2 * i + 1];
  const uint32_t q[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
  dist[i] = hamming256(q, b[2 * i], b[2 * i + 1]);
}

// ---------------------------------------------------------------------------------------------
// brute-force query x database top-2
// ---------------------------------------------------------------------------------------------
// Thread t of a block owns QPT queries in registers; the block streams its slice of the
// database through shared memory in tiles (cp.async double buffer); every row is an LDS.128
// broadcast shared by all threads.  Per (query,row): 8 LOP3 + 8 POPC (or 4 POPC in the
// carry-save form) + adds, then a branch-free top-2 update on packed keys (dist << 23 | row),
// which keeps the lowest row among equal distances exactly like the ascending scan with strict
// '<' of the reference (:216-225).  Partial results of the database slices are merged by
// merge_top2_kernel, the same kernel that merges per-GPU shards after the NCCL all-gather.
constexpr int kBfThreads = 128;
constexpr int kBfTile = 256;              // database rows per shared-memory tile (8 KB)
constexpr int kBfSliceMax = 1 << 22;      // rows per slice (row index must fit 23 bits)

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

template <int QPT, bool CSA>
__global__ void __launch_bounds__(kBfThreads)
bruteforce_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db, long long ndb,
                  long long slice_rows, int32_t idx_base, int32_t* __restrict__ o_idx,
                  int32_t* __restrict__ o_best, int32_t* __restrict__ o_second) {
  __shared__ __align__(16) uint4 tile[2][kBfTile * 2];
  const int slice = blockIdx.y;
  const long long row0 = (long long)slice * slice_rows;
  const long long rows = min(slice_rows, ndb - row0);
  uint32_t qr[QPT][8];
  unsigned best[QPT], second[QPT];
#pragma unroll
  for (int k = 0; k < QPT; ++k) {
    const int qi = min((blockIdx.x * QPT + k) * kBfThreads + threadIdx.x, nq - 1);
    const uint4 a0 = q[2 * qi], a1 = q[2 * qi + 1];
    qr[k][0] = a0.x; qr[k][1] = a0.y; qr[k][2] = a0.z; qr[k][3] = a0.w;
    qr[k][4] = a1.x; qr[k][5] = a1.y; qr[k][6] = a1.z; qr[k][7] = a1.w;
    best[k] = second[k] = (256u << 23) | 0x7fffffu;
  }
  const uint4* src = db + row0 * 2;
  const int ntiles = (int)((rows + kBfTile - 1) / kBfTile);
  auto issue = [&](int t, int buf) {
    const long long base = (long long)t * kBfTile * 2;
    const long long lim = rows * 2;
    for (int v = threadIdx.x; v < kBfTile * 2; v += kBfThreads)
      if (base + v < lim) cp_async16(&tile[buf][v], src + base + v);
    cp_async_commit();
  };
  if (ntiles > 0) issue(0, 0);
  for (int t = 0; t < ntiles; ++t) {
    if (t + 1 < ntiles) { issue(t + 1, (t + 1) & 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const uint4* tl = tile[t & 1];
    const int nr = (int)min((long long)kBfTile, rows - (long long)t * kBfTile);
    const unsigned rbase = (unsigned)(t * kBfTile);
#pragma unroll 4
    for (int r = 0; r < nr; ++r) {
      const uint4 lo = tl[2 * r], hi = tl[2 * r + 1];
#pragma unroll
      for (int k = 0; k < QPT; ++k) {
        const int d = CSA ? hamming256_csa(qr[k], lo, hi) : hamming256(qr[k], lo, hi);
        const unsigned key = ((unsigned)d << 23) | (rbase + r);
        second[k] = min(second[k], max(best[k], key));
        best[k] = min(best[k], key);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int k = 0; k < QPT; ++k) {
    const int qi = (blockIdx.x * QPT + k) * kBfThreads + threadIdx.x;
    if (qi >= nq) continue;
    const size_t o = (size_t)slice * nq + qi;
    const int bd = (int)(best[k] >> 23);
    o_best[o] = bd;
    o_second[o] = (int)(second[k] >> 23);
    o_idx[o] = bd < 256 ? (int32_t)(idx_base + row0 + (best[k] & 0x7fffffu)) : -1;  // 256 never beats the initial 256
  }
}

// ---------------------------------------------------------------------------------------------
// brute force on the tensor cores (legacy IMMA path of sm_100a: mma.sync m16n8k32 s8 x u8 -> s32)
// ---------------------------------------------------------------------------------------------
// With the query bits expanded to bytes a' in {-1,+1} and the database bits to bytes b in {0,1},
//   S = sum_k a'_k * b_k = |q & d| - |~q & d|   and   hamming(q, d) = popc(q) - S,
// so the 256-bit distance of 16 queries x 8 rows is eight m16n8k32 MMAs and the nearest row of a query is the
// one with the largest S.  A warp keeps 32 queries as A fragments in registers for the whole slice (64
// registers), streams the packed rows from L1/L2 (two LDG.128 per thread per 8 rows, prefetched one group
// ahead), expands nibbles to bytes through a 16-entry shared-memory table (conflict-free: equal nibbles
// broadcast, different nibbles are different banks) and keeps the top-2 of its four queries as packed keys
// (S + 256) << 22 | (0x3fffff - row): largest S first, lowest row among equals -- the order of the ascending
// scan with strict '<' (src/ORBmatcher.cc:216-225).  Measured IMMA rate: 1958 MAC/clk/SM
// (tools/ubench/imma_rate.cu) = 7.6 pairs/clk/SM, 3.9x the POPC pipe's.
constexpr int kMmaWarps = 8;                      // 256 queries per block
constexpr int kMmaQueries = kMmaWarps * 32;

__device__ __forceinline__ void imma_s8u8(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int kMmaTileRows = 32;                  // database rows expanded per barrier (groups of 8; multiple of 32)

__global__ void __launch_bounds__(kMmaWarps * 32, 2)
bruteforce_mma_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db, long long ndb,
                      long long slice_rows, int32_t idx_base, int32_t* __restrict__ o_idx,
                      int32_t* __restrict__ o_best, int32_t* __restrict__ o_second) {
  // expanded B tiles, double buffered: [buffer][group of 8 rows][j = k-step pair][lane] -> (b0,b1 of k-step 2j,
  // b0,b1 of k-step 2j+1), lane index XORed with j so that the expanding threads spread over the banks; a warp
  // reads its 16 fragment words of a group with four conflict-free LDS.128
  __shared__ uint4 s_b[2][kMmaTileRows / 8][4][32];
  const int slice = blockIdx.y;
  const long long row0 = (long long)slice * slice_rows;
  const int rows = (int)min(slice_rows, ndb - row0);
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31, gid = lane >> 2, tig = lane & 3;
  const int qbase = blockIdx.x * kMmaQueries + warp * 32;

  // A fragments of this thread: queries qbase + 16*mt + gid (+8), k-steps 0..7; bytes +1 / -1
  uint32_t a[2][8][4];
  int pq[4];  // popc of the thread's four queries: (mt, half) -> 2*mt + half
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int qi = min(qbase + 16 * mt + gid + 8 * half, nq - 1);
      const uint4 lo = __ldg(q + 2 * (size_t)qi), hi = __ldg(q + 2 * (size_t)qi + 1);
      const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
      int pc = 0;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        pc += __popc(w[ks]);
        const uint32_t e0 = (((w[ks] >> (4 * tig)) & 0xfu) * 0x00204081u) & 0x01010101u;
        const uint32_t e1 = (((w[ks] >> (16 + 4 * tig)) & 0xfu) * 0x00204081u) & 0x01010101u;
        a[mt][ks][half] = ((e0 ^ 0x01010101u) * 0xffu) | e0;      // a0 / a1: k = 4*tig .. +3
        a[mt][ks][2 + half] = ((e1 ^ 0x01010101u) * 0xffu) | e1;  // a2 / a3: k = 16 + 4*tig .. +3
      }
      pq[2 * mt + half] = pc;
    }

  // expansion duty of this thread: packed word e_ks of rows e_row and e_row + 32 of every tile -> 2 x 8 fragment words
  const int e_row = t >> 3, e_ks = t & 7;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(db + row0 * 2);
  const int ntiles = (rows + kMmaTileRows - 1) / kMmaTileRows;
  auto fetch = [&](int tile, int h) -> uint32_t {
    const int r = tile * kMmaTileRows + e_row + 32 * h;
    return (tile < ntiles && r < rows) ? __ldg(src + (size_t)r * 8 + e_ks) : 0u;
  };
  auto expand = [&](uint32_t w, int buf, int h) {
    const int er = e_row + 32 * h, j = e_ks >> 1;
    uint32_t* base = reinterpret_cast<uint32_t*>(&s_b[buf][er >> 3][j][0]) + (e_ks & 1) * 2;
#pragma unroll
    for (int g4 = 0; g4 < 4; ++g4) {  // fragment lane (row, tig = g4): nibbles g4 and 4 + g4
      uint2 v;
      v.x = (((w >> (4 * g4)) & 0xfu) * 0x00204081u) & 0x01010101u;
      v.y = (((w >> (16 + 4 * g4)) & 0xfu) * 0x00204081u) & 0x01010101u;
      *reinterpret_cast<uint2*>(base + 4 * ((((er & 7) * 4) + g4) ^ j)) = v;
    }
  };

  unsigned best[4] = {0, 0, 0, 0}, second[4] = {0, 0, 0, 0};  // key 0 = nothing seen
  int thr[4] = {-100000, -100000, -100000, -100000};           // S of the thread's second best per query
  constexpr int kHalves = kMmaTileRows / 32;
  uint32_t w_next[kHalves];
#pragma unroll
  for (int h = 0; h < kHalves; ++h) { w_next[h] = fetch(0, h); expand(w_next[h], 0, h); }
#pragma unroll
  for (int h = 0; h < kHalves; ++h) w_next[h] = fetch(1, h);
  for (int tile = 0; tile < ntiles; ++tile) {
    const int buf = tile & 1;
    __syncthreads();  // tile `tile` is expanded; everybody has finished reading the other buffer
    uint32_t w_cur[kHalves];
#pragma unroll
    for (int h = 0; h < kHalves; ++h) { w_cur[h] = w_next[h]; w_next[h] = fetch(tile + 2, h); }
#pragma unroll
    for (int g = 0; g < kMmaTileRows / 8; ++g) {
      int c0[4] = {0, 0, 0, 0}, c1[4] = {0, 0, 0, 0};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 v = s_b[buf][g][j][lane ^ j];
        imma_s8u8(c0, a[0][2 * j], v.x, v.y);
        imma_s8u8(c1, a[1][2 * j], v.x, v.y);
        imma_s8u8(c0, a[0][2 * j + 1], v.z, v.w);
        imma_s8u8(c1, a[1][2 * j + 1], v.z, v.w);
      }
      // quick reject: a row only matters if its S beats the S of the thread's current second best
      const bool hit = max(c0[0], c0[1]) > thr[0] || max(c0[2], c0[3]) > thr[1] || max(c1[0], c1[1]) > thr[2] ||
                       max(c1[2], c1[3]) > thr[3];
      if (hit) {
        const int col = tile * kMmaTileRows + g * 8 + 2 * tig;  // rows col, col + 1 of the slice
        const unsigned k0 = col < rows ? (256u << 22) + 0x3fffffu - (unsigned)col : 0u;
        const unsigned k1 = col + 1 < rows ? (256u << 22) + 0x3fffffu - (unsigned)(col + 1) : 0u;
        const int sv[8] = {c0[0], c0[1], c0[2], c0[3], c1[0], c1[1], c1[2], c1[3]};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int qslot = (i >> 2) * 2 + ((i >> 1) & 1);           // (mt, half)
          const unsigned kc = (i & 1) ? k1 : k0;
          const unsigned key = kc ? (unsigned)(sv[i] * (1 << 22)) + kc : 0u;  // rows past the slice end never count
          second[qslot] = max(second[qslot], min(best[qslot], key));
          best[qslot] = max(best[qslot], key);
        }
#pragma unroll
        for (int s4 = 0; s4 < 4; ++s4) thr[s4] = second[s4] ? (int)(second[s4] >> 22) - 256 : -100000;
      }
    }
    if (tile + 1 < ntiles) {
#pragma unroll
      for (int h = 0; h < kHalves; ++h) expand(w_cur[h], buf ^ 1, h);
    }
  }
  // merge the four lanes of a group (they saw different columns of the same queries)
#pragma unroll
  for (int o = 1; o <= 2; o <<= 1)
#pragma unroll
    for (int s4 = 0; s4 < 4; ++s4) {
      const unsigned ob = __shfl_xor_sync(0xffffffffu, best[s4], o), os = __shfl_xor_sync(0xffffffffu, second[s4], o);
      second[s4] = max(max(second[s4], os), min(best[s4], ob));
      best[s4] = max(best[s4], ob);
    }
  if (tig == 0) {
#pragma unroll
    for (int s4 = 0; s4 < 4; ++s4) {
      const int qi = qbase + 16 * (s4 >> 1) + gid + 8 * (s4 & 1);
      if (qi >= nq) continue;
      const size_t o = (size_t)slice * nq + qi;
      const int bd = best[s4] ? pq[s4] - ((int)(best[s4] >> 22) - 256) : 256;
      const int sd = second[s4] ? pq[s4] - ((int)(second[s4] >> 22) - 256) : 256;
      o_best[o] = bd;
      o_second[o] = sd;
      // a row at distance 256 never beats the initial best of the reference's scan (strict '<')
      o_idx[o] = (best[s4] && bd < 256) ? (int32_t)(idx_base + row0 + (0x3fffffu - (best[s4] & 0x3fffffu))) : -1;
    }
  }
}

// Exact merge of top-2 partials of disjoint ascending index ranges: scanning
// shards in ascending order with the reference's update rule applied to (best, second) pairs.
__global__ void merge_top2_kernel(const int32_t* __restrict__ idx, const int32_t* __restrict__ best,
                                  const int32_t* __restrict__ second, int nshards, int nq,
                                  int32_t* __restrict__ o_idx, int32_t* __restrict__ o_best,
                                  int32_t* __restrict__ o_second) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  int b = 256, s = 256, id = -1;
  for (int k = 0; k < nshards; ++k) {
    const int bk = best[(size_t)k * nq + i], sk = second[(size_t)k * nq + i], ik = idx[(size_t)k * nq + i];
    if (ik < 0) continue;  // empty shard
    if (bk < b) { s = min(b, sk); b = bk; id = ik; }
    else s = min(s, bk);
  }
  o_idx[i] = id; o_best[i] = b; o_second[i] = s;
}

__global__ void ratio_test_kernel(const int32_t* __restrict__ best, const int32_t* __restrict__ second,
                                  int nq, int th, float nnratio, uint8_t* __restrict__ accept) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  const int b = best[i], s = second[i];
  accept[i] = (b <= th) && ((float)b < __fmul_rn(nnratio, (float)s));  // ORBmatcher.cc:228-230
}

// ---------------------------------------------------------------------------------------------
// integer-pipe peak probe: dependent-free POPC stream
// ---------------------------------------------------------------------------------------------
__global__ void popc_peak_kernel(uint32_t seed, int iters, uint32_t* out) {
  uint32_t a[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = seed * (threadIdx.x + 1 + k) + blockIdx.x;
  uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      acc[k] += __popc(a[k]);
      a[k] += 0x9e3779b9u;  // keeps the operand changing so POPC cannot be hoisted
    }
  }
  uint32_t r = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) r += acc[k];
  if (r == 0xffffffffu) out[0] = r;
}

// tensor-pipe peak probe for the legacy int8 path: eight independent IMMA.16832 accumulation chains per warp
__global__ void imma_peak_kernel(int iters, int* out) {
  int c[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0;
  const uint32_t a[4] = {threadIdx.x * 0x01010101u, threadIdx.x * 0x01010101u + 1, 0x01ff01ffu, 0xff01ff01u};
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < 8; ++i) imma_s8u8(c[i], a, 0x01000101u, 0x00010100u);
  int r = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) r += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (r == 0x7fffffff) out[0] = r;
}

}  // namespace orbgpu

using namespace orbgpu;

// =============================================================================================
// C ABI (matcher part)
// =============================================================================================
extern "C" {

int orb_gpu_hamming_pairs_device(const uint8_t* d_a, const uint8_t* d_b, int n, int32_t* d_dist,
                                 void* stream) {
  if (n < 0 || (n > 0 && (!d_a || !d_b || !d_dist))) return ORB_GPU_ERR_ARG;
  if (n == 0) return ORB_GPU_OK;
  hamming_pairs_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>((const uint4*)d_a, (const uint4*)d_b, n, d_dist);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

int orb_gpu_merge_top2_device(const int32_t* d_idx, const int32_t* d_best, const int32_t* d_second,
                              int nshards, int nq, int32_t* d_out_idx, int32_t* d_out_best,
                              int32_t* d_out_second, void* stream) {
  if (nq < 0 || nshards < 1 || (nq > 0 && (!d_idx || !d_best || !d_second || !d_out_idx || !d_out_best || !d_out_second)))
    return ORB_GPU_ERR_ARG;
  if (nq == 0) return ORB_GPU_OK;
  merge_top2_kernel<<<(nq + 255) / 256, 256, 0, (cudaStream_t)stream>>>(d_idx, d_best, d_second, nshards, nq,
                                                                    d_out_idx, d_out_best, d_out_second);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

int orb_gpu_ratio_test_device(const int32_t* d_best, const int32_t* d_second, int nq, int th,
                              float nnratio, uint8_t* d_accept, void* stream) {
  if (nq < 0 || (nq > 0 && (!d_best || !d_second || !d_accept))) return ORB_GPU_ERR_ARG;
  if (nq == 0) return ORB_GPU_OK;
  ratio_test_kernel<<<(nq + 255) / 256, 256, 0, (cudaStream_t)stream>>>(d_best, d_second, nq, th, nnratio, d_accept);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

}  // extern "C"

// =============================================================================================
// brute force: launcher, host wrapper, POPC probe
// =============================================================================================
namespace orbgpu {

struct TmpDev {  // RAII for call-scoped device scratch
  void* p = nullptr;
  ~TmpDev() { if (p) cudaFree(p); }
  int alloc(size_t bytes) { ORB_CUDA_TRY(cudaMalloc(&p, bytes ? bytes : 16)); return ORB_GPU_OK; }
  template <class T> T* as() { return (T*)p; }
};

// Which brute-force form runs: 0: 8-POPC form, 1: carry-save form, 2: mma.sync IMMA form, 3: tcgen05 kind::i8 form
// (default), and the queries per thread of the POPC forms -- ONE atomic word (form | qpt << 8), so a call that races the
// test / bench setter still sees a consistent pair; all forms give identical results.
static std::atomic<int> g_bf_form{3 | (2 << 8)};

template <int QPT, bool CSA>
static void bf_launch(dim3 grid, cudaStream_t s, const uint4* q, int nq, const uint4* db, long long ndb,
                      long long slice_rows, int32_t base, int32_t* i, int32_t* b, int32_t* sec) {
  bruteforce_kernel<QPT, CSA><<<grid, kBfThreads, 0, s>>>(q, nq, db, ndb, slice_rows, base, i, b, sec);
}

}  // namespace orbgpu

extern "C" {

// test/bench knob: choose the inner-loop form (0 = 8 POPC, 1 = carry-save) and queries per thread
int orb_gpu_set_bruteforce_variant(int csa, int qpt) {
  if ((qpt != 1 && qpt != 2 && qpt != 4) || csa < 0 || csa > 3) return ORB_GPU_ERR_ARG;
  g_bf_form.store(csa | (qpt << 8), std::memory_order_relaxed);
  return ORB_GPU_OK;
}

int orb_gpu_match_bruteforce_device(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb,
                                    int32_t idx_base, int32_t* d_idx, int32_t* d_best, int32_t* d_second,
                                    void* stream) {
  int fused = 0;
  return orbgpu::bruteforce_device_impl(d_q, nq, d_db, ndb, idx_base, d_idx, d_best, d_second, (cudaStream_t)stream,
                                        nullptr, &fused);
}

}  // extern "C"

int orbgpu::bruteforce_device_impl(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb, int32_t idx_base,
                                   int32_t* d_idx, int32_t* d_best, int32_t* d_second, cudaStream_t s,
                                   const FusedExchange* fx, int* fused) {
  *fused = 0;
  if (nq < 0 || ndb < 0 || (nq > 0 && (!d_q || !d_idx || !d_best || !d_second)) || (ndb > 0 && !d_db))
    return ORB_GPU_ERR_ARG;
  if (nq == 0) return ORB_GPU_OK;
  int dev = 0, sms = 148;
  ORB_CUDA_TRY(cudaGetDevice(&dev));
  ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int form = g_bf_form.load(std::memory_order_relaxed);
  const int qpt = form >> 8, g_bf_variant = form & 0xff;
  const bool mma = g_bf_variant >= 2, tc5 = g_bf_variant == 3;  // both tensor-core forms take 256 queries per block
  const int qblocks = mma ? (nq + kMmaQueries - 1) / kMmaQueries : (nq + kBfThreads * qpt - 1) / (kBfThreads * qpt);
  // Database slices.  tcgen05 form (one block per SM): ONE wave of blocks -- every slice restarts the top-2
  // scan with no threshold, and the ~100 us a block spends in the exact merge before its threshold has
  // tightened is paid once per wave (measured: 5.56 -> 5.35 ms at 10 M rows, 0.99 -> 0.79 ms at 1.25 M);
  // more waves only when a slice would exceed the 22-bit row field of the packed keys.  Other forms: ~4 waves
  // of 4 blocks per SM.  Each slice is a whole number of tiles.
  static const int env_waves = [] { const char* e = getenv("ORB_GPU_BF_WAVES"); return e ? atoi(e) : 0; }();
  long long waves = env_waves > 0 ? env_waves : 1;
  while (tc5 && waves * sms / qblocks * kBfSliceMax < ndb && waves < 65535) ++waves;
  long long nslices = tc5 ? std::max<long long>(1, waves * sms / qblocks)
                          : std::max<long long>(1, (16LL * sms + qblocks - 1) / qblocks);
  long long slice_rows = (ndb + nslices - 1) / nslices;
  slice_rows = std::max<long long>(kBfTile, (slice_rows + kBfTile - 1) / kBfTile * kBfTile);
  slice_rows = std::min<long long>(slice_rows, kBfSliceMax);
  nslices = std::max<long long>(1, (ndb + slice_rows - 1) / slice_rows);
  if (nslices > 65535) return ORB_GPU_ERR_CAPACITY;
  // per-slice partials: stream-ordered scratch from the device's cached pool (no cudaMalloc / cudaFree and no
  // host synchronisation in the call; the merge kernel is the last reader and the free is ordered after it)
  void* part = nullptr;
  int32_t *pi = d_idx, *pb = d_best, *ps = d_second;
  unsigned *g_thr = nullptr, *g_ticket = nullptr;  // tcgen05 form: shared per-query threshold, slice tickets
  // the exchange across ranks rides in the same launch when the tcgen05 grid is at most one block per SM: the (at
  // most qblocks) blocks that wait for the peers then never keep a block of this grid from running
  const bool fuse = fx && fx->world > 1 && tc5 && (long long)qblocks * nslices <= sms && qblocks <= fx->qb_max &&
                    nq <= fx->nq_max;
  if (nslices > 1 || fuse) {
    int rc = use_cached_pool(dev);
    if (rc) return rc;
    const size_t part_bytes = (size_t)nslices * nq * 3 * sizeof(int32_t);
    const size_t extra = tc5 ? ((size_t)nq + (size_t)qblocks) * sizeof(unsigned) : 0;
    ORB_CUDA_TRY(cudaMallocAsync(&part, part_bytes + extra, s));
    pi = (int32_t*)part; pb = pi + (size_t)nslices * nq; ps = pb + (size_t)nslices * nq;
    if (tc5) {
      static const bool no_share = [] { const char* e = getenv("ORB_GPU_BF_NO_SHARE"); return e && atoi(e); }();
      g_thr = (unsigned*)((uint8_t*)part + part_bytes);
      g_ticket = g_thr + nq;
      ORB_CUDA_TRY(cudaMemsetAsync(g_thr, 0, extra, s));
      if (no_share) g_thr = nullptr;
    }
  }
  dim3 grid(qblocks, (unsigned)nslices);
  const uint4* q4 = (const uint4*)d_q; const uint4* db4 = (const uint4*)d_db;
  const int key = tc5 ? -2 : mma ? -1 : qpt * 2 + g_bf_variant;
  switch (key) {
    case -2: {
      int rc = launch_bruteforce_tc5(grid, s, q4, nq, db4, ndb, slice_rows, idx_base, pi, pb, ps, g_thr, g_ticket, d_idx,
                                     d_best, d_second, fuse ? fx : nullptr);
      *fused = fuse ? 1 : 0;
      if (rc) { if (part) cudaFreeAsync(part, s); return rc; }
      break;
    }
    case -1:
      bruteforce_mma_kernel<<<grid, kMmaWarps * 32, 0, s>>>(q4, nq, db4, ndb, slice_rows, idx_base, pi, pb, ps);
      break;
    case 2: bf_launch<1, false>(grid, s, q4, nq, db4, ndb, slice_rows, idx_base, pi, pb, ps); break;
    case 3: bf_launch<1, true>(grid, s, q4, nq, db4, ndb, slice_rows, idx_base, pi, pb, ps); break;
    case 4: bf_launch<2, false>(grid, s, q4, nq, db4, ndb, slice_rows, idx_base, pi, pb, ps); break;
    case 5: bf_launch<2, true>(grid, s, q4, nq, db4, ndb, slice_rows, idx_base, pi, pb, ps); break;
    case 8: bf_launch<4, false>(grid, s, q4, nq, db4, ndb, slice_rows, idx_base, pi, pb, ps); break;
    default: bf_launch<4, true>(grid, s, q4, nq, db4, ndb, slice_rows, idx_base, pi, pb, ps); break;
  }
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && nslices > 1 && !tc5) {  // the tcgen05 kernel merges its slices itself (last block per query block)
    merge_top2_kernel<<<(nq + 255) / 256, 256, 0, s>>>(pi, pb, ps, (int)nslices, nq, d_idx, d_best, d_second);
    count_launch();
    e = cudaGetLastError();
  }
  if (part) cudaFreeAsync(part, s);
  ORB_CUDA_TRY(e);
  return ORB_GPU_OK;
}

extern "C" {

int orb_gpu_match_bruteforce(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int device,
                             int32_t* idx, int32_t* best, int32_t* second) {
  if (nq < 0 || ndb < 0 || (nq > 0 && (!q || !idx || !best || !second)) || (ndb > 0 && !db))
    return ORB_GPU_ERR_ARG;
  if (nq == 0) return ORB_GPU_OK;
  ORB_CUDA_TRY(cudaSetDevice(device));
  TmpDev dq, ddb, dout;
  int rc;
  if ((rc = dq.alloc((size_t)nq * 32)) || (rc = ddb.alloc((size_t)ndb * 32)) ||
      (rc = dout.alloc((size_t)nq * 3 * sizeof(int32_t))))
    return rc;
  ORB_CUDA_TRY(cudaMemcpy(dq.p, q, (size_t)nq * 32, cudaMemcpyHostToDevice));
  if (ndb) ORB_CUDA_TRY(cudaMemcpy(ddb.p, db, (size_t)ndb * 32, cudaMemcpyHostToDevice));
  int32_t* o = dout.as<int32_t>();
  if ((rc = orb_gpu_match_bruteforce_device(dq.as<uint8_t>(), nq, ddb.as<uint8_t>(), ndb, 0, o, o + nq,
                                            o + 2 * (size_t)nq, nullptr)))
    return rc;
  ORB_CUDA_TRY(cudaDeviceSynchronize());
  ORB_CUDA_TRY(cudaMemcpy(idx, o, (size_t)nq * 4, cudaMemcpyDeviceToHost));
  ORB_CUDA_TRY(cudaMemcpy(best, o + nq, (size_t)nq * 4, cudaMemcpyDeviceToHost));
  ORB_CUDA_TRY(cudaMemcpy(second, o + 2 * (size_t)nq, (size_t)nq * 4, cudaMemcpyDeviceToHost));
  return ORB_GPU_OK;
}

// multiply-accumulates per second of mma.sync m16n8k32 s8 x u8 (the pipe bruteforce_mma_kernel runs on)
int orb_gpu_imma_peak(int device, double* mac_per_s) {
  if (!mac_per_s) return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(device));
  int sms = 148;
  ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  TmpDev out;
  int rc = out.alloc(16);
  if (rc) return rc;
  cudaEvent_t e0, e1;
  ORB_CUDA_TRY(cudaEventCreate(&e0));
  ORB_CUDA_TRY(cudaEventCreate(&e1));
  const int iters = 1 << 13, blocks = sms * 2, threads = 512;
  imma_peak_kernel<<<blocks, threads>>>(64, out.as<int>());  // warm-up
  double bestv = 0;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    imma_peak_kernel<<<blocks, threads>>>(iters, out.as<int>());
    cudaEventRecord(e1);
    ORB_CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    bestv = std::max(bestv, (double)blocks * (threads / 32) * iters * 8.0 * (16.0 * 8 * 32) / (ms * 1e-3));
  }
  count_launch(4);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  ORB_CUDA_TRY(cudaGetLastError());
  *mac_per_s = bestv;
  return ORB_GPU_OK;
}

int orb_gpu_popc_peak(int device, double* popc_per_s) {
  if (!popc_per_s) return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(device));
  int sms = 148;
  ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  TmpDev out;
  int rc = out.alloc(16);
  if (rc) return rc;
  cudaEvent_t e0, e1;
  ORB_CUDA_TRY(cudaEventCreate(&e0));
  ORB_CUDA_TRY(cudaEventCreate(&e1));
  const int iters = 1 << 14, blocks = sms * 8, threads = 256;
  popc_peak_kernel<<<blocks, threads>>>(12345u, 256, out.as<uint32_t>());  // warm-up
  double bestv = 0;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    popc_peak_kernel<<<blocks, threads>>>(12345u + rep, iters, out.as<uint32_t>());
    cudaEventRecord(e1);
    ORB_CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    bestv = std::max(bestv, (double)blocks * threads * iters * 8.0 / (ms * 1e-3));
  }
  count_launch(4);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  ORB_CUDA_TRY(cudaGetLastError());
  *popc_per_s = bestv;
  return ORB_GPU_OK;
}

}  // extern "C"

// =============================================================================================
// brute force without any allocation (orb_gpu_match_bruteforce_noalloc_device)
// =============================================================================================
// The per-slice partial results, the shared per-query thresholds and the slice tickets of the tcgen05 kernel live in
// this file-scope array instead of a stream-ordered allocation: 1 MiB that every GPU which loads the library holds
// until the process exits.  All calls share it and it keeps the last call's contents (the words that must start at
// zero are cleared on the caller's stream before the kernel): two calls that overlap on one GPU race on it.
static __device__ __align__(256) unsigned char g_bf_noalloc_scratch[ORB_GPU_NOALLOC_SCRATCH_BYTES];

namespace orbgpu {
// In host code `g_bf_noalloc_scratch` is only a stand-in for the device symbol: its device address is looked up once
// per device, with that device current, and that pointer is what the kernel gets.
static int bf_noalloc_scratch(int dev, unsigned char** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static unsigned char* ptr[kMaxDev] = {};
  if (dev < 0 || dev >= kMaxDev) return ORB_GPU_ERR_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_bf_noalloc_scratch);  // the current device's copy
    if (e != cudaSuccess || !p) {
      set_last_error("cudaGetSymbolAddress(g_bf_noalloc_scratch)", e);
      return ORB_GPU_ERR_CUDA;
    }
    ptr[dev] = (unsigned char*)p;
  }
  *out = ptr[dev];
  return ORB_GPU_OK;
}
}  // namespace orbgpu

extern "C" int orb_gpu_match_bruteforce_noalloc_device(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb,
                                                       int32_t idx_base, int32_t* d_idx, int32_t* d_best,
                                                       int32_t* d_second, void* stream) {
  if (nq < 0 || ndb < 0 || (nq > 0 && (!d_q || !d_idx || !d_best || !d_second)) || (ndb > 0 && !d_db))
    return ORB_GPU_ERR_ARG;
  if (nq == 0) return ORB_GPU_OK;
  cudaStream_t s = (cudaStream_t)stream;
  int dev = 0, sms = 148;
  ORB_CUDA_TRY(cudaGetDevice(&dev));
  ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // slices as in orb_gpu_match_bruteforce_device (tcgen05 form): one wave of blocks, more only when a slice would
  // exceed the row field of the packed keys
  const int qblocks = (nq + kMmaQueries - 1) / kMmaQueries;
  long long waves = 1;
  while (waves * sms / qblocks * kBfSliceMax < ndb && waves < 65535) ++waves;
  long long nslices = std::max<long long>(1, waves * sms / qblocks);
  long long slice_rows = (ndb + nslices - 1) / nslices;
  slice_rows = std::max<long long>(kBfTile, (slice_rows + kBfTile - 1) / kBfTile * kBfTile);
  slice_rows = std::min<long long>(slice_rows, kBfSliceMax);
  nslices = std::max<long long>(1, (ndb + slice_rows - 1) / slice_rows);
  if (nslices > 65535) {
    set_last_message("orb_gpu_match_bruteforce_noalloc_device: database needs more than 65535 slices");
    return ORB_GPU_ERR_CAPACITY;
  }
  int32_t *pi = d_idx, *pb = d_best, *ps = d_second;
  unsigned *g_thr = nullptr, *g_ticket = nullptr;
  if (nslices > 1) {
    const size_t part_bytes = (size_t)nslices * nq * 3 * sizeof(int32_t);
    const size_t extra = ((size_t)nq + (size_t)qblocks) * sizeof(unsigned);
    if (part_bytes + extra > (size_t)ORB_GPU_NOALLOC_SCRATCH_BYTES) {
      char msg[256];
      snprintf(msg, sizeof(msg), "orb_gpu_match_bruteforce_noalloc_device: %d queries x %lld rows need %zu bytes of scratch "
               "(%lld slices), the static scratch holds %d: use orb_gpu_match_bruteforce_device", nq, (long long)ndb,
               part_bytes + extra, nslices, (int)ORB_GPU_NOALLOC_SCRATCH_BYTES);
      set_last_message(msg);
      return ORB_GPU_ERR_CAPACITY;
    }
    unsigned char* scratch = nullptr;
    const int rc = bf_noalloc_scratch(dev, &scratch);
    if (rc) return rc;
    pi = (int32_t*)scratch; pb = pi + (size_t)nslices * nq; ps = pb + (size_t)nslices * nq;
    g_thr = (unsigned*)(scratch + part_bytes);
    g_ticket = g_thr + nq;
    ORB_CUDA_TRY(cudaMemsetAsync(g_thr, 0, extra, s));  // thresholds and tickets start at zero; the partials are overwritten
  }
  dim3 grid(qblocks, (unsigned)nslices);
  const int rc = launch_bruteforce_tc5(grid, s, (const uint4*)d_q, nq, (const uint4*)d_db, ndb, slice_rows, idx_base, pi, pb,
                                       ps, g_thr, g_ticket, d_idx, d_best, d_second);
  if (rc) return rc;
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}
