// GaussianBlur 7x7 sigma 2 (BORDER_REFLECT_101) on the 5th-generation tensor cores, tiles moved by TMA.
//
// Reference: GaussianBlur(workingMat, workingMat, Size(7, 7), 2, 2, BORDER_REFLECT_101), src/ORBextractor.cc:1085-1086
// (OpenCV's 8U fixed-point path: out = (sum_ij k_i k_j p + 32768) >> 16, k = [18 34 48 56 48 34 18]).  Every
// intermediate is an exact integer, so the two 1-D passes commute and each may run on any exact integer unit:
//
//   vertical pass   = banded Toeplitz matrix x pixel tile on tcgen05.mma kind::i8 (u8 x u8 -> s32, TMEM):
//                     D[r][c] = sum_k T[r][k] * X[k][c],  T[r][r + i] = k_i.  The frames of a level are stacked
//                     (rows of frame f start at f * h), a strip is 128 stacked rows, and REFLECT_101 at the top and
//                     bottom of a frame is folded into T (a tap that leaves the frame adds its weight to the row it
//                     reflects to, which always lies inside the 7-row window) -- T is rebuilt per strip (7 bytes per
//                     row).  X (136 rows x 256 columns, starting 16 columns left of the chunk: TMA wants 16-byte
//                     aligned box starts) comes straight from the pyramid by TMA (cp.async.bulk.tensor, 128-byte
//                     swizzle) and is the MN-major B operand; rows 136..159 of the K = 160 extent meet zero weights
//                     (they are whatever follows the tile in shared memory).
//   horizontal pass = in registers: the epilogue thread of TMEM lane r owns row r, reads its 16-bit column sums
//                     two to a register (tcgen05.ld pack::16b) and takes the seven taps from neighbouring registers
//                     (IDP.2A, the +32768 rounding folded into the accumulator).  REFLECT_101 at the left and right
//                     image border: the three halo byte columns of the tile are patched in shared memory before the
//                     MMA (the column sums of a reflected column are the reflected column sums).
//   output          = 32 rows x 112 columns per warp staged in shared memory, written by TMA (the hardware clips the
//                     columns past the image and the rows past the batch).
//
// One persistent block per SM walks strips (all column chunks of a strip share T).  Roles: warps 0..15 epilogue (two
// sets of eight, set i drains accumulator i: a warp owns 32 rows x 112 columns), warp 16 TMA loads, warp 17 MMA issue +
// TMEM allocation, warp 18 builds T and patches the halo columns.  Pipelines: x_full (TMA bytes) -> x_ready (patched)
// -> MMA -> x_empty / acc_full -> epilogue -> acc_empty; t_ready / t_empty for the two T buffers.
// Bring-up notes (tools/ubench/tma_probe.cu): a box whose first column is not a multiple of 16 bytes faults with
// "illegal instruction"; MN-major SWIZZLE_128B descriptors take LBO = bytes between the 128-column atoms, SBO = 1024.
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "umma.cuh"

namespace orbgpu {

constexpr int kBtRows = 128;                      // output rows per strip (M)
constexpr int kBtK = 160;                         // K extent of the MMA sequence (134 input rows used)
constexpr int kBtBoxRows = 136;                   // rows per TMA box
constexpr int kBtOutCols = 224;                   // output columns per chunk
constexpr int kBtLead = 16;                       // a tile starts 16 columns left of its first output column: TMA wants the
                                                  // box start 16-byte aligned (the halo itself is 3 columns)
constexpr int kBtAtom = kBtBoxRows * 128;          // bytes of one 128-column swizzle atom of X (17 KB, 1024-aligned); the K
                                                  // extent runs 24 rows into whatever follows it -- those meet zero weights
constexpr int kBtXStage = 2 * kBtAtom;            // one X stage: up to 256 input columns
constexpr int kBtStages = 3;                      // TMA -> patch -> MMA ring
constexpr int kBtT = kBtRows * kBtK;              // bytes of one T buffer
constexpr int kBtHalf = kBtOutCols / 2;              // output columns per epilogue warp
constexpr int kBtOutWarp = 32 * kBtHalf;            // staging bytes per epilogue warp
constexpr int kBtEpiWarps = 16;
constexpr int kBtMaxLevels = 8;
constexpr int kBtThreads = (kBtEpiWarps + 3) * 32;
constexpr int kBtOffT = kBtStages * kBtXStage;
constexpr int kBtOffOut = kBtOffT + 2 * kBtT;
constexpr int kBtOffBar = kBtOffOut + kBtEpiWarps * kBtOutWarp;
constexpr int kBtSmem = kBtOffBar + 256 + 1024;   // + alignment slack (the swizzle atoms need 1024-byte alignment)
// instruction descriptor without N: D = s32, A = u8 K-major, B = u8 MN-major, M = 128
constexpr uint32_t kBtIdesc = (2u << 4) | (0u << 7) | (0u << 10) | (1u << 16) | ((128u >> 4) << 24);

struct BlurTcLevel {
  int w, h;        // level size
  int rows;        // stacked rows of this launch: frames * h
  int nchunks;     // column chunks of kBtOutCols
  int strip0;      // first strip of this level in the launch's strip list
  int level;       // pyramid level (diagnostics)
};
struct BlurTcParams {
  CUtensorMap in[kBtMaxLevels];   // the parameter block stays below 4 KB: at most 8 levels per launch
  CUtensorMap out[kBtMaxLevels];
  BlurTcLevel lv[kBtMaxLevels];
  int nused, nstrips;
};

__device__ __forceinline__ void umma_u8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
// MN-major operand in 128-byte swizzle atoms: LBO = bytes between atoms along MN, SBO = bytes between 8-row groups along K
__device__ __forceinline__ uint64_t umma_desc_mn128(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
}

#define BT_R8(v, o) "=r"(v[o]), "=r"(v[o + 1]), "=r"(v[o + 2]), "=r"(v[o + 3]), "=r"(v[o + 4]), "=r"(v[o + 5]), "=r"(v[o + 6]), "=r"(v[o + 7])
#define BT_RW8(v, o) "+r"(v[o]), "+r"(v[o + 1]), "+r"(v[o + 2]), "+r"(v[o + 3]), "+r"(v[o + 4]), "+r"(v[o + 5]), "+r"(v[o + 6]), "+r"(v[o + 7])
// 64 (8) consecutive columns of this thread's TMEM lane, the low 16 bits of each, two to a register
__device__ __forceinline__ void bt_ld64p(uint32_t (&v)[32], uint32_t taddr) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : BT_R8(v, 0), BT_R8(v, 8), BT_R8(v, 16), BT_R8(v, 24)
      : "r"(taddr));
}
__device__ __forceinline__ void bt_ld8p(uint32_t (&v)[4], uint32_t taddr) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.pack::16b.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
               : "r"(taddr));
}
__device__ __forceinline__ void bt_wait(uint32_t (&v)[32], uint32_t (&h)[4]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : BT_RW8(v, 0), BT_RW8(v, 8), BT_RW8(v, 16), BT_RW8(v, 24), "+r"(h[0]), "+r"(h[1]), "+r"(h[2]), "+r"(h[3])
               :
               : "memory");
}

// 16 output pixels of a 64-column group.  The registers hold the column sums from one column left of the window of
// the group's first output (tile column 64 * grp + 12, an even column): output e = 2m of sub-block SB takes pairs
// 8 SB + m .. + 3 starting at their high halves, output 2m + 1 takes pairs 8 SB + m + 1 .. + 4 from their low halves.
// Pairs SPLIT and up are the halo registers.
template <int SB, int SPLIT>
__device__ __forceinline__ uint4 bt_row16(const uint32_t (&p)[32], const uint32_t (&hl)[4]) {
  // IDP2A weights: .lo takes bytes 0,1 and .hi bytes 2,3 of the second operand
  const unsigned W1 = (18u << 8) | (34u << 16) | (48u << 24);        // (0, k0) | (k1, k2)  window starting at an odd column
  const unsigned W2 = 56u | (48u << 8) | (34u << 16) | (18u << 24);  // (k3, k4) | (k5, k6)
  const unsigned W3 = 18u | (34u << 8) | (48u << 16) | (56u << 24);  // (k0, k1) | (k2, k3)  window starting at an even column
  const unsigned W4 = 48u | (34u << 8) | (18u << 16);                // (k4, k5) | (k6, 0)
  unsigned q[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) q[i] = (8 * SB + i < SPLIT) ? p[(8 * SB + i) & 31] : hl[(8 * SB + i - SPLIT) & 3];
  unsigned oe[8], oo[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    unsigned e = __dp2a_lo(q[m], W1, 32768u); e = __dp2a_hi(q[m + 1], W1, e); e = __dp2a_lo(q[m + 2], W2, e); e = __dp2a_hi(q[m + 3], W2, e);
    unsigned o = __dp2a_lo(q[m + 1], W3, 32768u); o = __dp2a_hi(q[m + 2], W3, o); o = __dp2a_lo(q[m + 3], W4, o); o = __dp2a_hi(q[m + 4], W4, o);
    oe[m] = e; oo[m] = o;
  }
  unsigned wv[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const unsigned lo = __byte_perm(oe[2 * k], oo[2 * k], 0x0062);          // (e.b2, o.b2, -, -)
    const unsigned hi = __byte_perm(oe[2 * k + 1], oo[2 * k + 1], 0x0062);
    wv[k] = __byte_perm(lo, hi, 0x5410);
  }
  return make_uint4(wv[0], wv[1], wv[2], wv[3]);
}

__global__ void __launch_bounds__(kBtThreads, 1) blur_tc_kernel(const __grid_constant__ BlurTcParams P) {
  extern __shared__ uint8_t bt_raw[];
  const uint32_t raw = smem_u32(bt_raw);
  uint8_t* sm = bt_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* s_x = sm;
  uint8_t* s_t = sm + kBtOffT;
  uint8_t* s_o = sm + kBtOffOut;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + kBtOffBar);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 24);
  const uint32_t bar0 = smem_u32(bars);
  // barrier slots (per X stage / T buffer / accumulator)
  constexpr int kXFull = 0, kXReady = kBtStages, kXEmpty = 2 * kBtStages, kTReady = 3 * kBtStages, kTEmpty = kTReady + 2,
                kAccFull = kTEmpty + 2, kAccEmpty = kAccFull + 2, kNumBars = kAccEmpty + 2;
  static_assert(kNumBars <= 24, "barrier slots");
  auto bar = [&](int i) { return bar0 + 8u * i; };
  const int t = threadIdx.x, lane = t & 31;
  const int warp = __shfl_sync(0xffffffffu, t >> 5, 0);  // provably warp-uniform: the role branches are uniform branches

  if (t == 0) {
    for (int i = 0; i < kAccEmpty; ++i) mbar_init(bar(i), 1);
    mbar_init(bar(kAccEmpty), 256); mbar_init(bar(kAccEmpty + 1), 256);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kBtEpiWarps + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // both T buffers start at zero: a strip only ever writes the 7 band entries of each row
  for (int i = t; i < 2 * kBtT / 16; i += kBtThreads) reinterpret_cast<uint4*>(s_t)[i] = make_uint4(0, 0, 0, 0);
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  // every role walks the same list: strips blockIdx.x, + gridDim.x, ...; the column chunks of a strip in order.
  // Tile number `tile` uses X stage tile % kBtStages and accumulator tile & 1; strip number sn uses T buffer sn & 1.
  // (strips are visited in ascending order, so the level index only moves forward: a running search from the last one)
  int lvl = 0;
  auto level_of = [&](int si) { while (lvl + 1 < P.nused && si >= P.lv[lvl + 1].strip0) ++lvl; return lvl; };

  if (warp < kBtEpiWarps) {
    // ---------------- epilogue: horizontal pass + store ----------------
    // warps 8 g + 4 hf + q: accumulator g, column half hf (112 output columns), TMEM lane quarter q (32 rows)
    const int g = warp >> 3, hf = (warp >> 2) & 1, q = warp & 3;
    const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16) + g * 256 + hf * kBtHalf + 12;
    uint8_t* stage = s_o + warp * kBtOutWarp;
    uint8_t* my_row = stage + lane * kBtHalf;
    int tile = 0;
    for (int si = blockIdx.x; si < P.nstrips; si += gridDim.x) {
      const int j = level_of(si);
      const BlurTcLevel L = P.lv[j];
      const int g0 = (si - L.strip0) * kBtRows;
      for (int c = 0; c < L.nchunks; ++c, ++tile) {
        if ((tile & 1) != g) continue;
        const int c0 = c * kBtOutCols + hf * kBtHalf;
        const int rem = min(kBtHalf, L.w - c0);   // output columns of this warp (<= 0: none)
        mbar_wait_suspend(bar(kAccFull + g), (tile >> 1) & 1);
        tc_fence_after();
        if (lane == 0) tma_store_wait_read();  // the previous store of this warp has left the staging rows
        __syncwarp();
        // The window of this warp's output i spans tile columns hf * 112 + i + 13 .. + 19; the registers hold the column
        // sums from tile column hf * 112 + 12 (first 64 outputs, 4 halo registers) and hf * 112 + 76 (last 48 outputs).
        if (rem > 0) {
          uint32_t p[32], hl[4];
          bt_ld64p(p, taddr);
          bt_ld8p(hl, taddr + 64);
          bt_wait(p, hl);
          uint4* dst = reinterpret_cast<uint4*>(my_row);
          if (rem <= 64) { tc_fence_before(); mbar_arrive(bar(kAccEmpty + g)); }
          dst[0] = bt_row16<0, 32>(p, hl);
          if (rem > 16) dst[1] = bt_row16<1, 32>(p, hl);
          if (rem > 32) dst[2] = bt_row16<2, 32>(p, hl);
          if (rem > 48) dst[3] = bt_row16<3, 32>(p, hl);
          if (rem > 64) {
            bt_ld64p(p, taddr + 64);
            bt_wait(p, hl);
            tc_fence_before();
            mbar_arrive(bar(kAccEmpty + g));  // the accumulator is in registers: the MMA of the tile after next may start
            dst[4] = bt_row16<0, 32>(p, hl);
            if (rem > 80) dst[5] = bt_row16<1, 32>(p, hl);
            if (rem > 96) dst[6] = bt_row16<2, 32>(p, hl);
          }
          fence_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&P.out[j], c0, g0 + 32 * q, smem_u32(stage));
            tma_store_commit();
          }
        } else {
          tc_fence_before();
          mbar_arrive(bar(kAccEmpty + g));
        }
      }
    }
    if (lane == 0) tma_store_wait_all();
  } else if (warp == kBtEpiWarps) {
    // ---------------- TMA loads ----------------
    if (lane == 0) {
      int tile = 0;
      for (int si = blockIdx.x; si < P.nstrips; si += gridDim.x) {
        const int j = level_of(si);
        const BlurTcLevel L = P.lv[j];
        const int g0 = (si - L.strip0) * kBtRows;
        for (int c = 0; c < L.nchunks; ++c, ++tile) {
          const int s = tile % kBtStages, u = tile / kBtStages;
          const int c0 = c * kBtOutCols;
          const int n = (min(kBtOutCols, L.w - c0) + kBtLead + 3 + 15) & ~15;
          const int natoms = n > 128 ? 2 : 1;
          mbar_wait_suspend(bar(kXEmpty + s), (u & 1) ^ 1);
          mbar_expect_tx(bar(kXFull + s), (uint32_t)(natoms * kBtBoxRows * 128));
          for (int a = 0; a < natoms; ++a)
            tma_load_2d(smem_u32(s_x + s * kBtXStage + a * kBtAtom), &P.in[j], c0 - kBtLead + 128 * a, g0 - 3,
                        bar(kXFull + s));
        }
      }
    }
  } else if (warp == kBtEpiWarps + 1) {
    // ---------------- MMA issue ----------------
    if (lane == 0) {
      int tile = 0, sn = 0;
      for (int si = blockIdx.x; si < P.nstrips; si += gridDim.x, ++sn) {
        const int j = level_of(si);
        const BlurTcLevel L = P.lv[j];
        const int b = sn & 1;
        mbar_wait_suspend(bar(kTReady + b), (sn >> 1) & 1);
        const uint64_t da = umma_desc(smem_u32(s_t + b * kBtT), 2048, 128);
        for (int c = 0; c < L.nchunks; ++c, ++tile) {
          const int s = tile % kBtStages, u = tile / kBtStages, acc = tile & 1;
          const int n = (min(kBtOutCols, L.w - c * kBtOutCols) + kBtLead + 3 + 15) & ~15;
          const uint64_t db = umma_desc_mn128(smem_u32(s_x + s * kBtXStage), kBtAtom, 1024);
          const uint32_t idesc = kBtIdesc | ((uint32_t)(n >> 3) << 17);
          mbar_wait_suspend(bar(kXReady + s), u & 1);
          mbar_wait_suspend(bar(kAccEmpty + acc), ((tile >> 1) & 1) ^ 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < kBtK / 32; ++k)  // 32 input rows per step: two 16-byte K chunks of T, four 8-row groups of X
            umma_u8(tmem_base + acc * 256, da + (uint64_t)(k * (2 * 2048 >> 4)), db + (uint64_t)(k * (4096 >> 4)), idesc, k > 0);
          umma_commit(bar(kXEmpty + s));
          umma_commit(bar(kAccFull + acc));
        }
        umma_commit(bar(kTEmpty + b));
      }
      // the last commits must have landed before the block (and its shared memory) goes away
      if (tile > 0) mbar_wait_suspend(bar(kXEmpty + (tile - 1) % kBtStages), ((tile - 1) / kBtStages) & 1);
      if (sn > 0) mbar_wait_suspend(bar(kTEmpty + ((sn - 1) & 1)), ((sn - 1) >> 1) & 1);
    }
  } else {
    // ---------------- T builder + halo patcher ----------------
    int tile = 0, sn = 0;
    for (int si = blockIdx.x; si < P.nstrips; si += gridDim.x, ++sn) {
      const int j = level_of(si);
      const BlurTcLevel L = P.lv[j];
      const int g0 = (si - L.strip0) * kBtRows;
      const int b = sn & 1;
      mbar_wait_suspend(bar(kTEmpty + b), ((sn >> 1) & 1) ^ 1);
      uint8_t* tb = s_t + b * kBtT;
      for (int r = lane; r < kBtRows; r += 32) {
        const int y = g0 + r, f = y / L.h, yf = y - f * L.h;
        unsigned v[7];
        if (yf >= 3 && yf + 3 < L.h) {
          v[0] = 18; v[1] = 34; v[2] = 48; v[3] = 56; v[4] = 48; v[5] = 34; v[6] = 18;
        } else {
          const unsigned kw[7] = {18, 34, 48, 56, 48, 34, 18};
#pragma unroll
          for (int kk = 0; kk < 7; ++kk) v[kk] = 0;
#pragma unroll
          for (int i = 0; i < 7; ++i) {
            int tt = yf + i - 3;
            tt = tt < 0 ? -tt : tt;
            tt = tt >= L.h ? 2 * (L.h - 1) - tt : tt;
            const int kk = tt - yf + 3;  // 0..6: the reflected row lies inside the window
#pragma unroll
            for (int m = 0; m < 7; ++m) v[m] += (kk == m) ? kw[i] : 0u;
          }
        }
#pragma unroll
        for (int kk = 0; kk < 7; ++kk) {
          const int k = r + kk;
          tb[(k >> 4) * 2048 + (r >> 3) * 128 + (r & 7) * 16 + (k & 15)] = (uint8_t)v[kk];
        }
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kTReady + b));
      for (int c = 0; c < L.nchunks; ++c, ++tile) {
        const int s = tile % kBtStages, u = tile / kBtStages;
        const int c0 = c * kBtOutCols;
        const bool left = c0 == 0, right = c0 + kBtOutCols + 3 > L.w;  // the tile reaches past column w - 1
        mbar_wait_suspend(bar(kXFull + s), u & 1);
        if (left || right) {
          // REFLECT_101 columns: tile columns 13..15 <- 19..17 (image columns -3..-1) and e..e+2 <- e-2..e-4 (image
          // columns w..w+2).  All source bytes of a lane's rows are read before any is written (the loads overlap).
          uint8_t* xs = s_x + s * kBtXStage;
          const int e = L.w - c0 + kBtLead;  // tile column of image column w
          constexpr int kRowsPerLane = (kBtBoxRows + 31) / 32;
          uint8_t v[kRowsPerLane][6];
          auto at = [&](int k, int n) -> uint8_t* {
            return xs + (n >> 7) * kBtAtom + k * 128 + (((((n & 127) >> 4) ^ (k & 7)) << 4) | (n & 15));
          };
#pragma unroll
          for (int i = 0; i < kRowsPerLane; ++i) {
            const int k = min(lane + 32 * i, kBtBoxRows - 1);
#pragma unroll
            for (int m = 0; m < 3; ++m) {
              v[i][m] = left ? *at(k, kBtLead + 3 - m) : (uint8_t)0;
              v[i][3 + m] = right ? *at(k, e - 2 - m) : (uint8_t)0;
            }
          }
#pragma unroll
          for (int i = 0; i < kRowsPerLane; ++i) {
            const int k = min(lane + 32 * i, kBtBoxRows - 1);
#pragma unroll
            for (int m = 0; m < 3; ++m) {
              if (left) *at(k, kBtLead - 3 + m) = v[i][m];
              if (right) *at(k, e + m) = v[i][3 + m];
            }
          }
          fence_async_smem();
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(kXReady + s));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kBtEpiWarps + 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ---------------- host side: tensor maps (cached) and launch ----------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static int tensor_map_2d(CUtensorMap* out, const void* base, int w, long long rows, int pitch, int box_w, int box_h,
                         bool swizzle128);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// u8 matrix [rows][w] with `pitch` bytes per row; box = box_w x box_h
bool tma_available() { return encode_tiled_fn() != nullptr; }
int tensor_map_2d_u8(void* out_map, const void* base, int w, long long rows, int pitch, int box_w, int box_h, bool swizzle128) {
  return tensor_map_2d(reinterpret_cast<CUtensorMap*>(out_map), base, w, rows, pitch, box_w, box_h, swizzle128);
}
static int tensor_map_2d(CUtensorMap* out, const void* base, int w, long long rows, int pitch, int box_w, int box_h,
                         bool swizzle128) {
  typedef std::tuple<const void*, int, long long, int, int, int> Key;
  static std::mutex mu;
  static std::map<Key, CUtensorMap> cache;
  const Key key(base, w, rows, pitch, box_w, box_h * 2 + (swizzle128 ? 1 : 0));
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return ORB_GPU_OK; }
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) { set_last_message("cuTensorMapEncodeTiled is not available"); return ORB_GPU_ERR_CUDA; }
  const cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch};
  const cuuint32_t box[2] = {(cuuint32_t)box_w, (cuuint32_t)box_h};
  const cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  const CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_last_message("cuTensorMapEncodeTiled failed"); return ORB_GPU_ERR_CUDA; }
  if (cache.size() > 8192) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return ORB_GPU_OK;
}

// Can level l of this launch go through the tensor-core kernel?  TMA wants 16-byte aligned bases and pitches, and the
// frames of the level must be stacked without gaps (always true for the handle's own buffers).
bool blur_tc_level_ok(const LevelGeom& g, int nframes) {
  if (!encode_tiled_fn()) return false;
  if ((reinterpret_cast<uintptr_t>(g.img) & 15) || (g.pitch & 15) || (reinterpret_cast<uintptr_t>(g.blur) & 15) ||
      (g.bpitch & 15))
    return false;
  if (nframes > 1 && (g.frame_stride != (long long)g.pitch * g.h || g.bframe_stride != (long long)g.bpitch * g.h))
    return false;
  return g.h >= 8 && g.w >= 8;
}

int launch_blur_tc(const Plan& p, int nframes, unsigned level_mask, cudaStream_t s) {
  int dev = 0, sms = 148;
  ORB_CUDA_TRY(cudaGetDevice(&dev));
  ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  ORB_CUDA_TRY(cudaFuncSetAttribute(blur_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBtSmem));
  for (int l = 0; l < p.nlevels;) {
    BlurTcParams P;
    std::memset(&P, 0, sizeof(P));
    int rc, strips = 0, j = 0;
    for (; l < p.nlevels && j < kBtMaxLevels; ++l) {
      if (!(level_mask >> l & 1u)) continue;
      const LevelGeom& g = p.lv[l];
      const long long rows = (long long)nframes * g.h;
      if ((rc = tensor_map_2d(&P.in[j], g.img, g.w, rows, g.pitch, 128, kBtBoxRows, true)) ||
          (rc = tensor_map_2d(&P.out[j], g.blur, g.w, rows, g.bpitch, kBtHalf, 32, false)))
        return rc;
      P.lv[j].w = g.w; P.lv[j].h = g.h; P.lv[j].rows = (int)rows; P.lv[j].level = l;
      P.lv[j].nchunks = (g.w + kBtOutCols - 1) / kBtOutCols;
      P.lv[j].strip0 = strips;
      strips += (int)((rows + kBtRows - 1) / kBtRows);
      ++j;
    }
    if (!j) break;
    P.nused = j; P.nstrips = strips;
    blur_tc_kernel<<<strips < sms ? strips : sms, kBtThreads, kBtSmem, s>>>(P);
    count_launch();
    ORB_CUDA_TRY(cudaGetLastError());
  }
  return ORB_GPU_OK;
}

}  // namespace orbgpu
