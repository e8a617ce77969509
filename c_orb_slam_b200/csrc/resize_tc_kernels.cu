// Pyramid level l from level l-1 (cv::resize INTER_LINEAR, 8U fixed point) with the horizontal pass on the 5th-generation
// tensor cores and the source tiles moved by TMA.
//
// Reference: resize(mvImagePyramid[level-1], mvImagePyramid[level], sz, 0, 0, INTER_LINEAR), src/ORBextractor.cc:1107-1132
// (OpenCV's 8U path):  T[y][dx] = s[y][x0]*a0 + s[y][x0+1]*a1 (11-bit weights, a0 + a1 = 2048), then
// dst[dy][dx] = (((b0 * (T[y0][dx] >> 4)) >> 16) + ((b1 * (T[y0+1][dx] >> 4)) >> 16) + 2) >> 2.
//   horizontal pass = T is linear in the pixels with exact integers, so it is a matrix product: D[y][dx] = sum_x X[y][x]
//                     W[dx][x] on tcgen05.mma kind::i8 (u8 x u8 -> s32).  The 11-bit weights are split into bytes:
//                     W_hi = 16 * (a >> 8) (at most 128) and W_lo = a & 255, two accumulators, and
//                     T >> 4 = D_hi + (D_lo >> 4) exactly (256 * (a >> 8) * s is a multiple of 16; D_lo <= 65280).
//                     X (128 stacked source rows x up to 256 source columns) comes from the level below by TMA (128-byte
//                     swizzle) and is the K-major A operand; W (128 destination columns x K) is written per column
//                     chunk into shared memory (two entries per destination column) and is the K-major B operand.
//   vertical pass   = per-term truncation ((b * U) >> 16) makes it non-linear: it stays on the CUDA cores.  The epilogue
//                     thread of TMEM lane r combines its row's sums into U = T >> 4 (u16) and parks them in shared memory;
//                     after a barrier the warps of the group walk the tile's source rows: a source row that is the
//                     upper tap of a destination row (yinv table) emits that row from U[r], U[r + 1] with two IMAD.HI
//                     per pixel and one 32-bit store per four pixels.
// The frames of a level are stacked (source row s of frame f at f * h_src + s); strips of 128 stacked source rows overlap
// by one row, so the pair (s, s + 1) of every destination row lies inside one strip; a destination row belongs to the
// strip that holds its upper source row at local row 0..126.
//
// One persistent block per SM takes a contiguous run of (column chunk, strip) items, chunk-major, so the weight matrix is
// rebuilt once or twice per block.  Roles: warps 0..15 epilogue (two sets of eight, set i drains accumulator pair i),
// warp 16 TMA loads, warp 17 MMA issue + TMEM allocation, warp 18 writes W.
#include <cuda.h>

#include <atomic>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "umma.cuh"

namespace orbgpu {

constexpr int kRtRows = 128;                   // stacked source rows per strip (M); consecutive strips share one row
constexpr int kRtStride = kRtRows - 1;
constexpr int kRtCols = 128;                   // destination columns per chunk (N)
constexpr int kRtKMax = 256;                   // source columns per tile (two TMA boxes of 128)
constexpr int kRtBox = kRtRows * 128;          // bytes of one TMA box (128 rows x 128 columns)
constexpr int kRtXStage = 2 * kRtBox;
constexpr int kRtStages = 2;
constexpr int kRtW = kRtCols * kRtKMax;        // bytes of one weight matrix
constexpr int kRtUPitch = 2 * kRtCols + 16;    // bytes per row of the U exchange buffer (u16 per column, + 16: bank spread)
constexpr int kRtU = kRtRows * kRtUPitch;
constexpr int kRtEpiWarps = 16;
constexpr int kRtThreads = (kRtEpiWarps + 3) * 32;
constexpr int kRtOffW = kRtStages * kRtXStage;
constexpr int kRtOffU = kRtOffW + 2 * kRtW;
constexpr int kRtOffBar = kRtOffU + 2 * kRtU;
constexpr int kRtSmem = kRtOffBar + 256 + 1024;
// instruction descriptor without N: D = s32, A = u8 K-major, B = u8 K-major, M = 128
constexpr uint32_t kRtIdesc = (2u << 4) | ((128u >> 4) << 24);

struct ResizeTcParams {
  CUtensorMap src_map;   // level l-1 as [frames * h_src][w_src], box 128 x 128, 128-byte swizzle
  LevelGeom src, dst;
  int src_rows, dst_rows;  // stacked rows: frames * h
  int nchunks, nstrips;
};

__device__ __forceinline__ void rt_umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major operand in 128-byte swizzle atoms (rows of 128 bytes, 8-row groups 1024 bytes apart)
__device__ __forceinline__ uint64_t rt_desc_k128(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(16 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}

#define RT_R8(v, o) "=r"(v[o]), "=r"(v[o + 1]), "=r"(v[o + 2]), "=r"(v[o + 3]), "=r"(v[o + 4]), "=r"(v[o + 5]), "=r"(v[o + 6]), "=r"(v[o + 7])
#define RT_RW8(v, o) "+r"(v[o]), "+r"(v[o + 1]), "+r"(v[o + 2]), "+r"(v[o + 3]), "+r"(v[o + 4]), "+r"(v[o + 5]), "+r"(v[o + 6]), "+r"(v[o + 7])
// 64 consecutive columns of this thread's TMEM lane, the low 16 bits of each, two to a register
__device__ __forceinline__ void rt_ld64p(uint32_t (&v)[32], uint32_t taddr) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : RT_R8(v, 0), RT_R8(v, 8), RT_R8(v, 16), RT_R8(v, 24)
      : "r"(taddr));
}
__device__ __forceinline__ void rt_wait(uint32_t (&a)[32], uint32_t (&b)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : RT_RW8(a, 0), RT_RW8(a, 8), RT_RW8(a, 16), RT_RW8(a, 24), RT_RW8(b, 0), RT_RW8(b, 8), RT_RW8(b, 16), RT_RW8(b, 24)
               :
               : "memory");
}

// geometry of a column chunk, from the level's tap table: first source column of its tile (16-byte aligned for TMA),
// K steps of 32 source columns, destination columns
struct RtChunk { int x0, ksteps, ncols; };
__device__ __forceinline__ RtChunk rt_chunk(const LevelGeom& dst, int c) {
  RtChunk r;
  const int d0 = c * kRtCols;
  r.ncols = min(kRtCols, dst.w - d0);
  r.x0 = __ldg(dst.xofs + d0) & ~15;
  r.ksteps = ((__ldg(dst.xofs + d0 + r.ncols - 1) + 1 - r.x0) >> 5) + 1;
  return r;
}

__global__ void __launch_bounds__(kRtThreads, 1) resize_tc_kernel(const __grid_constant__ ResizeTcParams P) {
  extern __shared__ uint8_t rt_raw[];
  const uint32_t raw = smem_u32(rt_raw);
  uint8_t* sm = rt_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* s_x = sm;
  uint8_t* s_w = sm + kRtOffW;   // [2][128 columns][256 K bytes]: hi, lo; (k >> 4) * 2048 + (n >> 3) * 128 + (n & 7) * 16 + (k & 15)
  uint8_t* s_u = sm + kRtOffU;   // [2][128 rows][kRtUPitch]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + kRtOffBar);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 16);
  const uint32_t bar0 = smem_u32(bars);
  constexpr int kXFull = 0, kXEmpty = kRtStages, kWReady = 2 * kRtStages, kWEmpty = kWReady + 1, kAccFull = kWEmpty + 1,
                kAccEmpty = kAccFull + 2;
  auto bar = [&](int i) { return bar0 + 8u * i; };
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;

  if (t == 0) {
    for (int i = 0; i < kAccEmpty; ++i) mbar_init(bar(i), 1);
    mbar_init(bar(kAccEmpty), 256); mbar_init(bar(kAccEmpty + 1), 256);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kRtEpiWarps + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // the weight matrices start at zero: a chunk only ever writes two entries per destination column (and clears them again)
  for (int i = t; i < 2 * kRtW / 16; i += kRtThreads) reinterpret_cast<uint4*>(s_w)[i] = make_uint4(0, 0, 0, 0);
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  // every role walks the same run of items (column chunk major, strips inside); item n of the run uses X stage n % 2 and
  // accumulator pair n & 1; the weight matrix is rewritten whenever the chunk changes (generation wg)
  const int total = P.nchunks * P.nstrips;
  const int per = (total + (int)gridDim.x - 1) / (int)gridDim.x;
  const int i0 = (int)blockIdx.x * per, i1 = min(total, i0 + per);
  const int h_src = P.src.h;

  if (warp < kRtEpiWarps) {
    // ---------------- epilogue ----------------
    const int g = warp >> 3, hf = (warp >> 2) & 1, q = warp & 3, w8 = warp & 7;
    const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16) + g * 256 + 64 * hf;
    uint8_t* ub = s_u + g * kRtU;
    uint8_t* my_row = ub + (32 * q + lane) * kRtUPitch + 128 * hf;
    int pc = -1;
    RtChunk ch = {0, 0, 0};
    for (int it = i0; it < i1; ++it) {
      const int n = it - i0;
      if ((n & 1) != g) continue;
      const int c = it / P.nstrips, k = it - c * P.nstrips;
      if (c != pc) { ch = rt_chunk(P.dst, c); pc = c; }
      mbar_wait_suspend(bar(kAccFull + g), (n >> 1) & 1);
      tc_fence_after();
      asm volatile("bar.sync %0, 256;" ::"r"(1 + g) : "memory");  // the group has finished reading U of its previous item
      if (64 * hf < ch.ncols) {
        uint32_t hi[32], lo[32];
        rt_ld64p(hi, taddr);
        rt_ld64p(lo, taddr + 128);
        rt_wait(hi, lo);
        tc_fence_before();
        mbar_arrive(bar(kAccEmpty + g));  // the accumulators are in registers: the MMAs of the item after next may start
        uint4* dst = reinterpret_cast<uint4*>(my_row);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          uint32_t u[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) u[j] = hi[4 * i + j] + ((lo[4 * i + j] >> 4) & 0x0fff0fffu);  // U = T >> 4, two u16
          dst[i] = make_uint4(u[0], u[1], u[2], u[3]);
        }
      } else {
        tc_fence_before();
        mbar_arrive(bar(kAccEmpty + g));
      }
      asm volatile("bar.sync %0, 256;" ::"r"(1 + g) : "memory");  // U of the whole tile is in shared memory
      // vertical pass: source row r of the tile (stacked row s) emits the destination row whose upper tap it is
      const int s0 = k * kRtStride;
      int r = w8, s = s0 + w8;
      int f = s / h_src, sl = s - f * h_src;
      const bool active = 4 * lane < ch.ncols;
      for (; r < kRtRows - 1 && s < P.src_rows; r += 8, s += 8, sl += 8) {
        while (sl >= h_src) { sl -= h_src; ++f; }
        const int d = __ldg(P.dst.yinv + sl);
        if (d < 0) continue;
        const unsigned b0 = (unsigned)(unsigned short)__ldg(P.dst.ay0 + d) << 16, b1 = (unsigned)(unsigned short)__ldg(P.dst.ay1 + d) << 16;
        if (active) {
          const uint2 ua = *reinterpret_cast<const uint2*>(ub + r * kRtUPitch + 8 * lane);
          const uint2 ud = *reinterpret_cast<const uint2*>(ub + (r + 1) * kRtUPitch + 8 * lane);
          const unsigned v0 = (__umulhi(b0, ua.x & 0xffffu) + __umulhi(b1, ud.x & 0xffffu) + 2u) >> 2;
          const unsigned v1 = (__umulhi(b0, ua.x >> 16) + __umulhi(b1, ud.x >> 16) + 2u) >> 2;
          const unsigned v2 = (__umulhi(b0, ua.y & 0xffffu) + __umulhi(b1, ud.y & 0xffffu) + 2u) >> 2;
          const unsigned v3 = (__umulhi(b0, ua.y >> 16) + __umulhi(b1, ud.y >> 16) + 2u) >> 2;
          uint8_t* out = P.dst.img + (size_t)f * P.dst.frame_stride + (size_t)d * P.dst.pitch + c * kRtCols + 4 * lane;
          *reinterpret_cast<unsigned*>(out) = v0 | (v1 << 8) | (v2 << 16) | (v3 << 24);
        }
      }
    }
  } else if (warp == kRtEpiWarps) {
    // ---------------- TMA loads ----------------
    if (lane == 0) {
      int pc = -1;
      RtChunk ch = {0, 0, 0};
      for (int it = i0; it < i1; ++it) {
        const int n = it - i0, s = n % kRtStages, u = n / kRtStages;
        const int c = it / P.nstrips, k = it - c * P.nstrips;
        if (c != pc) { ch = rt_chunk(P.dst, c); pc = c; }
        const int nboxes = ch.ksteps > 4 ? 2 : 1;
        mbar_wait_suspend(bar(kXEmpty + s), (u & 1) ^ 1);
        mbar_expect_tx(bar(kXFull + s), (uint32_t)(nboxes * kRtBox));
        for (int a = 0; a < nboxes; ++a)
          tma_load_2d(smem_u32(s_x + s * kRtXStage + a * kRtBox), &P.src_map, ch.x0 + 128 * a, k * kRtStride, bar(kXFull + s));
      }
    }
  } else if (warp == kRtEpiWarps + 1) {
    // ---------------- MMA issue ----------------
    if (lane == 0) {
      int pc = -1, wg = -1;
      RtChunk ch = {0, 0, 0};
      uint32_t idesc = 0;
      for (int it = i0; it < i1; ++it) {
        const int n = it - i0, s = n % kRtStages, u = n / kRtStages, acc = n & 1;
        const int c = it / P.nstrips;
        if (c != pc) {
          ch = rt_chunk(P.dst, c); pc = c; ++wg;
          idesc = kRtIdesc | ((uint32_t)(((ch.ncols + 15) & ~15) >> 3) << 17);
          mbar_wait_suspend(bar(kWReady), wg & 1);
        }
        mbar_wait_suspend(bar(kXFull + s), u & 1);
        mbar_wait_suspend(bar(kAccEmpty + acc), ((n >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t xa = smem_u32(s_x + s * kRtXStage);
        for (int half = 0; half < 2; ++half) {
          const uint32_t wa = smem_u32(s_w + half * kRtW);
          for (int ks = 0; ks < ch.ksteps; ++ks)  // 32 source columns per step: 32 bytes along the swizzled rows of X
            rt_umma(tmem_base + acc * 256 + half * 128, rt_desc_k128(xa + (ks >> 2) * kRtBox + (ks & 3) * 32),
                    umma_desc(wa + ks * 2 * 2048, 2048, 128), idesc, ks > 0);
        }
        umma_commit(bar(kXEmpty + s));
        umma_commit(bar(kAccFull + acc));
        if (it + 1 == i1 || (it + 1) / P.nstrips != c) umma_commit(bar(kWEmpty));  // the last item that reads this W
      }
      // the last commits must have landed before the block (and its shared memory) goes away
      if (i1 > i0) {
        const int n = i1 - 1 - i0;
        mbar_wait_suspend(bar(kXEmpty + n % kRtStages), (n / kRtStages) & 1);
        mbar_wait_suspend(bar(kWEmpty), wg & 1);
      }
    }
  } else {
    // ---------------- W builder ----------------
    int pc = -1, wg = -1;
    RtChunk ch = {0, 0, 0}, old = {0, 0, 0};
    for (int it = i0; it < i1; ++it) {
      const int c = it / P.nstrips;
      if (c == pc) continue;
      old = ch;
      const int oc = pc;
      ch = rt_chunk(P.dst, c); pc = c; ++wg;
      if (wg > 0) mbar_wait_suspend(bar(kWEmpty), (wg - 1) & 1);  // every MMA that read the previous W is done
      auto at = [&](int n, int k) { return (k >> 4) * 2048 + (n >> 3) * 128 + (n & 7) * 16 + (k & 15); };
      if (oc >= 0)
        for (int n = lane; n < old.ncols; n += 32) {
          const int k0 = __ldg(P.dst.xofs + oc * kRtCols + n) - old.x0;
          s_w[at(n, k0)] = 0; s_w[at(n, k0 + 1)] = 0;
          s_w[kRtW + at(n, k0)] = 0; s_w[kRtW + at(n, k0 + 1)] = 0;
        }
      __syncwarp();
      for (int n = lane; n < ch.ncols; n += 32) {
        const int dx = c * kRtCols + n;
        const int k0 = __ldg(P.dst.xofs + dx) - ch.x0;
        const unsigned a0 = (unsigned)(unsigned short)__ldg(P.dst.ax0 + dx), a1 = (unsigned)(unsigned short)__ldg(P.dst.ax1 + dx);
        s_w[at(n, k0)] = (uint8_t)((a0 >> 8) * 16); s_w[at(n, k0 + 1)] = (uint8_t)((a1 >> 8) * 16);
        s_w[kRtW + at(n, k0)] = (uint8_t)(a0 & 255u); s_w[kRtW + at(n, k0 + 1)] = (uint8_t)(a1 & 255u);
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kWReady));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kRtEpiWarps + 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ---------------- host side ----------------
static std::atomic<int> g_resize_variant{[] { const char* e = getenv("ORB_GPU_RESIZE"); return e && !strcmp(e, "tc") ? 1 : 0; }()};
int resize_variant() { return g_resize_variant.load(std::memory_order_relaxed); }
bool resize_tc_level_ok(const LevelGeom& src, const LevelGeom& dst, int nframes) {
  if (!tma_available()) return false;
  if ((reinterpret_cast<uintptr_t>(src.img) & 15) || (src.pitch & 15)) return false;
  if (nframes > 1 && src.frame_stride != (long long)src.pitch * src.h) return false;
  if ((dst.pitch & 3) || (dst.frame_stride & 3) || dst.w < 16 || src.w <= dst.w || src.h <= dst.h || src.h > 32767) return false;
  // a chunk of 128 destination columns must fit 256 source columns behind a 16-aligned start
  return 127.0 * (double)src.w / (double)dst.w + 20.0 <= (double)kRtKMax;
}

int launch_resize_tc(const LevelGeom& src, const LevelGeom& dst, int nframes, cudaStream_t s) {
  ResizeTcParams P;
  std::memset(&P, 0, sizeof(P));
  P.src = src; P.dst = dst;
  P.src_rows = nframes * src.h; P.dst_rows = nframes * dst.h;
  P.nchunks = (dst.w + kRtCols - 1) / kRtCols;
  P.nstrips = (P.src_rows - 1 + kRtStride - 1) / kRtStride;
  int rc = tensor_map_2d_u8(&P.src_map, src.img, src.w, P.src_rows, src.pitch, 128, kRtRows, true);
  if (rc) return rc;
  int dev = 0, sms = 148;
  ORB_CUDA_TRY(cudaGetDevice(&dev));
  ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  ORB_CUDA_TRY(cudaFuncSetAttribute(resize_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRtSmem));
  const int total = P.nchunks * P.nstrips;
  resize_tc_kernel<<<total < sms ? total : sms, kRtThreads, kRtSmem, s>>>(P);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

}  // namespace orbgpu

// test / bench knob: 0 = byte-SIMD resize (default), 1 = tensor-core horizontal pass for batches (identical pixels)
extern "C" int orb_gpu_set_resize_variant(int variant) {
  if (variant != 0 && variant != 1) return ORB_GPU_ERR_ARG;
  orbgpu::g_resize_variant.store(variant, std::memory_order_relaxed);
  return ORB_GPU_OK;
}
