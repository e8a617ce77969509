// Extraction kernels of the B200-native ORB front-end (sm_100a).
//
// Stages (reference: /root/reference/src/ORBextractor.cc):
//   resize_level_kernel   ComputePyramid :1107-1132   (cv::resize INTER_LINEAR, 11-bit fixed point)
//   fast_cells_kernel     per-cell FAST-9 + NMS + iniTh/minTh retry  :765-829
//   quadtree_kernel       DistributeOctTree :539-763 as sort + node-level list simulation
//   blur_kernel           GaussianBlur 7x7 sigma 2, REFLECT_101  :1085-1086
//   orient_desc_kernel    IC_Angle :77-104, computeOrbDescriptor :106-147, output fix-ups
// Integer stages are bit-exact with the CPU oracle; float stages use explicit _rn intrinsics
// (the TU is compiled with -fmad=false) so every operation rounds as in the oracle.
#include <cuda.h>

#include <cfloat>

#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "umma.cuh"

namespace orbgpu {

// rBRIEF tests (x0, y0, x1, y1) as floats, read by the descriptor kernel with coalesced 16-byte loads
__device__ float4 g_pattern[256];
static const int8_t h_pattern[256 * 4] = {
#include "brief_pattern.inc"
};

int upload_constants() {
  float4 pf[256];
  for (int k = 0; k < 256; ++k)
    pf[k] = make_float4((float)h_pattern[4 * k], (float)h_pattern[4 * k + 1], (float)h_pattern[4 * k + 2],
                        (float)h_pattern[4 * k + 3]);
  ORB_CUDA_TRY(cudaMemcpyToSymbol(g_pattern, pf, sizeof(pf)));
  return ORB_GPU_OK;
}

// =============================================================================================
// 1. pyramid level l from level l-1
// =============================================================================================
// cv::resize INTER_LINEAR, 8U fixed point:  T = s[x0]*a0 + s[x1]*a1 (11-bit weights),
// dst = (((b0 * (T0 >> 4)) >> 16) + ((b1 * (T1 >> 4)) >> 16) + 2) >> 2.
// A thread owns two adjacent destination columns of a band of kRzRows destination rows: its taps
// are fixed, so per source row it loads the two aligned words that hold them, one PRMT (lane-constant
// selector) puts the two tap pairs in one register and two IDP2A give T for both columns.  The 16 loads
// of kRzBatch destination rows are issued together; the second read of a source row (1.67 destination
// rows share one) hits L1.  No shared-memory staging, loads and stores are coalesced.
constexpr int kRzRows = 16, kRzBatch = 4;

__global__ void __launch_bounds__(512)
resize_level_kernel(LevelGeom src, LevelGeom dst) {
  __shared__ unsigned s_off0[kRzRows], s_off1[kRzRows];  // word offsets of the two source rows
  __shared__ unsigned s_b0[kRzRows], s_b1[kRzRows];      // vertical weights << 16
  const int f = blockIdx.y;
  const int dy0 = blockIdx.x * kRzRows;
  const int rstep = src.pitch >> 2;
  if (threadIdx.x < kRzRows) {
    const int dy = min(dy0 + (int)threadIdx.x, dst.h - 1);  // rows past the image repeat the last one (not stored)
    const int sy0 = dst.yofs[dy];
    s_off0[threadIdx.x] = (unsigned)(sy0 * rstep);
    s_off1[threadIdx.x] = (unsigned)(min(sy0 + 1, src.h - 1) * rstep);
    s_b0[threadIdx.x] = (unsigned)dst.ay0[dy] << 16;
    s_b1[threadIdx.x] = (unsigned)dst.ay1[dy] << 16;
  }
  __syncthreads();
  const int nt = (dst.w + 1) >> 1;
  const int last_word = rstep - 1;
  const unsigned* sbase = reinterpret_cast<const unsigned*>(src.img + (size_t)f * src.frame_stride);
  uint8_t* dbase = dst.img + (size_t)f * dst.frame_stride + (size_t)dy0 * dst.pitch;
  const int nrows = min(kRzRows, dst.h - dy0);
  for (int t = threadIdx.x; t < nt; t += blockDim.x) {
    const int dxa = 2 * t, dxb = min(dxa + 1, dst.w - 1);
    const int xa = dst.xofs[dxa], xb = dst.xofs[dxb];
    const unsigned wa = (unsigned)(unsigned short)dst.ax0[dxa] | ((unsigned)(unsigned short)dst.ax1[dxa] << 16);
    const unsigned wb = (unsigned)(unsigned short)dst.ax0[dxb] | ((unsigned)(unsigned short)dst.ax1[dxb] << 16);
    const int wi0 = xa >> 2;
    const unsigned* c0 = sbase + wi0;                       // this thread's two word columns
    const unsigned* c1 = sbase + min(wi0 + 1, last_word);
    const int oa = xa & 3, ob = xb - 4 * wi0;  // byte offsets of the two tap pairs in the 8-byte window (<= 6)
    const unsigned sel = (unsigned)oa | ((unsigned)(oa + 1) << 4) | ((unsigned)ob << 8) | ((unsigned)(ob + 1) << 12);
    uint8_t* drow = dbase + 2 * t;
#pragma unroll 1
    for (int i0 = 0; i0 < kRzRows; i0 += kRzBatch) {
      unsigned w[kRzBatch][4];
#pragma unroll
      for (int k = 0; k < kRzBatch; ++k) {
        const unsigned o0 = s_off0[i0 + k], o1 = s_off1[i0 + k];
        w[k][0] = c0[o0]; w[k][1] = c1[o0]; w[k][2] = c0[o1]; w[k][3] = c1[o1];
      }
#pragma unroll
      for (int k = 0; k < kRzBatch; ++k) {
        const unsigned p0 = __byte_perm(w[k][0], w[k][1], sel), p1 = __byte_perm(w[k][2], w[k][3], sel);
        const unsigned b0 = s_b0[i0 + k], b1 = s_b1[i0 + k];
        const unsigned va = (__umulhi(b0, __dp2a_lo(wa, p0, 0u) >> 4) + __umulhi(b1, __dp2a_lo(wa, p1, 0u) >> 4) + 2u) >> 2;
        const unsigned vb = (__umulhi(b0, __dp2a_hi(wb, p0, 0u) >> 4) + __umulhi(b1, __dp2a_hi(wb, p1, 0u) >> 4) + 2u) >> 2;
        if (i0 + k < nrows) *reinterpret_cast<unsigned short*>(drow) = (unsigned short)(va | (vb << 8));
        drow += dst.pitch;
      }
    }
  }
}

// Four destination columns per thread (scaleFactor <= 2: their taps span at most 11 bytes = three aligned words from
// the thread's first word).  Columns 0,1 take their tap pairs from words (0,1); columns 2,3 from words (1,2) when
// column 2 starts in word 1 or later, else from (0,1) as well -- a per-thread constant choice (two selects per row).
// The block walks the SOURCE rows of its band once: the horizontal pass T >> 4 of a source row (3 loads, 2 PRMT,
// 4 IDP.2A) is kept for the next row, and a destination row is emitted as soon as its lower source row is done --
// for 1 < scale <= 2 consecutive destination rows start at consecutive or next-but-one source rows, so every source
// row of the band is needed and each is loaded and filtered once (1.2 per destination row instead of 2).
//
// kStaged: the band's source rows (at most kRzSrcRows of them, whole pitch) come to shared memory by TMA first -- one
// elected thread issues a 256 x kRzSrcRows byte box per 256-byte column chunk, everybody waits on the mbarrier once --
// and the walk below reads shared memory.  The direct form waits for global loads once per batch of four source rows
// (30 % of its stall samples on the first use of a loaded word); staged, a block pays the memory latency once and the
// other blocks of the SM compute meanwhile.
constexpr int kRzSrcRows = 2 * kRzRows + 4;
constexpr int kRzChunk = 256;                              // bytes per column chunk = TMA box width
constexpr int kRzMaxChunks = 5;                            // source pitch <= 1280 bytes
struct RzStage { CUtensorMap map; int nchunks, src_h, box_h; };  // source level as [frames * h][pitch] bytes, box 256 x box_h
// box_h = the most source rows a band of kRzRows destination rows can need at this scale (<= kRzSrcRows)

template <bool kStaged>
__global__ void __launch_bounds__(256)
resize_level4_kernel(LevelGeom src, LevelGeom dst, int rows /*destination rows per block: 4, 8 or 16*/,
                     const __grid_constant__ RzStage S) {
  extern __shared__ __align__(128) uint8_t rz_band[];  // kStaged: [nchunks][box_h][256]
  __shared__ uint2 s_src[kRzSrcRows];  // per source row of the band: word offset, destination rows to emit (e1 | e2 << 8)
  __shared__ uint2 s_b[kRzRows];       // vertical weights << 16 of a destination row
  __shared__ int s_ns;
  __shared__ __align__(8) uint64_t s_bar;
  const int f = blockIdx.y;
  const int dy0 = blockIdx.x * rows;
  const int rstep = src.pitch >> 2;
  const int nrows = min(rows, dst.h - dy0);
  const int s_begin = dst.yofs[dy0];
  if (kStaged && threadIdx.x == 0) {
    const uint32_t bar = smem_u32(&s_bar);
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(bar, (uint32_t)(S.nchunks * S.box_h * kRzChunk));
    for (int k = 0; k < S.nchunks; ++k)  // rows past the frame (or the tensor) land too and are never read
      tma_load_2d(smem_u32(rz_band + k * S.box_h * kRzChunk), &S.map, kRzChunk * k, f * S.src_h + s_begin, bar);
  }
  if (threadIdx.x < kRzSrcRows) {
    const int row = min(s_begin + (int)threadIdx.x, src.h - 1);
    // staged: row of the band (rows past the staged box are table padding that is loaded but never emitted: clamp)
    s_src[threadIdx.x] = make_uint2(kStaged ? (unsigned)(min(row - s_begin, S.box_h - 1) * (kRzChunk / 4)) : (unsigned)(row * rstep), 0xffffu);
  }
  __syncthreads();
  if (threadIdx.x < nrows) {
    const int dy = dy0 + (int)threadIdx.x;
    const int sy0 = dst.yofs[dy], i0 = sy0 - s_begin;
    s_b[threadIdx.x] = make_uint2((unsigned)dst.ay0[dy] << 16, (unsigned)dst.ay1[dy] << 16);
    // e1: emitted after source row i0 + 1 from (previous, current); e2: the last source row serves as both taps
    // (sy0 == src.h - 1, the reference clamps sy0 + 1)
    unsigned char* e = reinterpret_cast<unsigned char*>(&s_src[sy0 + 1 <= src.h - 1 ? i0 + 1 : i0].y);
    e[sy0 + 1 <= src.h - 1 ? 0 : 1] = (unsigned char)threadIdx.x;
    if ((int)threadIdx.x == nrows - 1) s_ns = min(sy0 + 1, src.h - 1) - s_begin + 1;
  }
  __syncthreads();
  const int ns = s_ns;
  const int nt = (dst.w + 3) >> 2;
  const int last_word = rstep - 1;
  const unsigned* sbase = kStaged ? reinterpret_cast<const unsigned*>(rz_band)
                                  : reinterpret_cast<const unsigned*>(src.img + (size_t)f * src.frame_stride);
  // word column wi of the source row lives at chunk wi >> 6, word wi & 63 of the staged band
  const int chunk_words = kStaged ? S.box_h * (kRzChunk / 4) : 0;
  auto col = [&](int wi) { return kStaged ? (wi >> 6) * chunk_words + (wi & 63) : wi; };
  if (kStaged && ns > S.box_h) __trap();  // the host's bound on the band height (launch_resize_level) must hold
  uint8_t* dbase = dst.img + (size_t)f * dst.frame_stride + (size_t)dy0 * dst.pitch;
  if (kStaged) mbar_wait(smem_u32(&s_bar), 0);
  for (int t = threadIdx.x; t < nt; t += blockDim.x) {
    int x[4];
    unsigned wt[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int dx = min(4 * t + j, dst.w - 1);
      x[j] = dst.xofs[dx];
      wt[j] = (unsigned)(unsigned short)dst.ax0[dx] | ((unsigned)(unsigned short)dst.ax1[dx] << 16);
    }
    const int wi0 = x[0] >> 2;
    const bool hi = x[2] - 4 * wi0 >= 4;
    // word columns of the two tap windows: (ca, ca + 1) for destination columns 0,1 and (cb, cb + 1) for 2,3
    const unsigned* c0 = sbase + col(wi0);
    const unsigned* c1 = sbase + col(min(wi0 + 1, last_word));
    const unsigned* c2 = sbase + col(min(wi0 + 2, last_word));
    const int o0 = x[0] - 4 * wi0, o1 = x[1] - 4 * wi0;
    const int o2 = x[2] - 4 * wi0 - (hi ? 4 : 0), o3 = x[3] - 4 * wi0 - (hi ? 4 : 0);
    const unsigned sel01 = (unsigned)o0 | ((unsigned)(o0 + 1) << 4) | ((unsigned)o1 << 8) | ((unsigned)(o1 + 1) << 12);
    const unsigned sel23 = (unsigned)o2 | ((unsigned)(o2 + 1) << 4) | ((unsigned)o3 << 8) | ((unsigned)(o3 + 1) << 12);
    uint8_t* dcol = dbase + 4 * t;
    unsigned pv0 = 0, pv1 = 0, pv2 = 0, pv3 = 0;  // T >> 4 of the previous source row
    auto emit = [&](unsigned k, unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned h0, unsigned h1, unsigned h2,
                    unsigned h3) {
      const uint2 bw = s_b[k];
      const unsigned v0 = (__umulhi(bw.x, a0) + __umulhi(bw.y, h0) + 2u) >> 2;
      const unsigned v1 = (__umulhi(bw.x, a1) + __umulhi(bw.y, h1) + 2u) >> 2;
      const unsigned v2 = (__umulhi(bw.x, a2) + __umulhi(bw.y, h2) + 2u) >> 2;
      const unsigned v3 = (__umulhi(bw.x, a3) + __umulhi(bw.y, h3) + 2u) >> 2;
      *reinterpret_cast<unsigned*>(dcol + (size_t)k * dst.pitch) = v0 | (v1 << 8) | (v2 << 16) | (v3 << 24);
    };
#pragma unroll 1
    for (int j0 = 0; j0 < ns; j0 += kRzBatch) {
      unsigned w[kRzBatch][3], ev[kRzBatch];
#pragma unroll
      for (int k = 0; k < kRzBatch; ++k) {
        const uint2 sr = s_src[j0 + k];  // rows past the band repeat the last source row (never emitted)
        w[k][0] = c0[sr.x]; w[k][1] = c1[sr.x]; w[k][2] = c2[sr.x];
        ev[k] = sr.y;
      }
#pragma unroll
      for (int k = 0; k < kRzBatch; ++k) {
        const unsigned pa = __byte_perm(w[k][0], w[k][1], sel01);
        const unsigned pb = __byte_perm(hi ? w[k][1] : w[k][0], hi ? w[k][2] : w[k][1], sel23);
        const unsigned h0 = __dp2a_lo(wt[0], pa, 0u) >> 4, h1 = __dp2a_hi(wt[1], pa, 0u) >> 4;
        const unsigned h2 = __dp2a_lo(wt[2], pb, 0u) >> 4, h3 = __dp2a_hi(wt[3], pb, 0u) >> 4;
        const unsigned e1 = ev[k] & 0xffu, e2 = ev[k] >> 8;
        if (j0 + k < ns) {  // uniform: the emit lists are per block
          if (e1 != 0xffu) emit(e1, pv0, pv1, pv2, pv3, h0, h1, h2, h3);
          if (e2 != 0xffu) emit(e2, h0, h1, h2, h3, h0, h1, h2, h3);
        }
        pv0 = h0; pv1 = h1; pv2 = h2; pv3 = h3;
      }
    }
  }
}

int launch_resize_level(const Plan& p, int level, int nframes, cudaStream_t s) {
  const LevelGeom& d = p.lv[level];
  // selectable form (orb_gpu_set_resize_variant(1), not the default: measured slower): horizontal pass on
  // the tensor cores for batches of 8 frames and more when the level pair qualifies
  if (resize_variant() == 1 && nframes >= 8 && resize_tc_level_ok(p.lv[level - 1], d, nframes))
    return launch_resize_tc(p.lv[level - 1], d, nframes, s);
  static const bool two_col = [] { const char* e = getenv("ORB_GPU_RESIZE2"); return e && atoi(e); }();
  dim3 grid((d.h + kRzRows - 1) / kRzRows, nframes);
  // a few frames: bands of 4 rows instead of 16, so that one frame's level still spreads over the SMs
  const int rows4 = (long long)grid.x * nframes >= 2 * 148 ? kRzRows : kRzBatch;
  // the four-column form needs the three-word tap window (scaleFactor <= 2) and 4-byte aligned destination rows
  const bool four = !two_col && (long long)p.lv[level - 1].w <= 2LL * d.w + 1 && (d.pitch & 3) == 0 &&
                    (d.frame_stride & 3) == 0;
  if (four) {
    const int nt = (d.w + 3) >> 2;
    grid.x = (d.h + rows4 - 1) / rows4;
    const LevelGeom& sl = p.lv[level - 1];
    // TMA-staged source band: full bands only, 16-byte aligned rows, frames stacked without gaps, <= 5 column chunks
    static const bool no_stage = [] { const char* e = getenv("ORB_GPU_RESIZE_DIRECT"); return e && atoi(e); }();
    const int nchunks = (sl.pitch + kRzChunk - 1) / kRzChunk;
    const bool staged = !no_stage && rows4 == kRzRows && tma_available() && nchunks <= kRzMaxChunks &&
                        (reinterpret_cast<uintptr_t>(sl.img) & 15) == 0 && (sl.pitch & 15) == 0 &&
                        (nframes == 1 || sl.frame_stride == (long long)sl.pitch * sl.h);
    RzStage S;
    memset(&S, 0, sizeof(S));
    if (staged) {
      // source rows of a band: sy(dy0 + 15) - sy(dy0) <= floor(15 * scale) + 1, plus the lower tap row, plus one spare
      const int box_h = std::min(kRzSrcRows, (int)(15.0 * (double)sl.h / (double)d.h) + 4);
      int rc = tensor_map_2d_u8(&S.map, sl.img, sl.pitch, (long long)nframes * sl.h, sl.pitch, kRzChunk, box_h, false);
      if (rc) return rc;
      S.nchunks = nchunks; S.src_h = sl.h; S.box_h = box_h;
      const int smem = nchunks * box_h * kRzChunk;
      if (smem > 48 * 1024)
        ORB_CUDA_TRY(cudaFuncSetAttribute(resize_level4_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      resize_level4_kernel<true><<<grid, min(256, (nt + 31) & ~31), smem, s>>>(sl, d, rows4, S);
    } else {
      resize_level4_kernel<false><<<grid, min(256, (nt + 31) & ~31), 0, s>>>(sl, d, rows4, S);
    }
  } else {
    const int nt = (d.w + 1) >> 1;
    resize_level_kernel<<<grid, min(512, (nt + 31) & ~31), 0, s>>>(p.lv[level - 1], d);
  }
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

// Level-0 repack for frames whose rows are not 4-byte aligned (e.g. KITTI's 1241-byte rows): every kernel below
// reads rows as aligned 32-bit words, so such frames are copied once into the handle's pitched level-0 buffer.
// One warp per row, aligned word loads, the byte shift of the row done with a funnel shift; one launch per batch
// (a cudaMemcpy2DAsync per frame cost 4 us each: 1.05 ms for 256 KITTI frames).
__global__ void __launch_bounds__(256)
repack_rows_kernel(const uint8_t* __restrict__ src, size_t stride, size_t frame_stride, int width, int height,
                   LevelGeom dst) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= height) return;
  const uint8_t* sp = src + (size_t)blockIdx.y * frame_stride + (size_t)row * stride;
  const unsigned sh = (unsigned)(reinterpret_cast<uintptr_t>(sp) & 3) * 8;
  const unsigned* sw = reinterpret_cast<const unsigned*>(sp - (sh >> 3));
  unsigned* dw = reinterpret_cast<unsigned*>(dst.img + (size_t)blockIdx.y * dst.frame_stride + (size_t)row * dst.pitch);
  const int nw = (width + 3) >> 2;                     // destination words that hold pixels
  const int last = (int)((sh / 8 + width - 1) >> 2);   // last source word that holds a pixel of this row
  for (int j = lane; j < nw; j += 32) {
    const unsigned lo = sw[j], hi = sw[min(j + 1, last)];
    dw[j] = __funnelshift_r(lo, hi, sh);
  }
}

int launch_repack_rows(const uint8_t* d_frames, size_t stride, size_t frame_stride, int width, int height,
                       const LevelGeom& l0, int nframes, cudaStream_t s) {
  dim3 grid((height + 7) / 8, nframes);
  repack_rows_kernel<<<grid, 256, 0, s>>>(d_frames, stride, frame_stride, width, height, l0);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

// =============================================================================================
// 2. FAST-9/16 per cell, 3x3 NMS inside the cell, threshold retry
// =============================================================================================
// One warp per 30-px cell (:789-829).  score(p) = max over the 16 contiguous 9-arcs of
// max(min(d), -max(d)), d_k = I(p) - I(circle_k); p is a corner at t iff score > t and
// response = score - 1 (the largest threshold at which p is still a corner).
// Per threshold T (iniTh first, minTh only if the cell kept nothing, :809-816):
//   1. word-wide necessary test on 4 pixels at a time: a 9-arc contains one pixel of each antipodal
//      pair, so both (|v-c0|>T or |v-c8|>T) and (|v-c4|>T or |v-c12|>T) must hold (VABSDIFF4 +
//      SWAR compares).  Every lane scans a run of consecutive words and keeps the 4-bit results in
//      registers; one warp scan turns them into a row-major candidate list.
//   2. exact arc score of the candidates, two signs at once in s16x2 lanes (VIMNMX3.S16x2);
//   3. strict 3x3 NMS over the candidates, emitted in row-major order (cv::FAST's order).
constexpr int kFastWarps = 8;

__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}

// exact FAST score of the pixel at t (tile pitch in bytes): lo lane = v - c + 256, hi lane = c - v + 256
__device__ __forceinline__ int fast_score(const uint8_t* t, int pitch) {
  const unsigned vb = (unsigned)t[0] + 256u + ((256u - (unsigned)t[0]) << 16);
  unsigned e[16];
#define ORB_E(k, off) e[k] = (unsigned)t[off] * 0xffffu + vb
  ORB_E(0, 3 * pitch);       ORB_E(1, 3 * pitch + 1);   ORB_E(2, 2 * pitch + 2);   ORB_E(3, pitch + 3);
  ORB_E(4, 3);               ORB_E(5, -pitch + 3);      ORB_E(6, -2 * pitch + 2);  ORB_E(7, -3 * pitch + 1);
  ORB_E(8, -3 * pitch);      ORB_E(9, -3 * pitch - 1);  ORB_E(10, -2 * pitch - 2); ORB_E(11, -pitch - 3);
  ORB_E(12, -3);             ORB_E(13, pitch - 3);      ORB_E(14, 2 * pitch - 2);  ORB_E(15, 3 * pitch - 1);
#undef ORB_E
  unsigned m3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) m3[k] = __vimin3_s16x2(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
  unsigned m9[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) m9[k] = __vimin3_s16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
  unsigned a = __vimax3_s16x2(m9[0], m9[1], m9[2]), b = __vimax3_s16x2(m9[3], m9[4], m9[5]);
  unsigned c = __vimax3_s16x2(m9[6], m9[7], m9[8]), d = __vimax3_s16x2(m9[9], m9[10], m9[11]);
  unsigned f = __vimax3_s16x2(m9[12], m9[13], m9[14]);
  a = __vimax3_s16x2(a, b, c);
  d = __vimax3_s16x2(d, f, m9[15]);
  a = __vmaxs2(a, d);
  return max((int)(a & 0xffffu), (int)(a >> 16)) - 256;
}

// sign bits of (x > T) per byte for T < 127: x > T  <=>  x >= 128 or (x & 127) + (127 - T) >= 128
__device__ __forceinline__ unsigned gt_bytes(unsigned x, unsigned k127mT) {
  return ((x & 0x7f7f7f7fu) + k127mT) | x;
}

struct FastLayout { int tp4, tile_bytes, score_bytes, cand_cap, per_warp; unsigned inv_tp; };

__global__ void __launch_bounds__(kFastWarps * 32)
fast_cells_kernel(Plan p, Scratch sc, FastLayout L) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int lane = threadIdx.x & 31;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // provably warp-uniform: per-cell values stay in uniform registers
  const int cell = blockIdx.x * (blockDim.x >> 5) + warp;
  if (cell >= p.cells_per_frame) return;
  const int f = blockIdx.y;
  // cell ROI (:789-806), precomputed on the host: x | y << 16, cw | ch << 8 | level << 16 (cw == 0: skipped cell)
  const uint2 cd = __ldg(sc.cell_tab + cell);
  const int ini_x = (int)(cd.x & 0xffffu), ini_y = (int)(cd.x >> 16);
  const int cw = (int)(cd.y & 0xffu), ch = (int)((cd.y >> 8) & 0xffu);
  const LevelGeom& g = p.lv[(cd.y >> 16) & 0xffu];

  const int tp4 = L.tp4, tp = tp4 * 4;
  uint8_t* tile = smem + (size_t)warp * L.per_warp;
  uint8_t* score = tile + L.tile_bytes;  // same pitch as the tile, one ring of zeros around the interior
  uint16_t* cand = reinterpret_cast<uint16_t*>(score + L.score_bytes);
  uint16_t* corner = cand;  // compacted in place: a round only overwrites entries it has already read

  const size_t cell_slot = (size_t)f * p.cells_per_frame + cell;
  uint32_t* out = sc.cell_kp + cell_slot * p.cell_cap;
  if (cw == 0) {
    if (lane == 0) sc.cell_cnt[cell_slot] = 0;
    return;
  }
  const int iw = cw - 6, ih = ch - 6;  // detectable interior

  // tile load with aligned 32-bit words (rows are 4-byte aligned: pitch % 4 == 0, base % 4 == 0)
  const uint8_t* img = g.img + (size_t)f * g.frame_stride + (size_t)ini_y * g.pitch + ini_x;
  const int off = (int)(reinterpret_cast<uintptr_t>(img) & 3);
  uint32_t* tile32 = reinterpret_cast<uint32_t*>(tile);
  {
    const int nw = (off + cw + 3) >> 2;
    if (nw <= 16) {  // two rows per pass (16 lanes each), three when a row is at most 10 words (30 of 32 lanes busy)
      const bool three = nw <= 10;
      const int r0 = three ? (lane * 205) >> 11 : lane >> 4;  // lane / 10 for lane < 32
      const int c = three ? lane - 10 * r0 : lane & 15;
      const int rpp = three ? 3 : 2;
      const uint8_t* src = img - off + (size_t)r0 * g.pitch + 4 * c;
      unsigned dst = (unsigned)__cvta_generic_to_shared(tile32 + r0 * tp4 + c);
      const int sstep = rpp * g.pitch;
      const unsigned dstep = rpp * tp;
      if (c < nw && r0 < rpp)
#pragma unroll 4
        for (int r = r0; r < ch; r += rpp, src += sstep, dst += dstep)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst), "l"(src));
    } else {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(img - off);
      const int wpitch = g.pitch >> 2;
      const unsigned inv_nw = ((1u << 20) + nw - 1) / nw;
      for (int k = lane; k < nw * ch; k += 32) {
        const int r = (int)(((unsigned)k * inv_nw) >> 20), c = k - r * nw;
        cp_async4(tile32 + r * tp4 + c, src + (size_t)r * wpitch + c);
      }
    }
    uint4* z = reinterpret_cast<uint4*>(score);  // 16-byte aligned; zeroing a little past (ih+2) rows is harmless
    for (int k = lane; k < (tp4 * (ih + 2) + 3) >> 2; k += 32) z[k] = make_uint4(0, 0, 0, 0);
    cp_async_commit_wait_all();
  }
  __syncwarp();
  const int xs = off + 3;                         // tile byte column of interior pixel x = 0
  const uint8_t* t0 = tile + 3 * tp + xs;         // interior pixel (0,0); candidates are offsets y*tp + x from here
  uint8_t* s0 = score + tp + 1;                   // score of interior pixel (0,0)
  const int gw0 = xs >> 2;                        // first word group that holds interior pixels
  const int ng = ((xs + iw - 1) >> 2) - gw0 + 1;  // word groups per interior row
  const int lead = xs & 3;                        // bytes of the first group left of the interior
  const int tail = (xs + iw) - 4 * (gw0 + ng - 1);  // interior bytes of the last group (1..4)
  const int S = (ng + 15) >> 4;                   // row segments (1 unless the cell is wider than 64 px)
  const int seg_len = S == 1 ? ng : (ng + S - 1) / S;
  const int units = ih * S;

  int n_kept = 0;
  int T = p.ini_th;
  for (int pass = 0; pass < 2; ++pass) {
    const unsigned kT = (unsigned)(127 - min(T, 126)) * 0x01010101u;
    // ---- 1. necessary test, 4 pixels per word; a lane walks one row (segment) with a sliding window ----
    int ncand = 0;
    for (int u0 = 0; u0 < units; u0 += 32) {
      const int u = u0 + lane;
      const int uu = min(u, units - 1);  // idle lanes shadow a valid unit, their result is dropped
      int row = uu, seg = 0;
      if (S > 1) { row = uu / S; seg = uu - row * S; }
      const int g0 = seg * seg_len, len = min(seg_len, ng - g0);
      const uint32_t* rp = tile32 + (row + 3) * tp4 + gw0 + g0;
      const uint32_t* rup = rp - 3 * tp4;
      const uint32_t* rdn = rp + 3 * tp4;
      unsigned wm = rp[-1], wc = rp[0];
      unsigned acc_lo = 0, acc_hi = 0;
      for (int it = 0; it < seg_len; ++it) {
        const unsigned wp = rp[it + 1], up = rup[it], dn = rdn[it];
        const unsigned lf = __funnelshift_r(wm, wc, 8), rt = __funnelshift_r(wc, wp, 24);
        const unsigned v = gt_bytes(__vabsdiffu4(wc, up), kT) | gt_bytes(__vabsdiffu4(wc, dn), kT);
        const unsigned h = gt_bytes(__vabsdiffu4(wc, lf), kT) | gt_bytes(__vabsdiffu4(wc, rt), kT);
        const unsigned nib = __umulhi(v & h & 0x80808080u, 0x02040810u);  // sign bits 7,15,23,31 -> bits 0..3 (+ junk above)
        acc_lo = __funnelshift_r(acc_lo, acc_hi, 4);
        acc_hi = __funnelshift_r(acc_hi, nib, 4);
        wm = wc; wc = wp;
      }
      // right-align (group k of the segment at nibble k) and drop everything outside the interior
      unsigned long long acc = (((unsigned long long)acc_hi << 32) | acc_lo) >> (4 * (16 - seg_len));
      const int first_bit = g0 == 0 ? lead : 0;
      const int end_bit = g0 + len == ng ? 4 * (len - 1) + tail : 4 * len;
      acc &= (end_bit >= 64 ? ~0ull : (1ull << end_bit) - 1) & ~((1ull << first_bit) - 1);
      if (u >= units) acc = 0;
      unsigned a_lo = (unsigned)acc, a_hi = (unsigned)(acc >> 32);
      const int mine = __popc(a_lo) + __popc(a_hi);
      int incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int vv = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += vv;
      }
      int pos = ncand + incl - mine;
      ncand += __shfl_sync(0xffffffffu, incl, 31);
      // row-major emission: ascending bit order inside the lane, lanes in unit order
      const int cbase = row * tp + 4 * (gw0 + g0) - xs;  // bit b of acc is the pixel at offset cbase + b
      while (a_lo) {
        const int b = __ffs(a_lo) - 1;
        a_lo &= a_lo - 1;
        cand[pos++] = (uint16_t)(cbase + b);
      }
      while (a_hi) {
        const int b = __ffs(a_hi) - 1;
        a_hi &= a_hi - 1;
        cand[pos++] = (uint16_t)(cbase + 32 + b);
      }
      __syncwarp();
    }
    // ---- 2. exact score of the candidates; corners (score > T) go to a second ordered list ----
    int ncorner = 0;
    for (int q0 = 0; q0 < ncand; q0 += 32) {
      const int q = q0 + lane;
      int c = 0, s = 0;
      if (q < ncand) {
        c = cand[q];
        s = fast_score(t0 + c, tp);
      }
      __syncwarp();  // all reads of this round precede the in-place writes below
      const bool is_corner = s > T;
      const unsigned mk = __ballot_sync(0xffffffffu, is_corner);
      if (is_corner) {
        s0[c] = (uint8_t)s;
        corner[ncorner + __popc(mk & ((1u << lane) - 1))] = (uint16_t)c;
      }
      ncorner += __popc(mk);
    }
    __syncwarp();
    // ---- 3. NMS + ordered emission ----
    for (int q0 = 0; q0 < ncorner; q0 += 32) {
      const int q = q0 + lane;
      bool keep = false;
      int s = 0, c = 0;
      if (q < ncorner) {
        c = corner[q];
        const uint8_t* cs = s0 + c;
        s = cs[0];
        const int mx = max(max(max(cs[-tp - 1], cs[-tp]), max(cs[-tp + 1], cs[-1])),
                           max(max(cs[1], cs[tp - 1]), max(cs[tp], cs[tp + 1])));
        keep = s > mx;
      }
      const unsigned mk = __ballot_sync(0xffffffffu, keep);
      if (keep) {  // ROI coordinates as cv::FAST reports them (x+3, y+3), response = score-1
        const int y = (int)(((unsigned)c * L.inv_tp) >> 20), x = c - y * tp;
        out[n_kept + __popc(mk & ((1u << lane) - 1))] =
            (uint32_t)(x + 3) | ((uint32_t)(y + 3) << 8) | ((uint32_t)(s - 1) << 16);
      }
      n_kept += __popc(mk);
    }
    if (n_kept > 0 || T == p.min_th) break;
    T = p.min_th;  // empty at iniThFAST: rerun the cell at minThFAST (:812-816)
    // scores written at iniTh (none survived NMS, but some may be set) must not leak into the rerun
    for (int q = lane; q < ncorner; q += 32) s0[corner[q]] = 0;
    __syncwarp();
  }
  if (lane == 0) sc.cell_cnt[cell_slot] = (uint32_t)n_kept;
}

static FastLayout fast_smem_layout(const Plan& p) {
  int mw = 0, mh = 0;
  for (int l = 0; l < p.nlevels; ++l) {
    mw = max(mw, p.lv[l].wcell + 6);
    mh = max(mh, p.lv[l].hcell + 6);
  }
  FastLayout L;
  L.tp4 = ((mw + 3 + 3) / 4 + 1) | 1;  // + alignment offset (<= 3), + one word right of the last group; odd:
                                        // lanes walk different rows, an odd word pitch keeps them on different banks
  L.tile_bytes = (L.tp4 * 4 * mh + 15) & ~15;
  L.score_bytes = (L.tp4 * 4 * (mh - 4) + 15) & ~15;
  int area = 0;
  for (int l = 0; l < p.nlevels; ++l) area = max(area, p.lv[l].wcell * p.lv[l].hcell);
  L.cand_cap = (area + 7) & ~7;
  L.per_warp = L.tile_bytes + L.score_bytes + 2 * L.cand_cap;  // tile + score + candidate list
  L.inv_tp = ((1u << 20) + L.tp4 * 4 - 1) / (L.tp4 * 4);
  return L;
}

int launch_fast_cells(const Plan& p, const Scratch& sc, int nframes, cudaStream_t s) {
  const FastLayout L = fast_smem_layout(p);
  const size_t smem = (size_t)kFastWarps * L.per_warp;
  if (smem > 227 * 1024) return ORB_GPU_ERR_CAPACITY;
  // the opt-in belongs to the current device's context: set it on every launch rather than cache it process-wide
  if (smem > 48 * 1024)
    ORB_CUDA_TRY(cudaFuncSetAttribute(fast_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // 2 cells per block: a block's slot is held until its slowest cell is done and cells differ a lot in work, so small
  // blocks keep more warps resident (measured on 256 TUM frames: 8 warps 0.431 ms, 4: 0.418, 2: 0.414, 1: 0.416), and
  // the cells of a single frame still fill the SMs.  ORB_GPU_FAST_WPB overrides (1..8).
  static const int env_wpb = [] { const char* e = getenv("ORB_GPU_FAST_WPB"); return e ? atoi(e) : 0; }();
  const int wpb = env_wpb > 0 ? min(env_wpb, kFastWarps) : 2;
  const dim3 grid((p.cells_per_frame + wpb - 1) / wpb, nframes);
  fast_cells_kernel<<<grid, wpb * 32, smem * wpb / kFastWarps, s>>>(p, sc, L);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

// =============================================================================================
// 3. quadtree distribution
// =============================================================================================
// The reference splits std::list nodes 4 ways until it holds >= N nodes, then keeps the best
// response per node (:539-763).  Node boxes depend only on the path of quadrant choices from a
// root, so every candidate gets a PATH CODE (root id, then 2 bits per depth: bit0 = right of
// the split column, bit1 = below the split row, i.e. children n1..n4 = 0..3).  After sorting the
// keys by (code, arrival index) every node is a contiguous key range and a split is three
// binary searches.  The list itself (order, creation sequence, the "careful" last rounds that
// split the largest nodes first and stop at N) is then simulated on node records only, with
// scans instead of pointer chasing.  Equal-size ties use the creation sequence.
constexpr int kQtThreads = 256;
constexpr int kQtSmemKeys = 4096;   // keys sorted in shared memory up to this many, else global
constexpr int kQtBins = 4096;       // buckets of the shared-memory sort
constexpr int kQtKpt = kQtSmemKeys / kQtThreads;  // candidates per thread on the shared-memory path

struct QtNode { int start, end; uint32_t prefix; int depth_new; int seq; };  // depth | isnew<<8

__device__ __forceinline__ int block_exclusive_scan(int* a, int n, int* s_warp /*>=33 ints*/) {
  // in-place exclusive scan of a[0..n) by the whole block; returns the total
  constexpr int T = kQtThreads;  // compile-time block size: the chunk divisions below become shifts
  const int t = threadIdx.x;
  const int chunk = (n + T - 1) / T;
  const int lo = min(t * chunk, n), hi = min(lo + chunk, n);
  int sum = 0;
  for (int k = lo; k < hi; ++k) sum += a[k];
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((t & 31) >= o) incl += v;
  }
  if ((t & 31) == 31) s_warp[t >> 5] = incl;
  __syncthreads();
  if (t < 32) {
    const int nw = T >> 5;
    int w = t < nw ? s_warp[t] : 0, wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, wi, o);
      if (t >= o) wi += v;
    }
    if (t < nw) s_warp[t] = wi - w;
    if (t == nw - 1) s_warp[32] = wi;
  }
  __syncthreads();
  int run = s_warp[t >> 5] + incl - sum;
  for (int k = lo; k < hi; ++k) { const int v = a[k]; a[k] = run; run += v; }
  const int total = s_warp[32];
  __syncthreads();
  return total;
}

// the same for two arrays at once (one set of barriers): returns total of a, *total_b that of b
__device__ __forceinline__ int block_exclusive_scan2(int* a, int* b, int n, int* s_warp /*>=66 ints*/, int* total_b) {
  constexpr int T = kQtThreads;  // compile-time block size: the chunk divisions below become shifts
  const int t = threadIdx.x;
  const int chunk = (n + T - 1) / T;
  const int lo = min(t * chunk, n), hi = min(lo + chunk, n);
  int sa = 0, sb = 0;
  for (int k = lo; k < hi; ++k) { sa += a[k]; sb += b[k]; }
  int ia = sa, ib = sb;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int va = __shfl_up_sync(0xffffffffu, ia, o), vb = __shfl_up_sync(0xffffffffu, ib, o);
    if ((t & 31) >= o) { ia += va; ib += vb; }
  }
  if ((t & 31) == 31) { s_warp[t >> 5] = ia; s_warp[33 + (t >> 5)] = ib; }
  __syncthreads();
  if (t < 32) {
    const int nw = T >> 5;
    const int wa = t < nw ? s_warp[t] : 0, wb = t < nw ? s_warp[33 + t] : 0;
    int xa = wa, xb = wb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int va = __shfl_up_sync(0xffffffffu, xa, o), vb = __shfl_up_sync(0xffffffffu, xb, o);
      if (t >= o) { xa += va; xb += vb; }
    }
    if (t < nw) { s_warp[t] = xa - wa; s_warp[33 + t] = xb - wb; }
    if (t == nw - 1) { s_warp[32] = xa; s_warp[65] = xb; }
  }
  __syncthreads();
  int ra = s_warp[t >> 5] + ia - sa, rb = s_warp[33 + (t >> 5)] + ib - sb;
  for (int k = lo; k < hi; ++k) {
    const int va = a[k], vb = b[k];
    a[k] = ra; b[k] = rb;
    ra += va; rb += vb;
  }
  const int total = s_warp[32];
  *total_b = s_warp[65];
  __syncthreads();
  return total;
}

__device__ __forceinline__ void bitonic_step(uint64_t* a, int i, int j, int k) {
  const int l = i | j;
  const uint64_t x = a[i], y = a[l];
  const bool up = (i & k) == 0;
  if ((x > y) == up) { a[i] = y; a[l] = x; }
}

// Bitonic sort of a[0..n) (n a power of two) by the whole block.  Every warp owns an aligned chunk
// of n / nwarps elements: sub-stages whose partner distance stays inside a chunk only need a
// __syncwarp, so only the few widest sub-stages pay for a block barrier.
__device__ __forceinline__ void block_bitonic_sort(uint64_t* a, int n /*pow2*/) {
  constexpr int nwarps = kQtThreads >> 5;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int chunk = n / nwarps;  // nwarps is a power of two
  if (chunk < 2) {               // tiny input: plain block-wide network
    for (int k = 2; k <= n; k <<= 1)
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = threadIdx.x; t < (n >> 1); t += kQtThreads)
          bitonic_step(a, ((t & ~(j - 1)) << 1) | (t & (j - 1)), j, k);
        __syncthreads();
      }
    return;
  }
  const int half = chunk >> 1, base = warp * chunk;
  for (int k = 2; k <= n; k <<= 1) {
    int j = k >> 1;
    for (; 2 * j > chunk; j >>= 1) {  // partners in different chunks
      for (int t = threadIdx.x; t < (n >> 1); t += kQtThreads)
        bitonic_step(a, ((t & ~(j - 1)) << 1) | (t & (j - 1)), j, k);
      __syncthreads();
    }
    for (; j > 0; j >>= 1) {          // warp-local sub-stages
      for (int t = lane; t < half; t += 32)
        bitonic_step(a, base + (((t & ~(j - 1)) << 1) | (t & (j - 1))), j, k);
      __syncwarp();
    }
    if (2 * k > chunk) __syncthreads();  // the next k starts with (or ends the sort before) a cross-chunk read
  }
}

__device__ __forceinline__ int lower_bound_code(const uint64_t* keys, int lo, int hi, uint32_t code) {
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((uint32_t)(keys[mid] >> 32) < code) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// child boundaries of a node: b[0]=start, b[1..3] = first key of child 1..3, b[4]=end
__device__ __forceinline__ void child_bounds(const uint64_t* keys, const QtNode& nd, int b[5]) {
  const int depth = nd.depth_new & 0xff;
  const int shift = 24 - 2 * (depth + 1);
  b[0] = nd.start; b[4] = nd.end;
  b[2] = lower_bound_code(keys, nd.start, nd.end, nd.prefix + (2u << shift));
  b[1] = lower_bound_code(keys, nd.start, b[2], nd.prefix + (1u << shift));
  b[3] = lower_bound_code(keys, b[2], nd.end, nd.prefix + (3u << shift));
}

__host__ __device__ __forceinline__ int qt_ncap(int quota, int n_ini) {
  const int a = quota + 4, b = 4 * n_ini;
  return (a > b ? a : b) + 4;
}
__device__ __forceinline__ int pow2ceil_dev(int v) { return v <= 1 ? 1 : 1 << (32 - __clz(v - 1)); }

// Shared-memory layout of quadtree_kernel (bytes from the start of dynamic shared memory), the same on both sides:
//   A  [kQtSmemKeys] u64 sorted keys.  Until the keys are written (the last step of the sort) A holds the packed bucket
//      counters (s_bin) and, when they fit behind them, the per-cell arrival bases (s_cell): both are dead by then.
//   B  [kQtSmemKeys] u32 bucketed entries (s_bkt) during the sort; afterwards the node lists, the rank-ordered scratch
//      and the child boundaries (everything the list simulation needs) overlay it -- B is as large as the larger user.
//   C  72 ints of warp partials (+ s_cell when a level has too many cells for A).
// 75 KB of separate arrays (3 blocks per SM) became 51 KB (4 per SM) on the TUM configuration.
struct QtLayout { int off_b, off_warp, off_cell, total; };
__host__ __device__ __forceinline__ QtLayout qt_layout(int ncap, int sort_cap, int ncells) {
  QtLayout L;
  const int a_bytes = kQtSmemKeys * 8, bin_bytes = (kQtBins / 2 + 4) * 4, cell_bytes = (ncells + 1) * 4;
  const int overlay = 8 * sort_cap + (int)sizeof(int) * 19 * ncap;  // s_sort, 2 lists of 5 ints, nch/cb/sv, bnd[3], cbnd[3]
  int b_bytes = kQtSmemKeys * 4 > overlay ? kQtSmemKeys * 4 : overlay;
  b_bytes = (b_bytes + 15) & ~15;
  L.off_b = a_bytes;
  L.off_warp = a_bytes + b_bytes;
  const bool cell_in_a = bin_bytes + cell_bytes <= a_bytes;
  L.off_cell = cell_in_a ? bin_bytes : L.off_warp + 72 * 4;
  L.total = L.off_warp + 72 * 4 + (cell_in_a ? 0 : cell_bytes) + 16;
  return L;
}
static_assert(sizeof(int) * 5 == 20, "QtNode is five ints");

__global__ void __launch_bounds__(kQtThreads, 4)
quadtree_kernel(Plan p, Scratch sc, int nframes) {
  extern __shared__ __align__(16) uint8_t smem[];
  // one block per (level, frame); the big levels are scheduled first so the small ones fill the tail
  const int level = blockIdx.x / nframes;
  const LevelGeom& g = p.lv[level];
  const int f = blockIdx.x - level * nframes, t = threadIdx.x, T = kQtThreads;
  const int ncap = qt_ncap(g.quota, g.n_ini), sort_cap = pow2ceil_dev(ncap);
  const int ncells = g.nrows * g.ncols;
  const int N = g.quota;

  // shared memory carve-up (qt_layout)
  const QtLayout lay = qt_layout(ncap, sort_cap, ncells);
  uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem);                    // [kQtSmemKeys]
  uint32_t* s_bin = reinterpret_cast<uint32_t*>(smem);                     // [kQtBins / 2 + 4] packed u16 bucket counters / bases, in A until the keys are written
  int* s_cell = reinterpret_cast<int*>(smem + lay.off_cell);              // [ncells+1], in A behind s_bin when it fits
  uint32_t* s_bkt = reinterpret_cast<uint32_t*>(smem + lay.off_b);         // [kQtSmemKeys] bucketed (rest-of-code, arrival)
  uint64_t* s_sort = reinterpret_cast<uint64_t*>(smem + lay.off_b);        // [sort_cap]   -- from here on: after the sort
  QtNode* list_a = reinterpret_cast<QtNode*>(s_sort + sort_cap);          // [ncap]
  QtNode* list_b = list_a + ncap;                                          // [ncap]
  int* s_nch = reinterpret_cast<int*>(list_b + ncap);                      // [ncap] children per split node
  int* s_cb = s_nch + ncap;                                                // [ncap] creation base
  int* s_sv = s_cb + ncap;                                                 // [ncap] survivor flags/ranks
  int* s_bnd = s_sv + ncap;                                                // [3*ncap] rank-ordered scratch
  int* s_cbnd = s_bnd + 3 * ncap;                                          // [3*ncap] child boundaries found while counting
  int* s_warp = reinterpret_cast<int*>(smem + lay.off_warp);              // [72]
  __shared__ int s_misc[8];

  const uint32_t* __restrict__ cell_cnt = sc.cell_cnt + (size_t)f * p.cells_per_frame + g.cell_base;
  const uint32_t* __restrict__ cell_kp = sc.cell_kp + ((size_t)f * p.cells_per_frame + g.cell_base) * p.cell_cap;
  uint32_t* __restrict__ pay = sc.pay + (size_t)f * p.keys_per_frame + g.key_base;
  uint32_t* __restrict__ sel = sc.sel + (size_t)f * p.slots_per_frame + g.slot_base;

  // ---- gather candidates in arrival order (cell row-major, then FAST row-major) -------------
  for (int c = t; c < ncells; c += T) s_cell[c] = (int)cell_cnt[c];
  if (t == 0) s_cell[ncells] = 0;
  __syncthreads();
  const int C = block_exclusive_scan(s_cell, ncells, s_warp);
  if (t == 0) { sc.cand_cnt[f * p.nlevels + level] = C; s_cell[ncells] = C; }
  if (C > g.key_cap) {  // uniform branch
    if (t == 0) { *sc.status = ORB_GPU_ERR_CAPACITY; sc.sel_cnt[f * p.nlevels + level] = 0; }
    return;
  }
  int cpad = 1;
  while (cpad < C) cpad <<= 1;
  uint64_t* keys = (cpad <= kQtSmemKeys) ? s_keys
                                         : sc.keys + (size_t)f * p.keys_per_frame + g.key_base;
  __syncthreads();
  // One thread per candidate: its cell is the last one whose arrival base is <= the arrival index.
  // Candidates are distinct pixels, so their path codes are distinct: the order by code is found with a
  // bucket sort instead of a comparison network.  Bucket = root and the first d quadrant digits
  // (n_ini * 4^d <= kQtBins buckets); rank inside the bucket by counting smaller (rest-of-code, arrival).
  const bool in_smem = cpad <= kQtSmemKeys;
  int bdepth = 0;
  while ((g.n_ini << (2 * (bdepth + 1))) <= kQtBins && bdepth + 1 <= kCodeDepth) ++bdepth;
  const int bshift = 24 - 2 * bdepth;
  const int nbins = g.n_ini << (2 * bdepth);
  // shared-memory path: the owner cell of every arrival index, written cell by cell into the bucket array (idle until
  // the scatter below) -- one load per candidate instead of a 9-step binary search over the cell bases
  uint16_t* s_own = reinterpret_cast<uint16_t*>(s_bkt);
  if (in_smem) {
    for (int k = t; k < kQtBins / 2 + 1; k += T) s_bin[k] = 0;
    for (int c = t; c < ncells; c += T) {
      const int ke = s_cell[c + 1];
      for (int k = s_cell[c]; k < ke; ++k) s_own[k] = (uint16_t)c;
    }
    __syncthreads();
  }
  uint32_t my_code[kQtKpt], my_rank[kQtKpt];
#pragma unroll
  for (int q = 0; q < kQtKpt; ++q) {
    const int arrival = t + q * T;
    my_code[q] = 0; my_rank[q] = 0;
    if (in_smem ? arrival < C : q == 0) {
      for (int arr = arrival; arr < C; arr += in_smem ? C : T) {  // shared-memory path: exactly one candidate per slot
        int c;
        if (in_smem) {
          c = s_own[arr];
        } else {
          int lo = 0, hi = ncells - 1;
          while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s_cell[mid] <= arr) lo = mid; else hi = mid - 1;
          }
          c = lo;
        }
        // the cell's origin comes from the FAST cell table (ini = kBorder + index * cell size): no division by ncols
        const unsigned cxy = __ldg(&sc.cell_tab[g.cell_base + c].x);
        const uint32_t e = __ldg(&cell_kp[(size_t)c * p.cell_cap + (arr - s_cell[c])]);  // written by the FAST kernel: read-only here
        const int resp = (e >> 16) & 0xff;
        const int x = (int)(e & 0xff) + (int)(cxy & 0xffffu) - kBorder, y = (int)((e >> 8) & 0xff) + (int)(cxy >> 16) - kBorder;
        // root: vpIniNodes[kp.pt.x/hX] (:569), then 12 quadrant digits (DivideNode :481-525)
        int root = (int)__fdiv_rn((float)x, g.hx);
        root = min(root, g.n_ini - 1);
        int ulx = (int)__fmul_rn(g.hx, (float)root), brx = (int)__fmul_rn(g.hx, (float)(root + 1));
        int uly = 0, bry = g.box_h;
        uint32_t code = (uint32_t)root << 24;
#pragma unroll
        for (int d = 0; d < kCodeDepth; ++d) {
          const int mx = ulx + ((brx - ulx + 1) >> 1), my = uly + ((bry - uly + 1) >> 1);
          const int right = x >= mx, below = y >= my;
          code |= (uint32_t)(right | (below << 1)) << (22 - 2 * d);
          if (right) ulx = mx; else brx = mx;
          if (below) uly = my; else bry = my;
        }
        pay[arr] = (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)resp << 24);
        if (in_smem) {
          const uint32_t bin = code >> bshift;
          const uint32_t old = atomicAdd(&s_bin[bin >> 1], (bin & 1) ? 0x10000u : 1u);
          my_code[q] = code;
          my_rank[q] = ((bin & 1) ? old >> 16 : old & 0xffffu) | ((uint32_t)resp << 24);  // response rides along
        } else {
          keys[arr] = ((uint64_t)code << 32) | (uint32_t)arr | ((uint32_t)resp << 24);
        }
      }
    }
  }
  if (in_smem) {
    __syncthreads();
    // exclusive scan of the packed u16 counters -> bucket bases (entry nbins = C)
    {
      const int nwords = (nbins >> 1) + 1;
      const int chunk = (nwords + T - 1) / T;
      const int lo = min(t * chunk, nwords), hi = min(lo + chunk, nwords);
      int sum = 0;
      for (int k = lo; k < hi; ++k) { const uint32_t v = s_bin[k]; sum += (int)(v & 0xffffu) + (int)(v >> 16); }
      int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((t & 31) >= o) incl += v;
      }
      if ((t & 31) == 31) s_warp[t >> 5] = incl;
      __syncthreads();
      if (t < 32) {
        const int nw = T >> 5;
        int w = t < nw ? s_warp[t] : 0, wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int v = __shfl_up_sync(0xffffffffu, wi, o);
          if (t >= o) wi += v;
        }
        if (t < nw) s_warp[t] = wi - w;
      }
      __syncthreads();
      int run = s_warp[t >> 5] + incl - sum;
      for (int k = lo; k < hi; ++k) {
        const uint32_t v = s_bin[k];
        const int c0 = (int)(v & 0xffffu), c1 = (int)(v >> 16);
        s_bin[k] = (uint32_t)run | ((uint32_t)(run + c0) << 16);
        run += c0 + c1;
      }
    }
    __syncthreads();
    const uint32_t rem_mask = (1u << bshift) - 1u;
#pragma unroll
    for (int q = 0; q < kQtKpt; ++q) {
      const int arrival = t + q * T;
      if (arrival < C) {
        const uint32_t bin = my_code[q] >> bshift;
        const uint32_t base = (s_bin[bin >> 1] >> ((bin & 1) * 16)) & 0xffffu;
        s_bkt[base + (my_rank[q] & 0xffffu)] = ((my_code[q] & rem_mask) << 12) | (uint32_t)arrival;
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < kQtKpt; ++q) {
      const int arrival = t + q * T;
      if (arrival < C) {
        const uint32_t bin = my_code[q] >> bshift;
        const int bs = (int)((s_bin[bin >> 1] >> ((bin & 1) * 16)) & 0xffffu);
        const int be = (int)((s_bin[(bin + 1) >> 1] >> (((bin + 1) & 1) * 16)) & 0xffffu);
        const uint32_t mine = ((my_code[q] & rem_mask) << 12) | (uint32_t)arrival;
        int rank = 0;
        for (int k = bs; k < be; ++k) rank += s_bkt[k] < mine;
        my_rank[q] = (my_rank[q] & 0xff000000u) | (uint32_t)(bs + rank);  // final position (< kQtSmemKeys)
      }
    }
    __syncthreads();  // the keys overwrite s_bin (and s_cell): every read of them is behind this barrier
#pragma unroll
    for (int q = 0; q < kQtKpt; ++q) {
      const int arrival = t + q * T;
      // low word: arrival | response << 24 (codes are distinct, so the low word never decides the order)
      if (arrival < C)
        keys[my_rank[q] & 0xffffu] = ((uint64_t)my_code[q] << 32) | (uint32_t)arrival | (my_rank[q] & 0xff000000u);
    }
    __syncthreads();
  } else {
    for (int k = C + t; k < cpad; k += T) keys[k] = ~0ull;
    __syncthreads();
    block_bitonic_sort(keys, cpad);
  }

  // ---- initial list: non-empty roots (:543-585) --------------------------------------------
  QtNode* cur = list_a;
  QtNode* nxt = list_b;
  if (t == 0) {
    int n = 0;
    for (int r = 0; r < g.n_ini; ++r) {
      const int s = lower_bound_code(keys, 0, C, (uint32_t)r << 24);
      const int e = lower_bound_code(keys, s, C, (uint32_t)(r + 1) << 24);
      if (e > s) cur[n++] = QtNode{s, e, (uint32_t)r << 24, 0, r};
    }
    s_misc[0] = n;        // list size
    s_misc[1] = g.n_ini;  // next creation sequence number
  }
  __syncthreads();
  int n = s_misc[0];
  int seq_base = s_misc[1];
  bool careful = false;

  while (true) {
    const int prev_n = n;
    // -- choose the nodes to split and their creation bases --
    int K = 0, nsurv = 0;
    if (!careful) {
      for (int i = t; i < n; i += T) {
        const QtNode nd = cur[i];
        int nch = 0, sv = 1;
        if (nd.end - nd.start > 1 && (nd.depth_new & 0xff) < kCodeDepth) {
          int b[5];
          child_bounds(keys, nd, b);
          s_cbnd[3 * i] = b[1]; s_cbnd[3 * i + 1] = b[2]; s_cbnd[3 * i + 2] = b[3];  // reused when the node is split
          nch = (b[1] > b[0]) + (b[2] > b[1]) + (b[3] > b[2]) + (b[4] > b[3]);
          sv = 0;
        }
        s_nch[i] = nch; s_cb[i] = nch; s_sv[i] = sv;
      }
      __syncthreads();
      K = block_exclusive_scan2(s_cb, s_sv, n, s_warp, &nsurv);
    } else {
      // expandable nodes of the previous round, largest first, newest first among equals
      for (int i = t; i < n; i += T) { s_nch[i] = 0; s_cb[i] = 0; s_sv[i] = 1; }
      if (t == 0) s_misc[2] = 0;
      __syncthreads();
      for (int i = t; i < n; i += T) {
        const QtNode nd = cur[i];
        if ((nd.depth_new >> 8) && nd.end - nd.start > 1 && (nd.depth_new & 0xff) < kCodeDepth) {
          const int slot = atomicAdd(&s_misc[2], 1);
          const uint64_t key = ((uint64_t)(nd.end - nd.start) << 40) | ((uint64_t)nd.seq << 16) | (uint64_t)i;
          s_sort[slot] = ~key;  // ascending sort of ~key == descending (count, seq)
        }
      }
      __syncthreads();
      const int m = s_misc[2];
      int mpad = 1;
      while (mpad < m) mpad <<= 1;
      for (int k = m + t; k < mpad; k += T) s_sort[k] = ~0ull;
      __syncthreads();
      block_bitonic_sort(s_sort, mpad);
      // rank-ordered child counts; s_cb[j] (by rank) = children, s_sv reused later
      int* r_nch = s_bnd + 0;
      for (int j = t; j < m; j += T) {
        const int pos = (int)((~s_sort[j]) & 0xffff);
        int b[5];
        child_bounds(keys, cur[pos], b);
        s_cbnd[3 * pos] = b[1]; s_cbnd[3 * pos + 1] = b[2]; s_cbnd[3 * pos + 2] = b[3];
        r_nch[j] = (b[1] > b[0]) + (b[2] > b[1]) + (b[3] > b[2]) + (b[4] > b[3]);
      }
      __syncthreads();
      // inclusive growth scan -> first rank at which the list reaches N (:730-731)
      int* r_scan = s_bnd + ncap;
      for (int j = t; j < m; j += T) r_scan[j] = r_nch[j];
      __syncthreads();
      block_exclusive_scan(r_scan, m, s_warp);  // r_scan[j] = children created before rank j
      if (t == 0) s_misc[3] = m - 1;
      __syncthreads();
      for (int j = t; j < m; j += T) {
        const int size_after = n + (r_scan[j] + r_nch[j]) - (j + 1);
        if (size_after >= N) atomicMin(&s_misc[3], j);
      }
      __syncthreads();
      const int jstar = s_misc[3];
      int* r_pos = s_bnd + 2 * ncap;
      for (int j = t; j < m; j += T) r_pos[j] = (int)((~s_sort[j]) & 0xffff);
      __syncthreads();
      if (m > 0) K = r_scan[jstar] + r_nch[jstar];
      for (int j = t; j <= jstar && j < m; j += T) {
        const int pos = r_pos[j];
        s_nch[pos] = r_nch[j];
        s_cb[pos] = r_scan[j];
        s_sv[pos] = 0;
      }
      __syncthreads();
      nsurv = block_exclusive_scan(s_sv, n, s_warp);
    }
    // -- survivors keep their order behind the new children --
    if (t == 0) s_misc[4] = 0;
    __syncthreads();
    int n_exp_local = 0;
    for (int i = t; i < n; i += T) {
      const QtNode nd = cur[i];
      const bool split = (i + 1 < n ? s_sv[i + 1] : nsurv) == s_sv[i];  // flag was 0
      if (!split) {
        QtNode keep = nd;
        keep.depth_new &= 0xff;
        nxt[K + s_sv[i]] = keep;
      } else {
        const int b[5] = {nd.start, s_cbnd[3 * i], s_cbnd[3 * i + 1], s_cbnd[3 * i + 2], nd.end};
        const int depth = (nd.depth_new & 0xff) + 1;
        const int shift = 24 - 2 * depth;
        int m_idx = s_cb[i];
        for (int k = 0; k < 4; ++k) {
          if (b[k + 1] > b[k]) {
            QtNode ch{b[k], b[k + 1], nd.prefix + ((uint32_t)k << shift), depth | 0x100, seq_base + m_idx};
            nxt[K - 1 - m_idx] = ch;  // push_front: the last created child ends up first
            n_exp_local += (b[k + 1] - b[k]) > 1;
            ++m_idx;
          }
        }
      }
    }
    if (n_exp_local) atomicAdd(&s_misc[4], n_exp_local);
    __syncthreads();
    const int n_exp = s_misc[4];
    n = K + nsurv;
    seq_base += K;
    QtNode* tmp = cur; cur = nxt; nxt = tmp;
    __syncthreads();
    if (n >= N || n == prev_n) break;
    if (!careful && n + 3 * n_exp > N) careful = true;
  }

  // ---- best response per node, list order (:741-760) ----------------------------------------
  // Eight lanes per node: the final nodes hold ~C / N candidates each (18 on a TUM level 0), so a whole warp per node
  // left most lanes idle and paid five 64-bit shuffle steps per node; four nodes now share a warp's instructions.
  {
    const int grp = t >> 3, l8 = t & 7;
    for (int base = 0; base < n; base += T / 8) {  // uniform trip count: the shuffles below are executed by every lane
      const int i = base + grp;
      // max response, earliest arrival among equals: maximise (resp << 32) | ~arrival
      unsigned long long best = 0;
      if (i < n) {
        const QtNode nd = cur[i];
        for (int k = nd.start + l8; k < nd.end; k += 8) {
          const uint32_t low = (uint32_t)keys[k], arrival = low & 0xffffffu;
          const unsigned long long cand = ((unsigned long long)(low >> 24) << 32) | (uint32_t)(~arrival);
          best = cand > best ? cand : best;
        }
      }
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) {
        const unsigned long long v = __shfl_xor_sync(0xffffffffu, best, o);
        best = v > best ? v : best;
      }
      if (i < n && l8 == 0) s_sv[i] = (int)(~(uint32_t)best);  // arrival index of the node's keypoint
    }
  }
  // the payload gather in a pass of its own: one global load per node, all in flight at once (inside the loop above
  // each warp stalled on its load before it could store, ~27 dependent L2 round trips per warp)
  __syncthreads();
  for (int i = t; i < n; i += T) sel[i] = pay[s_sv[i]];
  if (t == 0) sc.sel_cnt[f * p.nlevels + level] = n;
}

static int pow2ceil(int v) { int p = 1; while (p < v) p <<= 1; return p; }

size_t quadtree_smem_bytes(const Plan& p, int level) {
  const LevelGeom& g = p.lv[level];
  const int ncap = qt_ncap(g.quota, g.n_ini);
  const int ncells = g.nrows * g.ncols;
  return (size_t)qt_layout(ncap, pow2ceil(ncap), ncells).total;
}

int launch_quadtree(const Plan& p, const Scratch& sc, int nframes, cudaStream_t s) {
  size_t smem = 0;
  for (int l = 0; l < p.nlevels; ++l) smem = max(smem, quadtree_smem_bytes(p, l));
  if (smem > 227 * 1024) return ORB_GPU_ERR_CAPACITY;
  if (smem > 48 * 1024)  // per device, see launch_fast_cells
    ORB_CUDA_TRY(cudaFuncSetAttribute(quadtree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  quadtree_kernel<<<nframes * p.nlevels, kQtThreads, smem, s>>>(p, sc, nframes);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

// =============================================================================================
// 4. GaussianBlur 7x7 sigma 2 (separable integer kernel, REFLECT_101 of the w x h image)
// =============================================================================================
// out = (sum_ij k_i k_j p + 32768) >> 16 with k = [18 34 48 56 48 34 18]; every intermediate is an exact
// integer, so the two 1-D passes commute: the vertical pass runs first, on packed bytes.
// One block = a band of kBlurR rows over the full width of one level of one frame; a thread owns
// 4 adjacent columns (one 32-bit word per row, coalesced loads straight from global memory).
//   vertical:   the word is split into even/odd bytes (2 x u16 lanes per register); with the 7-row
//               window in registers the 8.8 column sums are 3 IADD + 4 IMAD per register (sum <= 65280
//               never carries between the 16-bit lanes);
//   exchange:   (v0,v1),(v2,v3) u16 pairs go to a shared-memory row buffer (kBlurBatch rows per barrier
//               pair, double buffered); one small pass mirrors the 3 columns left and right (REFLECT_101);
//   horizontal: 4 IDP2A per pixel on the u16 pairs (16.16), +32768 folded into the accumulator, >>16,
//               one 32-bit store per thread and row.
constexpr int kBlurR = 32, kBlurBatch = 8;

__device__ __forceinline__ int reflect101(int p, int len) {
  // valid for -len < p < 2*len-1
  p = p < 0 ? -p : p;
  return p >= len ? 2 * (len - 1) - p : p;
}

__device__ __forceinline__ void bar_sync_n(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__global__ void __launch_bounds__(1024)
blur_kernel(Plan p, unsigned level_mask) {
  extern __shared__ __align__(16) uint16_t s_rows[];  // [2][kBlurBatch][pitch]
  int band = blockIdx.x;
  const int f = blockIdx.y;
  int level = 0;
  for (;; ++level) {  // the bands of the levels in level_mask, in level order
    if (!(level_mask >> level & 1u)) continue;
    const int nb = (p.lv[level].h + kBlurR - 1) / kBlurR;
    if (band < nb) break;
    band -= nb;
  }
  const LevelGeom& g = p.lv[level];
  const int t = threadIdx.x;
  const int nt = (g.w + 3) >> 2;             // threads that own columns of this level
  const int nactive = (nt + 31) & ~31;       // whole warps beyond leave before any barrier
  if (t >= nactive) return;
  const int pitch = 4 * (int)blockDim.x + 8; // u16 elements per row buffer: column x lives at index x + 4
  const int wi = min(t, nt - 1);
  const int y0 = band * kBlurR;
  const uint8_t* src = g.img + (size_t)f * g.frame_stride + 4 * wi;
  uint8_t* dst = g.blur + (size_t)f * g.bframe_stride + 4 * wi;
  const int h = g.h, w = g.w;
  auto load_row = [&](int yy) -> unsigned {  // word of input row yy (reflected; far rows of a partial band are clamped)
    const int sy = reflect101(min(yy, h + 2), h);
    return __ldg(reinterpret_cast<const unsigned*>(src + (size_t)sy * g.pitch));
  };

  unsigned e[7], o[7];
#pragma unroll
  for (int r = 0; r < 6; ++r) {
    const unsigned wd = load_row(y0 - 3 + r);
    e[r] = __byte_perm(wd, 0, 0x4240);  // (b0, 0, b2, 0)
    o[r] = __byte_perm(wd, 0, 0x4341);  // (b1, 0, b3, 0)
  }
  unsigned nxt[kBlurBatch];
#pragma unroll
  for (int k = 0; k < kBlurBatch; ++k) nxt[k] = load_row(y0 + 3 + k);

  // IDP2A weights: .lo uses bytes 0,1 and .hi bytes 2,3 of the second operand
  const unsigned W1 = (18u << 8) | (34u << 16) | (48u << 24);              // (0,k0) | (k1,k2)
  const unsigned W2 = 56u | (48u << 8) | (34u << 16) | (18u << 24);        // (k3,k4) | (k5,k6)
  const unsigned W3 = 18u | (34u << 8) | (48u << 16) | (56u << 24);        // (k0,k1) | (k2,k3)
  const unsigned W4 = 48u | (34u << 8) | (18u << 16);                      // (k4,k5) | (k6,0)

#pragma unroll
  for (int b = 0; b < kBlurR / kBlurBatch; ++b) {
    unsigned cur[kBlurBatch];
#pragma unroll
    for (int k = 0; k < kBlurBatch; ++k) cur[k] = nxt[k];
    if (b + 1 < kBlurR / kBlurBatch) {
#pragma unroll
      for (int k = 0; k < kBlurBatch; ++k) nxt[k] = load_row(y0 + 3 + kBlurBatch * (b + 1) + k);
    }
    uint16_t* buf = s_rows + (size_t)(b & 1) * kBlurBatch * pitch;
    // ---- vertical pass of kBlurBatch rows ----
#pragma unroll
    for (int k = 0; k < kBlurBatch; ++k) {
      const int r = kBlurBatch * b + k + 6;  // index of the newest input row of this output row
      e[r % 7] = __byte_perm(cur[k], 0, 0x4240);
      o[r % 7] = __byte_perm(cur[k], 0, 0x4341);
      const unsigned ve = 18u * (e[(r - 6) % 7] + e[r % 7]) + 34u * (e[(r - 5) % 7] + e[(r - 1) % 7]) +
                          48u * (e[(r - 4) % 7] + e[(r - 2) % 7]) + 56u * e[(r - 3) % 7];  // (v0, v2)
      const unsigned vo = 18u * (o[(r - 6) % 7] + o[r % 7]) + 34u * (o[(r - 5) % 7] + o[(r - 1) % 7]) +
                          48u * (o[(r - 4) % 7] + o[(r - 2) % 7]) + 56u * o[(r - 3) % 7];  // (v1, v3)
      uint2 pr;
      pr.x = __byte_perm(ve, vo, 0x5410);  // (v0, v1)
      pr.y = __byte_perm(ve, vo, 0x7632);  // (v2, v3)
      *reinterpret_cast<uint2*>(buf + k * pitch + 4 + 4 * t) = pr;
    }
    bar_sync_n(1, nactive);
    // ---- REFLECT_101 columns: x = -1-j <- 1+j and x = w+j <- w-2-j (j = 0..2) ----
    for (int q = t; q < kBlurBatch * 6; q += nactive) {
      const int k = q / 6, jj = q - 6 * k;
      uint16_t* row = buf + k * pitch + 4;
      if (jj < 3) row[-1 - jj] = row[1 + jj]; else row[w + jj - 3] = row[w + 1 - jj];
    }
    bar_sync_n(1, nactive);
    // ---- horizontal pass ----
    if (t < nt) {
#pragma unroll
      for (int k = 0; k < kBlurBatch; ++k) {
        const int y = y0 + kBlurBatch * b + k;
        if (y < h) {
          const uint2* rp = reinterpret_cast<const uint2*>(buf + k * pitch + 4 * t);
          const uint2 a = rp[0], c = rp[1], d = rp[2];  // (P0,P1) (P2,P3) (P4,P5): columns 4t-4 .. 4t+7
          unsigned o0 = __dp2a_lo(a.x, W1, 32768u); o0 = __dp2a_hi(a.y, W1, o0); o0 = __dp2a_lo(c.x, W2, o0); o0 = __dp2a_hi(c.y, W2, o0);
          unsigned o1 = __dp2a_lo(a.y, W3, 32768u); o1 = __dp2a_hi(c.x, W3, o1); o1 = __dp2a_lo(c.y, W4, o1); o1 = __dp2a_hi(d.x, W4, o1);
          unsigned o2 = __dp2a_lo(a.y, W1, 32768u); o2 = __dp2a_hi(c.x, W1, o2); o2 = __dp2a_lo(c.y, W2, o2); o2 = __dp2a_hi(d.x, W2, o2);
          unsigned o3 = __dp2a_lo(c.x, W3, 32768u); o3 = __dp2a_hi(c.y, W3, o3); o3 = __dp2a_lo(d.x, W4, o3); o3 = __dp2a_hi(d.y, W4, o3);
          const unsigned lo = __byte_perm(o0, o1, 0x0062);  // (o0.b2, o1.b2, -, -)
          const unsigned hi = __byte_perm(o2, o3, 0x0062);
          *reinterpret_cast<unsigned*>(dst + (size_t)y * g.bpitch) = __byte_perm(lo, hi, 0x5410);
        }
      }
    }
  }
}

// The tensor-core kernel (blur_tc_kernels.cu) takes every level whose buffers TMA can address; the byte-SIMD kernel
// above takes the rest (a caller's level-0 frames with rows that are not 16-byte aligned or with gaps between frames).
// ORB_GPU_BLUR=simd forces the byte-SIMD kernel for all levels (A/B measurements).
static bool blur_tc_enabled() {
  static const bool on = [] { const char* e = getenv("ORB_GPU_BLUR"); return !(e && !strcmp(e, "simd")); }();
  return on;
}
int launch_blur(const Plan& p, int nframes, cudaStream_t s) {
  const unsigned all = (p.nlevels >= 32) ? ~0u : ((1u << p.nlevels) - 1u);
  unsigned tc_mask = 0;
  if (blur_tc_enabled())
    for (int l = 0; l < p.nlevels; ++l)
      if (blur_tc_level_ok(p.lv[l], nframes)) tc_mask |= 1u << l;
  if (tc_mask) {
    const int rc = launch_blur_tc(p, nframes, tc_mask, s);
    if (rc) return rc;
  }
  const unsigned rest = all & ~tc_mask;
  if (!rest) return ORB_GPU_OK;
  int bands = 0, wmax = 0;
  for (int l = 0; l < p.nlevels; ++l)
    if (rest >> l & 1u) { bands += (p.lv[l].h + kBlurR - 1) / kBlurR; wmax = max(wmax, p.lv[l].w); }
  const int threads = (((wmax + 3) >> 2) + 31) & ~31;
  if (threads > 1024) return ORB_GPU_ERR_CAPACITY;
  const size_t smem = (size_t)2 * kBlurBatch * (4 * threads + 8) * sizeof(uint16_t);
  if (smem > 48 * 1024)  // per device, see launch_fast_cells
    ORB_CUDA_TRY(cudaFuncSetAttribute(blur_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  blur_kernel<<<dim3(bands, nframes), threads, smem, s>>>(p, rest);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

// =============================================================================================
// 5. orientation (IC_Angle), steered BRIEF, keypoint fix-ups and per-frame concatenation
// =============================================================================================
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  // cv::fastAtan2 (OpenCV 4.13): odd 7th-order polynomial, every op rounded to float32.
  const float s = 57.295779513082323f;  // (float)(180/pi)
  const float p1 = __fmul_rn(0.9997878412794807f, s), p3 = __fmul_rn(-0.3258083974640975f, s);
  const float p5 = __fmul_rn(0.1555786518463281f, s), p7 = __fmul_rn(-0.04432655554792128f, s);
  const float eps = 2.2204460492503131e-16f;  // (float)DBL_EPSILON
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

// sin and cos of x in [0, 2*pi] in double precision (error ~1e-16, the float rounding of the result is what
// the descriptor uses: identical to rounding libm's double sincos, without its large-argument slow path).
// Cody-Waite reduction by pi/2 and the fdlibm kernel polynomials on |r| <= pi/4.
__device__ __forceinline__ void sincos_0_2pi(double x, double* s, double* c) {
  const double k = rint(x * 0.63661977236758138);  // 2/pi
  double r = __fma_rn(-k, 1.5707963267948966, x);
  r = __fma_rn(-k, 6.123233995736766e-17, r);
  const double z = r * r;
  double ps = __fma_rn(z, 1.58969099521155010221e-10, -2.50507602534068634195e-08);
  ps = __fma_rn(z, ps, 2.75573137070700676789e-06);
  ps = __fma_rn(z, ps, -1.98412698298579493134e-04);
  ps = __fma_rn(z, ps, 8.33333333332248946124e-03);
  ps = __fma_rn(z, ps, -1.66666666666666324348e-01);
  const double sr = __fma_rn(r * z, ps, r);
  double pc = __fma_rn(z, -1.13596475577881948265e-11, 2.08757232129817482790e-09);
  pc = __fma_rn(z, pc, -2.75573143513906633035e-07);
  pc = __fma_rn(z, pc, 2.48015872894767294178e-05);
  pc = __fma_rn(z, pc, -1.38888888888741095749e-03);
  pc = __fma_rn(z, pc, 4.16666666666666019037e-02);
  const double cr = __fma_rn(z * z, pc, __fma_rn(z, -0.5, 1.0));
  const int q = (int)k & 3;
  const double s0 = (q & 1) ? cr : sr, c0 = (q & 1) ? sr : cr;
  *s = (q & 2) ? -s0 : s0;
  *c = ((q + 1) & 2) ? -c0 : c0;
}

constexpr int kOdWarps = 8, kOdSlotsPerWarp = 8, kOdSlotsPerBlock = kOdWarps * kOdSlotsPerWarp;
constexpr int kOdReach = 18;                      // |rotated tap| <= ceil(13 * sqrt(2)) = 18 < EDGE_THRESHOLD
constexpr int kOdRows = 2 * kOdReach + 1;         // 37 rows of the blurred level around the keypoint
constexpr int kOdWords = (kOdRows + 3 + 3) / 4;   // 10 words per row: 37 px + alignment offset (<= 3)

constexpr int kIcRows = 2 * kHalfPatch + 1;       // 31 rows of the level image around the keypoint (IC_Angle)
constexpr int kIcWords = 9;                       // 31 px + alignment offset (<= 3) -> 9 words; odd: conflict-free rows

__device__ __forceinline__ int round_half_even(float v) {
  // cvRound for |v| < 2^22: adding 1.5 * 2^23 leaves the rounded integer in the mantissa
  return __float_as_int(__fadd_rn(v, 12582912.f)) - 0x4B400000;
}

// One block = 32 consecutive output slots (level, rank in the level's selection) of one frame, one warp
// per slot, 8 slots per warp (the pattern copy and the level prefix of a block are shared by 64 keypoints).  IC_Angle: lane = patch column u, the rows of the disc that contain it
// are |v| <= umax[|u|] (the disc is symmetric), 31 coalesced byte loads, shuffle reduction.
// kTma: the two windows of a keypoint come by TMA (one 64 x 37 byte box of the blurred level, one 48 x 31 byte box of
// the level image; boxes start at the 16-byte boundary left of the window, which is what the wider pitches are for)
// instead of 24 rounds of 4-byte cp.async per keypoint -- a sixth of the kernel's instructions.  A level image that TMA
// cannot address (caller memory that is unaligned or has gaps between frames) keeps cp.async for its IC patch.
constexpr int kOdTilePitch = 64, kOdTileBytes = ((kOdTilePitch * kOdRows + 127) / 128) * 128;   // 2432
constexpr int kIcPitch = 48, kIcBytes = ((kIcPitch * kIcRows + 127) / 128) * 128;               // 1536
struct OdStage {
  CUtensorMap tile[kMaxLevels];  // blurred level l as [frames * h][bpitch], box 64 x 37
  CUtensorMap ic[kMaxLevels];    // level image l as [frames * h][pitch], box 48 x 31
  unsigned ic_mask;              // levels whose image TMA can address
};

template <bool kTma>
__global__ void __launch_bounds__(kOdWarps * 32, 5)
orient_desc_kernel(Plan p, Scratch sc, OrbKeyPoint* __restrict__ kps, uint8_t* __restrict__ desc,
                   int capacity, int* __restrict__ n_out, int slots_per_warp, const __grid_constant__ OdStage T) {
  __shared__ float4 s_pat[256];
  __shared__ __align__(128) uint32_t s_tile[kOdWarps][kTma ? kOdTileBytes / 4 : kOdRows * kOdWords];
  __shared__ __align__(128) uint32_t s_ic[kOdWarps][kTma ? kIcBytes / 4 : kIcRows * kIcWords];  // unblurred 31-row patch
  __shared__ __align__(8) uint64_t s_bar[kOdWarps][2];
  constexpr int kTileWords = kTma ? kOdTilePitch / 4 : kOdWords;  // words per staged row of the blurred window
  constexpr int kIcRowWords = kTma ? kIcPitch / 4 : kIcWords;     // words per staged row of the IC patch
  __shared__ int s_cnt[kMaxLevels], s_before[kMaxLevels + 1], s_base[kMaxLevels];
  const int f = blockIdx.y;
  s_pat[threadIdx.x] = g_pattern[threadIdx.x];
  if (threadIdx.x < p.nlevels) s_cnt[threadIdx.x] = sc.sel_cnt[f * p.nlevels + threadIdx.x];
  if (threadIdx.x < kMaxLevels)  // first slot of every level (unused levels never match)
    s_base[threadIdx.x] = threadIdx.x < p.nlevels ? p.lv[threadIdx.x].slot_base : 0x7fffffff;
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int l = 0; l < p.nlevels; ++l) { s_before[l] = run; run += s_cnt[l]; }
    s_before[p.nlevels] = run;
    if (blockIdx.x == 0) n_out[f] = min(run, capacity);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // provably warp-uniform
  // IC_Angle with lane = patch row v: the row's 31 pixels as 8 aligned words, columns |u| <= umax[|v|] kept by a
  // byte mask, m10 / row sum by two IDP.4A per word (weights u = 4k+j-15 as signed bytes / all ones)
  const int v_row = lane - kHalfPatch;
  const int dmax = lane < 31 ? p.umax[v_row < 0 ? -v_row : v_row] : -1;
  uint32_t ic_mask[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int uu = 4 * k + j - kHalfPatch;
      if ((uu < 0 ? -uu : uu) <= dmax) m |= 0xffu << (8 * j);
    }
    ic_mask[k] = m;
  }
  uint32_t* my_tile = s_tile[warp];
  const int tile_r = lane / kOdWords, tile_c = lane - tile_r * kOdWords;  // this lane's (row, word) inside a 3-row step
  const unsigned tile_dst = (unsigned)__cvta_generic_to_shared(my_tile + tile_r * kOdWords + tile_c);
  uint32_t* my_ic = s_ic[warp];
  const int ic_r = lane / kIcWords, ic_c = lane - ic_r * kIcWords;  // (row, word) inside a 3-row step, lanes 0..26
  const unsigned ic_dst = (unsigned)__cvta_generic_to_shared(my_ic + ic_r * kIcRowWords + ic_c);
  const uint32_t bar_ic = smem_u32(&s_bar[warp][0]), bar_tl = smem_u32(&s_bar[warp][1]);
  if (kTma) {
    if (lane == 0) {
      mbar_init(bar_ic, 1); mbar_init(bar_tl, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
  }
  uint32_t phase_tl = 0, phase_ic = 0;

  int level = 0;
  for (int it = 0; it < slots_per_warp; ++it) {
    const int slot = blockIdx.x * (slots_per_warp * kOdWarps) + it * kOdWarps + warp;
    if (slot >= p.slots_per_frame) break;
    // the slots a warp visits ascend, so the level only moves forward
    while (level + 1 < p.nlevels && slot >= s_base[level + 1]) ++level;
    const LevelGeom& g = p.lv[level];
    const int rank = slot - g.slot_base;
    if (rank >= s_cnt[level]) continue;
    const int out_idx = s_before[level] + rank;
    if (out_idx >= capacity) {
      if (lane == 0) *sc.status = ORB_GPU_ERR_CAPACITY;
      continue;
    }
    const uint32_t pv = sc.sel[(size_t)f * p.slots_per_frame + slot];
    const int cx = (int)(pv & 0xfff) + kBorder, cy = (int)((pv >> 12) & 0xfff) + kBorder;
    const int resp = (int)(pv >> 24);

    // The rotated BRIEF taps stay within 18 px of the keypoint (|x|,|y| <= 13): the 37 x 37 window of the
    // blurred level is staged in shared memory with coalesced word copies (10 words per row, issued now
    // so they fly during the orientation pass); the 512 byte gathers then hit shared memory, not L1/L2.
    const uint8_t* img = g.img + (size_t)f * g.frame_stride;
    const int ic_xal = (cx - kHalfPatch) & ~3;
    const bool ic_tma = kTma && ((T.ic_mask >> level) & 1u);
    if (kTma && lane == 0) {
      // rows f * h + y of the stacked levels; x rounded down to the 16-byte boundary (TMA's coordinate rule)
      mbar_expect_tx(bar_tl, (uint32_t)(kOdTilePitch * kOdRows));
      tma_load_2d(smem_u32(my_tile), &T.tile[level], (cx - kOdReach) & ~15, f * g.h + cy - kOdReach, bar_tl);
      if (ic_tma) {
        mbar_expect_tx(bar_ic, (uint32_t)(kIcPitch * kIcRows));
        tma_load_2d(smem_u32(my_ic), &T.ic[level], (cx - kHalfPatch) & ~15, f * g.h + cy - kHalfPatch, bar_ic);
      }
    }
    if (!ic_tma) {  // unblurred 31 x 36-byte patch for IC_Angle: 27 lanes copy 3 rows x 9 words per step, own commit group
      const uint8_t* src = img + (size_t)(cy - kHalfPatch + ic_r) * g.pitch + ic_xal + 4 * ic_c;
      unsigned dst = ic_dst;
      const int step = 3 * g.pitch;
#pragma unroll
      for (int r0 = 0; r0 < kIcRows; r0 += 3, src += step, dst += 3 * kIcRowWords * 4)
        if (lane < 27 && r0 + ic_r < kIcRows)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst), "l"(src));
      asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    if (!kTma) {
      const uint8_t* bl = g.blur + (size_t)f * g.bframe_stride;
      const int xal = (cx - kOdReach) & ~3;
      // 30 lanes copy 3 rows x 10 words per step: one pointer bump per step
      const uint8_t* src = bl + (size_t)(cy - kOdReach + tile_r) * g.bpitch + xal + 4 * tile_c;
      unsigned dst = tile_dst;
      const int step = 3 * g.bpitch;
#pragma unroll
      for (int r0 = 0; r0 < kOdRows; r0 += 3, src += step, dst += 3 * kOdWords * 4)
        if (lane < 30 && r0 + tile_r < kOdRows)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst), "l"(src));
      asm volatile("cp.async.commit_group;\n" ::: "memory");
    }

    // IC_Angle: m10 = sum u*I, m01 = sum v*I over the 31-px disc (:77-104)
    int m10 = 0, m01 = 0;
    // the unblurred patch has landed (the blurred one may still fly)
    if (ic_tma) { mbar_wait(bar_ic, phase_ic); phase_ic ^= 1u; }
    else if (kTma) asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    else asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    __syncwarp();
    if (lane < 31) {
      // TMA boxes start at the 16-byte boundary: the window begins ((cx - 15) & 15) bytes into the staged row
      const uint32_t* wp = my_ic + lane * kIcRowWords + (ic_tma ? ((cx - kHalfPatch) & 15) >> 2 : 0);
      const unsigned sh = (unsigned)((cx - kHalfPatch) & 3) * 8;
      uint32_t w[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) w[k] = wp[k];
      int srow = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t x = __funnelshift_r(w[k], w[k + 1], sh) & ic_mask[k];
        // signed weights (4k-15 .. 4k-12) as packed bytes
        const int wu = (int)(((uint32_t)(uint8_t)(int8_t)(4 * k - 15)) | ((uint32_t)(uint8_t)(int8_t)(4 * k - 14) << 8) |
                             ((uint32_t)(uint8_t)(int8_t)(4 * k - 13) << 16) | ((uint32_t)(uint8_t)(int8_t)(4 * k - 12) << 24));
        asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(x), "r"(wu));
        srow = (int)__dp4a(x, 0x01010101u, (unsigned)srow);
      }
      m01 = v_row * srow;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m10 += __shfl_xor_sync(0xffffffffu, m10, o);
      m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // steered BRIEF on the blurred level (:106-147)
    const float rad = __fmul_rn(angle, 0.017453292519943295f);  // (float)(CV_PI/180.f)
    double sd, cd;
    sincos_0_2pi((double)rad, &sd, &cd);
    const float a = (float)cd, b = (float)sd;
    if (kTma) { mbar_wait(bar_tl, phase_tl); phase_tl ^= 1u; }
    else asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncwarp();
    const uint8_t* tc = reinterpret_cast<const uint8_t*>(my_tile) + kOdReach * (4 * kTileWords) + kOdReach +
                        ((cx - kOdReach) & (kTma ? 15 : 3));
    int t0[8], t1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {  // 16 gathers per lane from the staged window
      const float4 t = s_pat[32 * k + lane];
      const int r0 = round_half_even(__fadd_rn(__fmul_rn(t.x, b), __fmul_rn(t.y, a)));
      const int q0 = round_half_even(__fsub_rn(__fmul_rn(t.x, a), __fmul_rn(t.y, b)));
      const int r1 = round_half_even(__fadd_rn(__fmul_rn(t.z, b), __fmul_rn(t.w, a)));
      const int q1 = round_half_even(__fsub_rn(__fmul_rn(t.z, a), __fmul_rn(t.w, b)));
      t0[k] = tc[r0 * (4 * kTileWords) + q0];
      t1[k] = tc[r1 * (4 * kTileWords) + q1];
    }
    __syncwarp();  // the window is rewritten by the next slot
    uint32_t mine = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t wk = __ballot_sync(0xffffffffu, t0[k] < t1[k]);  // bit = test 32k+lane
      mine = lane == k ? wk : mine;
    }
    uint32_t* drow = reinterpret_cast<uint32_t*>(desc + ((size_t)f * capacity + out_idx) * 32);
    if (lane < 8) drow[lane] = mine;  // one coalesced 32-byte store
    if (lane == 0) {
      OrbKeyPoint kp;
      const float fx = (float)cx, fy = (float)cy;
      kp.x = level ? __fmul_rn(fx, g.scale) : fx;  // keypoint->pt *= scale (:1095-1101)
      kp.y = level ? __fmul_rn(fy, g.scale) : fy;
      kp.size = g.patch_size;
      kp.angle = angle;
      kp.response = (float)resp;
      kp.octave = level;
      kp.class_id = -1;
      kps[(size_t)f * capacity + out_idx] = kp;
    }
  }
}

int launch_orient_desc(const Plan& p, const Scratch& sc, int nframes, OrbKeyPoint* d_kps,
                       uint8_t* d_desc, int capacity, int* d_n_out, cudaStream_t s) {
  // a few frames: one keypoint slot per warp instead of four (more, shorter blocks)
  static const int env_spw = [] { const char* e = getenv("ORB_GPU_OD_SPW"); return e ? atoi(e) : 0; }();
  const int spw = env_spw > 0 ? env_spw : (long long)p.slots_per_frame * nframes >= 4 * kOdSlotsPerBlock * 148 ? kOdSlotsPerWarp : 1;
  const dim3 grid((p.slots_per_frame + spw * kOdWarps - 1) / (spw * kOdWarps), nframes);
  // TMA staging of the keypoint windows: every blurred level must be addressable (16-byte aligned base and pitch, frames
  // stacked without gaps: always true for the handle's own buffers); a level image that is not keeps cp.async
  static const bool no_tma = [] { const char* e = getenv("ORB_GPU_OD_DIRECT"); return e && atoi(e); }();
  OdStage T;
  memset(&T, 0, sizeof(T));
  bool tma = !no_tma && tma_available();
  for (int l = 0; tma && l < p.nlevels; ++l) {
    const LevelGeom& g = p.lv[l];
    if ((reinterpret_cast<uintptr_t>(g.blur) & 15) || (g.bpitch & 15) ||
        (nframes > 1 && g.bframe_stride != (long long)g.bpitch * g.h) || g.h < kOdRows) { tma = false; break; }
    if (tensor_map_2d_u8(&T.tile[l], g.blur, g.bpitch, (long long)nframes * g.h, g.bpitch, kOdTilePitch, kOdRows, false)) { tma = false; break; }
    const bool ic_ok = (reinterpret_cast<uintptr_t>(g.img) & 15) == 0 && (g.pitch & 15) == 0 &&
                       (nframes == 1 || g.frame_stride == (long long)g.pitch * g.h);
    if (ic_ok && tensor_map_2d_u8(&T.ic[l], g.img, g.pitch, (long long)nframes * g.h, g.pitch, kIcPitch, kIcRows, false) == ORB_GPU_OK)
      T.ic_mask |= 1u << l;
  }
  if (tma) orient_desc_kernel<true><<<grid, kOdWarps * 32, 0, s>>>(p, sc, d_kps, d_desc, capacity, d_n_out, spw, T);
  else orient_desc_kernel<false><<<grid, kOdWarps * 32, 0, s>>>(p, sc, d_kps, d_desc, capacity, d_n_out, spw, T);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

// =============================================================================================
// test hook: candidates of one (frame, level) in arrival order as floats
// =============================================================================================
__global__ void export_candidates_kernel(Plan p, Scratch sc, int f, int level, float* xyr, int cap) {
  const LevelGeom& g = p.lv[level];
  const int n = min(sc.cand_cnt[f * p.nlevels + level], cap);
  const uint32_t* pay = sc.pay + (size_t)f * p.keys_per_frame + g.key_base;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
    const uint32_t v = pay[k];
    xyr[3 * k] = (float)(v & 0xfff);
    xyr[3 * k + 1] = (float)((v >> 12) & 0xfff);
    xyr[3 * k + 2] = (float)(v >> 24);
  }
}

int launch_export_candidates(const Plan& p, const Scratch& sc, int frame, int level, float* d_xyr,
                             int cap, cudaStream_t s) {
  export_candidates_kernel<<<32, 256, 0, s>>>(p, sc, frame, level, d_xyr, cap);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

}  // namespace orbgpu
