// Brute-force Hamming matching on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in TMEM).
//
// Reference: the query x database scan with the top-2 update rule of src/ORBmatcher.cc:216-225 over
// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663).  Same identity as bruteforce_mma_kernel
// (match_kernels.cu): with query bits expanded to signed bytes a' (+w set, -w clear) and database bits to
// unsigned bytes b (0 or 64/w), sum_k a'_k * b_k = 64 * (|q & d| - |~q & d|) = 64 * S and
// hamming(q, d) = popc(q) - S.  The weights let the expanding threads skip every shift: a database byte is
// replicated into the four bytes of a word by one PRMT and masked, so bit j of the byte stays at bit j of
// output byte j (weights 1,2,4,8 for bits 0..3; for bits 4..7 the word is first shifted right by one, giving
// 8,16,32,64), and the query side carries 64,32,16,8 and 8,4,2,1.
//
// One block = 256 queries (two M=128 A tiles, expanded once into shared memory) x one slice of the database.
// Roles (one warp each unless noted):
//   warps 9..12  expanders: two database rows per thread and tile of 256 rows -> 32 STS.128 into the K-major,
//                unswizzled core-matrix layout (8 rows x 16 bytes contiguous; LBO = 4096, SBO = 128)
//   warp 8       MMA issuer (the warp stays converged, one elected lane issues): per tile 2 x 8 tcgen05.mma
//                M128 N256 K32 into two TMEM accumulators, every descriptor in its own uniform register pair
//   warps 0..7   epilogue (set i = warps 4i..4i+3 drains accumulator i): the 256-column row of the thread's
//                query comes to registers in four packed tcgen05.ld (three in flight), the accumulator is
//                handed back as soon as the last one lands, and the rest of the scan (max filter against the
//                running second best, packed (S, row) keys) overlaps the next MMA into it
// Pipelines: b_full/b_empty (expanders <-> MMA, 2 stages), acc_full/acc_empty (MMA <-> epilogue, the two
// accumulators alternate so the epilogue of one overlaps the MMAs of the other).
#include <cstring>
#include <type_traits>

#include "common.cuh"
#include "umma.cuh"

namespace orbgpu {

constexpr int kT5Queries = 256;
constexpr int kT5TileRows = 256;
constexpr int kT5EpiWarps = 8;                           // two sets of four: set i drains accumulator i
constexpr int kT5ExpWarps = 4;
constexpr int kT5ExpThreads = kT5ExpWarps * 32;
constexpr int kT5RowsPerExp = kT5TileRows / kT5ExpThreads;
constexpr int kT5Threads = (kT5EpiWarps + 1 + kT5ExpWarps) * 32;
constexpr uint32_t kT5ATile = 128 * 256;                 // bytes of one expanded A tile
constexpr uint32_t kT5BStage = kT5TileRows * 256;        // bytes of one expanded B stage
constexpr uint32_t kT5SmemBytes = 2 * kT5ATile + 2 * kT5BStage + 128;
constexpr uint32_t kT5TmemCols = 512;
// instruction descriptor (cute/arch/mma_sm100_desc.hpp layout): D = s32, A = signed 8 bit, B = unsigned 8 bit,
// both K-major, N = 256, M = 128
constexpr uint32_t kT5Idesc = (2u << 4) | (1u << 7) | (0u << 10) | ((256u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(kT5Idesc), "r"(accumulate)
      : "memory");
}

#define T5_R8(v, o) "=r"(v[o]), "=r"(v[o + 1]), "=r"(v[o + 2]), "=r"(v[o + 3]), "=r"(v[o + 4]), "=r"(v[o + 5]), "=r"(v[o + 6]), "=r"(v[o + 7])
#define T5_RW8(v, o) "+r"(v[o]), "+r"(v[o + 1]), "+r"(v[o + 2]), "+r"(v[o + 3]), "+r"(v[o + 4]), "+r"(v[o + 5]), "+r"(v[o + 6]), "+r"(v[o + 7])
// The issuing warp runs its loop converged and one elected lane issues (elect.sync inside the asm): the control flow
// around the MMAs stays warp-uniform, so the descriptors are computed and kept in uniform registers.
__device__ __forceinline__ void umma_i8_elect(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(kT5Idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
      ::"r"(bar)
      : "memory");
}
// 64 consecutive accumulator columns of this thread's TMEM lane, the low 16 bits of each packed two to a
// register (column 2r in the low half of register r): |64 * S| <= 16384 fits
__device__ __forceinline__ void tmem_ld64p(uint32_t (&v)[32], uint32_t taddr) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : T5_R8(v, 0), T5_R8(v, 8), T5_R8(v, 16), T5_R8(v, 24)
      : "r"(taddr));
}
// wait for the outstanding tcgen05.ld; the registers are operands so that no use of them moves above the wait
__device__ __forceinline__ void tmem_wait(uint32_t (&v)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : T5_RW8(v, 0), T5_RW8(v, 8), T5_RW8(v, 16), T5_RW8(v, 24)
               :
               : "memory");
}

// packed maximum of 64 accumulator columns (two s16 lanes), a depth-4 tree of three-input maxima
__device__ __forceinline__ uint32_t t5_rowmax(const uint32_t (&v)[32]) {
  uint32_t l1[11], l2[4];
#pragma unroll
  for (int i = 0; i < 10; ++i) l1[i] = __vmaxs2(__vmaxs2(v[3 * i], v[3 * i + 1]), v[3 * i + 2]);
  l1[10] = __vmaxs2(v[30], v[31]);
#pragma unroll
  for (int i = 0; i < 3; ++i) l2[i] = __vmaxs2(__vmaxs2(l1[3 * i], l1[3 * i + 1]), l1[3 * i + 2]);
  l2[3] = __vmaxs2(l1[9], l1[10]);
  return __vmaxs2(__vmaxs2(__vmaxs2(l2[0], l2[1]), l2[2]), l2[3]);
}
// does either half exceed the threshold held in both halves of thr2?
__device__ __forceinline__ bool t5_hit(uint32_t packed, uint32_t thr2) { return __vmaxs2(packed, thr2) != thr2; }

// exact top-2 merge of 64 accumulator columns into (best, second): keys of the two halves of a register are
// ordered first, four independent (best, second) pairs take the registers round robin, and the pairs are merged
// at the end -- straight-line code with a short dependency chain (all lanes run it; it is reached for ~5% of the
// calls).  CHECK masks the columns past the end of the slice (last tile only).
__device__ __forceinline__ void t5_merge2(unsigned& b, unsigned& s2, unsigned hi, unsigned lo) {
  const unsigned t = min(b, hi);
  s2 = max(max(s2, t), lo);
  b = max(b, hi);
}
template <bool CHECK>
__device__ __forceinline__ void t5_merge64(const uint32_t (&v)[32], int col0, int rows, unsigned& best,
                                           unsigned& second) {
  const unsigned kbase = (256u << 22) + 0x3fffffu - (unsigned)col0;
  const int lim = rows - col0;
  unsigned b[4] = {0, 0, 0, 0}, s2[4] = {0, 0, 0, 0};
#pragma unroll
  for (int r = 0; r < 32; ++r) {
    // (64 * S) << 16 = S << 22: the packed halves are already the key's distance field
    unsigned k0 = (v[r] << 16) + (kbase - 2 * r), k1 = (v[r] & 0xffff0000u) + (kbase - (2 * r + 1));
    if (CHECK) {
      k0 = 2 * r < lim ? k0 : 0u;
      k1 = 2 * r + 1 < lim ? k1 : 0u;
    }
    t5_merge2(b[r & 3], s2[r & 3], max(k0, k1), min(k0, k1));
  }
  t5_merge2(b[0], s2[0], b[1], s2[1]);
  t5_merge2(b[2], s2[2], b[3], s2[3]);
  t5_merge2(b[0], s2[0], b[2], s2[2]);
  t5_merge2(best, second, b[0], s2[0]);
}
// exact top-2 merge of the 16 columns held by registers 8 * J .. 8 * J + 7 (same keys as t5_merge64)
template <int J>
__device__ __forceinline__ void t5_merge16(const uint32_t (&v)[32], int col0, unsigned& best, unsigned& second) {
  const unsigned kbase = (256u << 22) + 0x3fffffu - (unsigned)col0;
  unsigned b[2] = {0, 0}, s2[2] = {0, 0};
#pragma unroll
  for (int r = 8 * J; r < 8 * J + 8; ++r) {
    const unsigned k0 = (v[r] << 16) + (kbase - 2 * r), k1 = (v[r] & 0xffff0000u) + (kbase - (2 * r + 1));
    t5_merge2(b[r & 1], s2[r & 1], max(k0, k1), min(k0, k1));
  }
  t5_merge2(b[0], s2[0], b[1], s2[1]);
  t5_merge2(best, second, b[0], s2[0]);
}
template <int J>
__device__ __forceinline__ void t5_scan16(const uint32_t (&v)[32], int col0, unsigned& best, unsigned& second,
                                          uint32_t thr2) {
  uint32_t m = __vmaxs2(__vmaxs2(v[8 * J], v[8 * J + 1]), __vmaxs2(v[8 * J + 2], v[8 * J + 3]));
  m = __vmaxs2(m, __vmaxs2(__vmaxs2(v[8 * J + 4], v[8 * J + 5]), __vmaxs2(v[8 * J + 6], v[8 * J + 7])));
  if (__any_sync(0xffffffffu, t5_hit(m, thr2))) t5_merge16<J>(v, col0, best, second);
}
// A 64-column chunk in which some query of the warp beats its running second best: after the first few thousand
// rows of a slice that is one column in 64, so the quarter that holds it is found first and only it is merged.
__device__ __forceinline__ uint32_t t5_thr_of(unsigned field) {  // packed 64 * S of the distance field S + 256
  return ((field * 64u - 16384u) & 0xffffu) * 0x10001u;
}
__device__ __forceinline__ void t5_scan64(const uint32_t (&v)[32], int col0, int rows, unsigned& best,
                                          unsigned& second, uint32_t& thr2, unsigned floor_f) {
  if (rows - col0 >= 64) {
    t5_scan16<0>(v, col0, best, second, thr2);
    t5_scan16<1>(v, col0, best, second, thr2);
    t5_scan16<2>(v, col0, best, second, thr2);
    t5_scan16<3>(v, col0, best, second, thr2);
  } else {
    t5_merge64<true>(v, col0, rows, best, second);
  }
  // the filter is the running second best of this slice or, when it is tighter, what the other slices of the
  // query block have published (floor_f, already one step looser: a tie with another slice must still get through)
  if (second || floor_f) thr2 = t5_thr_of(max(second >> 22, floor_f));
}

__global__ void __launch_bounds__(kT5Threads, 1)
bruteforce_tc5_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db, long long ndb,
                      long long slice_rows, int32_t idx_base, int32_t* __restrict__ o_idx,
                      int32_t* __restrict__ o_best, int32_t* __restrict__ o_second,
                      unsigned* __restrict__ g_thr /*[nq] shared second-best field of a query over the slices, or null*/,
                      unsigned* __restrict__ g_ticket /*[query blocks] slices finished*/, int nslices,
                      int32_t* __restrict__ f_idx, int32_t* __restrict__ f_best, int32_t* __restrict__ f_second,
                      FusedExchange fx /*world > 1: finish the step across the ranks here*/) {
  extern __shared__ __align__(128) uint8_t t5_smem[];
  uint8_t* s_a = t5_smem;
  uint8_t* s_b = t5_smem + 2 * kT5ATile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(t5_smem + 2 * kT5ATile + 2 * kT5BStage);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 8);
  const uint32_t bar0 = smem_u32(bars);
  // barrier slots: 0,1 b_full  2,3 b_empty  4,5 acc_full  6,7 acc_empty
  auto bar = [&](int i) { return bar0 + 8u * i; };

  const int t = threadIdx.x, lane = t & 31;
  const int warp = __shfl_sync(0xffffffffu, t >> 5, 0);  // provably warp-uniform: the role branches stay uniform
  const int slice = blockIdx.y;
  const long long row0 = (long long)slice * slice_rows;
  const int rows = (int)min(slice_rows, ndb - row0);
  const int ntiles = (rows + kT5TileRows - 1) / kT5TileRows;
  const int qbase = blockIdx.x * kT5Queries;

  if (t == 0) {
    mbar_init(bar(0), kT5ExpThreads); mbar_init(bar(1), kT5ExpThreads);
    mbar_init(bar(2), 1); mbar_init(bar(3), 1);
    mbar_init(bar(4), 1); mbar_init(bar(5), 1);
    mbar_init(bar(6), 128); mbar_init(bar(7), 128);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kT5EpiWarps) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)),
                 "r"(kT5TmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp > kT5EpiWarps) {
    // A operand: query qbase + e as signed weighted bytes (see the header), K-major core matrices
    for (int e = t - (kT5EpiWarps + 1) * 32; e < kT5Queries; e += kT5ExpThreads) {
      const int qi = min(qbase + e, nq - 1);
      const uint4 lo = __ldg(q + 2 * (size_t)qi), hi = __ldg(q + 2 * (size_t)qi + 1);
      const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
      uint8_t* dst = s_a + (e >> 7) * kT5ATile + (e & 127) * 16;
#pragma unroll 2
      for (int c = 0; c < 16; ++c) {
        const uint32_t h = (w[c >> 1] >> (16 * (c & 1))) & 0xffffu;
        uint32_t o[4] = {0, 0, 0, 0};
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int wx = 64 >> j, wy = 8 >> j;  // query weights of bits j and 4 + j
            const uint32_t x = (h >> (8 * b + j)) & 1u ? (uint32_t)wx : (uint32_t)(256 - wx);
            const uint32_t y = (h >> (8 * b + 4 + j)) & 1u ? (uint32_t)wy : (uint32_t)(256 - wy);
            o[2 * b] |= x << (8 * j);
            o[2 * b + 1] |= y << (8 * j);
          }
        *reinterpret_cast<uint4*>(dst + c * 2048) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp < kT5EpiWarps) {
    // ---- epilogue: thread = TMEM lane = query qbase + t (warps 0..3: accumulator 0, warps 4..7: accumulator 1)
    const int acc = warp >> 2;
    unsigned best = 0, second = 0;
    uint32_t thr2 = 0x80008000u;  // 64 * S of the running second best in both halves; -32768 = none yet
    // Every slice of a query block starts its scan cold; the slices publish their running second best per query
    // (atomicMax on its distance field) and adopt the others' -- less one step, so that a row that only TIES what
    // another slice holds still arrives (the lowest row wins ties in the final merge).  A row filtered this way is
    // below two rows of the same query elsewhere: the merged top-2 is unchanged, and the warm-up is shared.
    unsigned floor_f = 0, published = 0;
    const bool share = g_thr != nullptr && qbase + t < nq;
    const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + acc * 256;
    // three buffers of 64 packed columns: 13 warps put four on one scheduler, which caps a thread at 128 registers
    uint32_t va[32], vb[32], vc[32];
    for (int tile = 0; tile < ntiles; ++tile) {
      mbar_wait(bar(4 + acc), tile & 1);
      tc_fence_after();
      const int col0 = tile * kT5TileRows;
      tmem_ld64p(va, taddr);
      tmem_ld64p(vb, taddr + 64);
      tmem_ld64p(vc, taddr + 128);
      tmem_wait(va);
      tmem_wait(vb);
      tmem_wait(vc);
      if (__any_sync(0xffffffffu, t5_hit(t5_rowmax(va), thr2))) t5_scan64(va, col0, rows, best, second, thr2, floor_f);
      tmem_ld64p(va, taddr + 192);  // the fourth quarter lands while the second and third are filtered
      const bool hb = __any_sync(0xffffffffu, t5_hit(t5_rowmax(vb), thr2));
      const bool hc = __any_sync(0xffffffffu, t5_hit(t5_rowmax(vc), thr2));
      tmem_wait(va);
      tc_fence_before();
      mbar_arrive(bar(6 + acc));  // the accumulator is in registers: the next MMA into it overlaps the rest of the scan
      if (hb) t5_scan64(vb, col0 + 64, rows, best, second, thr2, floor_f);
      if (hc) t5_scan64(vc, col0 + 128, rows, best, second, thr2, floor_f);
      if (__any_sync(0xffffffffu, t5_hit(t5_rowmax(va), thr2))) t5_scan64(va, col0 + 192, rows, best, second, thr2, floor_f);
      if (share && (tile < 32 || (tile & 7) == 7)) {
        const unsigned mine = second >> 22;
        unsigned seen;
        if (mine > published) { seen = atomicMax(g_thr + qbase + t, mine); published = mine; }
        else seen = __ldcg(g_thr + qbase + t);
        if (seen > floor_f + 1) {
          floor_f = seen - 1;
          if (floor_f > mine) thr2 = t5_thr_of(floor_f);
        }
      }
    }
    const int qi = qbase + t;
    if (qi < nq) {
      const uint4 lo = __ldg(q + 2 * (size_t)qi), hi = __ldg(q + 2 * (size_t)qi + 1);
      const int pq = __popc(lo.x) + __popc(lo.y) + __popc(lo.z) + __popc(lo.w) + __popc(hi.x) + __popc(hi.y) +
                     __popc(hi.z) + __popc(hi.w);
      const size_t o = (size_t)slice * nq + qi;
      const int bd = best ? pq - ((int)(best >> 22) - 256) : 256;
      const int sd = second ? pq - ((int)(second >> 22) - 256) : 256;
      o_best[o] = bd;
      o_second[o] = sd;
      // a row at distance 256 never beats the initial best of the reference's scan (strict '<')
      o_idx[o] = (best && bd < 256) ? (int32_t)(idx_base + row0 + (0x3fffffu - (best & 0x3fffffu))) : -1;
      if (nslices > 1 || fx.world > 1) __threadfence();  // the slice's result is visible before this block takes its ticket
    }
  } else if (warp == kT5EpiWarps) {
    // ---- MMA issuer: the whole warp, converged (see umma_i8_elect)
    {
      // Every descriptor is base + constant with compile-time stage, accumulator and K step (the tile loop is
      // unrolled by the two stages).  A descriptor selected at run time goes through ONE uniform register pair
      // that is rewritten before every MMA, and the rewrite waits until the previous MMA has read it: the tensor
      // pipe then idles ~40 clk per MMA (tools/ubench/tmem_a_probe.cu).
      const uint64_t da0 = umma_desc(smem_u32(s_a), 2048, 128), db0 = umma_desc(smem_u32(s_b), 4096, 128);
      auto issue_tile = [&](auto stage, int tile) {
        constexpr int s = decltype(stage)::value;
        mbar_wait(bar(s), (tile >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int acc = 0; acc < 2; ++acc) {
          mbar_wait(bar(6 + acc), s ^ 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 8; ++k)  // K = 32 bytes = two 16-byte chunks per step
            umma_i8_elect(tmem_base + acc * 256, da0 + (uint64_t)((acc * kT5ATile + k * 2 * 2048) >> 4),
                          db0 + (uint64_t)((s * kT5BStage + k * 2 * 4096) >> 4), k > 0);
          umma_commit_elect(bar(4 + acc));
        }
        umma_commit_elect(bar(2 + s));
      };
      for (int tile = 0; tile < ntiles; tile += 2) {
        issue_tile(std::integral_constant<int, 0>{}, tile);
        if (tile + 1 < ntiles) issue_tile(std::integral_constant<int, 1>{}, tile + 1);
      }
      // the last commit must have landed before the block (and its shared memory) goes away
      if (ntiles > 0) mbar_wait(bar(2 + ((ntiles - 1) & 1)), ((ntiles - 1) >> 1) & 1);
      __syncwarp();
    }
  } else {
    // ---- expanders: rows e, e + 128 of every tile
    const int e = t - (kT5EpiWarps + 1) * 32;
    const uint4* src = db + row0 * 2;
    auto fetch = [&](int tile, uint4 (&lo)[kT5RowsPerExp], uint4 (&hi)[kT5RowsPerExp]) {
#pragma unroll
      for (int i = 0; i < kT5RowsPerExp; ++i) {
        const int r = tile * kT5TileRows + e + i * kT5ExpThreads;
        if (tile < ntiles && r < rows) { lo[i] = __ldg(src + 2 * (size_t)r); hi[i] = __ldg(src + 2 * (size_t)r + 1); }
        else { lo[i] = make_uint4(0, 0, 0, 0); hi[i] = lo[i]; }
      }
    };
    uint4 lo[kT5RowsPerExp], hi[kT5RowsPerExp], nlo[kT5RowsPerExp], nhi[kT5RowsPerExp];
    fetch(0, lo, hi);
    for (int tile = 0; tile < ntiles; ++tile) {
      const int s = tile & 1, u = tile >> 1;
      fetch(tile + 1, nlo, nhi);
      mbar_wait(bar(2 + s), (u & 1) ^ 1);
#pragma unroll
      for (int i = 0; i < kT5RowsPerExp; ++i) {
        uint8_t* dst = s_b + s * kT5BStage + (e + i * kT5ExpThreads) * 16;
        const uint32_t w[8] = {lo[i].x, lo[i].y, lo[i].z, lo[i].w, hi[i].x, hi[i].y, hi[i].z, hi[i].w};
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          const uint32_t x = w[c >> 1], x1 = x >> 1;
          const uint32_t s0 = (c & 1) ? 0x2222u : 0x0000u, s1 = (c & 1) ? 0x3333u : 0x1111u;  // byte replication
          *reinterpret_cast<uint4*>(dst + c * 4096) =
              make_uint4(__byte_perm(x, 0, s0) & 0x08040201u, __byte_perm(x1, 0, s0) & 0x40201008u,
                         __byte_perm(x, 0, s1) & 0x08040201u, __byte_perm(x1, 0, s1) & 0x40201008u);
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_arrive(bar(s));
#pragma unroll
      for (int i = 0; i < kT5RowsPerExp; ++i) { lo[i] = nlo[i]; hi[i] = nhi[i]; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kT5EpiWarps) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kT5TmemCols) : "memory");
  }
  // ---- the block that finishes a query block's last slice merges the slices (ascending rows, the reference's
  // update rule on (best, second) pairs: same as merge_top2_kernel) -- no separate merge launch
  if (nslices > 1 || fx.world > 1) {
    if (t == 0) s_tmem[1] = atomicAdd(g_ticket + blockIdx.x, 1u) == (unsigned)(nslices - 1);
    __syncthreads();
    if (s_tmem[1] && t < kT5Queries) {  // the eight epilogue warps of the last block of this query block
      __threadfence();
      const int qi = qbase + t;
      const bool live = qi < nq;
      int b = 256, sd = 256, id = -1;
      if (live)
        for (int k = 0; k < nslices; ++k) {
          const size_t o = (size_t)k * nq + qi;
          const int ik = __ldcg(o_idx + o), bk = __ldcg(o_best + o), sk = __ldcg(o_second + o);
          if (ik < 0) continue;
          if (bk < b) { sd = min(b, sk); b = bk; id = ik; }
          else sd = min(sd, bk);
        }
      if (fx.world > 1) {
        // ---- exchange across the ranks, fused: this rank's merged triples go into slot `rank` of every
        // rank's gather buffer over NVLink, the query block's sequence number is released to every rank, and once
        // every rank's number for this query block has arrived the `world` slots are merged in ascending row order.
        const size_t nqm = (size_t)fx.nq_max;
        const size_t slot = ((size_t)fx.parity * fx.world + fx.rank) * 3 * nqm;
        if (live)
          for (int p = 0; p < fx.world; ++p) {
            int32_t* dst = fx.peers.buf[p] + slot;
            dst[qi] = id; dst[nqm + qi] = b; dst[2 * nqm + qi] = sd;
          }
        __threadfence_system();
        asm volatile("bar.sync 2, 256;" ::: "memory");
        if (t == 0) {
          const size_t fo = (size_t)kExMaxWorld + (size_t)fx.rank * fx.qb_max + blockIdx.x;
          for (int p = 0; p < fx.world; ++p)
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(fx.peers.flag[p] + fo), "r"(fx.seq) : "memory");
          const unsigned* mine = fx.peers.flag[fx.rank] + kExMaxWorld + blockIdx.x;
          const long long t0 = clock64();
          for (int r = 0; r < fx.world; ++r) {
            unsigned v;
            do {
              asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine + (size_t)r * fx.qb_max) : "memory");
              if (clock64() - t0 > (1ll << 35)) __trap();  // ~17 s: a rank never made the call
            } while ((int)(v - fx.seq) < 0);
          }
          __threadfence_system();
        }
        asm volatile("bar.sync 2, 256;" ::: "memory");
        if (live) {
          const int32_t* g = fx.peers.buf[fx.rank] + (size_t)fx.parity * fx.world * 3 * nqm;
          b = 256; sd = 256; id = -1;
          for (int r = 0; r < fx.world; ++r) {
            const int32_t* e = g + (size_t)r * 3 * nqm;
            const int ik = __ldcg(e + qi), bk = __ldcg(e + nqm + qi), sk = __ldcg(e + 2 * nqm + qi);
            if (ik < 0) continue;
            if (bk < b) { sd = min(b, sk); b = bk; id = ik; }
            else sd = min(sd, bk);
          }
        }
      }
      if (live) { f_idx[qi] = id; f_best[qi] = b; f_second[qi] = sd; }
    }
    if (s_tmem[1] && t == 0) g_ticket[blockIdx.x] = 0;  // ready for the next call
  }
}

int launch_bruteforce_tc5(dim3 grid, cudaStream_t s, const uint4* q, int nq, const uint4* db, long long ndb,
                          long long slice_rows, int32_t idx_base, int32_t* o_idx, int32_t* o_best,
                          int32_t* o_second, unsigned* g_thr, unsigned* g_ticket, int32_t* f_idx, int32_t* f_best,
                          int32_t* f_second, const FusedExchange* fx) {
  ORB_CUDA_TRY(cudaFuncSetAttribute(bruteforce_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kT5SmemBytes));
  FusedExchange f;
  if (fx) f = *fx;
  else { memset(&f, 0, sizeof(f)); f.world = 1; }
  bruteforce_tc5_kernel<<<grid, kT5Threads, kT5SmemBytes, s>>>(q, nq, db, ndb, slice_rows, idx_base, o_idx, o_best,
                                                              o_second, g_thr, g_ticket, (int)grid.y, f_idx, f_best,
                                                              f_second, f);
  return ORB_GPU_OK;
}

// Peak probe: thread 0 of one block per SM issues `n` M128 N256 K32 kind::i8 MMAs back to back on fixed operands
// (the shape and operand layout of bruteforce_tc5_kernel) and waits for the commit.
__global__ void __launch_bounds__(128, 1) umma_peak_kernel(int n) {
  extern __shared__ __align__(128) uint8_t t5_smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(t5_smem + kT5ATile + kT5BStage);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 1);
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < (int)(kT5ATile + kT5BStage) / 4; i += 128) reinterpret_cast<uint32_t*>(t5_smem)[i] = 0x01010101u;
  if (t == 0) {
    mbar_init(smem_u32(bars), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)),
                 "r"(kT5TmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  if (t == 0) {
    const uint64_t da = umma_desc(smem_u32(t5_smem), 2048, 128), db = umma_desc(smem_u32(t5_smem + kT5ATile), 4096, 128);
    for (int i = 0; i < n; ++i) {
      const int k = i & 7;
      umma_i8(tmem_base + ((i >> 3) & 1) * 256, da + (uint64_t)(k * 256), db + (uint64_t)(k * 512), i > 15);
    }
    umma_commit(smem_u32(bars));
    mbar_wait(smem_u32(bars), 0);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kT5TmemCols) : "memory");
  }
}

}  // namespace orbgpu

extern "C" int orb_gpu_umma_peak(int device, double* mac_per_s) {
  using namespace orbgpu;
  if (!mac_per_s) return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(device));
  int sms = 148;
  ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  const int smem = (int)(kT5ATile + kT5BStage) + 64, n = 8192;
  ORB_CUDA_TRY(cudaFuncSetAttribute(umma_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  ORB_CUDA_TRY(cudaEventCreate(&e0));
  ORB_CUDA_TRY(cudaEventCreate(&e1));
  umma_peak_kernel<<<sms, 128, smem>>>(n);
  float best = 1e30f;
  for (int r = 0; r < 3; ++r) {
    ORB_CUDA_TRY(cudaEventRecord(e0));
    umma_peak_kernel<<<sms, 128, smem>>>(n);
    ORB_CUDA_TRY(cudaEventRecord(e1));
    ORB_CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    ORB_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    best = ms < best ? ms : best;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  ORB_CUDA_TRY(cudaGetLastError());
  *mac_per_s = (double)sms * n * 128.0 * 256.0 * 32.0 / (best * 1e-3);
  return ORB_GPU_OK;
}
