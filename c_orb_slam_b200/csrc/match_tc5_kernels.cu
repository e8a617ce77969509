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
//   warps 9..16  expanders: one database row per thread and tile of 256 rows -> 16 STS.128 into the K-major,
//                unswizzled core-matrix layout (8 rows x 16 bytes contiguous; LBO = 4096, SBO = 128)
//   warp 8       MMA issuer (lane 0): per tile 2 x 8 tcgen05.mma M128 N256 K32 into two TMEM accumulators
//   warps 0..7   epilogue (set i = warps 4i..4i+3 drains accumulator i): tcgen05.ld 32 columns at a time, max filter against the running second best,
//                packed (S, row) keys; one query per thread and accumulator
// Pipelines: b_full/b_empty (expanders <-> MMA, 2 stages), acc_full/acc_empty (MMA <-> epilogue, the two
// accumulators alternate so the epilogue of one overlaps the MMAs of the other).
#include <cstdlib>

#include "common.cuh"

namespace orbgpu {

constexpr int kT5Queries = 256;
constexpr int kT5TileRows = 256;
constexpr int kT5EpiWarps = 8;                           // two sets of four: set i drains accumulator i
constexpr int kT5ExpWarps = 8;
constexpr int kT5Threads = (kT5EpiWarps + 1 + kT5ExpWarps) * 32;
constexpr uint32_t kT5ATile = 128 * 256;                 // bytes of one expanded A tile
constexpr uint32_t kT5BStage = kT5TileRows * 256;        // bytes of one expanded B stage
constexpr uint32_t kT5SmemBytes = 2 * kT5ATile + 2 * kT5BStage + 128;
constexpr uint32_t kT5TmemCols = 512;
// instruction descriptor (cute/arch/mma_sm100_desc.hpp layout): D = s32, A = signed 8 bit, B = unsigned 8 bit,
// both K-major, N = 256, M = 128
constexpr uint32_t kT5Idesc = (2u << 4) | (1u << 7) | (0u << 10) | ((256u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Spin on a phase parity; a wait that lasts seconds is a protocol bug and traps instead of hanging the device.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  long long t0 = 0;
  for (;;) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    const long long now = clock64();
    if (t0 == 0) t0 = now;
    else if (now - t0 > 6000000000LL) __trap();
  }
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// shared-memory matrix descriptor, no swizzle, K-major: start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 | version 1
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         (1ull << 46);
}
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
// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t (&v)[32], uint32_t taddr) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
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

// top-2 update over 32 accumulator columns (database rows col0 .. col0 + 31 of the slice) of one query
__device__ __forceinline__ void t5_update(const uint32_t (&v)[32], int col0, int rows, unsigned& best,
                                          unsigned& second, int& thr) {
  int m4[4];  // four independent chains: the warp is alone on its scheduler, so the chain depth is what it pays
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m4[i] = max((int)v[i], (int)v[4 + i]);
#pragma unroll
    for (int j = 8; j < 32; j += 8) m4[i] = max(m4[i], max((int)v[j + i], (int)v[j + 4 + i]));
  }
  const int m = max(max(m4[0], m4[1]), max(m4[2], m4[3]));
  // a row only matters if 64 * S beats the current second best; the branch is warp-uniform because the
  // tcgen05.ld that follows is .aligned
  if (__any_sync(0xffffffffu, m > thr)) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int col = col0 + j;
      const unsigned key =
          col < rows ? (unsigned)((int)v[j] * (1 << 16)) + ((256u << 22) + 0x3fffffu - (unsigned)col) : 0u;
      second = max(second, min(best, key));
      best = max(best, key);
    }
    thr = second ? ((int)(second >> 22) - 256) * 64 : -100000;
  }
}

__global__ void __launch_bounds__(kT5Threads, 1)
bruteforce_tc5_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db, long long ndb,
                      long long slice_rows, int32_t idx_base, int32_t* __restrict__ o_idx,
                      int32_t* __restrict__ o_best, int32_t* __restrict__ o_second, int dbg) {
  extern __shared__ __align__(128) uint8_t t5_smem[];
  uint8_t* s_a = t5_smem;
  uint8_t* s_b = t5_smem + 2 * kT5ATile;
  uint64_t* bars = reinterpret_cast<uint64_t*>(t5_smem + 2 * kT5ATile + 2 * kT5BStage);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 8);
  const uint32_t bar0 = smem_u32(bars);
  // barrier slots: 0,1 b_full  2,3 b_empty  4,5 acc_full  6,7 acc_empty
  auto bar = [&](int i) { return bar0 + 8u * i; };

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int slice = blockIdx.y;
  const long long row0 = (long long)slice * slice_rows;
  const int rows = (int)min(slice_rows, ndb - row0);
  const int ntiles = (rows + kT5TileRows - 1) / kT5TileRows;
  const int qbase = blockIdx.x * kT5Queries;

  if (t == 0) {
    mbar_init(bar(0), kT5ExpWarps * 32); mbar_init(bar(1), kT5ExpWarps * 32);
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
    const int e = t - (kT5EpiWarps + 1) * 32;
    const int qi = min(qbase + e, nq - 1);
    const uint4 lo = __ldg(q + 2 * (size_t)qi), hi = __ldg(q + 2 * (size_t)qi + 1);
    const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    uint8_t* dst = s_a + (e >> 7) * kT5ATile + (e & 127) * 16;
#pragma unroll
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
    int thr = -100000;
    const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + acc * 256;
    uint32_t va[32], vb[32];
    for (int tile = 0; tile < ntiles; ++tile) {
      mbar_wait(bar(4 + acc), tile & 1);
      tc_fence_after();
      const int col0 = tile * kT5TileRows;
      tmem_ld32(va, taddr);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_wait(va);
        tmem_ld32(vb, taddr + (2 * c + 1) * 32);
        if (!(dbg & 1)) t5_update(va, col0 + 2 * c * 32, rows, best, second, thr);
        tmem_wait(vb);
        if (c < 3) {
          tmem_ld32(va, taddr + (2 * c + 2) * 32);
        } else {
          tc_fence_before();
          mbar_arrive(bar(6 + acc));  // the accumulator is in registers: the next MMA may overwrite it
        }
        if (!(dbg & 1)) t5_update(vb, col0 + (2 * c + 1) * 32, rows, best, second, thr);
        else best += va[3] + vb[5];
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
    }
  } else if (warp == kT5EpiWarps) {
    // ---- MMA issuer
    if (lane == 0) {
      const uint64_t da[2] = {umma_desc(smem_u32(s_a), 2048, 128), umma_desc(smem_u32(s_a + kT5ATile), 2048, 128)};
      const uint64_t dbs[2] = {umma_desc(smem_u32(s_b), 4096, 128), umma_desc(smem_u32(s_b + kT5BStage), 4096, 128)};
      for (int tile = 0; tile < ntiles; ++tile) {
        const int s = tile & 1, u = tile >> 1;
        mbar_wait(bar(s), u & 1);
        tc_fence_after();
#pragma unroll
        for (int acc = 0; acc < 2; ++acc) {
          mbar_wait(bar(6 + acc), (tile & 1) ^ 1);
          tc_fence_after();
          if (!(dbg & 4))
#pragma unroll
          for (int k = 0; k < 8; ++k)  // K = 32 bytes = two 16-byte chunks per step
            umma_i8(tmem_base + acc * 256, da[acc] + (uint64_t)(k * (2 * 2048 >> 4)),
                    dbs[s] + (uint64_t)(k * (2 * 4096 >> 4)), k > 0);
          umma_commit(bar(4 + acc));
        }
        umma_commit(bar(2 + s));
      }
      // the last commit must have landed before the block (and its shared memory) goes away
      if (ntiles > 0) mbar_wait(bar(2 + ((ntiles - 1) & 1)), ((ntiles - 1) >> 1) & 1);
    }
  } else {
    // ---- expanders: row e of every tile
    const int e = t - (kT5EpiWarps + 1) * 32;
    const uint4* src = db + row0 * 2;
    auto fetch = [&](int tile, uint4& lo, uint4& hi) {
      const int r = tile * kT5TileRows + e;
      if (tile < ntiles && r < rows) { lo = __ldg(src + 2 * (size_t)r); hi = __ldg(src + 2 * (size_t)r + 1); }
      else { lo = make_uint4(0, 0, 0, 0); hi = lo; }
    };
    uint4 lo, hi, nlo, nhi;
    fetch(0, lo, hi);
    for (int tile = 0; tile < ntiles; ++tile) {
      const int s = tile & 1, u = tile >> 1;
      fetch(tile + 1, nlo, nhi);
      mbar_wait(bar(2 + s), (u & 1) ^ 1);
      uint8_t* dst = s_b + s * kT5BStage + e * 16;
      const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
      if (!(dbg & 2))
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const uint32_t x = w[c >> 1], x1 = x >> 1;
        const uint32_t s0 = (c & 1) ? 0x2222u : 0x0000u, s1 = (c & 1) ? 0x3333u : 0x1111u;  // byte replication
        *reinterpret_cast<uint4*>(dst + c * 4096) =
            make_uint4(__byte_perm(x, 0, s0) & 0x08040201u, __byte_perm(x1, 0, s0) & 0x40201008u,
                       __byte_perm(x, 0, s1) & 0x08040201u, __byte_perm(x1, 0, s1) & 0x40201008u);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_arrive(bar(s));
      lo = nlo; hi = nhi;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kT5EpiWarps) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kT5TmemCols) : "memory");
  }
}

int launch_bruteforce_tc5(dim3 grid, cudaStream_t s, const uint4* q, int nq, const uint4* db, long long ndb,
                          long long slice_rows, int32_t idx_base, int32_t* o_idx, int32_t* o_best,
                          int32_t* o_second) {
  static const int dbg = getenv("ORB_GPU_TC5_DEBUG") ? atoi(getenv("ORB_GPU_TC5_DEBUG")) : 0;  // timing experiments only
  ORB_CUDA_TRY(cudaFuncSetAttribute(bruteforce_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kT5SmemBytes));
  bruteforce_tc5_kernel<<<grid, kT5Threads, kT5SmemBytes, s>>>(q, nq, db, ndb, slice_rows, idx_base, o_idx, o_best,
                                                              o_second, dbg);
  return ORB_GPU_OK;
}

}  // namespace orbgpu
