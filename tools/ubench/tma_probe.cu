// Bring-up probe for blur_tc_kernels.cu: (1) one TMA box load (128 x 136 bytes, 128-byte swizzle) of a u8 image with
// negative / out-of-range coordinates, checked against the expected shared-memory layout; (2) the vertical blur pass
// as one tcgen05.mma kind::i8 sequence: A = banded Toeplitz matrix (K-major, unswizzled), B = the TMA tile (MN-major,
// 128-byte swizzle, two atoms), compared with the host column sums.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tma_probe tma_probe.cu      usage: tma_probe <mode> [lbo sbo]
//   mode 0: TMA load, tensor map in global memory     mode 1: tensor map inside a __grid_constant__ parameter struct
//   (both work; what faults with 'illegal instruction' is a box whose first column is not a multiple of 16 bytes)
//   mode 2: TMA load + MMA (MN-major B)
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ volatile int* g_flags;  // host-mapped progress markers (readable after a fault)
__device__ __forceinline__ void mark(int i, int v) { if (g_flags) { g_flags[i] = v; __threadfence_system(); } }
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  long long t0 = clock64();
  for (;;) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return true;
    if (clock64() - t0 > 400000000LL) return false;  // 0.2 s
  }
}

struct Maps { CUtensorMap in[4]; int pad[5]; };

constexpr int kAtom = 160 * 128;

__device__ void load_tile(uint8_t* sm, const void* tmap, int x, int y, uint32_t bar, int natoms) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(natoms * 136 * 128) : "memory");
  for (int a = 0; a < natoms; ++a)
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(sm + a * kAtom)), "l"(tmap), "r"(bar), "r"(x + 128 * a), "r"(y) : "memory");
}

__global__ void __launch_bounds__(128, 1) k_direct(const CUtensorMap* m, int x, int y, uint8_t* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = raw + (((smem_u32(raw) + 1023u) & ~1023u) - smem_u32(raw));
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 2 * kAtom);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * kAtom; i += blockDim.x) sm[i] = 0xcd;
  __syncthreads();
  if (threadIdx.x == 0) { mark(0, 1); load_tile(sm, m, x, y, smem_u32(bar), 2); mark(1, 1); }
  const bool ok = mbar_wait(smem_u32(bar), 0);
  if (threadIdx.x == 0) mark(2, ok ? 1 : -1);
  for (int i = threadIdx.x; i < 2 * kAtom; i += blockDim.x) out[i] = sm[i];
  if (threadIdx.x == 0) mark(3, 1);
}

__global__ void __launch_bounds__(128, 1) k_struct(const __grid_constant__ Maps P, int j, int x, int y, uint8_t* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = raw + (((smem_u32(raw) + 1023u) & ~1023u) - smem_u32(raw));
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 2 * kAtom);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) load_tile(sm, &P.in[j], x, y, smem_u32(bar), 2);
  mbar_wait(smem_u32(bar), 0);
  for (int i = threadIdx.x; i < 2 * kAtom; i += blockDim.x) out[i] = sm[i];
}

// TMA tile + Toeplitz A -> D[128][256] s32
__global__ void __launch_bounds__(128, 1) k_mma(const CUtensorMap* m, int x, int y, uint32_t lbo, uint32_t sbo,
                                                uint32_t idesc, int32_t* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = raw + (((smem_u32(raw) + 1023u) & ~1023u) - smem_u32(raw));
  uint8_t* s_t = sm + 2 * kAtom;  // 128 x 160, K-major core matrices: (k>>4)*2048 + (r>>3)*128 + (r&7)*16 + (k&15)
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_t + 128 * 160);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar + 2);
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < 128 * 160; i += 128) s_t[i] = 0;
  __syncthreads();
  const int kw[7] = {18, 34, 48, 56, 48, 34, 18};
  for (int i = 0; i < 7; ++i) { const int r = t, k = r + i; s_t[(k >> 4) * 2048 + (r >> 3) * 128 + (r & 7) * 16 + (k & 15)] = kw[i]; }
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + 1)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;
  if (t == 0) {
    load_tile(sm, m, x, y, smem_u32(bar), 2);
    mbar_wait(smem_u32(bar), 0);
    const uint64_t da = (uint64_t)((smem_u32(s_t) & 0x3ffffu) >> 4) | ((uint64_t)(2048 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
    const uint64_t db = (uint64_t)((smem_u32(sm) & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) | (2ull << 61);
    for (int k = 0; k < 5; ++k)
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                   ::"r"(tmem), "l"(da + (uint64_t)(k * 256)), "l"(db + (uint64_t)(k * 256)), "r"(idesc), "r"(k > 0 ? 1u : 0u) : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar + 1)) : "memory");
  }
  mbar_wait(smem_u32(bar + 1), 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c = 0; c < 256; c += 8) {
    uint32_t v[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(tmem + ((uint32_t)(warp * 32) << 16) + c));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 8; ++i) out[t * 256 + c + i] = (int32_t)v[i];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

// mode 3: 1-D bulk copy (no tensor map)
__global__ void __launch_bounds__(128, 1) k_bulk1d(const uint8_t* src, uint8_t* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = raw + (((smem_u32(raw) + 1023u) & ~1023u) - smem_u32(raw));
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 2 * kAtom);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mark(0, 1);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(4096) : "memory");
    mark(4, 1);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(sm)), "l"(src), "r"(4096), "r"(smem_u32(bar)) : "memory");
    mark(1, 1);
  }
  const bool ok = mbar_wait(smem_u32(bar), 0);
  if (threadIdx.x == 0) mark(2, ok ? 1 : -1);
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) out[i] = sm[i];
  if (threadIdx.x == 0) mark(3, 1);
}
// modes 4, 5: tensor map read from global memory / small unswizzled box; `bytes` = box bytes
__global__ void __launch_bounds__(128, 1) k_gmap(const CUtensorMap* gm, int x, int y, int bytes, uint8_t* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = raw + (((smem_u32(raw) + 1023u) & ~1023u) - smem_u32(raw));
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 2 * kAtom);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mark(0, 1);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    mark(4, 1);
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(sm)), "l"(gm), "r"(smem_u32(bar)), "r"(x), "r"(y) : "memory");
    mark(1, 1);
  }
  const bool ok = mbar_wait(smem_u32(bar), 0);
  if (threadIdx.x == 0) mark(2, ok ? 1 : -1);
  for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = sm[i];
  if (threadIdx.x == 0) mark(3, 1);
}

// mode 8: horizontal resize pass as MMAs: A = image tile from TMA (K-major, SWIZZLE_128B, two atoms of 128 columns),
// B = weight matrices [N = 128 destination columns][K = 192 source columns] (K-major, unswizzled core matrices; hi
// weights x 16 and lo weights), D_hi / D_lo in two 128-column accumulators.  Host check: D_hi + (D_lo >> 4) == T >> 4.
__global__ void __launch_bounds__(128, 1) k_hres(const CUtensorMap* m, int x, int y, const uint8_t* w_hi, const uint8_t* w_lo,
                                                 uint32_t a_lbo, int32_t* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = raw + (((smem_u32(raw) + 1023u) & ~1023u) - smem_u32(raw));
  uint8_t* s_a = sm;                       // 2 atoms x 128 rows x 128 B
  uint8_t* s_w = sm + 2 * 16384;           // 2 x [128][192]: (k>>4)*2048 + (n>>3)*128 + (n&7)*16 + (k&15)
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_w + 2 * 128 * 192);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar + 2);
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < 128 * 192; i += 128) { s_w[i] = w_hi[i]; s_w[128 * 192 + i] = w_lo[i]; }
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + 1)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;
  if (t == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(2 * 16384) : "memory");
    for (int a = 0; a < 2; ++a)
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                   ::"r"(smem_u32(s_a + a * 16384)), "l"(m), "r"(smem_u32(bar)), "r"(x + 128 * a), "r"(y) : "memory");
    mbar_wait(smem_u32(bar), 0);
    // instruction descriptor: D s32, A u8 K-major, B u8 K-major, N = 128, M = 128
    const uint32_t idesc = (2u << 4) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
    for (int half = 0; half < 2; ++half)
      for (int k = 0; k < 6; ++k) {
        const uint32_t a_addr = smem_u32(s_a + (k >> 2) * 16384) + (k & 3) * 32;
        const uint64_t da = (uint64_t)((a_addr & 0x3ffffu) >> 4) | ((uint64_t)(a_lbo >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
        const uint32_t b_addr = smem_u32(s_w + half * 128 * 192) + k * 2 * 2048;
        const uint64_t db = (uint64_t)((b_addr & 0x3ffffu) >> 4) | ((uint64_t)(2048 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(tmem + half * 128), "l"(da), "l"(db), "r"(idesc), "r"(k > 0 ? 1u : 0u) : "memory");
      }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar + 1)) : "memory");
  }
  mbar_wait(smem_u32(bar + 1), 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c = 0; c < 256; c += 8) {
    uint32_t v[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(tmem + ((uint32_t)(warp * 32) << 16) + c));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 8; ++i) out[t * 256 + c + i] = (int32_t)v[i];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  const int mode = argc > 1 ? atoi(argv[1]) : 0;
  const int W = 640, H = 480, x0 = argc > 4 ? atoi(argv[4]) : -3, y0 = argc > 5 ? atoi(argv[5]) : -3;
  std::vector<uint8_t> img((size_t)W * H);
  srand(7);
  for (auto& v : img) v = (uint8_t)(rand() >> 5);
  uint8_t* d_img; CK(cudaMalloc(&d_img, img.size()));
  CK(cudaMemcpy(d_img, img.data(), img.size(), cudaMemcpyHostToDevice));
  int* h_flags; CK(cudaHostAlloc(&h_flags, 64, cudaHostAllocMapped));
  memset(h_flags, 0, 64);
  { int* d_flags; CK(cudaHostGetDevicePointer(&d_flags, h_flags, 0)); CK(cudaMemcpyToSymbol(g_flags, &d_flags, sizeof(d_flags))); }
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  if (!fp || q != cudaDriverEntryPointSuccess) { printf("no cuTensorMapEncodeTiled (%d)\n", (int)q); return 1; }
  const cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H};
  const cuuint64_t strides[1] = {(cuuint64_t)W};
  const cuuint32_t box[2] = {128, 136};
  const cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = ((EncodeTiledFn)fp)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d_img, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode rc=%d  sizeof(CUtensorMap)=%zu alignof=%zu\n", (int)r, sizeof(CUtensorMap), alignof(CUtensorMap));
  if (r != CUDA_SUCCESS) return 1;
  const int smem = 2 * kAtom + 128 * 160 + 1024 + 256;
  CUtensorMap* d_map; CK(cudaMalloc(&d_map, sizeof(m)));
  CK(cudaMemcpy(d_map, &m, sizeof(m), cudaMemcpyHostToDevice));
  auto expect = [&](int k, int n) -> int { const int y = y0 + k, x = x0 + n; return (y < 0 || y >= H || x < 0 || x >= W) ? 0 : img[(size_t)y * W + x]; };
  if (mode == 8) {
    // box 128 columns x 128 rows for the A tile
    const cuuint32_t boxa[2] = {128, 128};
    CUtensorMap ma;
    r = ((EncodeTiledFn)fp)(&ma, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d_img, dims, strides, boxa, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode A rc=%d\n", (int)r); return 1; }
    CUtensorMap* d_ma; CK(cudaMalloc(&d_ma, sizeof(ma))); CK(cudaMemcpy(d_ma, &ma, sizeof(ma), cudaMemcpyHostToDevice));
    // destination column n takes source columns xo[n], xo[n] + 1 of the tile with 11-bit weights a0 + a1 = 2048
    std::vector<int> xo(128), a0(128), a1(128);
    std::vector<uint8_t> whi(128 * 192, 0), wlo(128 * 192, 0);
    auto at = [](int n, int k) { return (k >> 4) * 2048 + (n >> 3) * 128 + (n & 7) * 16 + (k & 15); };
    for (int n = 0; n < 128; ++n) {
      xo[n] = (int)(n * 1.2) + 5; a1[n] = (n * 37) % 2049; a0[n] = 2048 - a1[n];
      whi[at(n, xo[n])] = (uint8_t)((a0[n] >> 8) * 16); wlo[at(n, xo[n])] = (uint8_t)(a0[n] & 255);
      whi[at(n, xo[n] + 1)] = (uint8_t)((a1[n] >> 8) * 16); wlo[at(n, xo[n] + 1)] = (uint8_t)(a1[n] & 255);
    }
    uint8_t *d_whi, *d_wlo; CK(cudaMalloc(&d_whi, whi.size())); CK(cudaMalloc(&d_wlo, wlo.size()));
    CK(cudaMemcpy(d_whi, whi.data(), whi.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(d_wlo, wlo.data(), wlo.size(), cudaMemcpyHostToDevice));
    int32_t* d_out; CK(cudaMalloc(&d_out, 128 * 256 * 4));
    const int smem8 = 2 * 16384 + 2 * 128 * 192 + 1024 + 256;
    CK(cudaFuncSetAttribute(k_hres, cudaFuncAttributeMaxDynamicSharedMemorySize, smem8));
    const uint32_t a_lbo = argc > 2 ? atoi(argv[2]) : 16;
    k_hres<<<1, 128, smem8>>>(d_ma, x0, y0, d_whi, d_wlo, a_lbo, d_out);
    CK(cudaGetLastError());
    { cudaError_t e = cudaDeviceSynchronize(); if (e) { printf("mode 8 sync: %s\n", cudaGetErrorString(e)); return 1; } }
    std::vector<int32_t> out(128 * 256);
    CK(cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost));
    long bad = 0;
    for (int rr = 0; rr < 128; ++rr)
      for (int n = 0; n < 128; ++n) {
        const int T = a0[n] * expect(rr, xo[n]) + a1[n] * expect(rr, xo[n] + 1);
        const int got = out[rr * 256 + n] + (out[rr * 256 + 128 + n] >> 4);
        if (got != (T >> 4)) { if (bad < 6) printf("  (row=%d, n=%d) hi %d lo %d -> %d want %d\n", rr, n, out[rr * 256 + n], out[rr * 256 + 128 + n], got, T >> 4); ++bad; }
      }
    printf("mode 8 (A lbo %u): horizontal-pass mismatches: %ld of %d\n", a_lbo, bad, 128 * 128);
    return 0;
  }
  if (mode >= 3 && mode <= 6) {
    uint8_t* d_out; CK(cudaMalloc(&d_out, 2 * kAtom));
    int bytes = 4096;
    if (mode == 3) {
      CK(cudaFuncSetAttribute(k_bulk1d, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      k_bulk1d<<<1, 128, smem>>>(d_img, d_out);
    } else {
      CUtensorMap mm = m;
      if (mode >= 5) {
        const cuuint32_t box2[2] = {64, 8};
        r = ((EncodeTiledFn)fp)(&mm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d_img, dims, strides, box2, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                mode == 6 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("encode small rc=%d\n", (int)r);
        bytes = 512;
      } else bytes = 136 * 128;
      CUtensorMap* d_m; CK(cudaMalloc(&d_m, sizeof(mm)));
      CK(cudaMemcpy(d_m, &mm, sizeof(mm), cudaMemcpyHostToDevice));
      CK(cudaFuncSetAttribute(k_gmap, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      k_gmap<<<1, 128, smem>>>(d_m, x0, y0, bytes, d_out);
    }
    CK(cudaGetLastError());
    cudaError_t e = cudaDeviceSynchronize();
    printf("mode %d sync: %s; flags start=%d expect_tx=%d issued=%d wait=%d dumped=%d\n", mode, cudaGetErrorString(e), h_flags[0], h_flags[4], h_flags[1], h_flags[2], h_flags[3]);
    if (e) return 1;
    std::vector<uint8_t> out(bytes);
    CK(cudaMemcpy(out.data(), d_out, bytes, cudaMemcpyDeviceToHost));
    printf("  first bytes: %d %d %d %d (image: %d %d %d %d)\n", out[0], out[1], out[2], out[3], img[0], img[1], img[2], img[3]);
    return 0;
  }
  if (mode == 0 || mode == 1) {
    uint8_t* d_out; CK(cudaMalloc(&d_out, 2 * kAtom));
    CK(cudaMemset(d_out, 0xee, 2 * kAtom));
    if (mode == 0) {
      CK(cudaFuncSetAttribute(k_direct, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      k_direct<<<1, 128, smem>>>(d_map, x0, y0, d_out);
    } else {
      Maps P; memset(&P, 0, sizeof(P)); for (int i = 0; i < 4; ++i) P.in[i] = m;
      CK(cudaFuncSetAttribute(k_struct, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      k_struct<<<1, 128, smem>>>(P, 2, x0, y0, d_out);
    }
    CK(cudaGetLastError());
    { cudaError_t e = cudaDeviceSynchronize(); printf("sync: %s; flags issue=%d issued=%d wait=%d dumped=%d\n", cudaGetErrorString(e), h_flags[0], h_flags[1], h_flags[2], h_flags[3]); if (e) return 1; }
    std::vector<uint8_t> out(2 * kAtom);
    CK(cudaMemcpy(out.data(), d_out, out.size(), cudaMemcpyDeviceToHost));
    long bad = 0;
    for (int k = 0; k < 136; ++k)
      for (int n = 0; n < 256; ++n) {
        const int got = out[(n >> 7) * kAtom + k * 128 + (((((n & 127) >> 4) ^ (k & 7)) << 4) | (n & 15))];
        if (k == 0 && n < 3) printf("  sample (k=%d,n=%d) got %d\n", k, n, got);
        if (got != expect(k, n)) { if (bad < 5) printf("  (k=%d, n=%d) got %d want %d\n", k, n, got, expect(k, n)); ++bad; }
      }
    printf("mode %d: TMA tile layout mismatches: %ld of %d\n", mode, bad, 136 * 256);
  } else {
    const uint32_t lbo = argc > 2 ? atoi(argv[2]) : kAtom, sbo = argc > 3 ? atoi(argv[3]) : 1024;
    const uint32_t idesc = (2u << 4) | (1u << 16) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
    int32_t* d_out; CK(cudaMalloc(&d_out, 128 * 256 * 4));
    CK(cudaFuncSetAttribute(k_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    k_mma<<<1, 128, smem>>>(d_map, x0, y0, lbo, sbo, idesc, d_out);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<int32_t> out(128 * 256);
    CK(cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost));
    const int kw[7] = {18, 34, 48, 56, 48, 34, 18};
    long bad = 0;
    for (int rr = 0; rr < 128; ++rr)
      for (int n = 0; n < 256; ++n) {
        int want = 0;
        for (int i = 0; i < 7; ++i) want += kw[i] * expect(rr + i, n);
        if (out[rr * 256 + n] != want) { if (bad < 6) printf("  (r=%d, n=%d) got %d want %d\n", rr, n, out[rr * 256 + n], want); ++bad; }
      }
    printf("mode 2 (lbo %u sbo %u): vertical-pass mismatches: %ld of %d\n", lbo, sbo, bad, 128 * 256);
  }
  return 0;
}
