// Layout probe for tcgen05.mma kind::i8 with unswizzled K-major shared-memory descriptors: one M128 N256 K32 MMA
// on random s8 x u8 operands, compared with the host product for several (LBO, SBO) readings.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tc5_layout tc5_layout.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1)
probe(const uint8_t* a_img, const uint8_t* b_img, uint32_t lboA, uint32_t sboA, uint32_t lboB, uint32_t sboB,
      uint32_t idesc, int32_t* d_out) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint8_t* s_a = sm;            // 4 KB
  uint8_t* s_b = sm + 4096;     // 8 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 12288);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(sm + 12288 + 8);
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < 4096; i += 128) s_a[i] = a_img[i];
  for (int i = t; i < 8192; i += 128) s_b[i] = b_img[i];
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
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
    const uint64_t da = (uint64_t)((smem_u32(s_a) & 0x3ffffu) >> 4) | ((uint64_t)(lboA >> 4) << 16) | ((uint64_t)(sboA >> 4) << 32) | (1ull << 46);
    const uint64_t db = (uint64_t)((smem_u32(s_b) & 0x3ffffu) >> 4) | ((uint64_t)(lboB >> 4) << 16) | ((uint64_t)(sboB >> 4) << 32) | (1ull << 46);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(0u) : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
  }
  uint32_t done = 0;
  long long t0 = clock64();
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)), "r"(0u) : "memory");
    if (!done && clock64() - t0 > 2000000000LL) __trap();
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c = 0; c < 256; c += 8) {
    uint32_t v[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(tmem + ((uint32_t)(warp * 32) << 16) + c));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 8; ++j) d_out[t * 256 + c + j] = (int32_t)v[j];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
  }
}

static size_t off(int r, int k, uint32_t lbo, uint32_t sbo) { return (size_t)(k / 16) * lbo + (size_t)(r / 8) * sbo + (r % 8) * 16 + k % 16; }

int main() {
  std::vector<int8_t> A(128 * 32);
  std::vector<uint8_t> B(256 * 32);
  srand(5);
  for (auto& x : A) x = (int8_t)(rand() % 33 - 16);
  for (auto& x : B) x = (uint8_t)(rand() % 17);
  std::vector<int32_t> ref(128 * 256);
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < 256; ++n) {
      int s = 0;
      for (int k = 0; k < 32; ++k) s += (int)A[m * 32 + k] * (int)B[n * 32 + k];
      ref[m * 256 + n] = s;
    }
  uint8_t *da, *db; int32_t* dd;
  cudaMalloc(&da, 4096); cudaMalloc(&db, 8192); cudaMalloc(&dd, 128 * 256 * 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  const uint32_t idesc_sa = (2u << 4) | (1u << 7) | (0u << 10) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
  const uint32_t idesc_sb = (2u << 4) | (0u << 7) | (1u << 10) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
  struct Lay { const char* name; uint32_t lboA, sboA, lboB, sboB; } lays[] = {
      {"chunk-major (LBO=rows*16, SBO=128)", 2048, 128, 4096, 128},
      {"group-major (LBO=128, SBO=256)", 128, 256, 128, 256},
  };
  for (auto& L : lays)
    for (int swap = 0; swap < 2; ++swap)
      for (int fmt = 0; fmt < 2; ++fmt) {
        std::vector<uint8_t> ai(4096, 0), bi(8192, 0);
        for (int r = 0; r < 128; ++r) for (int k = 0; k < 32; ++k) ai[off(r, k, L.lboA, L.sboA)] = (uint8_t)A[r * 32 + k];
        for (int r = 0; r < 256; ++r) for (int k = 0; k < 32; ++k) bi[off(r, k, L.lboB, L.sboB)] = B[r * 32 + k];
        cudaMemcpy(da, ai.data(), 4096, cudaMemcpyHostToDevice);
        cudaMemcpy(db, bi.data(), 8192, cudaMemcpyHostToDevice);
        cudaMemset(dd, 0xff, 128 * 256 * 4);
        if (!swap) probe<<<1, 128, 16384>>>(da, db, L.lboA, L.sboA, L.lboB, L.sboB, fmt ? idesc_sb : idesc_sa, dd);
        else probe<<<1, 128, 16384>>>(da, db, L.sboA, L.lboA, L.sboB, L.lboB, fmt ? idesc_sb : idesc_sa, dd);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s swap=%d fmt=%d: CUDA error %s\n", L.name, swap, fmt, cudaGetErrorString(e)); return 1; }
        std::vector<int32_t> out(128 * 256);
        cudaMemcpy(out.data(), dd, out.size() * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (size_t i = 0; i < out.size(); ++i) bad += out[i] != ref[i];
        printf("%s  descriptor fields %s  idesc %s: mismatches %d / %zu   D[0][0..3] = %d %d %d %d (ref %d %d %d %d)  D[1][0]=%d (ref %d)\n",
               L.name, swap ? "(SBO,LBO) swapped" : "(LBO,SBO)", fmt ? "A=u8,B=s8" : "A=s8,B=u8", bad, out.size(), out[0], out[1], out[2], out[3],
               ref[0], ref[1], ref[2], ref[3], out[256], ref[256]);
      }
  return 0;
}
