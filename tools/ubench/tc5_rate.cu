// Issue rate of tcgen05.mma kind::i8 (M128 N256 K32, unswizzled K-major operands in shared memory): every SM runs one
// CTA whose thread 0 issues `n` MMAs back to back into one accumulator and waits for the commit.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tc5_rate tc5_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1) rate(int n, int nmma_n, long long* cycles) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 98304);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(sm + 98304 + 8);
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < 98304 / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0x01010101u;
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;
  if (t == 0) {
    const uint32_t idesc = (2u << 4) | (1u << 7) | (((uint32_t)nmma_n >> 3) << 17) | ((128u >> 4) << 24);
    // A: 32 KB at 0 (128 rows x 256 bytes, LBO 2048), B: 64 KB at 32768 (256 rows x 256 bytes, LBO 4096)
    const uint64_t da = (uint64_t)((smem_u32(sm) & 0x3ffffu) >> 4) | ((uint64_t)(2048 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
    const uint64_t db = (uint64_t)((smem_u32(sm + 32768) & 0x3ffffu) >> 4) | ((uint64_t)(4096 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
    const long long c0 = clock64();
    for (int i = 0; i < n; ++i) {
      const int k = i & 7;
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem + ((i >> 3) & 1) * 256),
                   "l"(da + (uint64_t)(k * 256)), "l"(db + (uint64_t)(k * 512)), "r"(idesc), "r"((uint32_t)(i > 15))
                   : "memory");
    }
    const long long c1 = clock64();
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
    uint32_t done = 0;
    while (!done) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)), "r"(0u) : "memory");
      if (!done && clock64() - c0 > 4000000000LL) __trap();
    }
    const long long c2 = clock64();
    cycles[2 * blockIdx.x] = c1 - c0;
    cycles[2 * blockIdx.x + 1] = c2 - c0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

int main() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  long long* d; cudaMalloc(&d, sms * 16);
  cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304 + 64);
  for (int nn : {256, 128})
    for (int grid : {1, sms})
      for (int n : {8, 64, 512, 4096}) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        rate<<<grid, 128, 98304 + 64>>>(n, nn, d);  // warm
        cudaEventRecord(e0);
        rate<<<grid, 128, 98304 + 64>>>(n, nn, d);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        const double mac = (double)grid * n * 128.0 * nn * 32.0;
        printf("N=%d grid=%3d n=%4d: issue %8lld clk, done %8lld clk -> %.1f clk/MMA, %.0f MAC/clk/SM; kernel %.3f ms -> %.1f TMAC/s\n", nn, grid, n, h[0], h[1],
               (double)h[1] / n, 128.0 * nn * 32.0 * n / (double)h[1], ms, mac / ms / 1e9);
      }
  return 0;
}
