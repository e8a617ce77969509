// Probes around tcgen05.mma kind::i8 issue (B200):
//  1. A operand in tensor memory: A[128][32] (s8) is stored by tcgen05.st (thread = TMEM lane = row, register j = bytes
//     K = 4j .. 4j+3), B[64][32] (u8) sits in shared memory as unswizzled K-major core matrices; D = A * B^T is compared
//     with the host (bit-exact).
//  2. rate(): one lane per SM issues M128 x N x K32 MMAs back to back, A from shared memory or from tensor memory, with
//     an optional tcgen05.commit every c MMAs: both operand sources run at N/2 clk per MMA (N >= 192), a commit after
//     every MMA costs nothing, and any extra instructions in the issuing loop show up directly (the loop, not the tensor
//     pipe, sets the rate below ~75-135 clk per MMA).
//  3. rate2<FORM>(): 16 MMAs unrolled with descriptors as base + constant, issued by lane 0 alone (FORM 0) or by the
//     converged warp with elect.sync inside the asm (FORM 1): 48 clk per MMA at N = 64, 128 at N = 256 -- every
//     descriptor sits in its own uniform register pair.  A loop that rewrites ONE pair before each MMA (what a
//     run-time selected descriptor compiles to) measured N/2 + 40 clk per MMA.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o tmem_a_probe tmem_a_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1) probe(const uint8_t* a_rows /*[128][64]*/, const uint8_t* b_img /*[2 ksteps][core layout 64 rows]*/, uint32_t idesc, int ksteps, int32_t* d_out) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint8_t* s_b = sm;  // K-major: (k >> 4) * 1024 + (n >> 3) * 128 + (n & 7) * 16 + (k & 15) for 64 rows
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 8192);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(sm + 8192 + 8);
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < 64 * 64; i += 128) s_b[i] = b_img[i];
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;
  // A: this thread's row, 64 bytes = 16 registers, into columns 64..79
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(a_rows + t * 64);
    uint32_t v[16];
    for (int j = 0; j < 16; ++j) v[j] = src[j];
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 ::"r"(tmem + ((uint32_t)(warp * 32) << 16) + 64), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]),
                   "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (t == 0) {
    for (int k = 0; k < ksteps; ++k) {
      const uint32_t baddr = smem_u32(s_b) + k * 2 * 1024;
      const uint64_t db = (uint64_t)((baddr & 0x3ffffu) >> 4) | ((uint64_t)(1024 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
                   ::"r"(tmem), "r"(tmem + 64 + 8 * k), "l"(db), "r"(idesc), "r"(k > 0 ? 1u : 0u) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
  }
  {
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c = 0; c < 64; c += 8) {
    uint32_t v[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(tmem + ((uint32_t)(warp * 32) << 16) + c));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 8; ++i) d_out[t * 64 + c + i] = (int32_t)v[i];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
}

// Issue rate: lane 0 of one block per SM issues n x 8 MMAs (M128, N, K32) back to back, A from shared memory (mode 0)
// or from tensor memory (mode 1), two accumulators alternating when 2N + 64 <= 512 columns.
__global__ void __launch_bounds__(128, 1) rate(int mode, int N, int n, long long* clk, int commit_every) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint8_t* s_a = sm;                 // 128 x 256 bytes
  uint8_t* s_b = sm + 32768;         // N x 256 bytes
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 32768 + 65536);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar + 2);
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < (32768 + 65536) / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0x01010101u;
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + 1)) : "memory");
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
  const bool two = 2 * N + (mode ? 64 : 0) <= 512;
  const uint32_t acol = two ? 2 * N : N;
  if (t == 0) {
    const uint32_t idesc = (2u << 4) | (1u << 7) | (((uint32_t)N >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t da = (uint64_t)((smem_u32(s_a) & 0x3ffffu) >> 4) | ((uint64_t)(2048 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
    const uint64_t db = (uint64_t)((smem_u32(s_b) & 0x3ffffu) >> 4) | ((uint64_t)(16 * N >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
    const long long c0 = clock64();
    // flat loop, one MMA per iteration (an unrolled loop with the mode test inside issues no faster than one MMA per 93 clk)
    const uint32_t kb = (uint32_t)(2 * 16 * N >> 4);
    const int cmask = commit_every ? commit_every - 1 : 0x7fffffff;
    if (mode == 0) {
      for (int i = 0; i < 8 * n; ++i) {
        const int k = i & 7;
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(tmem + (two ? ((i >> 3) & 1) * N : 0)), "l"(da + (uint64_t)(k * 256)), "l"(db + (uint64_t)(k * kb)), "r"(idesc), "r"(i > 15 ? 1u : 0u) : "memory");
        if ((i & cmask) == cmask)  // a commit nobody waits for (the phase just flips)
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar + 1)) : "memory");
      }
    } else {
      for (int i = 0; i < 8 * n; ++i) {
        const int k = i & 7;
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
                     ::"r"(tmem + (two ? ((i >> 3) & 1) * N : 0)), "r"(tmem + acol + 8 * k), "l"(db + (uint64_t)(k * kb)), "r"(idesc), "r"(i > 15 ? 1u : 0u) : "memory");
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)) : "memory");
    if (blockIdx.x == 0) *clk = clock64() - c0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

static void run_rate(int mode, int N, int n = 8192, int commit_every = 0) {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int smem = 32768 + 65536 + 64;
  cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  static long long* d_clk = nullptr;
  if (!d_clk) cudaMalloc(&d_clk, 8);
  for (int r = 0; r < 40; ++r) rate<<<sms, 128, smem>>>(mode, N, n, d_clk, commit_every);
  float best = 1e9f;
  for (int r = 0; r < 12; ++r) {
    cudaEventRecord(e0);
    rate<<<sms, 128, smem>>>(mode, N, n, d_clk, commit_every);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  const cudaError_t e = cudaDeviceSynchronize();
  const double mac = (double)sms * n * 8 * 128.0 * N * 32.0 / (best * 1e-3);
  long long clk = 0;
  cudaMemcpy(&clk, d_clk, 8, cudaMemcpyDeviceToHost);
  printf("rate (%5d x 8 MMAs, commit every %d): A from %s, N = %3d: %.3f ms, %.1f TMAC/s (%.0f Gpairs/s of 256-bit rows), %.1f clk per MMA (ideal %d), %.0f MHz  [%s]\n",
         n, commit_every, mode ? "TMEM" : "smem", N, best, mac * 1e-12, mac / 256 * 1e-9, (double)clk / (8.0 * n), N / 2, (double)clk / (best * 1e3),
         cudaGetErrorString(e));
}

// Issue-cost variants (A in shared memory, two accumulators): FORM 0 = lane 0 alone, 8 MMAs unrolled, descriptors as
// base + k * step; FORM 1 = the whole warp runs the loop and elect.sync picks the issuing lane inside the asm.
__device__ __forceinline__ void mma_plain(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_elect(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t@e tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
template <int FORM, int N>
__global__ void __launch_bounds__(128, 1) rate2(int n, long long* clk) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint8_t* s_a = sm;
  uint8_t* s_b = sm + 32768;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 32768 + 65536);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar + 2);
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < (32768 + 65536) / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0x01010101u;
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
  constexpr uint32_t idesc = (2u << 4) | (1u << 7) | (((uint32_t)N >> 3) << 17) | ((128u >> 4) << 24);
  const uint64_t da = (uint64_t)((smem_u32(s_a) & 0x3ffffu) >> 4) | ((uint64_t)(2048 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
  const uint64_t db = (uint64_t)((smem_u32(s_b) & 0x3ffffu) >> 4) | ((uint64_t)(16 * N >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
  constexpr uint32_t kb = 2 * 16 * N >> 4;
  if (FORM == 0 ? t == 0 : warp == 0) {
    const long long c0 = clock64();
    for (int i = 0; i < n; i += 2) {
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (FORM == 0) mma_plain(tmem + h * N, da + (uint64_t)(k * 256), db + (uint64_t)(k * kb), idesc, (i > 0 || k > 0) ? 1u : 0u);
          else mma_elect(tmem + h * N, da + (uint64_t)(k * 256), db + (uint64_t)(k * kb), idesc, (i > 0 || k > 0) ? 1u : 0u);
        }
    }
    if (t == 0) {
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
      uint32_t done = 0;
      while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)) : "memory");
      if (blockIdx.x == 0) *clk = clock64() - c0;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}
template <int FORM, int N>
static void run_rate2() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int smem = 32768 + 65536 + 64, n = 8192;
  cudaFuncSetAttribute(rate2<FORM, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  long long* d_clk; cudaMalloc(&d_clk, 8);
  for (int r = 0; r < 10; ++r) rate2<FORM, N><<<sms, 128, smem>>>(n, d_clk);
  const cudaError_t e = cudaDeviceSynchronize();
  long long clk = 0;
  cudaMemcpy(&clk, d_clk, 8, cudaMemcpyDeviceToHost);
  printf("issue form %d (%s), N = %3d: %.1f clk per MMA (tensor time %d)  [%s]\n", FORM, FORM ? "whole warp + elect.sync" : "lane 0, unrolled x16",
         N, (double)clk / (8.0 * n), N / 2, cudaGetErrorString(e));
  cudaFree(d_clk);
}

int main() {
  const int KS = 2, K = 32 * KS;
  std::vector<int8_t> A(128 * 64);
  std::vector<uint8_t> B(64 * 64), Bimg(64 * 64, 0);
  srand(3);
  for (auto& v : A) v = (int8_t)(rand() % 255 - 127);
  for (auto& v : B) v = (uint8_t)(rand() & 255);
  for (int n = 0; n < 64; ++n)
    for (int k = 0; k < K; ++k) Bimg[(k >> 4) * 1024 + (n >> 3) * 128 + (n & 7) * 16 + (k & 15)] = B[n * 64 + k];
  uint8_t *dA, *dB; int32_t* dD;
  cudaMalloc(&dA, A.size()); cudaMalloc(&dB, Bimg.size()); cudaMalloc(&dD, 128 * 64 * 4);
  cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice); cudaMemcpy(dB, Bimg.data(), Bimg.size(), cudaMemcpyHostToDevice);
  // D s32, A s8 K-major, B u8 K-major, N = 64, M = 128
  const uint32_t idesc = (2u << 4) | (1u << 7) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
  probe<<<1, 128, 8192 + 64>>>(dA, dB, idesc, KS, dD);
  cudaError_t e = cudaDeviceSynchronize();
  printf("sync: %s\n", cudaGetErrorString(e));
  if (e) return 1;
  std::vector<int32_t> D(128 * 64);
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  long bad = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < 64; ++n) {
      int want = 0;
      for (int k = 0; k < K; ++k) want += (int)A[m * 64 + k] * (int)B[n * 64 + k];
      if (D[m * 64 + n] != want) { if (bad < 6) printf("  (m=%d, n=%d) got %d want %d\n", m, n, D[m * 64 + n], want); ++bad; }
    }
  printf("A-from-TMEM kind::i8: mismatches %ld of %d\n", bad, 128 * 64);
  for (int N : {256, 192})
    for (int mode = 0; mode < 2; ++mode) run_rate(mode, N);
  for (int ce : {8, 1}) run_rate(0, 256, 8192, ce);
  run_rate2<0, 64>(); run_rate2<1, 64>(); run_rate2<0, 256>(); run_rate2<1, 256>();
  return 0;
}
