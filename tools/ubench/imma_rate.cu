// Micro-benchmark: throughput of legacy mma.sync int8 (IMMA m16n8k32) and b1 and.popc (m16n8k256) on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o imma_rate imma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int ILP>
__global__ void imma_kernel(int iters, int* out) {
  int c[ILP][4];
#pragma unroll
  for (int i = 0; i < ILP; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0;
  unsigned a0 = threadIdx.x * 0x01010101u, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = 0x01ff01ffu, b1 = 0xff01ff01u;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i)
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  int s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void bmma_kernel(int iters, int* out) {
  int c[ILP][4];
#pragma unroll
  for (int i = 0; i < ILP; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0;
  unsigned a0 = threadIdx.x * 0x9e3779b9u, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 ^ 0x55555555u, b1 = a1 ^ 0x33333333u;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i)
      asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.and.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  int s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 0, clk = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  int* out; cudaMalloc(&out, (size_t)sms * 8 * 1024 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int warps = 4; warps <= 32; warps *= 2) {
    for (int mode = 0; mode < 2; ++mode) {
      const int blocks = sms * (warps >= 16 ? 2 : 1), threads = 32 * warps / (warps >= 16 ? 2 : 1);
      if (mode == 0) imma_kernel<8><<<blocks, threads>>>(100, out); else bmma_kernel<8><<<blocks, threads>>>(100, out);
      cudaEventRecord(e0);
      if (mode == 0) imma_kernel<8><<<blocks, threads>>>(iters, out); else bmma_kernel<8><<<blocks, threads>>>(iters, out);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
      const double mmas = (double)blocks * (threads / 32) * iters * 8;
      const double macs = mmas * (mode == 0 ? 16.0 * 8 * 32 : 16.0 * 8 * 256);
      printf("%s warps/SM %2d: %.3f ms  %.1f T(mac|bit-and)/s  = %.0f per clk per SM (at %d MHz)  %s\n", mode ? "b1 and.popc m16n8k256" : "s8 m16n8k32        ",
             warps, ms, macs / ms / 1e9, macs / (ms * 1e-3) / sms / (clk * 1e3), clk / 1000, cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
