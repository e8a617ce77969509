// MapPoint::ComputeDistinctiveDescriptors (reference src/MapPoint.cc:242-307) for a batch of map points:
// all-pairs Hamming distances between the observed descriptors of a point, the median of every row
// (sorted[0.5*(N-1)], :292-294) and the first row with the smallest median (:296-300).
//
// One block per map point: the point's descriptors are staged in shared memory (32 B each), each warp takes
// rows i, i+4, ...; a row's N distances live in a per-warp shared array and the k-th smallest is found by a
// 9-step bitwise search over the value range [0, 256] with ballot-free per-lane counts + one shuffle
// reduction per step.  The block minimum of (median << 16 | row) is the answer.
#include <climits>

#include "common.cuh"

namespace orbgpu {

constexpr int kDistinctMaxObs = 1024;  // observations per map point handled in shared memory
constexpr int kDistinctWarps = 4;

__global__ void __launch_bounds__(kDistinctWarps * 32)
distinctive_kernel(const uint4* __restrict__ desc, const int* __restrict__ offsets, int npoints, int nmax,
                   int32_t* __restrict__ best_idx, int* __restrict__ status) {
  extern __shared__ uint4 s_desc[];  // [N][2]
  __shared__ unsigned s_best;
  const int p = blockIdx.x, t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int beg = offsets[p], N = offsets[p + 1] - beg;
  if (N <= 0) { if (t == 0) best_idx[p] = -1; return; }
  if (N > nmax) {  // nmax <= kDistinctMaxObs: what the launch sized shared memory for
  if (t == 0) { best_idx[p] = -1; atomicExch(status, ORB_GPU_ERR_CAPACITY); } return; }
  unsigned short* s_row = (unsigned short*)(s_desc + 2 * (size_t)N) + (size_t)warp * kDistinctMaxObs;
  for (int i = t; i < 2 * N; i += blockDim.x) s_desc[i] = desc[2 * (size_t)beg + i];
  if (t == 0) s_best = 0xffffffffu;
  __syncthreads();
  const int k = (int)(0.5 * (double)(N - 1));  // vDists[0.5*(N-1)] (:294): index truncated from double
  for (int i = warp; i < N; i += kDistinctWarps) {
    const uint4 a0 = s_desc[2 * i], a1 = s_desc[2 * i + 1];
    for (int j = lane; j < N; j += 32) {
      const uint4 b0 = s_desc[2 * j], b1 = s_desc[2 * j + 1];
      s_row[j] = (unsigned short)(__popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) +
                                  __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) +
                                  __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w));
    }
    __syncwarp();
    int lo = 0;  // largest v with #{d < v} <= k  ==  k-th smallest
    for (int bit = 8; bit >= 0; --bit) {
      const int cand = lo | (1 << bit);
      int c = 0;
      for (int j = lane; j < N; j += 32) c += s_row[j] < cand;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
      if (c <= k) lo = cand;
    }
    if (lane == 0) atomicMin(&s_best, ((unsigned)lo << 16) | (unsigned)i);  // strict '<' keeps the first row
    __syncwarp();
  }
  __syncthreads();
  if (t == 0) best_idx[p] = (int)(s_best & 0xffffu);
}

int launch_distinctive(const uint8_t* d_desc, const int* d_offsets, int npoints, int max_obs, int32_t* d_best,
                       int* d_status, cudaStream_t s) {
  if (npoints <= 0) return ORB_GPU_OK;
  const int nmax = max_obs < 1 ? 1 : (max_obs > kDistinctMaxObs ? kDistinctMaxObs : max_obs);
  const size_t smem = (size_t)nmax * 32 + (size_t)kDistinctWarps * kDistinctMaxObs * 2;
  if (smem > 48 * 1024)
    ORB_CUDA_TRY(cudaFuncSetAttribute(distinctive_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  distinctive_kernel<<<npoints, kDistinctWarps * 32, smem, s>>>((const uint4*)d_desc, d_offsets, npoints, nmax,
                                                               d_best, d_status);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

}  // namespace orbgpu

using namespace orbgpu;

extern "C" {

int orb_gpu_distinctive_descriptors_device(const uint8_t* d_desc, const int* d_offsets, int npoints, int max_obs,
                                           int32_t* d_best_idx, int* d_status, void* stream) {
  if (npoints < 0 || (npoints > 0 && (!d_desc || !d_offsets || !d_best_idx || !d_status))) return ORB_GPU_ERR_ARG;
  return launch_distinctive(d_desc, d_offsets, npoints, max_obs, d_best_idx, d_status, (cudaStream_t)stream);
}

int orb_gpu_distinctive_descriptors(const uint8_t* desc, const int32_t* offsets, int npoints, int device,
                                    int32_t* best_idx) {
  if (npoints < 0 || (npoints > 0 && (!offsets || !best_idx))) return ORB_GPU_ERR_ARG;
  if (npoints == 0) return ORB_GPU_OK;
  int max_obs = 0;
  for (int p = 0; p < npoints; ++p) {
    if (offsets[p + 1] < offsets[p] || offsets[0] != 0) return ORB_GPU_ERR_ARG;
    max_obs = offsets[p + 1] - offsets[p] > max_obs ? offsets[p + 1] - offsets[p] : max_obs;
  }
  const int total = offsets[npoints];
  if (total > 0 && !desc) return ORB_GPU_ERR_ARG;
  if (max_obs > kDistinctMaxObs) return ORB_GPU_ERR_CAPACITY;
  { const int rc_pool = use_cached_pool(device); if (rc_pool) return rc_pool; }
  TmpDevW bd, boff, bbest, bst;
  int rc0;
  if ((rc0 = bd.alloc((size_t)(total > 0 ? total : 1) * 32)) || (rc0 = boff.alloc((size_t)(npoints + 1) * 4)) ||
      (rc0 = bbest.alloc((size_t)npoints * 4)) || (rc0 = bst.alloc(4)))
    return rc0;
  void *dd = bd.p, *doff = boff.p, *dbest = bbest.p, *dst = bst.p;
  if (total > 0) ORB_CUDA_TRY(cudaMemcpy(dd, desc, (size_t)total * 32, cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemcpy(doff, offsets, (size_t)(npoints + 1) * 4, cudaMemcpyHostToDevice));
  ORB_CUDA_TRY(cudaMemset(dst, 0, 4));
  int rc = launch_distinctive((const uint8_t*)dd, (const int*)doff, npoints, max_obs, (int32_t*)dbest, (int*)dst, 0);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaMemcpy(best_idx, dbest, (size_t)npoints * 4, cudaMemcpyDeviceToHost));
  int st = 0;
  ORB_CUDA_TRY(cudaMemcpy(&st, dst, 4, cudaMemcpyDeviceToHost));
  return st;
}

}  // extern "C"
