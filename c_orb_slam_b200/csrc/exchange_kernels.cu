// Exchange step of the database-sharded brute-force matcher: every rank holds the exact
// top-2 of the replicated queries against ITS rows; the global result is the merge of the per-rank results in
// ascending row order (lowest index wins ties, as in the ascending scan of src/ORBmatcher.cc:216-225).
//
// One kernel does the whole exchange over NVLink peer memory instead of an NCCL all-gather followed by a merge:
//   1. a thread stores its query's (idx, best, second) into slot `rank` of EVERY rank's gather buffer (peer stores),
//   2. system-scope fence, block barrier, one system-scope atomic per peer on that peer's arrival counter [rank],
//   3. thread 0 spins (acquire, system scope) until its own counters of all ranks reach this call's target,
//   4. the block merges the `world` slots of the local gather buffer (ld.cg: written by peers, never through L1).
// No rank waits before it has delivered its own data, so the exchange cannot deadlock as long as every rank makes
// the call; the gather buffer is double-buffered by call parity (a rank can be at most one call ahead of a peer,
// because completing a call needs that peer's arrival of the same call).  A wait of ~17 s traps instead of hanging.
//
// Peer pointers come from cudaIpcOpenMemHandle (one process per GPU: the handles travel through
// torch.distributed, c_orb_slam_b200/parallel.py) or are wired directly when all ranks live in one process.
#include <algorithm>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace orbgpu {

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// At most kExMaxBlocks blocks, each striding over the queries: every block's thread 0 waits for ALL blocks of ALL
// ranks, so the whole grid has to be resident at once -- 64 blocks of 256 threads without shared memory always are
// (148 SMs), whatever nq_max is.
constexpr int kExMaxBlocks = 64;

__global__ void __launch_bounds__(256)
exchange_top2_kernel(const int32_t* __restrict__ l_idx, const int32_t* __restrict__ l_best,
                     const int32_t* __restrict__ l_second, int nq, int nq_max, int rank, int world, int parity,
                     unsigned target, ExPeers peers, int32_t* __restrict__ o_idx, int32_t* __restrict__ o_best,
                     int32_t* __restrict__ o_second) {
  const int stride = gridDim.x * blockDim.x;
  const size_t slot = ((size_t)parity * world + rank) * 3 * (size_t)nq_max;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nq; i += stride) {
    const int32_t a = l_idx[i], b = l_best[i], c = l_second[i];
    for (int p = 0; p < world; ++p) {
      int32_t* dst = peers.buf[p] + slot;
      dst[i] = a; dst[nq_max + i] = b; dst[2 * (size_t)nq_max + i] = c;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int p = 0; p < world; ++p) atomicAdd_system(peers.flag[p] + rank, 1u);
    const unsigned* mine = peers.flag[rank];
    const long long t0 = clock64();
    for (int r = 0; r < world; ++r)
      while ((int)(ld_acquire_sys(mine + r) - target) < 0)
        if (clock64() - t0 > (1ll << 35)) __trap();  // ~17 s: a rank never made the call
    __threadfence_system();
  }
  __syncthreads();
  const int32_t* g = peers.buf[rank] + (size_t)parity * world * 3 * (size_t)nq_max;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nq; i += stride) {
    int b = 256, s = 256, id = -1;
    for (int r = 0; r < world; ++r) {  // ranks hold ascending row ranges: same rule as merge_top2_kernel
      const int32_t* e = g + (size_t)r * 3 * nq_max;
      const int ik = __ldcg(e + i), bk = __ldcg(e + nq_max + i), sk = __ldcg(e + 2 * (size_t)nq_max + i);
      if (ik < 0) continue;
      if (bk < b) { s = min(b, sk); b = bk; id = ik; }
      else s = min(s, bk);
    }
    o_idx[i] = id; o_best[i] = b; o_second[i] = s;
  }
}

}  // namespace orbgpu

using namespace orbgpu;

struct OrbGpuExchange {
  int device = 0, rank = 0, world = 1, nq_max = 0;
  int32_t* buf = nullptr;    // [2][world][3][nq_max]
  unsigned* flag = nullptr;  // [world]
  ExPeers peers{};
  std::vector<void*> opened;  // IPC mappings to close
  bool connected = false;
  unsigned long long arrivals = 0;  // blocks that have signalled per source rank so far (same on every rank)
  unsigned calls = 0;               // exchanges so far, fused or not (the parity of the gather buffer)
  unsigned fused_calls = 0;         // sequence number of the fused form
  int qb_max = 0;                   // query blocks of 256 in nq_max
  size_t flag_words() const { return (size_t)kExMaxWorld + (size_t)world * qb_max; }
};

extern "C" {

int orb_gpu_exchange_create(int device, int rank, int world, int nq_max, OrbGpuExchange** out) {
  if (!out || world < 1 || world > kExMaxWorld || rank < 0 || rank >= world || nq_max < 1) return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(device));
  OrbGpuExchange* x = new OrbGpuExchange();
  x->device = device; x->rank = rank; x->world = world; x->nq_max = nq_max;
  x->qb_max = (nq_max + 255) / 256;
  const size_t bytes = (size_t)2 * world * 3 * nq_max * sizeof(int32_t);
  // plain cudaMalloc: IPC handles cannot be taken from the stream-ordered pool
  if (cudaMalloc(&x->buf, bytes) != cudaSuccess || cudaMalloc(&x->flag, x->flag_words() * sizeof(unsigned)) != cudaSuccess) {
    set_last_error("cudaMalloc (exchange)", cudaGetLastError());
    if (x->buf) cudaFree(x->buf);
    delete x;
    return ORB_GPU_ERR_CUDA;
  }
  ORB_CUDA_TRY(cudaMemset(x->buf, 0xff, bytes));
  ORB_CUDA_TRY(cudaMemset(x->flag, 0, x->flag_words() * sizeof(unsigned)));
  ORB_CUDA_TRY(cudaDeviceSynchronize());
  x->peers.buf[rank] = x->buf;
  x->peers.flag[rank] = x->flag;
  x->connected = world == 1;
  *out = x;
  return ORB_GPU_OK;
}

int orb_gpu_exchange_handle_bytes(void) { return 2 * (int)sizeof(cudaIpcMemHandle_t); }

int orb_gpu_exchange_local_handle(OrbGpuExchange* x, uint8_t* handle) {
  if (!x || !handle) return ORB_GPU_ERR_ARG;
  ORB_CUDA_TRY(cudaSetDevice(x->device));
  cudaIpcMemHandle_t a, b;
  ORB_CUDA_TRY(cudaIpcGetMemHandle(&a, x->buf));
  ORB_CUDA_TRY(cudaIpcGetMemHandle(&b, x->flag));
  memcpy(handle, &a, sizeof(a));
  memcpy(handle + sizeof(a), &b, sizeof(b));
  return ORB_GPU_OK;
}

int orb_gpu_exchange_connect(OrbGpuExchange* x, const uint8_t* handles) {
  if (!x || !handles) return ORB_GPU_ERR_ARG;
  if (x->connected) return ORB_GPU_OK;
  ORB_CUDA_TRY(cudaSetDevice(x->device));
  const size_t hb = (size_t)orb_gpu_exchange_handle_bytes();
  for (int r = 0; r < x->world; ++r) {
    if (r == x->rank) continue;
    cudaIpcMemHandle_t a, b;
    memcpy(&a, handles + r * hb, sizeof(a));
    memcpy(&b, handles + r * hb + sizeof(a), sizeof(b));
    void *pa = nullptr, *pb = nullptr;
    ORB_CUDA_TRY(cudaIpcOpenMemHandle(&pa, a, cudaIpcMemLazyEnablePeerAccess));
    x->opened.push_back(pa);
    ORB_CUDA_TRY(cudaIpcOpenMemHandle(&pb, b, cudaIpcMemLazyEnablePeerAccess));
    x->opened.push_back(pb);
    x->peers.buf[r] = (int32_t*)pa;
    x->peers.flag[r] = (unsigned*)pb;
  }
  x->connected = true;
  return ORB_GPU_OK;
}

// all ranks in one process (tests; a single-process multi-GPU caller enables peer access itself)
int orb_gpu_exchange_connect_local(OrbGpuExchange* const* all, int world) {
  if (!all || world < 1 || world > kExMaxWorld) return ORB_GPU_ERR_ARG;
  for (int a = 0; a < world; ++a) {
    if (!all[a] || all[a]->world != world || all[a]->rank != a || all[a]->nq_max != all[0]->nq_max) return ORB_GPU_ERR_ARG;
    for (int b = 0; b < world; ++b) { all[a]->peers.buf[b] = all[b]->buf; all[a]->peers.flag[b] = all[b]->flag; }
    all[a]->connected = true;
  }
  return ORB_GPU_OK;
}

int orb_gpu_exchange_destroy(OrbGpuExchange* x) {
  if (!x) return ORB_GPU_ERR_ARG;
  cudaSetDevice(x->device);
  cudaDeviceSynchronize();
  for (void* p : x->opened) cudaIpcCloseMemHandle(p);
  cudaFree(x->buf);
  cudaFree(x->flag);
  delete x;
  return ORB_GPU_OK;
}

int orb_gpu_exchange_top2_device(OrbGpuExchange* x, const int32_t* d_idx, const int32_t* d_best,
                                 const int32_t* d_second, int nq, int32_t* d_out_idx, int32_t* d_out_best,
                                 int32_t* d_out_second, void* stream) {
  if (!x || nq < 1 || nq > x->nq_max || !d_idx || !d_best || !d_second || !d_out_idx || !d_out_best || !d_out_second)
    return ORB_GPU_ERR_ARG;
  if (!x->connected) return ORB_GPU_ERR_STATE;
  ORB_CUDA_TRY(cudaSetDevice(x->device));
  const int blocks = std::min((nq + 255) / 256, kExMaxBlocks);
  x->arrivals += (unsigned long long)blocks;  // every rank makes the same calls: the target is the same everywhere
  const int parity = (int)(x->calls++ & 1u);
  exchange_top2_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_idx, d_best, d_second, nq, x->nq_max, x->rank, x->world,
                                                                 parity, (unsigned)x->arrivals, x->peers, d_out_idx,
                                                                 d_out_best, d_out_second);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

int orb_gpu_exchange_match_bruteforce_device(OrbGpuExchange* x, const uint8_t* d_q, int nq, const uint8_t* d_db,
                                             int64_t ndb, int32_t idx_base, int32_t* d_idx, int32_t* d_best,
                                             int32_t* d_second, void* stream) {
  if (!x || nq < 1 || nq > x->nq_max || !d_q || !d_idx || !d_best || !d_second) return ORB_GPU_ERR_ARG;
  if (!x->connected) return ORB_GPU_ERR_STATE;
  ORB_CUDA_TRY(cudaSetDevice(x->device));
  if (x->world == 1) {
    int fused = 0;
    return bruteforce_device_impl(d_q, nq, d_db, ndb, idx_base, d_idx, d_best, d_second, (cudaStream_t)stream, nullptr, &fused);
  }
  static const bool no_fuse = [] { const char* e = getenv("ORB_GPU_NO_FUSED_EXCHANGE"); return e && atoi(e); }();
  FusedExchange fx;
  fx.peers = x->peers;
  fx.rank = x->rank; fx.world = x->world; fx.nq_max = x->nq_max; fx.qb_max = x->qb_max;
  fx.parity = (int)(x->calls & 1u);
  fx.seq = x->fused_calls + 1;
  int fused = 0;
  int rc = bruteforce_device_impl(d_q, nq, d_db, ndb, idx_base, d_idx, d_best, d_second, (cudaStream_t)stream,
                                  no_fuse ? nullptr : &fx, &fused);
  if (rc) return rc;
  if (fused) {  // the match kernel stored, signalled, waited and merged: the step was one launch
    ++x->calls;
    ++x->fused_calls;
    return ORB_GPU_OK;
  }
  // shapes the fused form does not take (another kernel form, a grid of more than one wave): match, then the
  // stand-alone exchange kernel in place.  Every rank takes the same branch: it depends on nq, the database size
  // class and the SM count only through the grid shape, which the ranks of one box share.
  return orb_gpu_exchange_top2_device(x, d_idx, d_best, d_second, nq, d_idx, d_best, d_second, stream);
}

}  // extern "C"
