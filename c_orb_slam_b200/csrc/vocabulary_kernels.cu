// DBoW2 vocabulary-tree descent for a batch of descriptors: TemplatedVocabulary::transform(feature, word_id,
// weight, nid, levelsup) (reference Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1213-1259), the Hamming part of
// Frame::ComputeBoW (src/Frame.cc:395-402).  The tree lives in HBM for the life of the handle (the shipped ORB
// vocabulary is k = 10, L = 6: 1.1 M nodes x 32 B = 35 MB, L2-resident on a B200 after the first frames).
//
// One warp per descriptor: at every level the lanes take the children of the current node (any branching
// factor), the warp minimum of (distance << 20 | child position) picks the first closest child (strict '<'
// in the reference's scan, :1238-1246).  The walk ends at a node without children.
#include "common.cuh"

struct OrbGpuVocabulary {
  int device = 0;
  int n_nodes = 0, depth = 0;
  int* child_begin = nullptr;   // [n_nodes + 1]
  int* children = nullptr;      // [n_children]
  uint4* desc = nullptr;        // [n_nodes][2]
  int* word_id = nullptr;       // [n_nodes]
  double* weight = nullptr;     // [n_nodes]
};

namespace orbgpu {

__global__ void __launch_bounds__(256)
voc_transform_kernel(const int* __restrict__ child_begin, const int* __restrict__ children,
                     const uint4* __restrict__ node_desc, const int* __restrict__ word_id,
                     const double* __restrict__ weight, int n_nodes, int nid_level,
                     const uint4* __restrict__ feat, int n, uint32_t* __restrict__ out_word,
                     double* __restrict__ out_weight, uint32_t* __restrict__ out_node) {
  const int f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (f >= n) return;
  const uint4 a0 = feat[2 * (size_t)f], a1 = feat[2 * (size_t)f + 1];
  int node = 0, level = 0;
  uint32_t nid = 0;  // root when nid_level <= 0 (:1221)
  int beg = child_begin[0], end = child_begin[1];
  while (end > beg && level < n_nodes) {
    ++level;
    unsigned best = 0xffffffffu;
    for (int c = beg + lane; c < end; c += 32) {
      const int id = children[c];
      const uint4 b0 = node_desc[2 * (size_t)id], b1 = node_desc[2 * (size_t)id + 1];
      const unsigned d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
      best = min(best, (d << 20) | (unsigned)(c - beg));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    node = children[beg + (int)(best & 0xfffffu)];
    if (level == nid_level) nid = (uint32_t)node;
    beg = child_begin[node];
    end = child_begin[node + 1];
  }
  if (lane == 0) {
    out_word[f] = (uint32_t)word_id[node];
    out_weight[f] = weight[node];
    out_node[f] = nid;
  }
}

}  // namespace orbgpu

using namespace orbgpu;

extern "C" {

int orb_gpu_vocabulary_create(int n_nodes, const int32_t* child_begin, const int32_t* children,
                              const uint8_t* node_desc, const int32_t* word_id, const double* weight, int depth,
                              int device, OrbGpuVocabulary** out) {
  if (!out || n_nodes < 2 || !child_begin || !children || !node_desc || !word_id || !weight || depth < 1)
    return ORB_GPU_ERR_ARG;
  *out = nullptr;
  if (child_begin[0] != 0) return ORB_GPU_ERR_ARG;
  const int n_children = child_begin[n_nodes];
  if (n_children < 1 || n_children >= (1 << 20) * 32) return ORB_GPU_ERR_ARG;
  for (int i = 0; i < n_nodes; ++i)
    if (child_begin[i + 1] < child_begin[i] || child_begin[i + 1] - child_begin[i] >= (1 << 20)) return ORB_GPU_ERR_ARG;
  for (int i = 0; i < n_children; ++i)
    if (children[i] <= 0 || children[i] >= n_nodes) return ORB_GPU_ERR_ARG;  // the root is nobody's child
  ORB_CUDA_TRY(cudaSetDevice(device));
  OrbGpuVocabulary* v = new OrbGpuVocabulary();
  v->device = device; v->n_nodes = n_nodes; v->depth = depth;
  cudaError_t e = cudaSuccess;
  auto up = [&](void** dst, const void* src, size_t bytes) {
    if (e == cudaSuccess) e = cudaMalloc(dst, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
  };
  up((void**)&v->child_begin, child_begin, (size_t)(n_nodes + 1) * 4);
  up((void**)&v->children, children, (size_t)n_children * 4);
  up((void**)&v->desc, node_desc, (size_t)n_nodes * 32);
  up((void**)&v->word_id, word_id, (size_t)n_nodes * 4);
  up((void**)&v->weight, weight, (size_t)n_nodes * 8);
  if (e != cudaSuccess) {
    set_last_error("orb_gpu_vocabulary_create", e);
    orb_gpu_vocabulary_destroy(v);
    return ORB_GPU_ERR_CUDA;
  }
  *out = v;
  return ORB_GPU_OK;
}

void orb_gpu_vocabulary_destroy(OrbGpuVocabulary* v) {
  if (!v) return;
  cudaSetDevice(v->device);
  cudaFree(v->child_begin); cudaFree(v->children); cudaFree(v->desc); cudaFree(v->word_id); cudaFree(v->weight);
  delete v;
}

int orb_gpu_vocabulary_transform_device(const OrbGpuVocabulary* v, const uint8_t* d_desc, int n, int levelsup,
                                        uint32_t* d_word_id, double* d_weight, uint32_t* d_node_id, void* stream) {
  if (!v || n < 0 || (n > 0 && (!d_desc || !d_word_id || !d_weight || !d_node_id))) return ORB_GPU_ERR_ARG;
  if (n == 0) return ORB_GPU_OK;
  voc_transform_kernel<<<(n * 32 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      v->child_begin, v->children, v->desc, v->word_id, v->weight, v->n_nodes, v->depth - levelsup,
      (const uint4*)d_desc, n, d_word_id, d_weight, d_node_id);
  count_launch();
  ORB_CUDA_TRY(cudaGetLastError());
  return ORB_GPU_OK;
}

int orb_gpu_vocabulary_transform(const OrbGpuVocabulary* v, const uint8_t* desc, int n, int levelsup,
                                 uint32_t* word_id, double* weight, uint32_t* node_id) {
  if (!v || n < 0 || (n > 0 && (!desc || !word_id || !weight || !node_id))) return ORB_GPU_ERR_ARG;
  if (n == 0) return ORB_GPU_OK;
  { const int rc_pool = use_cached_pool(v->device); if (rc_pool) return rc_pool; }
  TmpDevW bd, bw, bv, bn;
  int rc0;
  if ((rc0 = bd.alloc((size_t)n * 32)) || (rc0 = bw.alloc((size_t)n * 4)) || (rc0 = bv.alloc((size_t)n * 8)) ||
      (rc0 = bn.alloc((size_t)n * 4)))
    return rc0;
  void *dd = bd.p, *dw = bw.p, *dv = bv.p, *dn = bn.p;
  ORB_CUDA_TRY(cudaMemcpy(dd, desc, (size_t)n * 32, cudaMemcpyHostToDevice));
  int rc = orb_gpu_vocabulary_transform_device(v, (const uint8_t*)dd, n, levelsup, (uint32_t*)dw, (double*)dv,
                                               (uint32_t*)dn, nullptr);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaMemcpy(word_id, dw, (size_t)n * 4, cudaMemcpyDeviceToHost));
  ORB_CUDA_TRY(cudaMemcpy(weight, dv, (size_t)n * 8, cudaMemcpyDeviceToHost));
  ORB_CUDA_TRY(cudaMemcpy(node_id, dn, (size_t)n * 4, cudaMemcpyDeviceToHost));
  return ORB_GPU_OK;
}

}  // extern "C"
