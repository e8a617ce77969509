/* orb_gpu.h -- C ABI of the B200-native ORB front-end (liborb_gpu.so).
 *
 * Drop-in boundary for ORB-SLAM2's (reference: junejunejune/c_orb_slam)
 *   ORB_SLAM2::ORBextractor::operator()   include/ORBextractor.h:59-61, src/ORBextractor.cc:1043-1105
 *   ORB_SLAM2::ORBmatcher Hamming matching include/ORBmatcher.h:37-102, src/ORBmatcher.cc
 *   Frame::ComputeStereoMatches (Hamming) src/Frame.cc:466-549
 * The reference has no FFI layer of its own; the C++ adapter classes in
 * c_orb_slam_b200/csrc/adapter/ reproduce the two reference classes on top of these calls
 * (INTEGRATION.md shows the binding a maintainer adds).
 *
 * Conventions: every function returns ORB_GPU_OK (0) or a negative OrbGpuStatus; nothing throws;
 * there is NO CPU fallback -- without a CUDA device every compute call fails with
 * ORB_GPU_ERR_CUDA.  Pointers named d_* are device pointers, all others are host pointers.
 * `stream` arguments are cudaStream_t values passed as void* (NULL = the handle's own stream).
 * A handle is single-caller; different handles may be used concurrently from different threads
 * (the reference runs two extractors from two std::threads, src/Frame.cc:78-81).
 */
#ifndef ORB_GPU_H_
#define ORB_GPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  ORB_GPU_OK = 0,
  ORB_GPU_ERR_ARG = -1,      /* null pointer, bad size, unsupported shape (e.g. portrait: nIni == 0) */
  ORB_GPU_ERR_CAPACITY = -2, /* caller buffer or an internal candidate pool too small */
  ORB_GPU_ERR_CUDA = -3,     /* CUDA runtime error (no device, launch failure, OOM) */
  ORB_GPU_ERR_STATE = -4     /* call order violated (e.g. pyramid requested before any extract) */
} OrbGpuStatus;

/* Layout-compatible with cv::KeyPoint (28 bytes): pt.x, pt.y, size, angle, response, octave,
 * class_id.  Replaces std::vector<cv::KeyPoint> at include/ORBextractor.h:60. */
typedef struct {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} OrbKeyPoint;

typedef struct OrbGpuExtractor OrbGpuExtractor; /* owns one device, one stream, pyramids, scratch */

#define ORB_GPU_MAX_LEVELS 16
/* upper bound of keypoints per frame: nfeatures + 3 per level (the quadtree may overshoot its
 * per-level quota by up to 3, src/ORBextractor.cc:730-731). */
#define ORB_GPU_CAPACITY(nfeatures, nlevels) ((nfeatures) + 3 * (nlevels))

/* ---- extractor ------------------------------------------------------------------------- */

/* ORBextractor::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 * (include/ORBextractor.h:51-52, src/ORBextractor.cc:410-470).  max_width/max_height/max_batch
 * size the device buffers once; device = CUDA ordinal. */
int orb_gpu_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                   int max_width, int max_height, int max_batch, int device,
                   OrbGpuExtractor** out);
int orb_gpu_destroy(OrbGpuExtractor* h);

/* ORBextractor::operator()(image, mask, keypoints, descriptors) (src/ORBextractor.cc:1043-1105);
 * the mask is ignored by the reference (include/ORBextractor.h:58) and has no parameter here.
 * gray: host, 8-bit single channel, `stride` bytes per row.  kps[capacity], desc[capacity*32].
 * An empty image (width or height 0) returns OK with *n_out = 0 (:1046). */
int orb_gpu_extract(OrbGpuExtractor* h, const uint8_t* gray, int width, int height, size_t stride,
                    OrbKeyPoint* kps, uint8_t* desc, int capacity, int* n_out);

/* Batched form (config 3 of BASELINE.json; no reference function -- defined as the call above
 * applied independently to each frame).  Host frames: frame f starts at frames + f*frame_stride.
 * Outputs are [nframes][capacity] / [nframes][capacity*32] / n_out[nframes]. */
int orb_gpu_extract_batch(OrbGpuExtractor* h, const uint8_t* frames, int nframes, int width,
                          int height, size_t stride, size_t frame_stride, OrbKeyPoint* kps,
                          uint8_t* desc, int capacity, int* n_out);

/* Same with every buffer resident in device memory; asynchronous on `stream`.  Level 0 of the
 * pyramid aliases d_frames (no copy), so d_frames must stay valid until the stream has drained.
 * d_status (optional, int[1]) receives 0 or ORB_GPU_ERR_CAPACITY from the device side. */
int orb_gpu_extract_batch_device(OrbGpuExtractor* h, const uint8_t* d_frames, int nframes,
                                 int width, int height, size_t stride, size_t frame_stride,
                                 OrbKeyPoint* d_kps, uint8_t* d_desc, int capacity, int* d_n_out,
                                 int* d_status, void* stream);

/* Getters of include/ORBextractor.h:63-83 (values bit-identical to the reference's float/double
 * mix) plus mnFeaturesPerLevel and umax for tests.  Arrays hold nlevels entries (umax: 16). */
int orb_gpu_get_levels(const OrbGpuExtractor* h);
float orb_gpu_get_scale_factor(const OrbGpuExtractor* h);
int orb_gpu_get_scale_tables(const OrbGpuExtractor* h, float* scale, float* inv_scale,
                             float* sigma2, float* inv_sigma2);
int orb_gpu_get_features_per_level(const OrbGpuExtractor* h, int* quota, int* umax16);

/* mvImagePyramid[level] of the last extract call (include/ORBextractor.h:85; read by stereo
 * matching, src/Frame.cc:473,563,575,580).  frame = index inside the last batch.  blurred != 0
 * returns the 7x7 sigma-2 blurred level used for the descriptors (:1085-1086) instead.
 * dst may be NULL to query the size only. */
int orb_gpu_get_pyramid(OrbGpuExtractor* h, int frame, int level, int blurred, uint8_t* dst,
                        size_t dst_stride, int* width, int* height);

/* Stage access for parity tests: FAST candidates of one level of one frame in the reference's
 * arrival order (vToDistributeKeys, src/ORBextractor.cc:778-829): xyr[3*i] = x, y (relative to
 * the 16-px border), response.  Returns count through *n (only min(*n, cap) are written). */
int orb_gpu_get_candidates(OrbGpuExtractor* h, int frame, int level, float* xyr, int cap, int* n);

/* ---- matcher --------------------------------------------------------------------------- */

#define ORB_GPU_TH_HIGH 100 /* ORBmatcher::TH_HIGH, src/ORBmatcher.cc:37 */
#define ORB_GPU_TH_LOW 50   /* ORBmatcher::TH_LOW,  src/ORBmatcher.cc:38 */
#define ORB_GPU_HISTO_LENGTH 30

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663) for n pairs: dist[i] =
 * hamming(a[i], b[i]), 32-byte rows, device pointers. */
int orb_gpu_hamming_pairs_device(const uint8_t* d_a, const uint8_t* d_b, int n, int32_t* d_dist,
                                 void* stream);

/* Brute-force query x database top-2 (config 5; defined from ORBmatcher.cc:216-225,228-230):
 * for each query, ascending scan of the database keeping best / lowest best index / second.
 * idx = index + idx_base (int32; -1 if ndb == 0), best/second in 0..256.  Device pointers. */
int orb_gpu_match_bruteforce_device(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb,
                                    int32_t idx_base, int32_t* d_idx, int32_t* d_best,
                                    int32_t* d_second, void* stream);
/* Host-pointer form (copies in, computes on `device`, copies out). */
int orb_gpu_match_bruteforce(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int device,
                             int32_t* idx, int32_t* best, int32_t* second);
/* The same search for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * arguments and bit-identical results; the default (tcgen05) kernel only.  It calls no cudaMalloc / cudaFree /
 * cudaMallocAsync or any allocating routine: its scratch (per-slice partial results, shared thresholds, slice tickets)
 * lives in a file-scope device array of ORB_GPU_NOALLOC_SCRATCH_BYTES (1 MiB) that every GPU which loads the library
 * holds until the process exits.  Needed: slices * nq * 12 + (nq + ceil(nq / 256)) * 4 bytes when the database is cut
 * into more than one slice, slices = floor(SMs / ceil(nq / 256)) on a database of up to slices * 2^22 rows (e.g.
 * 4096 queries x 10 M rows on 148 SMs: 9 slices, 0.46 MB); a call that needs more returns ORB_GPU_ERR_CAPACITY with
 * the sizes in orb_gpu_last_error() and launches nothing.  Every call on a GPU shares that one array, which keeps the
 * last call's contents (what must start at zero is cleared on `stream` before the kernel): calls that overlap on the
 * same GPU -- two streams, two threads -- RACE on it; serialise them or use orb_gpu_match_bruteforce_device. */
#define ORB_GPU_NOALLOC_SCRATCH_BYTES (1 << 20)
int orb_gpu_match_bruteforce_noalloc_device(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb,
                                            int32_t idx_base, int32_t* d_idx, int32_t* d_best,
                                            int32_t* d_second, void* stream);
/* Exact merge of per-shard results of disjoint ascending index ranges (shard s holds rows
 * [base_s, base_s + n_s) with base_s increasing in s): inputs are [nshards][nq].  This is the
 * step that follows the NCCL all-gather when the database is sharded across GPUs. */
int orb_gpu_merge_top2_device(const int32_t* d_idx, const int32_t* d_best, const int32_t* d_second,
                              int nshards, int nq, int32_t* d_out_idx, int32_t* d_out_best,
                              int32_t* d_out_second, void* stream);
/* The same exchange + merge as ONE kernel over NVLink peer memory (no NCCL on the data path): every rank stores
 * its [3][nq] result into slot `rank` of every rank's gather buffer, signals the peers' arrival counters
 * (system-scope atomics), waits for its own counters and merges the `world` slots (lowest index wins ties).
 * One OrbGpuExchange per rank; one process per GPU exchanges the handles of orb_gpu_exchange_local_handle
 * (orb_gpu_exchange_handle_bytes() bytes each, cudaIpcMemHandle_t inside) by any host-side means and passes all of
 * them, in rank order, to orb_gpu_exchange_connect; ranks living in one process use orb_gpu_exchange_connect_local.
 * Every rank must make the same sequence of orb_gpu_exchange_top2_device calls (same nq <= nq_max); a rank that
 * never arrives makes the others trap after ~17 s instead of hanging.  The reference has no counterpart (it
 * relocalises through the BoW inverted file, src/KeyFrameDatabase.cc:199-309); config 5 of BASELINE.json defines it. */
typedef struct OrbGpuExchange OrbGpuExchange;
int orb_gpu_exchange_create(int device, int rank, int world, int nq_max, OrbGpuExchange** out);
int orb_gpu_exchange_handle_bytes(void);
int orb_gpu_exchange_local_handle(OrbGpuExchange* x, uint8_t* handle);
int orb_gpu_exchange_connect(OrbGpuExchange* x, const uint8_t* handles /* [world][handle_bytes] */);
int orb_gpu_exchange_connect_local(OrbGpuExchange* const* all, int world);
int orb_gpu_exchange_destroy(OrbGpuExchange* x);
int orb_gpu_exchange_top2_device(OrbGpuExchange* x, const int32_t* d_idx, const int32_t* d_best,
                                 const int32_t* d_second, int nq, int32_t* d_out_idx, int32_t* d_out_best,
                                 int32_t* d_out_second, void* stream);
/* Fused form: the brute-force search over this rank's rows AND the exchange in ONE launch.  The block
 * of the tcgen05 kernel that finishes the last database slice of a block of 256 queries merges the slices, stores the
 * triples into every rank's gather buffer over NVLink, releases a per-query-block sequence number to every rank, waits
 * for the peers' numbers of the same query block and merges the `world` slots -- the exchange of one query block
 * overlaps the matching of the others, and there is no merge or exchange launch.  Arguments as
 * orb_gpu_match_bruteforce_device (idx_base = global index of this rank's row 0; ranks hold ascending row ranges);
 * the outputs are the GLOBAL top-2, identical on every rank.  Every rank makes the same sequence of calls; it may be
 * mixed with orb_gpu_exchange_top2_device calls.  Shapes the fused form does not take (a non-default kernel form, a
 * grid of more than one block per SM) run the search and then the stand-alone exchange kernel, same results. */
int orb_gpu_exchange_match_bruteforce_device(OrbGpuExchange* x, const uint8_t* d_q, int nq, const uint8_t* d_db,
                                             int64_t ndb, int32_t idx_base, int32_t* d_idx, int32_t* d_best,
                                             int32_t* d_second, void* stream);
/* accept flags of ORBmatcher.cc:228-230: best <= th && (float)best < nnratio*(float)second */
int orb_gpu_ratio_test_device(const int32_t* d_best, const int32_t* d_second, int nq, int th,
                              float nnratio, uint8_t* d_accept, void* stream);

/* The MapPoint fields ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) reads
 * (src/ORBmatcher.cc:52-66). */
typedef struct {
  float u, v, ur;   /* mTrackProjX, mTrackProjY, mTrackProjXR */
  int32_t level;    /* mnTrackScaleLevel */
  float view_cos;   /* mTrackViewCos */
  uint8_t valid;    /* mbTrackInView && !isBad() */
  uint8_t has_obs;  /* Observations() > 0 */
  uint8_t pad_[2];
} OrbProjQuery;

/* Frame grid parameters (src/Frame.cc:230-245,327-392; include/Frame.h:37-38: 64 x 48 cells). */
typedef struct {
  float min_x, min_y, inv_w, inv_h; /* mnMinX, mnMinY, mfGridElementWidthInv, ...HeightInv */
} OrbFrameGrid;

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:45-129)
 * including GetFeaturesInArea candidate order and the in-order claim of keypoints.  Host
 * pointers.  u_right may be NULL (monocular).  kp_blocked[nk] in/out: keypoint already holds a
 * MapPoint with Observations()>0.  match_of_kp[nk] out: query index assigned or -1. */
int orb_gpu_search_by_projection(const OrbProjQuery* q, const uint8_t* qdesc, int nq,
                                 const OrbKeyPoint* kps, const uint8_t* kdesc, const float* u_right,
                                 int nk, OrbFrameGrid grid, const float* scale_factors, int nlevels,
                                 float th, float nnratio, int th_high, int device,
                                 uint8_t* kp_blocked, int32_t* match_of_kp, int* nmatches);

/* What the best-only projection searches know about a source point before their candidate loop
 * (ORBmatcher.cc:1355-1390 / :1497-1527 / :337-362; computed by the caller exactly as the reference does). */
typedef struct {
  float u, v;        /* projection into the target frame */
  float radius;      /* th * mvScaleFactors[level] */
  float ur;          /* u - mbf*invzc, used by the stereo gate (:1407-1413) */
  int32_t min_level, max_level; /* GetFeaturesInArea(u, v, radius, minLevel, maxLevel) */
  float angle;       /* source keypoint angle (rotation histogram, :1432-1441) */
  uint8_t valid;     /* passed the caller's geometric tests */
  uint8_t blocks;    /* a match hides the keypoint from later points (Observations()>0 in :1403-1405; 1 elsewhere) */
  uint8_t pad_[2];
} OrbWindowQuery;

/* Best-only windowed search with in-order claims and optional rotation consistency:
 *   ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono)          src/ORBmatcher.cc:1328-1470
 *   ORBmatcher::SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)     :1472-1599
 *   ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th)           :290-403
 *   ORBmatcher::Fuse(KeyFrame*, vpMapPoints, th) / Fuse(KeyFrame*, Scw, ...) search loops  :825-975, :977-1100
 *   ORBmatcher::SearchBySim3: each of its two passes                                       :1102-1326
 * (candidate order of GetFeaturesInArea, bestDist from 256 with strict '<', accept best <= th_accept, claim,
 * ComputeThreeMaxima :1601-1642 when check_orientation).  gate: 0 none; 1 stereo radius gate (:1407-1413:
 * u_right[i] > 0 and |ur - u_right[i]| > radius -> skip); 2 chi-square gate of Fuse (:904-931:
 * e2 * inv_level_sigma2[octave] > 5.99, or 7.8 with the right coordinate when u_right[i] >= 0).
 * Host pointers.  kp_blocked in/out; match_of_kp[nk] = query index, -1 (untouched) or -2 (matched, then reset
 * to NULL by the rotation check, :1461); match_of_query[nq] (optional) = keypoint chosen by query i or -1. */
int orb_gpu_search_windowed(const OrbWindowQuery* q, const uint8_t* qdesc, int nq,
                            const OrbKeyPoint* kps, const uint8_t* kdesc, const float* u_right,
                            int nk, OrbFrameGrid grid, int th_accept, int gate,
                            const float* inv_level_sigma2, int nlevels, int check_orientation,
                            int device, uint8_t* kp_blocked, int32_t* match_of_kp,
                            int32_t* match_of_query, int* nmatches);

/* ---- frames resident in device memory (matcher side) -----------------------------------------------------------
 * The reference keeps mvKeysUn, mDescriptors, mvuRight and the 64 x 48 feature grid inside the Frame / KeyFrame
 * (src/Frame.cc:230-253, include/Frame.h:130-161) and every ORBmatcher search reads them there.  OrbGpuFrame is that
 * state in device memory: filled once per frame -- from host arrays, or from the device arrays orb_gpu_extract_batch_device
 * left (no host round trip) -- the grid of AssignFeaturesToGrid built once, then any number of searches.  A search
 * moves only its queries in and its results out (one copy each way through pinned staging blocks owned by the frame,
 * no allocation, one synchronisation).  A frame is single-caller like an extractor handle; it owns one stream. */
typedef struct OrbGpuFrame OrbGpuFrame;
int orb_gpu_frame_create(int device, OrbGpuFrame** out);
int orb_gpu_frame_destroy(OrbGpuFrame* f);
int orb_gpu_frame_size(const OrbGpuFrame* f);
/* kps = mvKeysUn, desc = mDescriptors [n][32], u_right = mvuRight or NULL (monocular); host pointers. */
int orb_gpu_frame_upload(OrbGpuFrame* f, const OrbKeyPoint* kps, const uint8_t* desc, const float* u_right, int n,
                         OrbFrameGrid grid);
/* The same from device arrays (e.g. frame i of the [frame][capacity] outputs of orb_gpu_extract_batch_device and the
 * d_u_right row of orb_gpu_stereo_matches_device; the keypoints are taken as undistorted, i.e. a camera without
 * distortion coefficients as in src/Frame.cc:406-412).  `stream` = the stream that produced them (NULL: already
 * complete).  The arrays are copied; the caller may reuse them after the call returns. */
int orb_gpu_frame_from_device(OrbGpuFrame* f, const OrbKeyPoint* d_kps, const uint8_t* d_desc, const float* d_u_right,
                              int n, OrbFrameGrid grid, void* stream);
/* orb_gpu_search_by_projection / orb_gpu_search_windowed (below) against a resident frame: same arguments without
 * the target arrays, same results. */
int orb_gpu_frame_search_by_projection(OrbGpuFrame* f, const OrbProjQuery* q, const uint8_t* qdesc, int nq,
                                       const float* scale_factors, int nlevels, float th, float nnratio, int th_high,
                                       uint8_t* kp_blocked, int32_t* match_of_kp, int* nmatches);
int orb_gpu_frame_search_windowed(OrbGpuFrame* f, const OrbWindowQuery* q, const uint8_t* qdesc, int nq, int th_accept,
                                  int gate, const float* inv_level_sigma2, int nlevels, int check_orientation,
                                  uint8_t* kp_blocked, int32_t* match_of_kp, int32_t* match_of_query, int* nmatches);

/* Replaces Frame::ComputeStereoMatches (src/Frame.cc:466-640) whole: Hamming stage (:476-549), sub-pixel
 * refinement (11x11 SAD over 11 shifts + parabola, :551-620) and the median outlier filter (:623-639), run on
 * the unblurred pyramids (mvImagePyramid, include/ORBextractor.h:85) that `left` and `right` kept in device
 * memory from their last extract call -- frame_l / frame_r are indices inside those batches (0 for
 * orb_gpu_extract); left == right is allowed.  kps / desc are the host arrays those calls returned (mvKeys,
 * mDescriptors, mvKeysRight, mDescriptorsRight).  bf = Frame::mbf, b = Frame::mb (minD = 0, maxD = mbf/mb,
 * :496-498).  u_right / depth [nl] = mvuRight / mvDepth (-1 where there is no match).  A keypoint whose 11x11
 * window would leave its level image (cv::Mat::rowRange would throw in the reference) gets no match. */
int orb_gpu_compute_stereo_matches(OrbGpuExtractor* left, int frame_l, OrbGpuExtractor* right,
                                   int frame_r, const OrbKeyPoint* kps_l, const uint8_t* desc_l, int nl,
                                   const OrbKeyPoint* kps_r, const uint8_t* desc_r, int nr, float bf,
                                   float b, float* u_right, float* depth);

/* The same for `npairs` stereo pairs with everything resident in device memory, asynchronous on `stream`
 * (NULL = left's stream).  Pair i uses frame frame_l0 + i*frame_l_step of left's last batch and frame
 * frame_r0 + i*frame_r_step of right's, for the images and for the keypoint arrays alike: d_kps_* / d_desc_* /
 * d_n_* are the [frame][capacity] outputs of orb_gpu_extract_batch_device.  Outputs: d_u_right / d_depth
 * [npairs][capacity_l]; d_sad [npairs][capacity_l] int scratch (best SAD or -1). */
int orb_gpu_stereo_matches_device(OrbGpuExtractor* left, int frame_l0, int frame_l_step,
                                  OrbGpuExtractor* right, int frame_r0, int frame_r_step, int npairs,
                                  const OrbKeyPoint* d_kps_l, const uint8_t* d_desc_l, const int* d_n_l,
                                  int capacity_l, const OrbKeyPoint* d_kps_r, const uint8_t* d_desc_r,
                                  const int* d_n_r, int capacity_r, float bf, float b, float* d_u_right,
                                  float* d_depth, int* d_sad, void* stream);

/* Replaces the arithmetic of MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307) for a batch of
 * map points: point p owns the descriptors desc[offsets[p] .. offsets[p+1]) (its observations in the iteration
 * order of mObservations, bad key frames already dropped, :261-267).  best_idx[p] = position inside that list
 * of the descriptor with the least median distance to the others -- median = sorted row [0.5*(N-1)], first
 * minimum wins (:288-301) -- or -1 for a point without descriptors.  At most 1024 observations per point
 * (ORB_GPU_ERR_CAPACITY beyond).  Host pointers. */
int orb_gpu_distinctive_descriptors(const uint8_t* desc, const int32_t* offsets, int npoints, int device,
                                    int32_t* best_idx);
/* Device pointers, asynchronous on `stream`; max_obs >= the longest list; *d_status is set to
 * ORB_GPU_ERR_CAPACITY when a list is longer than 1024 (initialise it to 0). */
int orb_gpu_distinctive_descriptors_device(const uint8_t* d_desc, const int* d_offsets, int npoints,
                                           int max_obs, int32_t* d_best_idx, int* d_status, void* stream);

/* ---- vocabulary (DBoW2) ------------------------------------------------------------------- */

/* A DBoW2 vocabulary tree resident in device memory.  Node 0 is the root; node i owns the children
 * children[child_begin[i] .. child_begin[i+1]) in the order of Node::children
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:148-176); a node without children is a word with word_id /
 * weight (Node::word_id, Node::weight); depth = m_L. */
typedef struct OrbGpuVocabulary OrbGpuVocabulary;
int orb_gpu_vocabulary_create(int n_nodes, const int32_t* child_begin, const int32_t* children,
                              const uint8_t* node_desc, const int32_t* word_id, const double* weight,
                              int depth, int device, OrbGpuVocabulary** out);
void orb_gpu_vocabulary_destroy(OrbGpuVocabulary* v);

/* Replaces TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup)
 * (TemplatedVocabulary.h:1213-1259) for n descriptors -- the Hamming work of Frame::ComputeBoW
 * (src/Frame.cc:395-402, levelsup = 4).  node_id[i] = ancestor of the word at level depth - levelsup (0 = root
 * when that is <= 0).  The caller folds (word_id, weight) into the BowVector and (node_id, i) into the
 * FeatureVector in feature order (:1148-1165).  Host pointers / device pointers + stream. */
int orb_gpu_vocabulary_transform(const OrbGpuVocabulary* v, const uint8_t* desc, int n, int levelsup,
                                 uint32_t* word_id, double* weight, uint32_t* node_id);
int orb_gpu_vocabulary_transform_device(const OrbGpuVocabulary* v, const uint8_t* d_desc, int n,
                                        int levelsup, uint32_t* d_word_id, double* d_weight,
                                        uint32_t* d_node_id, void* stream);

/* Frame::GetFeaturesInArea / KeyFrame::GetFeaturesInArea (src/Frame.cc:327-380, src/KeyFrame.cc:569-608) over
 * the 64x48 grid of AssignFeaturesToGrid (:230-245) for nq windows at once: the candidates of window i are
 * indices[offsets[i] .. offsets[i+1]) in the reference's order (grid column, then row, then insertion order).
 * Only u, v, radius, min_level, max_level and valid of OrbWindowQuery are read.  offsets has nq+1 entries and is
 * always filled; ORB_GPU_ERR_CAPACITY when offsets[nq] > capacity.  Host pointers. */
int orb_gpu_features_in_area(const OrbWindowQuery* q, int nq, const OrbKeyPoint* kps, int nk,
                             OrbFrameGrid grid, int device, int32_t* offsets, int32_t* indices,
                             int capacity);

/* One source feature of a search over equal vocabulary nodes: feature `index` of frame 1 is compared with the
 * features cand[begin .. end) of frame 2 (the index list of the same DBoW2 node, in its insertion order).  The
 * caller emits the rows in the reference's visiting order (nodes ascending, then the node's list order) and
 * drops the features the reference skips before the inner loop (no / bad map point, :194-200, :559-563,
 * :699-709). */
typedef struct OrbBucketQuery {
  int32_t index, begin, end;
} OrbBucketQuery;

/* Replaces the loops of
 *   ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)       src/ORBmatcher.cc:159-288
 *   ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&)    src/ORBmatcher.cc:522-655
 * (top-2 inside the node, accept best <= th_accept and (float)best < nnratio * (float)second, second = 256
 * when alone; the chosen feature of frame 2 is claimed, :209/:576; rotation histogram :268-285).
 * th_accept = ORB_GPU_TH_LOW for the first form and ORB_GPU_TH_LOW - 1 for the second (strict '<', :598).
 * skip2[n2] (optional): features of frame 2 that are never candidates (no / bad map point, :576-580).
 * match12[n1] = feature of frame 2 matched to feature i of frame 1 or -1; match21[n2] the inverse.  Host
 * pointers. */
int orb_gpu_search_by_bow(const OrbBucketQuery* q, int nq, const int32_t* cand2, int ncand2,
                          const OrbKeyPoint* kps1, const uint8_t* desc1, int n1,
                          const OrbKeyPoint* kps2, const uint8_t* desc2, const uint8_t* skip2, int n2,
                          float nnratio, int th_accept, int check_orientation, int device,
                          int32_t* match12, int32_t* match21, int* nmatches);

/* Replaces the loops of ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-823) including
 * CheckDistEpipolarLine (:138-157).  u_right1/2 optional (NULL = monocular, all < 0); skip2[n2] = feature of
 * frame 2 already has a map point (:722-726); F12 row-major 3x3; (ex, ey) the epipole in image 2 (:663-670);
 * scale_factors2 / level_sigma2_2 of key frame 2.  The caller drops sources with a map point and, with
 * only_stereo, the monocular ones (:699-709).  match12[n1] = feature of frame 2 or -1. */
int orb_gpu_search_for_triangulation(const OrbBucketQuery* q, int nq, const int32_t* cand2, int ncand2,
                                     const OrbKeyPoint* kps1, const uint8_t* desc1,
                                     const float* u_right1, int n1, const OrbKeyPoint* kps2,
                                     const uint8_t* desc2, const float* u_right2,
                                     const uint8_t* skip2, int n2, const float* F12, float ex, float ey,
                                     const float* scale_factors2, const float* level_sigma2_2,
                                     int nlevels, int only_stereo, int check_orientation, int device,
                                     int32_t* match12, int* nmatches);

/* Replaces ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520) up to :511: level-0 sources, a
 * `window`-pixel box around prev_matched[i] = (x, y) in frame 2's grid, top-2 from INT_MAX, per-target best
 * distance table with steal-back, rotation histogram.  match12[n1] = index in frame 2 or -1; the caller updates
 * vbPrevMatched from it (:514-517). */
int orb_gpu_search_for_initialization(const OrbKeyPoint* kps1, const uint8_t* desc1, int n1,
                                      const float* prev_matched, const OrbKeyPoint* kps2,
                                      const uint8_t* desc2, int n2, OrbFrameGrid grid2, int window,
                                      float nnratio, int check_orientation, int device,
                                      int32_t* match12, int* nmatches);

/* Hamming stage of Frame::ComputeStereoMatches (src/Frame.cc:466-549): for each left keypoint
 * the best right keypoint in its row band within the disparity and octave gates.
 * best_idx_r[nl] = right index or -1, best_dist[nl] = distance (th_high when none). */
int orb_gpu_stereo_match(const OrbKeyPoint* kl, const uint8_t* dl, int nl, const OrbKeyPoint* kr,
                         const uint8_t* dr, int nr, int n_rows, const float* scale_factors,
                         int nlevels, float min_d, float max_d, int th_high, int device,
                         int32_t* best_idx_r, int32_t* best_dist);

/* test/bench knob for the brute-force kernel: csa = 0 (8 POPC per pair), 1 (carry-save, 4 POPC per pair),
 * 2 (tensor cores through mma.sync m16n8k32 s8 x u8) or 3 (tcgen05.mma kind::i8 with TMEM accumulators; the
 * default); qpt = queries per thread of forms 0 and 1 (1, 2 or 4).  All forms return identical results. */
int orb_gpu_set_bruteforce_variant(int csa, int qpt);
/* Test / bench knob for A2 (ComputePyramid, src/ORBextractor.cc:1107-1132): 0 = the byte-SIMD resize kernel (default),
 * 1 = the form with the horizontal pass on the tensor cores (tcgen05.mma + TMA) for batches of 8 frames and more.
 * Identical pixels; the second form is measured slower and kept as evidence. */
int orb_gpu_set_resize_variant(int variant);

/* Per-stage device time of the last extract call on this handle (CUDA events on the stream the
 * kernels ran on): ms5 = pyramid, FAST, quadtree, blur, orientation+descriptors. */
int orb_gpu_set_stage_timing(OrbGpuExtractor* h, int enable);
int orb_gpu_get_stage_ms(OrbGpuExtractor* h, float* ms5);

/* ---- diagnostics ----------------------------------------------------------------------- */
const char* orb_gpu_version(void);
/* last CUDA error string seen by this thread inside the library ("" if none) */
const char* orb_gpu_last_error(void);
/* number of kernels this library has launched in this process (for bench.py gpu_launches) */
int64_t orb_gpu_launch_count(void);
/* dependent-free POPC stream on all SMs: returns achieved 32-bit POPC/s (integer-pipe peak
 * denominator for the Hamming roofline; MEASURED_PEAKS.json has no integer figure). */
int orb_gpu_popc_peak(int device, double* popc_per_s);
/* Measured multiply-accumulate rate (MAC/s) of mma.sync m16n8k32 s8 x u8 -> s32 (IMMA), the tensor-core path of
 * brute-force variant 2: 256 MACs per descriptor pair. */
int orb_gpu_imma_peak(int device, double* mac_per_s);
/* Measured multiply-accumulate rate (MAC/s) of tcgen05.mma kind::i8 (M128 N256 K32, operands in shared memory,
 * accumulators in TMEM), the path of the default brute-force kernel (variant 3): 256 MACs per descriptor pair. */
int orb_gpu_umma_peak(int device, double* mac_per_s);

#ifdef __cplusplus
}
#endif
#endif /* ORB_GPU_H_ */
