"""Python host side of the C ABI (include/orb_gpu.h), mirroring the reference's two classes:

  ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)   include/ORBextractor.h:45-111
      __call__(image) -> (keypoints, descriptors)                        src/ORBextractor.cc:1043-1105
      GetLevels / GetScaleFactor / GetScaleFactors / GetInverseScaleFactors /
      GetScaleSigmaSquares / GetInverseScaleSigmaSquares, mvImagePyramid
  ORBmatcher(nnratio=0.6, checkOri=True)                                  include/ORBmatcher.h:37-102
      DescriptorDistance, SearchByProjection (Frame, MapPoints, th), TH_HIGH/TH_LOW/HISTO_LENGTH

Everything computes in liborb_gpu.so (hand-written sm_100a kernels).  There is no CPU path: if the
library cannot be loaded or no CUDA device is present, calls raise OrbGpuError.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liborb_gpu.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
PROJ_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("ur", "<f4"), ("level", "<i4"),
                       ("view_cos", "<f4"), ("valid", "u1"), ("has_obs", "u1"), ("pad", "u1", (2,))])
BUCKET_DTYPE = np.dtype([("index", "<i4"), ("begin", "<i4"), ("end", "<i4")])
WINDOW_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("radius", "<f4"), ("ur", "<f4"), ("min_level", "<i4"),
                         ("max_level", "<i4"), ("angle", "<f4"), ("valid", "u1"), ("blocks", "u1"),
                         ("pad", "u1", (2,))])
assert KP_DTYPE.itemsize == 28 and PROJ_DTYPE.itemsize == 24 and WINDOW_DTYPE.itemsize == 32

EXPORTS = [
    "orb_gpu_create", "orb_gpu_destroy", "orb_gpu_extract", "orb_gpu_extract_batch",
    "orb_gpu_extract_batch_device", "orb_gpu_get_levels", "orb_gpu_get_scale_factor",
    "orb_gpu_get_scale_tables", "orb_gpu_get_features_per_level", "orb_gpu_get_pyramid",
    "orb_gpu_get_candidates", "orb_gpu_hamming_pairs_device", "orb_gpu_match_bruteforce_device",
    "orb_gpu_match_bruteforce", "orb_gpu_merge_top2_device", "orb_gpu_ratio_test_device",
    "orb_gpu_search_by_projection", "orb_gpu_search_windowed", "orb_gpu_features_in_area", "orb_gpu_search_by_bow",
    "orb_gpu_search_for_triangulation", "orb_gpu_search_for_initialization", "orb_gpu_stereo_match",
    "orb_gpu_compute_stereo_matches", "orb_gpu_stereo_matches_device", "orb_gpu_distinctive_descriptors",
    "orb_gpu_distinctive_descriptors_device", "orb_gpu_vocabulary_create", "orb_gpu_vocabulary_destroy",
    "orb_gpu_vocabulary_transform", "orb_gpu_vocabulary_transform_device", "orb_gpu_version", "orb_gpu_last_error",
    "orb_gpu_launch_count", "orb_gpu_popc_peak", "orb_gpu_imma_peak", "orb_gpu_umma_peak",
    "orb_gpu_frame_create", "orb_gpu_frame_destroy", "orb_gpu_frame_size", "orb_gpu_frame_upload",
    "orb_gpu_frame_from_device", "orb_gpu_frame_search_by_projection", "orb_gpu_frame_search_windowed",
    "orb_gpu_match_bruteforce_noalloc_device", "orb_gpu_exchange_match_bruteforce_device", "orb_gpu_set_resize_variant",
]

STATUS = {0: "OK", -1: "ERR_ARG", -2: "ERR_CAPACITY", -3: "ERR_CUDA", -4: "ERR_STATE"}


class OrbGpuError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        msg = "%s failed: %s (%d)" % (where, STATUS.get(code, "?"), code)
        try:
            detail = lib().orb_gpu_last_error().decode()
            if detail:
                msg += " -- " + detail
        except Exception:
            pass
        super().__init__(msg)


class _Grid(C.Structure):
    _fields_ = [("min_x", C.c_float), ("min_y", C.c_float), ("inv_w", C.c_float), ("inv_h", C.c_float)]


_lib = None


def lib():
    """The loaded C-ABI library.  Raises if liborb_gpu.so has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OrbGpuError(-3, "loading %s (run `python -m c_orb_slam_b200.build` / __graft_entry__.build())" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.orb_gpu_version.restype = C.c_char_p
        L.orb_gpu_last_error.restype = C.c_char_p
        L.orb_gpu_launch_count.restype = C.c_int64
        L.orb_gpu_get_scale_factor.restype = C.c_float
        L.orb_gpu_get_scale_factor.argtypes = [C.c_void_p]
        L.orb_gpu_get_levels.argtypes = [C.c_void_p]
        L.orb_gpu_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.orb_gpu_destroy.argtypes = [C.c_void_p]
        L.orb_gpu_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int,
                                      C.POINTER(C.c_int)]
        L.orb_gpu_extract_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                            C.c_size_t, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orb_gpu_extract_batch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                                   C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int,
                                                   C.c_void_p, C.c_void_p, C.c_void_p]
        L.orb_gpu_get_scale_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.orb_gpu_get_features_per_level.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orb_gpu_get_pyramid.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t,
                                          C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orb_gpu_get_candidates.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.orb_gpu_hamming_pairs_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orb_gpu_match_bruteforce_device.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int32,
                                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orb_gpu_match_bruteforce_noalloc_device.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int32,
                                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orb_gpu_match_bruteforce.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int,
                                               C.c_void_p, C.c_void_p, C.c_void_p]
        L.orb_gpu_merge_top2_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orb_gpu_ratio_test_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                                C.c_void_p, C.c_void_p]
        L.orb_gpu_exchange_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.orb_gpu_exchange_local_handle.argtypes = [C.c_void_p, C.c_void_p]
        L.orb_gpu_exchange_connect.argtypes = [C.c_void_p, C.c_void_p]
        L.orb_gpu_exchange_connect_local.argtypes = [C.POINTER(C.c_void_p), C.c_int]
        L.orb_gpu_exchange_destroy.argtypes = [C.c_void_p]
        L.orb_gpu_exchange_top2_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orb_gpu_exchange_match_bruteforce_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64,
                                                               C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orb_gpu_search_by_projection.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                   C.c_void_p, C.c_int, _Grid, C.c_void_p, C.c_int, C.c_float,
                                                   C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                   C.POINTER(C.c_int)]
        L.orb_gpu_search_windowed.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_int, _Grid, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                              C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]
        vp, ci, cf = C.c_void_p, C.c_int, C.c_float
        L.orb_gpu_frame_create.argtypes = [ci, C.POINTER(vp)]
        L.orb_gpu_frame_destroy.argtypes = [vp]
        L.orb_gpu_frame_size.argtypes = [vp]
        L.orb_gpu_frame_upload.argtypes = [vp, vp, vp, vp, ci, _Grid]
        L.orb_gpu_frame_from_device.argtypes = [vp, vp, vp, vp, ci, _Grid, vp]
        L.orb_gpu_frame_search_by_projection.argtypes = [vp, vp, vp, ci, vp, ci, cf, cf, ci, vp, vp, C.POINTER(ci)]
        L.orb_gpu_frame_search_windowed.argtypes = [vp, vp, vp, ci, ci, ci, vp, ci, ci, vp, vp, vp, C.POINTER(ci)]
        L.orb_gpu_search_by_bow.argtypes = [vp, ci, vp, ci, vp, vp, ci, vp, vp, vp, ci, cf, ci, ci, ci, vp, vp,
                                            C.POINTER(ci)]
        L.orb_gpu_search_for_triangulation.argtypes = [vp, ci, vp, ci, vp, vp, vp, ci, vp, vp, vp, vp, ci, vp, cf, cf,
                                                       vp, vp, ci, ci, ci, ci, vp, C.POINTER(ci)]
        L.orb_gpu_search_for_initialization.argtypes = [vp, vp, ci, vp, vp, vp, ci, _Grid, ci, cf, ci, ci, vp,
                                                        C.POINTER(ci)]
        L.orb_gpu_compute_stereo_matches.argtypes = [vp, ci, vp, ci, vp, vp, ci, vp, vp, ci, cf, cf, vp, vp]
        L.orb_gpu_stereo_matches_device.argtypes = [vp, ci, ci, vp, ci, ci, ci, vp, vp, vp, ci, vp, vp, vp, ci, cf, cf,
                                                    vp, vp, vp, vp]
        L.orb_gpu_distinctive_descriptors.argtypes = [vp, vp, ci, ci, vp]
        L.orb_gpu_distinctive_descriptors_device.argtypes = [vp, vp, ci, ci, vp, vp, vp]
        L.orb_gpu_vocabulary_create.argtypes = [ci, vp, vp, vp, vp, vp, ci, ci, C.POINTER(vp)]
        L.orb_gpu_vocabulary_destroy.argtypes = [vp]
        L.orb_gpu_vocabulary_destroy.restype = None
        L.orb_gpu_vocabulary_transform.argtypes = [vp, vp, ci, ci, vp, vp, vp]
        L.orb_gpu_vocabulary_transform_device.argtypes = [vp, vp, ci, ci, vp, vp, vp, vp]
        L.orb_gpu_features_in_area.argtypes = [vp, ci, vp, ci, _Grid, ci, vp, vp, ci]
        L.orb_gpu_stereo_match.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                           C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int,
                                           C.c_void_p, C.c_void_p]
        L.orb_gpu_set_stage_timing.argtypes = [C.c_void_p, C.c_int]
        L.orb_gpu_get_stage_ms.argtypes = [C.c_void_p, C.c_void_p]
        L.orb_gpu_popc_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
        L.orb_gpu_imma_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
        L.orb_gpu_umma_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
        L.orb_gpu_set_bruteforce_variant.argtypes = [C.c_int, C.c_int]
        _lib = L
    return _lib


def _check(code, where):
    if code != 0:
        raise OrbGpuError(code, where)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def launch_count():
    return int(lib().orb_gpu_launch_count())


def imma_peak(device=0):
    """Measured MAC/s of the legacy int8 tensor path (mma.sync m16n8k32) on `device`."""
    v = C.c_double()
    _check(lib().orb_gpu_imma_peak(device, C.byref(v)), "orb_gpu_imma_peak")
    return v.value


def umma_peak(device=0):
    """Measured MAC/s of tcgen05.mma kind::i8 (M128 N256 K32, the default brute-force kernel's path) on `device`."""
    v = C.c_double()
    _check(lib().orb_gpu_umma_peak(device, C.byref(v)), "orb_gpu_umma_peak")
    return v.value


def popc_peak(device=0):
    v = C.c_double()
    _check(lib().orb_gpu_popc_peak(device, C.byref(v)), "orb_gpu_popc_peak")
    return v.value


class ORBextractor:
    """Drop-in for ORB_SLAM2::ORBextractor (include/ORBextractor.h:45-111)."""

    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                 max_width=1920, max_height=1200, max_batch=1, device=0):
        self._h = C.c_void_p()
        self.nfeatures, self.nlevels, self.device = nfeatures, nlevels, device
        self.capacity = nfeatures + 3 * nlevels
        _check(lib().orb_gpu_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_width,
                                    max_height, max_batch, device, C.byref(self._h)), "orb_gpu_create")

    def close(self):
        if getattr(self, "_h", None):
            lib().orb_gpu_destroy(self._h)
            self._h = None

    __del__ = close

    # -- getters (ORBextractor.h:63-83) --
    def GetLevels(self):
        return lib().orb_gpu_get_levels(self._h)

    def GetScaleFactor(self):
        return lib().orb_gpu_get_scale_factor(self._h)

    def _tables(self):
        t = [np.empty(self.nlevels, np.float32) for _ in range(4)]
        _check(lib().orb_gpu_get_scale_tables(self._h, *[_p(a) for a in t]), "orb_gpu_get_scale_tables")
        return t

    def GetScaleFactors(self):
        return self._tables()[0]

    def GetInverseScaleFactors(self):
        return self._tables()[1]

    def GetScaleSigmaSquares(self):
        return self._tables()[2]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()[3]

    def features_per_level(self):
        q = np.empty(self.nlevels, np.int32)
        u = np.empty(16, np.int32)
        _check(lib().orb_gpu_get_features_per_level(self._h, _p(q), _p(u)), "orb_gpu_get_features_per_level")
        return q, u

    # -- operator() --
    def __call__(self, image, mask=None):
        """-> (keypoints[K] as KP_DTYPE, descriptors[K,32] uint8).  `mask` is ignored, as in the
        reference (ORBextractor.h:58).  An empty image returns empty outputs (:1046)."""
        image = np.asarray(image)
        if image.size == 0:
            return np.empty(0, KP_DTYPE), np.empty((0, 32), np.uint8)
        n, kps, desc = self.extract_batch(image[None])
        return kps[0, :n[0]].copy(), desc[0, :n[0]].copy()

    def extract_batch(self, frames):
        """frames uint8 [B,H,W] (host) -> counts[B], kps[B,cap], desc[B,cap,32]."""
        frames = np.asarray(frames)
        if frames.dtype != np.uint8 or frames.ndim != 3:
            raise TypeError("expected uint8 [B,H,W] (the reference asserts CV_8UC1, ORBextractor.cc:1050)")
        if frames.strides[2] != 1 or frames.strides[1] < frames.shape[2]:
            frames = np.ascontiguousarray(frames)
        B, H, W = frames.shape
        cap = self.capacity
        n = np.zeros(B, np.int32)
        kps = np.zeros((B, cap), KP_DTYPE)
        desc = np.zeros((B, cap, 32), np.uint8)
        _check(lib().orb_gpu_extract_batch(self._h, _p(frames), B, W, H, frames.strides[1], frames.strides[0],
                                           _p(kps), _p(desc), cap, _p(n)), "orb_gpu_extract_batch")
        return n, kps, desc

    def extract_batch_device(self, d_frames_ptr, B, W, H, stride, frame_stride, d_kps_ptr, d_desc_ptr,
                             d_n_ptr, d_status_ptr=None, stream=None, capacity=None):
        """All pointers are device addresses (ints); asynchronous on `stream` (cudaStream_t int)."""
        _check(lib().orb_gpu_extract_batch_device(self._h, d_frames_ptr, B, W, H, stride, frame_stride,
                                                  d_kps_ptr, d_desc_ptr, capacity or self.capacity, d_n_ptr,
                                                  d_status_ptr, stream), "orb_gpu_extract_batch_device")

    def set_stage_timing(self, enable=True):
        _check(lib().orb_gpu_set_stage_timing(self._h, int(enable)), "orb_gpu_set_stage_timing")

    def stage_ms(self):
        """Device ms of the last call: dict(pyramid, fast, quadtree, blur, orient_desc)."""
        ms = np.empty(5, np.float32)
        _check(lib().orb_gpu_get_stage_ms(self._h, _p(ms)), "orb_gpu_get_stage_ms")
        return dict(zip(["pyramid", "fast", "quadtree", "blur", "orient_desc"], ms.tolist()))

    def pyramid_level(self, level, frame=0, blurred=False):
        """mvImagePyramid[level] of the last call (ORBextractor.h:85)."""
        w, h = C.c_int(), C.c_int()
        _check(lib().orb_gpu_get_pyramid(self._h, frame, level, int(blurred), None, 0, C.byref(w), C.byref(h)),
               "orb_gpu_get_pyramid")
        out = np.empty((h.value, w.value), np.uint8)
        _check(lib().orb_gpu_get_pyramid(self._h, frame, level, int(blurred), _p(out), w.value, C.byref(w),
                                         C.byref(h)), "orb_gpu_get_pyramid")
        return out

    @property
    def mvImagePyramid(self):
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    def candidates(self, level, frame=0):
        n = C.c_int()
        _check(lib().orb_gpu_get_candidates(self._h, frame, level, None, 0, C.byref(n)), "orb_gpu_get_candidates")
        out = np.empty((n.value, 3), np.float32)
        if n.value:
            _check(lib().orb_gpu_get_candidates(self._h, frame, level, _p(out), n.value, C.byref(n)),
                   "orb_gpu_get_candidates")
        return out


class ORBVocabulary:
    """Device-resident DBoW2 vocabulary (ORB_SLAM2::ORBVocabulary = TemplatedVocabulary<FORB::TDescriptor, FORB>,
    include/ORBVocabulary.h) for the transform of Frame::ComputeBoW / KeyFrame::ComputeBoW
    (src/Frame.cc:395-402): TF_IDF weighting, L1 norm (the header line of ORBvoc.txt)."""

    def __init__(self, tree, device=0):
        self.depth = int(tree["depth"])
        self._keep = [np.ascontiguousarray(tree["child_begin"], np.int32), np.ascontiguousarray(tree["children"], np.int32),
                      np.ascontiguousarray(tree["desc"], np.uint8), np.ascontiguousarray(tree["word_id"], np.int32),
                      np.ascontiguousarray(tree["weight"], np.float64)]
        self._h = C.c_void_p()
        _check(lib().orb_gpu_vocabulary_create(len(self._keep[3]), *[_p(a) for a in self._keep], self.depth, device,
                                               C.byref(self._h)), "orb_gpu_vocabulary_create")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orb_gpu_vocabulary_destroy(self._h)
            self._h = None

    def transform_features(self, desc, levelsup=4):
        """-> word_id uint32 [n], weight float64 [n], node_id uint32 [n] (TemplatedVocabulary.h:1213-1259)"""
        d = np.ascontiguousarray(desc, np.uint8)
        n = len(d)
        w = np.empty(n, np.uint32); v = np.empty(n, np.float64); nd = np.empty(n, np.uint32)
        _check(lib().orb_gpu_vocabulary_transform(self._h, _p(d), n, int(levelsup), _p(w), _p(v), _p(nd)),
               "orb_gpu_vocabulary_transform")
        return w, v, nd

    def transform(self, desc, levelsup=4):
        """TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1127-1192).
        -> BowVector as (word ids ascending uint32, values float64), FeatureVector {node id: [feature indices]}.
        The descent runs on the GPU; the two std::map fills are host bookkeeping in the reference's order."""
        words, weights, nodes = self.transform_features(desc, levelsup)
        bow, fv = {}, {}
        for i in range(len(words)):
            w = float(weights[i])
            if w > 0:                                   # not stopped (:1155)
                k = int(words[i])
                bow[k] = bow[k] + w if k in bow else w   # BowVector::addWeight
                fv.setdefault(int(nodes[i]), []).append(i)
        ids = np.array(sorted(bow), np.uint32)
        vals = np.array([bow[int(k)] for k in ids], np.float64)
        norm = 0.0
        for x in vals:                                  # BowVector::normalize(L1): ascending key order
            norm += abs(float(x))
        if norm > 0.0:
            vals = vals / norm
        return (ids, vals), fv


class GpuFrame:
    """The matcher-side data of a Frame / KeyFrame resident in device memory (mvKeysUn, mDescriptors, mvuRight and
    the 64 x 48 grid of Frame::AssignFeaturesToGrid, src/Frame.cc:230-253): filled once, searched many times."""

    def __init__(self, device=0):
        self.device = device
        self._h = C.c_void_p()
        _check(lib().orb_gpu_frame_create(device, C.byref(self._h)), "orb_gpu_frame_create")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orb_gpu_frame_destroy(self._h)
            self._h = None

    def __len__(self):
        return lib().orb_gpu_frame_size(self._h)

    def upload(self, kps, desc, u_right, grid):
        k = np.ascontiguousarray(kps, KP_DTYPE); d = np.ascontiguousarray(desc, np.uint8)
        ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        _check(lib().orb_gpu_frame_upload(self._h, _p(k), _p(d), _p(ur), len(k), _Grid(*[float(v) for v in grid])),
               "orb_gpu_frame_upload")
        return self

    def from_device(self, d_kps_ptr, d_desc_ptr, d_u_right_ptr, n, grid, stream=None):
        """Device addresses (ints), e.g. one frame of the outputs of ORBextractor.extract_batch_device."""
        _check(lib().orb_gpu_frame_from_device(self._h, d_kps_ptr, d_desc_ptr, d_u_right_ptr, int(n),
                                               _Grid(*[float(v) for v in grid]), stream), "orb_gpu_frame_from_device")
        return self


class ORBmatcher:
    """Drop-in for the Hamming-matching surface of ORB_SLAM2::ORBmatcher (include/ORBmatcher.h)."""

    TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30  # src/ORBmatcher.cc:37-39

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self.mfNNratio, self.mbCheckOrientation, self.device = float(nnratio), bool(checkOri), device

    @staticmethod
    def DescriptorDistance(a, b):
        """src/ORBmatcher.cc:1647-1663 for one pair (host inline: one 256-bit pair is not worth a
        launch); the batched device form is orb_gpu_hamming_pairs_device."""
        a = np.ascontiguousarray(a, np.uint8).view(np.uint64)
        b = np.ascontiguousarray(b, np.uint8).view(np.uint64)
        return int(sum(bin(int(x) ^ int(y)).count("1") for x, y in zip(a, b)))

    def match_bruteforce(self, queries, database):
        """query x database top-2 (BASELINE config 5): -> idx, best, second (int32 arrays)."""
        q = np.ascontiguousarray(queries, np.uint8)
        db = np.ascontiguousarray(database, np.uint8)
        nq = len(q)
        idx = np.empty(nq, np.int32); best = np.empty(nq, np.int32); second = np.empty(nq, np.int32)
        _check(lib().orb_gpu_match_bruteforce(_p(q), nq, _p(db), len(db), self.device, _p(idx), _p(best),
                                              _p(second)), "orb_gpu_match_bruteforce")
        return idx, best, second

    def accept(self, best, second, th=None):
        """best <= th && best < nnratio*second (ORBmatcher.cc:228-230), evaluated like the kernel."""
        th = self.TH_LOW if th is None else th
        b = np.asarray(best); s = np.asarray(second)
        return (b <= th) & (b.astype(np.float32) < np.float32(self.mfNNratio) * s.astype(np.float32))

    def accept_device(self, best, second, th=None):
        """The same flags from the device kernel (orb_gpu_ratio_test_device); best / second: host int32 arrays."""
        import torch
        th = self.TH_LOW if th is None else th
        b = torch.from_numpy(np.ascontiguousarray(best, np.int32)).cuda(self.device)
        s = torch.from_numpy(np.ascontiguousarray(second, np.int32)).cuda(self.device)
        out = torch.empty(len(b), dtype=torch.uint8, device=b.device)
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream().cuda_stream
            _check(lib().orb_gpu_ratio_test_device(b.data_ptr(), s.data_ptr(), len(b), int(th), self.mfNNratio,
                                                   out.data_ptr(), st), "orb_gpu_ratio_test_device")
            torch.cuda.synchronize()
        return out.cpu().numpy().astype(bool)

    def SearchByProjection(self, frame, map_points, th=3.0):
        """ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) (ORBmatcher.cc:45-129).
        frame: dict(kps KP_DTYPE (undistorted), desc [N,32], u_right [N] or None, grid (minX,minY,invW,invH),
                    scale_factors [L], blocked uint8 [N] (keypoint holds a MapPoint with observations)).
        map_points: dict(proj PROJ_DTYPE [M], desc [M,32]).
        -> nmatches, match_of_kp [N] (map-point index or -1); frame['blocked'] is updated in place."""
        res = frame.get("resident")      # a GpuFrame: the target stays on the device, only the queries move
        nk = len(res) if res is not None else len(frame["kps"])
        sf = np.ascontiguousarray(frame["scale_factors"], np.float32)
        blocked = frame.get("blocked")
        if blocked is None:
            blocked = frame["blocked"] = np.zeros(nk, np.uint8)
        proj = np.ascontiguousarray(map_points["proj"], PROJ_DTYPE)
        qd = np.ascontiguousarray(map_points["desc"], np.uint8)
        if res is not None:
            out = np.empty(nk, np.int32)
            n = C.c_int()
            _check(lib().orb_gpu_frame_search_by_projection(res._h, _p(proj), _p(qd), len(proj), _p(sf), len(sf), th,
                                                            self.mfNNratio, self.TH_HIGH, _p(blocked), _p(out),
                                                            C.byref(n)), "orb_gpu_frame_search_by_projection")
            return n.value, out
        kps = np.ascontiguousarray(frame["kps"], KP_DTYPE)
        kd = np.ascontiguousarray(frame["desc"], np.uint8)
        ur = frame.get("u_right")
        ur = None if ur is None else np.ascontiguousarray(ur, np.float32)
        g = _Grid(*[float(v) for v in frame["grid"]])
        out = np.empty(len(kps), np.int32)
        n = C.c_int()
        _check(lib().orb_gpu_search_by_projection(_p(proj), _p(qd), len(proj), _p(kps), _p(kd), _p(ur), len(kps),
                                                  g, _p(sf), len(sf), th, self.mfNNratio, self.TH_HIGH,
                                                  self.device, _p(blocked), _p(out), C.byref(n)),
               "orb_gpu_search_by_projection")
        return n.value, out

    def SearchWindowed(self, frame, queries, qdesc, th_accept=None, gate=0, inv_level_sigma2=None):
        """Best-only windowed search shared by ORBmatcher::SearchByProjection(Cur, Last, th, bMono)
        (ORBmatcher.cc:1328-1470; th_accept=TH_HIGH, stereo_gate=True), (Cur, KF, sAlreadyFound, th, ORBdist)
        (:1472-1599; th_accept=ORBdist) and (KF, Scw, vpPoints, vpMatched, th) (:290-403; th_accept=TH_LOW,
        no orientation check).  queries: WINDOW_DTYPE [M] computed by the caller like the reference does
        (projection, radius, level window, source angle).  frame as in SearchByProjection.
        gate: 0 none, 1 stereo radius gate (:1407-1413), 2 chi-square gate of Fuse (:904-931, needs
        inv_level_sigma2).  With blocks = 0 in every query nothing is claimed (the two passes of SearchBySim3).
        -> nmatches, match_of_kp [N], match_of_query [M]; frame['blocked'] is updated in place."""
        res = frame.get("resident")
        nk = len(res) if res is not None else len(frame["kps"])
        blocked = frame.get("blocked")
        if blocked is None:
            blocked = frame["blocked"] = np.zeros(nk, np.uint8)
        wq = np.ascontiguousarray(queries, WINDOW_DTYPE)
        sig = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
        outq = np.empty(len(wq), np.int32)
        qd = np.ascontiguousarray(qdesc, np.uint8)
        th_acc = self.TH_HIGH if th_accept is None else int(th_accept)
        if res is not None:
            out = np.empty(nk, np.int32)
            n = C.c_int()
            _check(lib().orb_gpu_frame_search_windowed(res._h, _p(wq), _p(qd), len(wq), th_acc, int(gate), _p(sig),
                                                       0 if sig is None else len(sig), int(self.mbCheckOrientation),
                                                       _p(blocked), _p(out), _p(outq), C.byref(n)),
                   "orb_gpu_frame_search_windowed")
            return n.value, out, outq
        kps = np.ascontiguousarray(frame["kps"], KP_DTYPE)
        kd = np.ascontiguousarray(frame["desc"], np.uint8)
        ur = frame.get("u_right")
        ur = None if ur is None else np.ascontiguousarray(ur, np.float32)
        g = _Grid(*[float(v) for v in frame["grid"]])
        out = np.empty(len(kps), np.int32)
        n = C.c_int()
        th_accept = self.TH_HIGH if th_accept is None else int(th_accept)
        _check(lib().orb_gpu_search_windowed(_p(wq), _p(qd), len(wq), _p(kps), _p(kd), _p(ur), len(kps), g,
                                             th_accept, int(gate), _p(sig), 0 if sig is None else len(sig),
                                             int(self.mbCheckOrientation), self.device, _p(blocked), _p(out),
                                             _p(outq), C.byref(n)),
               "orb_gpu_search_windowed")
        return n.value, out, outq

    @staticmethod
    def bucket_queries(fv1, fv2, keep1):
        """Walk two DBoW2 feature vectors ({node id: [feature indices]}) the way the reference's two iterators
        do (src/ORBmatcher.cc:178-266: nodes ascending, equal ids only) and emit one BUCKET_DTYPE row per source
        feature with keep1[index] set, plus the flat candidate list of frame 2.  Host bookkeeping only."""
        rows, cand = [], []
        for node in sorted(set(fv1) & set(fv2)):
            begin = len(cand)
            cand.extend(int(j) for j in fv2[node])
            rows.extend((int(i), begin, len(cand)) for i in fv1[node] if keep1[i])
        q = np.array(rows, BUCKET_DTYPE) if rows else np.zeros(0, BUCKET_DTYPE)
        return q, np.asarray(cand, np.int32)

    def SearchByBoW(self, kf, frame, good_mp1, good_mp2=None):
        """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) (ORBmatcher.cc:159-288) when good_mp2 is None, else
        SearchByBoW(KeyFrame*, KeyFrame*, ...) (:522-655, strict best < TH_LOW).  kf / frame: dict(kps, desc,
        feat_vec {node: [indices]}); good_mp*: feature holds a map point that is not bad.
        -> nmatches, match12 [n1], match21 [n2]"""
        k1 = np.ascontiguousarray(kf["kps"], KP_DTYPE); d1 = np.ascontiguousarray(kf["desc"], np.uint8)
        k2 = np.ascontiguousarray(frame["kps"], KP_DTYPE); d2 = np.ascontiguousarray(frame["desc"], np.uint8)
        q, cand = self.bucket_queries(kf["feat_vec"], frame["feat_vec"], np.asarray(good_mp1, bool))
        skip2 = None if good_mp2 is None else np.ascontiguousarray(~np.asarray(good_mp2, bool), np.uint8)
        th = self.TH_LOW if good_mp2 is None else self.TH_LOW - 1
        m12 = np.empty(len(k1), np.int32); m21 = np.empty(len(k2), np.int32)
        n = C.c_int()
        _check(lib().orb_gpu_search_by_bow(_p(q), len(q), _p(cand), len(cand), _p(k1), _p(d1), len(k1), _p(k2),
                                           _p(d2), _p(skip2), len(k2), self.mfNNratio, th,
                                           int(self.mbCheckOrientation), self.device, _p(m12), _p(m21),
                                           C.byref(n)), "orb_gpu_search_by_bow")
        return n.value, m12, m21

    def SearchForTriangulation(self, kf1, kf2, F12, epipole, has_mp1, has_mp2, only_stereo=False):
        """ORBmatcher::SearchForTriangulation (ORBmatcher.cc:657-823).  kf*: dict(kps, desc, feat_vec, u_right or
        None); kf2 also scale_factors, level_sigma2.  -> nmatches, match12 [n1]"""
        k1 = np.ascontiguousarray(kf1["kps"], KP_DTYPE); d1 = np.ascontiguousarray(kf1["desc"], np.uint8)
        k2 = np.ascontiguousarray(kf2["kps"], KP_DTYPE); d2 = np.ascontiguousarray(kf2["desc"], np.uint8)
        u1 = kf1.get("u_right"); u2 = kf2.get("u_right")
        u1 = None if u1 is None else np.ascontiguousarray(u1, np.float32)
        u2 = None if u2 is None else np.ascontiguousarray(u2, np.float32)
        keep1 = ~np.asarray(has_mp1, bool)
        if only_stereo:
            keep1 &= (u1 >= 0) if u1 is not None else False
        q, cand = self.bucket_queries(kf1["feat_vec"], kf2["feat_vec"], keep1)
        skip2 = np.ascontiguousarray(has_mp2, np.uint8)
        F = np.ascontiguousarray(F12, np.float32).reshape(9)
        sf = np.ascontiguousarray(kf2["scale_factors"], np.float32)
        sg = np.ascontiguousarray(kf2["level_sigma2"], np.float32)
        m12 = np.empty(len(k1), np.int32)
        n = C.c_int()
        _check(lib().orb_gpu_search_for_triangulation(
            _p(q), len(q), _p(cand), len(cand), _p(k1), _p(d1), _p(u1), len(k1), _p(k2), _p(d2), _p(u2), _p(skip2),
            len(k2), _p(F), float(epipole[0]), float(epipole[1]), _p(sf), _p(sg), len(sf), int(bool(only_stereo)),
            int(self.mbCheckOrientation), self.device, _p(m12), C.byref(n)), "orb_gpu_search_for_triangulation")
        return n.value, m12

    def SearchForInitialization(self, f1, f2, prev_matched, window=100):
        """ORBmatcher::SearchForInitialization (ORBmatcher.cc:405-520).  f1 / f2: dict(kps, desc), f2 also grid;
        prev_matched float32 [n1, 2] is updated in place from the matches (:514-517).
        -> nmatches, vnMatches12 [n1]"""
        k1 = np.ascontiguousarray(f1["kps"], KP_DTYPE); d1 = np.ascontiguousarray(f1["desc"], np.uint8)
        k2 = np.ascontiguousarray(f2["kps"], KP_DTYPE); d2 = np.ascontiguousarray(f2["desc"], np.uint8)
        if prev_matched.dtype != np.float32 or prev_matched.shape != (len(k1), 2) or not prev_matched.flags.c_contiguous:
            raise ValueError("prev_matched must be a C-contiguous float32 [n1, 2] array")
        g = _Grid(*[float(v) for v in f2["grid"]])
        m12 = np.empty(len(k1), np.int32)
        n = C.c_int()
        _check(lib().orb_gpu_search_for_initialization(_p(k1), _p(d1), len(k1), _p(prev_matched), _p(k2), _p(d2),
                                                       len(k2), g, int(window), self.mfNNratio,
                                                       int(self.mbCheckOrientation), self.device, _p(m12),
                                                       C.byref(n)), "orb_gpu_search_for_initialization")
        hit = m12 >= 0
        prev_matched[hit, 0] = k2["x"][m12[hit]]
        prev_matched[hit, 1] = k2["y"][m12[hit]]
        return n.value, m12

    def GetFeaturesInArea(self, frame, windows):
        """Frame::GetFeaturesInArea (src/Frame.cc:327-380) for a batch of windows (WINDOW_DTYPE rows: u, v, radius,
        min_level, max_level, valid).  -> list of int32 index arrays in the reference's candidate order."""
        kps = np.ascontiguousarray(frame["kps"], KP_DTYPE)
        wq = np.ascontiguousarray(windows, WINDOW_DTYPE)
        g = _Grid(*[float(v) for v in frame["grid"]])
        offs = np.zeros(len(wq) + 1, np.int32)
        idx = np.empty(max(1, 64 * len(wq)), np.int32)
        rc = lib().orb_gpu_features_in_area(_p(wq), len(wq), _p(kps), len(kps), g, self.device, _p(offs), _p(idx),
                                            len(idx))
        if rc == -2 and offs[-1] > len(idx):
            idx = np.empty(int(offs[-1]), np.int32)
            rc = lib().orb_gpu_features_in_area(_p(wq), len(wq), _p(kps), len(kps), g, self.device, _p(offs), _p(idx),
                                                len(idx))
        _check(rc, "orb_gpu_features_in_area")
        return [idx[offs[i]:offs[i + 1]].copy() for i in range(len(wq))]

    def ComputeDistinctiveDescriptors(self, desc_lists):
        """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307) for many map points at once.
        desc_lists: one [N_p, 32] uint8 array per map point (its observations in mObservations order).
        -> best index per point (int32; -1 for an empty list); the caller clones that row into mDescriptor."""
        offs = np.zeros(len(desc_lists) + 1, np.int32)
        for i, d in enumerate(desc_lists):
            offs[i + 1] = offs[i] + len(d)
        flat = (np.concatenate([np.ascontiguousarray(d, np.uint8).reshape(-1, 32) for d in desc_lists])
                if len(desc_lists) and offs[-1] else np.zeros((0, 32), np.uint8))
        best = np.empty(len(desc_lists), np.int32)
        _check(lib().orb_gpu_distinctive_descriptors(_p(flat), _p(offs), len(desc_lists), self.device, _p(best)),
               "orb_gpu_distinctive_descriptors")
        return best

    def ComputeStereoMatches(self, ex_left, ex_right, kps_l, desc_l, kps_r, desc_r, bf, b, frame_l=0, frame_r=0):
        """Frame::ComputeStereoMatches (src/Frame.cc:466-640) on the pyramids the two ORBextractor objects kept on
        the device from their last call.  -> mvuRight [nl], mvDepth [nl] (float32, -1 = no match)"""
        kl = np.ascontiguousarray(kps_l, KP_DTYPE); dl = np.ascontiguousarray(desc_l, np.uint8)
        kr = np.ascontiguousarray(kps_r, KP_DTYPE); dr = np.ascontiguousarray(desc_r, np.uint8)
        ur = np.empty(len(kl), np.float32); dp = np.empty(len(kl), np.float32)
        _check(lib().orb_gpu_compute_stereo_matches(ex_left._h, frame_l, ex_right._h, frame_r, _p(kl), _p(dl), len(kl),
                                                    _p(kr), _p(dr), len(kr), float(bf), float(b), _p(ur), _p(dp)),
               "orb_gpu_compute_stereo_matches")
        return ur, dp

    def stereo_match(self, kps_l, desc_l, kps_r, desc_r, n_rows, scale_factors, min_d, max_d):
        """Hamming stage of Frame::ComputeStereoMatches (Frame.cc:466-549) -> best_idx_r, best_dist."""
        kl = np.ascontiguousarray(kps_l, KP_DTYPE); kr = np.ascontiguousarray(kps_r, KP_DTYPE)
        dl = np.ascontiguousarray(desc_l, np.uint8); dr = np.ascontiguousarray(desc_r, np.uint8)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        bi = np.empty(len(kl), np.int32); bd = np.empty(len(kl), np.int32)
        _check(lib().orb_gpu_stereo_match(_p(kl), _p(dl), len(kl), _p(kr), _p(dr), len(kr), n_rows, _p(sf),
                                          len(sf), min_d, max_d, self.TH_HIGH, self.device, _p(bi), _p(bd)),
               "orb_gpu_stereo_match")
        return bi, bd
