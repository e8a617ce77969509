"""ctypes front of the CPU oracle (oracle/liborb_oracle.so) and, when built, of the reference's
own ORBextractor.cc compiled verbatim (oracle/_ref/libref_orb.so).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never imported by c_orb_slam_b200.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liborb_oracle.so")
REF_PATH = os.path.join(HERE, "_ref", "libref_orb.so")
REF_SEQ_PATH = os.path.join(HERE, "_ref", "libref_orb_seq.so")
REF_MATCHER_PATH = os.path.join(HERE, "_ref", "libref_matcher.so")   # src/ORBmatcher.cc verbatim over oracle/slam_shim

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
PROJ_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("ur", "<f4"), ("level", "<i4"),
                       ("view_cos", "<f4"), ("valid", "u1"), ("has_obs", "u1")], align=True)
WINDOW_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("radius", "<f4"), ("ur", "<f4"), ("min_level", "<i4"),
                         ("max_level", "<i4"), ("angle", "<f4"), ("valid", "u1"), ("blocks", "u1")], align=True)
assert KP_DTYPE.itemsize == 28 and PROJ_DTYPE.itemsize == 24 and WINDOW_DTYPE.itemsize == 32


def build(force=False):
    """(Re)build the oracle with oracle/Makefile (gcc only; no GPU, no reference build system)."""
    # make is incremental (sources vs .so timestamps), so it is simply always consulted
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    subprocess.check_call(["make", "-s", "-C", HERE, "all"])


_lib = None
_ref = None
_refm = None


def have_ref_matcher():
    return os.path.exists(REF_MATCHER_PATH)


def ref_matcher():
    """The reference's own src/ORBmatcher.cc, compiled verbatim (oracle/Makefile target `refmatch`)."""
    global _refm
    if _refm is None:
        if not have_ref_matcher():
            raise RuntimeError("oracle/_ref/libref_matcher.so is not built (needs /root/reference)")
        _refm = C.CDLL(REF_MATCHER_PATH)
    return _refm


import contextlib


@contextlib.contextmanager
def matcher_library(path):
    """Inside the block every refm_* call (impl="reference", ref_search_*, ref_fuse*) goes to the library at `path`
    instead of the verbatim reference build.  tests/cpp/Makefile builds such a library from the same harness
    (oracle/ref_matcher_wrapper.cc) around the PRODUCT's drop-in ORB_SLAM2::ORBmatcher class, so a test can hand the
    same flat arrays to the reference's matcher and to the drop-in one."""
    global _refm
    keep = _refm
    _refm = C.CDLL(path)
    try:
        yield _refm
    finally:
        _refm = keep


def _m(impl, name):
    """Entry point `name` of the restated oracle (impl="oracle") or of the verbatim reference build
    (impl="reference"): both take the same flat arrays."""
    if impl == "oracle":
        return getattr(lib(), "orbo_" + name)
    if impl == "reference":
        return getattr(ref_matcher(), "refm_" + name)
    raise ValueError(impl)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orbo_fast_atan2.restype = C.c_float
        L.orbo_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orbo_extractor_create.restype = C.c_void_p
        L.orbo_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orbo_extract_batch.restype = C.c_long
        L.orbo_cv_round.argtypes = [C.c_double]
        _lib = L
    return _lib


def u8c(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


def resize_linear(src, dw, dh):
    src = u8c(src)
    dst = np.empty((dh, dw), np.uint8)
    lib().orbo_resize_linear_u8(_p(src), src.shape[1], src.shape[0], C.c_size_t(src.strides[0]),
                                _p(dst), dw, dh, C.c_size_t(dw))
    return dst


def resize_axis_table(S, D):
    ofs = np.empty(D, np.int32)
    a0 = np.empty(D, np.int16)
    a1 = np.empty(D, np.int16)
    lib().orbo_resize_axis_table(S, D, _p(ofs), _p(a0), _p(a1))
    return ofs, a0, a1


def gaussian_blur7(src):
    src = u8c(src)
    dst = np.empty_like(src)
    h, w = src.shape
    lib().orbo_gaussian_blur7_u8(_p(src), w, h, C.c_size_t(w), _p(dst), C.c_size_t(w))
    return dst


def copy_make_border(src, t, b, l, r):
    src = u8c(src)
    h, w = src.shape
    dst = np.empty((h + t + b, w + l + r), np.uint8)
    lib().orbo_copy_make_border_u8(_p(src), w, h, C.c_size_t(w), _p(dst), C.c_size_t(dst.shape[1]),
                                   t, b, l, r)
    return dst


def fast9_nms(img, threshold):
    """-> int32 array [n,3] of (x, y, score) in row-major emission order."""
    img = u8c(img)
    h, w = img.shape
    cap = max(16, w * h // 4 + 16)
    out = np.empty((cap, 3), np.int32)
    n = lib().orbo_fast9_nms(_p(img), w, h, C.c_size_t(w), threshold, _p(out), cap)
    return out[:n].copy()


def fast_atan2(y, x):
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(y)
    lib().orbo_fast_atan2_n(_p(y), _p(x), _p(out), int(y.size))
    return out


class Extractor:
    """Restated ORBextractor (oracle/orb_oracle.cc)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nlevels = nlevels
        self.args = (nfeatures, float(scale_factor), nlevels, ini_th, min_th)
        self._h = C.c_void_p(lib().orbo_extractor_create(nfeatures, C.c_float(scale_factor), nlevels,
                                                         ini_th, min_th))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orbo_extractor_destroy(self._h)
            self._h = None

    def tables(self):
        n = self.nlevels
        f = [np.empty(n, np.float32) for _ in range(4)]
        quota = np.empty(n, np.int32)
        umax = np.empty(16, np.int32)
        lib().orbo_extractor_tables(self._h, _p(f[0]), _p(f[1]), _p(f[2]), _p(f[3]), _p(quota), _p(umax))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3], quota=quota, umax=umax)

    def __call__(self, img):
        img = u8c(img)
        h, w = img.shape
        n = lib().orbo_extractor_run(self._h, _p(img), w, h, C.c_size_t(img.strides[0]))
        kps = np.empty(n, KP_DTYPE)
        desc = np.empty((n, 32), np.uint8)
        lib().orbo_extractor_results(self._h, _p(kps), _p(desc))
        return kps, desc

    def level_image(self, level, blurred=False):
        w = C.c_int()
        h = C.c_int()
        lib().orbo_extractor_level_size(self._h, level, C.byref(w), C.byref(h))
        if blurred and not lib().orbo_extractor_level_has_blur(self._h, level):
            return None
        out = np.empty((h.value, w.value), np.uint8)
        lib().orbo_extractor_level_image(self._h, level, int(blurred), _p(out))
        return out

    def candidates(self, level):
        n = lib().orbo_extractor_num_candidates(self._h, level)
        out = np.empty((n, 3), np.float32)
        lib().orbo_extractor_candidates(self._h, level, _p(out))
        return out

    def level_kps(self, level):
        n = lib().orbo_extractor_num_level_kps(self._h, level)
        out = np.empty(n, KP_DTYPE)
        lib().orbo_extractor_level_kps(self._h, level, _p(out))
        return out

    def stage_seconds(self):
        out = np.empty(6, np.float64)
        lib().orbo_extractor_stage_seconds(self._h, _p(out))
        return dict(zip(["pyramid", "fast", "quadtree", "orient", "blur", "desc"], out.tolist()))


def distribute_quadtree(xyr, minX, maxX, minY, maxY, N):
    xyr = np.ascontiguousarray(xyr, np.float32)
    cap = max(16, len(xyr))
    out = np.empty(cap, np.int32)
    n = lib().orbo_distribute_quadtree(_p(xyr), len(xyr), minX, maxX, minY, maxY, N, _p(out), cap)
    return out[:n].copy()


def extract_batch(frames, nfeatures, scale_factor, nlevels, ini_th, min_th, threads=1, cap=None,
                  want_output=True):
    """frames: uint8 [B,H,W] contiguous.  -> counts[B], kps[B,cap], desc[B,cap,32]"""
    frames = u8c(frames)
    B, H, W = frames.shape
    cap = cap or (nfeatures + 3 * nlevels)
    counts = np.zeros(B, np.int32)
    kps = np.zeros((B, cap), KP_DTYPE) if want_output else None
    desc = np.zeros((B, cap, 32), np.uint8) if want_output else None
    lib().orbo_extract_batch(nfeatures, C.c_float(scale_factor), nlevels, ini_th, min_th, _p(frames),
                             B, W, H, C.c_size_t(W), C.c_size_t(W * H), threads, _p(counts),
                             _p(kps), _p(desc), cap)
    return counts, kps, desc


def descriptor_distance(a, b, impl="oracle"):
    a = u8c(a)
    b = u8c(b)
    return _m(impl, "descriptor_distance")(_p(a), _p(b))


def match_bruteforce(q, db, threads=1):
    q = u8c(q)
    db = u8c(db)
    nq = len(q)
    idx = np.empty(nq, np.int32)
    best = np.empty(nq, np.int32)
    second = np.empty(nq, np.int32)
    lib().orbo_match_bruteforce(_p(q), nq, _p(db), C.c_int64(len(db)), _p(idx), _p(best), _p(second),
                                threads)
    return idx, best, second


def search_by_projection(queries, qdesc, kps, kdesc, u_right, grid, scale_factors, th, nnratio,
                         th_high=100, blocked=None, impl="oracle"):
    """grid = (minX, minY, invW, invH).  -> nmatches, match_of_kp[nk], blocked_after[nk]"""
    queries = np.ascontiguousarray(queries, PROJ_DTYPE)
    qdesc = u8c(qdesc)
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    kdesc = u8c(kdesc)
    nk = len(kps)
    blocked = np.zeros(nk, np.uint8) if blocked is None else np.array(blocked, np.uint8)
    ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    out = np.empty(nk, np.int32)
    n = _m(impl, "search_by_projection")(_p(queries), _p(qdesc), len(queries), _p(kps), _p(kdesc),
                                        _p(ur), nk,
                                        C.c_float(grid[0]), C.c_float(grid[1]), C.c_float(grid[2]),
                                        C.c_float(grid[3]), _p(sf), C.c_float(th), C.c_float(nnratio),
                                        th_high, _p(blocked), _p(out))
    return n, out, blocked


def search_windowed(queries, qdesc, kps, kdesc, u_right, grid, th_accept, gate, check_orientation,
                    blocked=None, inv_level_sigma2=None):
    """M3/M4/M5 (best-only windowed search + rotation histogram); gate 0 none / 1 stereo radius / 2 chi-square.
    -> nmatches, match_of_kp[nk], blocked_after[nk], match_of_query[nq]"""
    queries = np.ascontiguousarray(queries, WINDOW_DTYPE)
    qdesc = u8c(qdesc)
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    kdesc = u8c(kdesc)
    nk = len(kps)
    blocked = np.zeros(nk, np.uint8) if blocked is None else np.array(blocked, np.uint8)
    ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
    out = np.empty(nk, np.int32)
    outq = np.empty(len(queries), np.int32)
    sig = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
    n = lib().orbo_search_windowed(_p(queries), _p(qdesc), len(queries), _p(kps), _p(kdesc), _p(ur), nk,
                                   C.c_float(grid[0]), C.c_float(grid[1]), C.c_float(grid[2]),
                                   C.c_float(grid[3]), int(th_accept), int(gate), _p(sig),
                                   int(bool(check_orientation)), _p(blocked), _p(out), _p(outq))
    return n, out, blocked, outq


def flatten_feat_vec(fv):
    """{node id: [feature indices]} -> (node ids ascending, start[n+1], flat indices), int32: the iteration order of
    the std::map behind DBoW2::FeatureVector."""
    nodes = sorted(fv)
    start = np.zeros(len(nodes) + 1, np.int32)
    for i, nd in enumerate(nodes):
        start[i + 1] = start[i] + len(fv[nd])
    idx = np.fromiter((j for nd in nodes for j in fv[nd]), np.int32, count=int(start[-1]))
    return np.asarray(nodes, np.int32), start, idx


def _fv_args(fv):
    node, start, idx = flatten_feat_vec(fv)
    return (node, start, idx), (_p(node), _p(start), _p(idx), len(node))


def search_by_bow_frame(fv1, fv2, kps1, d1, good_mp1, kps2, d2, nnratio, check_orientation, impl="oracle"):
    """M6 SearchByBoW(KF, Frame) -> nmatches, match_of_2[n2] (key-frame feature index or -1)"""
    k1, a1 = _fv_args(fv1)
    k2, a2 = _fv_args(fv2)
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    d1 = u8c(d1); d2 = u8c(d2)
    g1 = np.ascontiguousarray(good_mp1, np.uint8)
    out = np.empty(len(kps2), np.int32)
    n = _m(impl, "search_by_bow_frame")(*a1, *a2, _p(kps1), _p(d1), _p(g1), len(kps1), _p(kps2), _p(d2), len(kps2),
                                       C.c_float(nnratio), int(bool(check_orientation)), _p(out))
    return n, out


def search_by_bow_keyframes(fv1, fv2, kps1, d1, good_mp1, kps2, d2, good_mp2, nnratio, check_orientation,
                            impl="oracle"):
    """M6 SearchByBoW(KF, KF) -> nmatches, match12[n1]"""
    k1, a1 = _fv_args(fv1)
    k2, a2 = _fv_args(fv2)
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    d1 = u8c(d1); d2 = u8c(d2)
    g1 = np.ascontiguousarray(good_mp1, np.uint8); g2 = np.ascontiguousarray(good_mp2, np.uint8)
    out = np.empty(len(kps1), np.int32)
    n = _m(impl, "search_by_bow_keyframes")(*a1, *a2, _p(kps1), _p(d1), _p(g1), len(kps1), _p(kps2), _p(d2), _p(g2),
                                           len(kps2), C.c_float(nnratio), int(bool(check_orientation)), _p(out))
    return n, out


def search_for_triangulation(fv1, fv2, kps1, d1, ur1, has_mp1, kps2, d2, ur2, has_mp2, F12, ex, ey, scale_factors2,
                             level_sigma2_2, only_stereo, check_orientation, impl="oracle"):
    """M6 SearchForTriangulation -> nmatches, match12[n1]"""
    k1, a1 = _fv_args(fv1)
    k2, a2 = _fv_args(fv2)
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    d1 = u8c(d1); d2 = u8c(d2)
    h1 = np.ascontiguousarray(has_mp1, np.uint8); h2 = np.ascontiguousarray(has_mp2, np.uint8)
    u1 = None if ur1 is None else np.ascontiguousarray(ur1, np.float32)
    u2 = None if ur2 is None else np.ascontiguousarray(ur2, np.float32)
    F = np.ascontiguousarray(F12, np.float32).reshape(9)
    sf = np.ascontiguousarray(scale_factors2, np.float32)
    sg = np.ascontiguousarray(level_sigma2_2, np.float32)
    out = np.empty(len(kps1), np.int32)
    n = _m(impl, "search_for_triangulation")(*a1, *a2, _p(kps1), _p(d1), _p(u1), _p(h1), len(kps1), _p(kps2), _p(d2),
                                            _p(u2), _p(h2), len(kps2), _p(F), C.c_float(ex), C.c_float(ey), _p(sf),
                                            _p(sg), int(bool(only_stereo)), int(bool(check_orientation)), _p(out))
    return n, out


def search_for_initialization(kps1, d1, prev_matched, kps2, d2, grid, window, nnratio, check_orientation,
                              impl="oracle"):
    """M7 SearchForInitialization -> nmatches, match12[n1] (vbPrevMatched is left to the caller)"""
    kps1 = np.ascontiguousarray(kps1, KP_DTYPE); kps2 = np.ascontiguousarray(kps2, KP_DTYPE)
    d1 = u8c(d1); d2 = u8c(d2)
    # the verbatim reference updates vbPrevMatched in place (:514-517): it gets the caller's array when that is
    # a contiguous float32 array, so the caller sees the update
    pm = np.ascontiguousarray(prev_matched, np.float32).reshape(-1)
    assert len(pm) == 2 * len(kps1)
    out = np.empty(len(kps1), np.int32)
    n = _m(impl, "search_for_initialization")(_p(kps1), _p(d1), len(kps1), _p(pm), _p(kps2), _p(d2), len(kps2),
                                             C.c_float(grid[0]), C.c_float(grid[1]), C.c_float(grid[2]),
                                             C.c_float(grid[3]), int(window), C.c_float(nnratio),
                                             int(bool(check_orientation)), _p(out))
    return n, out


def features_in_area(kps, grid, x, y, r, min_level, max_level):
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    out = np.empty(len(kps) + 1, np.int32)
    n = lib().orbo_features_in_area(_p(kps), len(kps), C.c_float(grid[0]), C.c_float(grid[1]),
                                    C.c_float(grid[2]), C.c_float(grid[3]), C.c_float(x), C.c_float(y),
                                    C.c_float(r), min_level, max_level, _p(out), len(out))
    return out[:n].copy()


def stereo_match(kl, dl, kr, dr, n_rows, scale_factors, min_d, max_d, th_high=100):
    kl = np.ascontiguousarray(kl, KP_DTYPE)
    kr = np.ascontiguousarray(kr, KP_DTYPE)
    dl = u8c(dl)
    dr = u8c(dr)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    bi = np.empty(len(kl), np.int32)
    bd = np.empty(len(kl), np.int32)
    lib().orbo_stereo_match(_p(kl), _p(dl), len(kl), _p(kr), _p(dr), len(kr), n_rows, _p(sf),
                            C.c_float(min_d), C.c_float(max_d), th_high, _p(bi), _p(bd))
    return bi, bd


# ------------------------------------------------------------------------------------------------
# The reference's own ORBextractor.cc, compiled verbatim over oracle/opencv_shim (oracle/_ref).
# ------------------------------------------------------------------------------------------------
def stereo_refine(kl, kr, best_idx_r, best_dist, levels_l, levels_r, scale_factors, inv_scale_factors, min_d, max_d,
                  bf):
    """N1 (Frame.cc:551-640) -> u_right[nl], depth[nl], n_out_of_image.  levels_*: list of 2-D uint8 level images."""
    kl = np.ascontiguousarray(kl, KP_DTYPE)
    kr = np.ascontiguousarray(kr, KP_DTYPE)
    bi = np.ascontiguousarray(best_idx_r, np.int32)
    bd = np.ascontiguousarray(best_dist, np.int32)
    ll = [u8c(a) for a in levels_l]
    lr = [u8c(a) for a in levels_r]
    n = len(ll)
    pl = (C.c_void_p * n)(*[a.ctypes.data for a in ll])
    pr = (C.c_void_p * n)(*[a.ctypes.data for a in lr])
    sl = np.array([(a.shape[1], a.shape[0]) for a in ll], np.int32)
    sr = np.array([(a.shape[1], a.shape[0]) for a in lr], np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    isf = np.ascontiguousarray(inv_scale_factors, np.float32)
    ur = np.empty(len(kl), np.float32)
    dp = np.empty(len(kl), np.float32)
    bad = lib().orbo_stereo_refine(_p(kl), len(kl), _p(kr), _p(bi), _p(bd), pl, pr, _p(sl), _p(sr), n, _p(sf), _p(isf),
                                   C.c_float(min_d), C.c_float(max_d), C.c_float(bf), _p(ur), _p(dp))
    return ur, dp, bad


def distinctive_descriptors(desc_lists):
    """N4 (MapPoint.cc:242-307) -> best index per map point"""
    offs = np.zeros(len(desc_lists) + 1, np.int32)
    for i, d in enumerate(desc_lists):
        offs[i + 1] = offs[i] + len(d)
    flat = (np.concatenate([u8c(d).reshape(-1, 32) for d in desc_lists]) if offs[-1] else np.zeros((0, 32), np.uint8))
    best = np.empty(len(desc_lists), np.int32)
    lib().orbo_distinctive_descriptors(_p(flat), _p(offs), len(desc_lists), _p(best))
    return best


def voc_transform(tree, features, levelsup=4):
    """N3: DBoW2 transform (TemplatedVocabulary.h:1127-1259, TF_IDF + L1).  tree: dict(child_begin, children, desc,
    word_id, weight, depth).  -> per-feature (word, weight, node), BowVector (ids, values), FeatureVector dict"""
    cb = np.ascontiguousarray(tree["child_begin"], np.int32); ch = np.ascontiguousarray(tree["children"], np.int32)
    nd = u8c(tree["desc"]); wid = np.ascontiguousarray(tree["word_id"], np.int32)
    wt = np.ascontiguousarray(tree["weight"], np.float64)
    f = u8c(features)
    n = len(f)
    fw = np.empty(n, np.uint32); fv = np.empty(n, np.float64); fn = np.empty(n, np.uint32)
    bid = np.empty(max(n, 1), np.uint32); bval = np.empty(max(n, 1), np.float64); node = np.empty(n, np.uint32)
    k = lib().orbo_voc_transform(len(wid), _p(cb), _p(ch), _p(nd), _p(wid), _p(wt), int(tree["depth"]), _p(f), n,
                                 int(levelsup), _p(fw), _p(fv), _p(fn), _p(bid), _p(bval), _p(node))
    featvec = {}
    for i in range(n):
        if node[i] != 0xffffffff:
            featvec.setdefault(int(node[i]), []).append(i)
    return (fw, fv, fn), (bid[:k].copy(), bval[:k].copy()), featvec


def have_ref():
    if not os.path.exists(REF_PATH) and os.path.isdir("/root/reference"):
        try:
            build()
        except Exception:
            return False
    return os.path.exists(REF_PATH)


_refs = {}


def ref(seq=False):
    """seq=False: the reference as built (heap-address tie-break).  seq=True: same objects with a
    bump arena for list nodes, so address order == creation order (see ref_wrapper.cc)."""
    if seq not in _refs:
        R = C.CDLL(REF_SEQ_PATH if seq else REF_PATH)
        R.ref_extractor_create.restype = C.c_void_p
        R.ref_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        R.ref_extract_batch.restype = C.c_long
        _refs[seq] = R
    return _refs[seq]


class RefExtractor:
    """ORB_SLAM2::ORBextractor from /root/reference/src/ORBextractor.cc, unmodified."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, seq=False):
        self.nlevels = nlevels
        self._R = ref(seq)
        self._h = C.c_void_p(self._R.ref_extractor_create(nfeatures, C.c_float(scale_factor), nlevels,
                                                        ini_th, min_th))

    def __del__(self):
        if getattr(self, "_h", None):
            self._R.ref_extractor_destroy(self._h)
            self._h = None

    def __call__(self, img):
        img = u8c(img)
        h, w = img.shape
        n = self._R.ref_extractor_run(self._h, _p(img), w, h, C.c_size_t(img.strides[0]))
        kps = np.empty(n, KP_DTYPE)
        desc = np.empty((n, 32), np.uint8)
        self._R.ref_extractor_results(self._h, _p(kps), _p(desc))
        return kps, desc

    def tables(self):
        n = self.nlevels
        f = [np.empty(n, np.float32) for _ in range(4)]
        self._R.ref_extractor_tables(self._h, _p(f[0]), _p(f[1]), _p(f[2]), _p(f[3]))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3])

    def level_image(self, level):
        w = C.c_int()
        h = C.c_int()
        self._R.ref_extractor_level_size(self._h, level, C.byref(w), C.byref(h))
        out = np.empty((h.value, w.value), np.uint8)
        self._R.ref_extractor_level_image(self._h, level, _p(out))
        return out


def ref_extract_batch(frames, nfeatures, scale_factor, nlevels, ini_th, min_th, threads=1):
    frames = u8c(frames)
    B, H, W = frames.shape
    counts = np.zeros(B, np.int32)
    ref().ref_extract_batch(nfeatures, C.c_float(scale_factor), nlevels, ini_th, min_th, _p(frames),
                            B, W, H, C.c_size_t(W), C.c_size_t(W * H), threads, _p(counts))
    return counts


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, np.float32)


def ref_matcher_last_call_seconds():
    """Seconds the last refm_* call of this thread spent inside the matcher method itself (stand-in objects are built
    outside the timed part): the CPU baseline bench.py puts next to a GPU matcher call."""
    f = ref_matcher().refm_last_call_seconds
    f.restype = C.c_double
    return float(f())


def ref_bruteforce_top2(q, db, threads=1):
    """Brute-force top-2 (config 5) around the reference's own DescriptorDistance (src/ORBmatcher.cc:1647-1663,
    verbatim build), `threads` host threads over the queries.  -> idx, best, second, seconds inside the loop."""
    q = u8c(q); db = u8c(db)
    nq = len(q)
    idx = np.empty(nq, np.int32); best = np.empty(nq, np.int32); second = np.empty(nq, np.int32)
    ref_matcher().refm_bruteforce_top2(_p(q), nq, _p(db), len(db), int(threads), _p(idx), _p(best), _p(second))
    return idx, best, second, ref_matcher_last_call_seconds()


def ref_search_last_frame(cam, Tcw_cur, Tcw_last, kps_cur, d_cur, ur_cur, scale_factors, blocked, kps_last, has_mp,
                          outlier, world_pos, mp_desc, mp_obs, th, mono, check_orientation):
    """The reference's SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1328-1470), verbatim.
    cam = (fx, fy, cx, cy, mbf, mb, minX, minY, maxX, maxY, invW, invH); poses row-major 3x4.
    -> nmatches, match_of_kp[n_cur] (last-frame feature or -1), windows[n_last] (WINDOW_DTYPE: what it searched)."""
    cam = _f32(cam); Tc = _f32(Tcw_cur).reshape(12); Tl = _f32(Tcw_last).reshape(12)
    kc = np.ascontiguousarray(kps_cur, KP_DTYPE); kl = np.ascontiguousarray(kps_last, KP_DTYPE)
    dc = u8c(d_cur); md = u8c(mp_desc)
    ur = _f32(ur_cur); sf = _f32(scale_factors)
    bl = None if blocked is None else np.ascontiguousarray(blocked, np.uint8)
    hm = np.ascontiguousarray(has_mp, np.uint8); ol = np.ascontiguousarray(outlier, np.uint8)
    wp = _f32(world_pos); ob = np.ascontiguousarray(mp_obs, np.int32)
    out = np.empty(len(kc), np.int32)
    win = np.zeros(len(kl), WINDOW_DTYPE)
    n = ref_matcher().refm_search_last_frame(_p(cam), _p(Tc), _p(Tl), _p(kc), _p(dc), _p(ur), len(kc), _p(sf), len(sf),
                                             _p(bl), _p(kl), len(kl), _p(hm), _p(ol), _p(wp), _p(md), _p(ob),
                                             C.c_float(th), int(bool(mono)), int(bool(check_orientation)), _p(out),
                                             _p(win))
    return n, out, win


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, np.int32)


def _u8(a):
    return None if a is None else np.ascontiguousarray(a, np.uint8)


def ref_search_keyframe_points(cam, Tcw_cur, kps_cur, d_cur, scale_factors, cur_has_mp, kps_kf, has_mp, world_pos,
                               desc, flags, dist_inv, pred_level, th, orb_dist, check_orientation):
    """The reference's SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (:1472-1599), verbatim.
    flags: bit 0 bad, bit 1 in sAlreadyFound.  -> nmatches, match_of_kp[n_cur], windows[n]"""
    cam = _f32(cam); Tc = _f32(Tcw_cur).reshape(12)
    kc = np.ascontiguousarray(kps_cur, KP_DTYPE); kk = np.ascontiguousarray(kps_kf, KP_DTYPE)
    dc = u8c(d_cur); sf = _f32(scale_factors); ch = _u8(cur_has_mp); hm = _u8(has_mp)
    wp = _f32(world_pos); ds = u8c(desc); fl = _u8(flags); di = _f32(dist_inv); pl = _i32(pred_level)
    out = np.empty(len(kc), np.int32)
    win = np.zeros(len(kk), WINDOW_DTYPE)
    n = ref_matcher().refm_search_keyframe_points(_p(cam), _p(Tc), _p(kc), _p(dc), len(kc), _p(sf), len(sf), _p(ch),
                                                  _p(kk), _p(hm), len(kk), _p(wp), _p(ds), _p(fl), _p(di), _p(pl),
                                                  C.c_float(th), int(orb_dist), int(bool(check_orientation)),
                                                  _p(out), _p(win))
    return n, out, win


def ref_search_sim3_projection(cam, Scw, kps_kf, d_kf, scale_factors, kf_matched, world_pos, normal, desc, flags,
                               dist_inv, pred_level, th):
    """The reference's SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:290-403), verbatim.
    -> nmatches, match_of_kp[n_kf] (cloud point newly written into vpMatched, or -1), windows[n]"""
    cam = _f32(cam); S = _f32(Scw).reshape(12)
    kk = np.ascontiguousarray(kps_kf, KP_DTYPE); dk = u8c(d_kf); sf = _f32(scale_factors); km = _u8(kf_matched)
    wp = _f32(world_pos); nm = _f32(normal); ds = u8c(desc); fl = _u8(flags); di = _f32(dist_inv); pl = _i32(pred_level)
    out = np.empty(len(kk), np.int32)
    win = np.zeros(len(ds), WINDOW_DTYPE)
    n = ref_matcher().refm_search_sim3_projection(_p(cam), _p(S), _p(kk), _p(dk), len(kk), _p(sf), len(sf), _p(km),
                                                  len(ds), _p(wp), _p(nm), _p(ds), _p(fl), _p(di), _p(pl), int(th),
                                                  _p(out), _p(win))
    return n, out, win


def ref_fuse(cam, Tcw, kps_kf, d_kf, ur_kf, scale_factors, level_sigma2, kf_obs, world_pos, normal, desc, flags, obs,
             dist_inv, pred_level, th):
    """The reference's Fuse(pKF, vpMapPoints, th) (:825-975), verbatim.  flags: bit 0 bad, bit 1 IsInKeyFrame.
    -> nFused, match_of_query[n], action[n] (1 add, 2 point replaced by the key frame's, 3 the reverse), windows[n]"""
    cam = _f32(cam); T = _f32(Tcw).reshape(12)
    kk = np.ascontiguousarray(kps_kf, KP_DTYPE); dk = u8c(d_kf); ur = _f32(ur_kf); sf = _f32(scale_factors)
    sg = _f32(level_sigma2); ko = _i32(kf_obs)
    wp = _f32(world_pos); nm = _f32(normal); ds = u8c(desc); fl = _u8(flags); ob = _i32(obs); di = _f32(dist_inv)
    pl = _i32(pred_level)
    mq = np.empty(len(ds), np.int32); act = np.empty(len(ds), np.int32)
    win = np.zeros(len(ds), WINDOW_DTYPE)
    n = ref_matcher().refm_fuse(_p(cam), _p(T), _p(kk), _p(dk), _p(ur), len(kk), _p(sf), _p(sg), len(sf), _p(ko),
                                len(ds), _p(wp), _p(nm), _p(ds), _p(fl), _p(ob), _p(di), _p(pl), C.c_float(th),
                                _p(mq), _p(act), _p(win))
    return n, mq, act, win


def ref_search_by_sim3(cam, T1w, T2w, s12, R12, t12, scale_factors, kf1, kf2, pre12, th):
    """The reference's SearchBySim3 (:1102-1326), verbatim.  kf1 / kf2: dict(kps, desc, has_mp, pos, flags, dist_inv,
    pred_level).  -> nFound, match12[n1], windows12[n1], windows21[n2]"""
    cam = _f32(cam); a = _f32(T1w).reshape(12); b = _f32(T2w).reshape(12)
    R = _f32(R12).reshape(9); t = _f32(t12).reshape(3); sf = _f32(scale_factors)

    def side(k):
        return (np.ascontiguousarray(k["kps"], KP_DTYPE), u8c(k["desc"]), _u8(k["has_mp"]), _f32(k["pos"]),
                _u8(k["flags"]), _f32(k["dist_inv"]), _i32(k["pred_level"]))
    s1, s2 = side(kf1), side(kf2)
    n1, n2 = len(s1[0]), len(s2[0])
    pre = _i32(pre12)
    out = np.empty(n1, np.int32)
    w12 = np.zeros(n1, WINDOW_DTYPE); w21 = np.zeros(n2, WINDOW_DTYPE)
    n = ref_matcher().refm_search_by_sim3(_p(cam), _p(a), _p(b), C.c_float(s12), _p(R), _p(t), _p(sf), len(sf),
                                          _p(s1[0]), _p(s1[1]), _p(s1[2]), n1, _p(s1[3]), _p(s1[4]), _p(s1[5]), _p(s1[6]),
                                          _p(s2[0]), _p(s2[1]), _p(s2[2]), n2, _p(s2[3]), _p(s2[4]), _p(s2[5]), _p(s2[6]),
                                          _p(pre), C.c_float(th), _p(out), _p(w12), _p(w21))
    return n, out, w12, w21


def ref_fuse_sim3(cam, Scw, kps_kf, d_kf, scale_factors, kf_obs, world_pos, normal, desc, flags, dist_inv, pred_level,
                  th):
    """The reference's Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:977-1100), verbatim.
    -> nFused, match_of_query[n], replace[n], windows[n]"""
    cam = _f32(cam); S = _f32(Scw).reshape(12)
    kk = np.ascontiguousarray(kps_kf, KP_DTYPE); dk = u8c(d_kf); sf = _f32(scale_factors); ko = _i32(kf_obs)
    wp = _f32(world_pos); nm = _f32(normal); ds = u8c(desc); fl = _u8(flags); di = _f32(dist_inv); pl = _i32(pred_level)
    mq = np.empty(len(ds), np.int32); rp = np.empty(len(ds), np.int32)
    win = np.zeros(len(ds), WINDOW_DTYPE)
    n = ref_matcher().refm_fuse_sim3(_p(cam), _p(S), _p(kk), _p(dk), len(kk), _p(sf), len(sf), _p(ko), len(ds), _p(wp),
                                     _p(nm), _p(ds), _p(fl), _p(di), _p(pl), C.c_float(th), _p(mq), _p(rp), _p(win))
    return n, mq, rp, win


# ------------------------------------------------------------------------------------------------
# The reference's own Frame.cc, compiled verbatim over oracle/slam_shim/frame_shim.h (oracle/_ref/libref_frame.so).
# ------------------------------------------------------------------------------------------------
REF_FRAME_PATH = os.path.join(HERE, "_ref", "libref_frame.so")
_reff = None


def have_ref_frame():
    return os.path.exists(REF_FRAME_PATH)


def ref_frame():
    global _reff
    if _reff is None:
        if not have_ref_frame():
            raise RuntimeError("oracle/_ref/libref_frame.so is not built (needs /root/reference)")
        _reff = C.CDLL(REF_FRAME_PATH)
    return _reff


def ref_compute_stereo_matches(kl, dl, kr, dr, levels_l, levels_r, scale_factors, inv_scale_factors, mb, mbf):
    """The reference's Frame::ComputeStereoMatches (src/Frame.cc:466-640), verbatim -> mvuRight[nl], mvDepth[nl], rc
    (rc = -1: a correlation window left its level image, where OpenCV asserts)."""
    kl = np.ascontiguousarray(kl, KP_DTYPE); kr = np.ascontiguousarray(kr, KP_DTYPE)
    dl = u8c(dl); dr = u8c(dr)
    ll = [u8c(a) for a in levels_l]; lr = [u8c(a) for a in levels_r]
    n = len(ll)
    pl = (C.c_void_p * n)(*[a.ctypes.data for a in ll]); pr = (C.c_void_p * n)(*[a.ctypes.data for a in lr])
    sl = np.array([[a.shape[1], a.shape[0]] for a in ll], np.int32); sr = np.array([[a.shape[1], a.shape[0]] for a in lr], np.int32)
    sf = _f32(scale_factors); isf = _f32(inv_scale_factors)
    ur = np.empty(len(kl), np.float32); dp = np.empty(len(kl), np.float32)
    rc = ref_frame().reff_compute_stereo_matches(_p(kl), _p(dl), len(kl), _p(kr), _p(dr), len(kr), pl, pr, _p(sl), _p(sr), n,
                                                 _p(sf), _p(isf), C.c_float(mb), C.c_float(mbf), _p(ur), _p(dp))
    return ur, dp, rc


def ref_frame_last_call_seconds():
    f = ref_frame().reff_last_call_seconds
    f.restype = C.c_double
    return float(f())


def ref_features_in_area(kps, grid, x, y, r, min_level, max_level):
    """The reference's Frame::AssignFeaturesToGrid + GetFeaturesInArea (src/Frame.cc:230-245, :327-392), verbatim."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    out = np.empty(len(kps) + 1, np.int32)
    n = ref_frame().reff_features_in_area(_p(kps), len(kps), C.c_float(grid[0]), C.c_float(grid[1]), C.c_float(grid[2]),
                                          C.c_float(grid[3]), C.c_float(x), C.c_float(y), C.c_float(r), min_level,
                                          max_level, _p(out), len(out))
    return out[:n].copy()


# ------------------------------------------------------------------------------------------------
# DBoW2 as vendored by the reference, compiled verbatim (oracle/_ref/libref_vocabulary.so).
# ------------------------------------------------------------------------------------------------
REF_VOC_PATH = os.path.join(HERE, "_ref", "libref_vocabulary.so")
_refv = None


def have_ref_vocabulary():
    return os.path.exists(REF_VOC_PATH)


def ref_vocabulary_lib():
    global _refv
    if _refv is None:
        if not have_ref_vocabulary():
            raise RuntimeError("oracle/_ref/libref_vocabulary.so is not built (needs /root/reference)")
        _refv = C.CDLL(REF_VOC_PATH)
        _refv.refv_load_text.restype = C.c_void_p
        _refv.refv_load_text.argtypes = [C.c_char_p]
        _refv.refv_destroy.argtypes = [C.c_void_p]
        _refv.refv_size.argtypes = [C.c_void_p]
    return _refv


def tree_as_loaded(tree):
    """The synthetic tree as TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1338-1430) rebuilds it from
    a text file: node ids = line numbers with every parent before its children (the loader appends to m_nodes line by
    line and indexes the parent, so the file is in top-down order -- breadth first here), children of a node in
    ascending id order, word ids in ascending node id order."""
    cb, ch = tree["child_begin"], tree["children"]
    n = len(tree["word_id"])
    order = [0]                                   # old ids in breadth-first order
    for old in order:
        order.extend(int(c) for c in ch[cb[old]:cb[old + 1]])
    assert len(order) == n
    new_of = np.empty(n, np.int64)
    new_of[order] = np.arange(n)
    child_begin = np.zeros(n + 1, np.int32)
    children = []
    for new, old in enumerate(order):
        children.extend(sorted(int(new_of[c]) for c in ch[cb[old]:cb[old + 1]]))
        child_begin[new + 1] = len(children)
    order = np.asarray(order)
    leaf = tree["word_id"][order] >= 0
    word_id = np.full(n, -1, np.int32)
    word_id[leaf] = np.arange(int(leaf.sum()), dtype=np.int32)
    return dict(tree, child_begin=child_begin, children=np.asarray(children, np.int32), desc=tree["desc"][order],
                weight=tree["weight"][order], word_id=word_id)


def write_vocabulary_text(tree, k, path):
    """ORBvoc.txt format: "k L scoring weighting", then one line per node id 1..n-1: parent, is-leaf, 32 bytes, weight.
    No trailing newline (the reference's eof loop would read an empty record as one more child of the root)."""
    cb, ch = tree["child_begin"], tree["children"]
    n = len(tree["word_id"])
    parent = np.zeros(n, np.int64)
    for i in range(n):
        parent[ch[cb[i]:cb[i + 1]]] = i
    lines = ["%d %d 0 0" % (k, tree["depth"])]      # L1_NORM, TF_IDF (what ORBvoc.txt declares)
    for i in range(1, n):
        lines.append("%d %d %s %r" % (parent[i], int(tree["word_id"][i] >= 0),
                                      " ".join(str(int(b)) for b in tree["desc"][i]), float(tree["weight"][i])))
    with open(path, "w") as f:
        f.write("\n".join(lines))


class RefVocabulary:
    """ORBVocabulary (include/ORBVocabulary.h:31-32) of the verbatim DBoW2 build, loaded from a text file."""

    def __init__(self, path):
        self._L = ref_vocabulary_lib()
        self._v = self._L.refv_load_text(path.encode())
        if not self._v:
            raise RuntimeError("loadFromTextFile failed: " + path)

    def __del__(self):
        if getattr(self, "_v", None):
            self._L.refv_destroy(self._v)
            self._v = None

    def words(self):
        return self._L.refv_size(self._v)

    def transform(self, features, levelsup=4):
        """-> BowVector (ids, values) in key order, FeatureVector dict"""
        f = u8c(features)
        n = len(f)
        bid = np.empty(max(n, 1), np.uint32); bval = np.empty(max(n, 1), np.float64); node = np.empty(n, np.uint32)
        k = self._L.refv_transform(C.c_void_p(self._v), _p(f), n, int(levelsup), _p(bid), _p(bval), _p(node))
        featvec = {}
        for i in range(n):
            if node[i] != 0xffffffff:
                featvec.setdefault(int(node[i]), []).append(i)
        return (bid[:k].copy(), bval[:k].copy()), featvec


# ------------------------------------------------------------------------------------------------
# The reference's own MapPoint.cc, compiled verbatim with its own MapPoint.h (oracle/_ref/libref_mappoint.so).
# ------------------------------------------------------------------------------------------------
REF_MAPPOINT_PATH = os.path.join(HERE, "_ref", "libref_mappoint.so")
_refp = None


def have_ref_mappoint():
    return os.path.exists(REF_MAPPOINT_PATH)


def ref_distinctive_descriptors(desc_lists, bad_lists=None):
    """The reference's MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307), verbatim, for every list of
    observed descriptors -> index of the descriptor kept (-1: none).  bad_lists: per observation, its key frame isBad()."""
    global _refp
    if _refp is None:
        if not have_ref_mappoint():
            raise RuntimeError("oracle/_ref/libref_mappoint.so is not built (needs /root/reference)")
        _refp = C.CDLL(REF_MAPPOINT_PATH)
    offs = np.zeros(len(desc_lists) + 1, np.int32)
    for i, d in enumerate(desc_lists):
        offs[i + 1] = offs[i] + len(d)
    flat = (np.concatenate([u8c(d).reshape(-1, 32) for d in desc_lists]) if offs[-1] else np.zeros((0, 32), np.uint8))
    bad = None
    if bad_lists is not None:
        bad = (np.concatenate([np.asarray(b, np.uint8).reshape(-1) for b in bad_lists]) if offs[-1]
               else np.zeros(0, np.uint8))
    best = np.empty(len(desc_lists), np.int32)
    _refp.refp_distinctive_descriptors(_p(flat), _p(bad), _p(offs), len(desc_lists), _p(best))
    return best
