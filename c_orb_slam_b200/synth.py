"""Deterministic synthetic inputs (datasets are not available offline).

Frames are "textured grayscale": five octaves of bilinearly upsampled uniform noise, filled
rectangles (true corners at every pyramid level) and +-6 pixel noise.  Descriptor sets are i.i.d.
256-bit rows with queries derived from database rows by bit flips, plus adversarial duplicates.
Pure numpy (PCG64), so the CPU oracle and the CUDA path read identical bytes everywhere.
"""
import numpy as np

SHAPES = {"tum": (640, 480), "euroc": (752, 480), "kitti": (1241, 376)}


def _upsample_bilinear(grid, h, w):
    gh, gw = grid.shape
    ys = np.linspace(0, gh - 1, h, dtype=np.float32)
    xs = np.linspace(0, gw - 1, w, dtype=np.float32)
    y0 = np.minimum(ys.astype(np.int32), gh - 2)
    x0 = np.minimum(xs.astype(np.int32), gw - 2)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = grid[y0][:, x0]
    b = grid[y0][:, x0 + 1]
    c = grid[y0 + 1][:, x0]
    d = grid[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def make_frame(width, height, seed, n_rects=1200):
    """One uint8 [height, width] textured frame."""
    rng = np.random.default_rng(seed)
    img = np.zeros((height, width), np.float32)
    for cell, wgt in ((64, 0.3), (32, 0.25), (16, 0.2), (8, 0.15), (4, 0.1)):
        gh, gw = height // cell + 2, width // cell + 2
        img += wgt * _upsample_bilinear(rng.uniform(0, 255, (gh, gw)).astype(np.float32), height, width)
    for _ in range(n_rects):
        rw, rh = rng.integers(6, 61, 2)
        x0 = int(rng.integers(-20, width - 4))
        y0 = int(rng.integers(-20, height - 4))
        val = float(rng.integers(0, 256))
        img[max(y0, 0):max(y0 + rh, 0), max(x0, 0):max(x0 + rw, 0)] = val
    img += rng.integers(-6, 7, (height, width)).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def make_frames(width, height, n, seed0=1000):
    return np.stack([make_frame(width, height, seed0 + i) for i in range(n)])


def make_frames_fast(width, height, n, seed0=1000, n_base=8):
    """n frames derived from n_base generated frames by cyclic shifts + per-frame noise (cheap to
    make in bulk for the benchmark; each frame is still a distinct textured image)."""
    base = [make_frame(width, height, seed0 + i) for i in range(min(n_base, n))]
    out = np.empty((n, height, width), np.uint8)
    rng = np.random.default_rng(seed0 + 77)
    for i in range(n):
        b = base[i % len(base)]
        sx, sy = (int(v) for v in rng.integers(0, 64, 2))
        f = np.roll(b, (sy, sx), (0, 1)).astype(np.int16)
        f += rng.integers(-3, 4, f.shape, dtype=np.int16)
        out[i] = np.clip(f, 0, 255).astype(np.uint8)
    return out


def make_stereo_pair(width, height, seed):
    """left, right: right = left resampled with an integer disparity per 32-row band in [4,64]."""
    left = make_frame(width, height, seed)
    rng = np.random.default_rng(seed + 555)
    right = np.empty_like(left)
    for y0 in range(0, height, 32):
        d = int(rng.integers(4, 65))
        band = left[y0:y0 + 32]
        right[y0:y0 + 32] = np.roll(band, -d, axis=1)
        right[y0:y0 + 32, width - d:] = band[:, -1:]
    noise = rng.integers(-2, 3, right.shape)
    return left, np.clip(right.astype(np.int16) + noise, 0, 255).astype(np.uint8)


def make_stereo_pair_subpixel(width, height, seed):
    """left, right: right = left shifted by a fractional disparity per 32-row band in [3, 60] px (linear
    interpolation between the two neighbouring integer shifts) plus noise: exercises the parabola fit of the
    sub-pixel stereo refinement."""
    left = make_frame(width, height, seed)
    rng = np.random.default_rng(seed + 777)
    right = np.empty_like(left)
    for y0 in range(0, height, 32):
        d = float(rng.uniform(3, 60))
        d0 = int(np.floor(d))
        f = d - d0
        band = left[y0:y0 + 32].astype(np.float32)
        a = np.roll(band, -d0, axis=1)
        b = np.roll(band, -(d0 + 1), axis=1)
        mix = (1 - f) * a + f * b
        mix[:, width - d0 - 1:] = band[:, -1:]
        right[y0:y0 + 32] = np.clip(np.rint(mix), 0, 255).astype(np.uint8)
    noise = rng.integers(-2, 3, right.shape)
    return left, np.clip(right.astype(np.int16) + noise, 0, 255).astype(np.uint8)


def make_vocabulary(k=10, depth=4, seed=1, shuffle_ids=True, stop_frac=0.02):
    """Synthetic DBoW2 vocabulary tree (the shipped ORBvoc.txt is k = 10, L = 6; it is not in the reference
    tree): every node's descriptor is its parent's with ~256/(level+2) random bits flipped, so descents are
    decisive near the root and close calls (equal distances) appear near the leaves.  Branching is k, a few
    nodes get fewer children.  Node ids are shuffled (children of a node are not contiguous, as in a loaded
    vocabulary); words are numbered in creation order; weights mimic idf values, a fraction is 0 (stopped).
    -> dict(child_begin, children, desc, word_id, weight, depth)"""
    rng = np.random.default_rng(seed)
    descs = [rng.integers(0, 256, 32, dtype=np.uint8)]
    kids = [[]]
    level_nodes = [0]
    for lv in range(depth):
        nxt = []
        for nd in level_nodes:
            nk = k if rng.random() < 0.9 else int(rng.integers(1, k + 1))
            for _ in range(nk):
                d = descs[nd].copy()
                bits = rng.choice(256, size=max(4, 256 // (lv + 2) // 2), replace=False)
                for b in bits:
                    d[b >> 3] ^= np.uint8(1 << (b & 7))
                descs.append(d)
                kids.append([])
                kids[nd].append(len(descs) - 1)
                nxt.append(len(descs) - 1)
        level_nodes = nxt
    n = len(descs)
    perm = np.arange(n)
    if shuffle_ids:
        perm[1:] = rng.permutation(np.arange(1, n))   # new id of old node i; the root stays 0
    inv = np.argsort(perm)
    child_begin = np.zeros(n + 1, np.int32)
    children = []
    for new in range(n):
        old = inv[new]
        children.extend(int(perm[c]) for c in kids[old])
        child_begin[new + 1] = len(children)
    desc = np.stack(descs)[inv]
    word_id = np.full(n, -1, np.int32)
    weight = np.zeros(n, np.float64)
    words = 0
    for old in range(n):                      # creation order
        if not kids[old]:
            word_id[perm[old]] = words
            words += 1
            weight[perm[old]] = 0.0 if rng.random() < stop_frac else float(np.log(rng.uniform(1.5, 400.0)))
    return dict(child_begin=child_begin, children=np.asarray(children, np.int32), desc=desc, word_id=word_id,
                weight=weight, depth=depth, words=words)


def random_descriptors(n, seed=7):
    return np.random.default_rng(seed).integers(0, 256, (n, 32), dtype=np.uint8)


def flip_bits(desc, nflips, rng):
    """Return copies of desc rows with nflips[i] distinct random bits flipped."""
    out = desc.copy()
    for i, k in enumerate(nflips):
        if k <= 0:
            continue
        bits = rng.choice(256, size=int(k), replace=False)
        for b in bits:
            out[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return out


def make_match_set(nq, ndb, seed=7, n_dup=64):
    """database [ndb,32], queries [nq,32]: queries = db rows with U{0..80} flips, 5% pure random,
    plus duplicate database rows far apart (equal-distance ties across shards)."""
    rng = np.random.default_rng(seed)
    db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
    n_dup = min(n_dup, ndb // 4)
    if n_dup:
        src = rng.choice(ndb // 2, n_dup, replace=False)
        dst = ndb - 1 - rng.choice(ndb // 2, n_dup, replace=False)
        db[dst] = db[src]
    perm = rng.integers(0, ndb, nq)
    q = flip_bits(db[perm], rng.integers(0, 81, nq), rng)
    rnd = rng.random(nq) < 0.05
    q[rnd] = rng.integers(0, 256, (int(rnd.sum()), 32), dtype=np.uint8)
    if n_dup and nq >= n_dup:
        q[:n_dup] = db[src]          # exact duplicates: best == second == 0, lowest index must win
    return q, db
