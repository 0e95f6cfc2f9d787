"""Seeded synthetic EM-like inputs: jittered-grid Voronoi superpixels and a
dyadic boundary-probability map.  Pure integer hashing, so the CUDA generator in
csrc/synth.cu (used by bench.py for the 1024^3 volume) produces the same labels.

labels: uint32, cell index + 1 (label 0 is reserved by the reference as a scratch key,
util/struct_merge_bc.hxx:19-20).  pb: float32 q/256, q in [0,255], ridge ~1 on superpixel faces.
"""
import numpy as np

_M = np.uint64(0xFFFFFFFF)


def _hash32(x):
    """murmur3 finaliser on uint32 arrays (held in uint64 to avoid overflow warnings)."""
    x = x.astype(np.uint64) & _M
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x85EBCA6B)) & _M
    x ^= x >> np.uint64(13)
    x = (x * np.uint64(0xC2B2AE35)) & _M
    x ^= x >> np.uint64(16)
    return x


def seed_positions(shape_xyz, cell, seed):
    """Fixed-point (1/256 voxel) seed position of every grid cell: [ncell, ndim] int64."""
    nd = len(shape_xyz)
    nc = [(s + cell - 1) // cell for s in shape_xyz]
    idx = np.arange(int(np.prod(nc)), dtype=np.uint64)
    coords = []
    rem = idx.copy()
    for d in range(nd):  # x fastest
        coords.append(rem % np.uint64(nc[d]))
        rem = rem // np.uint64(nc[d])
    pos = np.zeros((len(idx), nd), np.int64)
    for d in range(nd):
        u = _hash32((np.uint64(seed) * np.uint64(0x9E3779B9) + idx * np.uint64(4) + np.uint64(d)) & _M) >> np.uint64(8)
        pos[:, d] = (coords[d] * np.uint64(cell * 256) + ((u * np.uint64(cell * 256)) >> np.uint64(24))).astype(np.int64)
    return pos, nc


def voronoi_volume(shape, cell, seed=1234):
    """shape in numpy order ([z,] y, x).  Returns (labels uint32, pb float32)."""
    nd = len(shape)
    sx = list(shape[::-1])
    pos, nc = seed_positions(sx, cell, seed)
    N = int(np.prod(shape))
    lin = np.arange(N, dtype=np.int64)
    xyz = []
    rem = lin.copy()
    for d in range(nd):
        xyz.append(rem % sx[d])
        rem //= sx[d]
    c = [v // cell for v in xyz]
    best1 = np.full(N, np.iinfo(np.int64).max, np.int64)
    best2 = best1.copy()
    lab = np.zeros(N, np.int64)
    offs = np.stack(np.meshgrid(*[[-1, 0, 1]] * nd, indexing="ij"), -1).reshape(-1, nd)
    # visit neighbour cells in ascending cell-index order so ties keep the lower index
    offs = offs[np.lexsort([offs[:, d] for d in range(nd)])]
    for o in offs:
        cc = [c[d] + o[d] for d in range(nd)]
        ok = np.ones(N, bool)
        for d in range(nd):
            ok &= (cc[d] >= 0) & (cc[d] < nc[d])
        ci = np.zeros(N, np.int64)
        mul = 1
        for d in range(nd):
            ci += np.where(ok, cc[d], 0) * mul
            mul *= nc[d]
        d2 = np.zeros(N, np.int64)
        for d in range(nd):
            dd = xyz[d] * 256 - pos[ci, d]
            d2 += dd * dd
        d2 = np.where(ok, d2, np.iinfo(np.int64).max)
        closer = d2 < best1
        best2 = np.where(closer, best1, np.minimum(best2, d2))
        lab = np.where(closer, ci, lab)
        best1 = np.where(closer, d2, best1)
    d1 = np.sqrt(best1.astype(np.float64)) / 256.0
    has2 = best2 < np.iinfo(np.int64).max
    d2v = np.where(has2, np.sqrt(np.where(has2, best2, 0).astype(np.float64)) / 256.0, d1 + 1e9)
    u = _hash32((np.uint64(seed) * np.uint64(0x85EBCA77) + np.uint64(0x1234567) + lin.astype(np.uint64)) & _M)
    u = u.astype(np.float64) / 4294967296.0
    val = 0.8 * np.exp(-(d2v - d1) / 2.0) + 0.2 * u
    q = np.floor(255.0 * np.clip(val, 0.0, 1.0))
    pb = (q / 256.0).astype(np.float32)
    return (lab + 1).astype(np.uint32).reshape(shape), pb.reshape(shape)


def speckle_volume(shape, nlabels, seed=0, dyadic=True):
    """Random-speckle labels: many tiny/disconnected regions, non-mutual edges, equal areas —
    the adversarial case for the RAG rules."""
    rng = np.random.default_rng(seed)
    lab = rng.integers(1, nlabels + 1, size=shape, dtype=np.uint32)
    # smooth a little so that regions have some extent
    for ax in range(len(shape)):
        keep = rng.random(shape) < 0.6
        lab = np.where(keep, np.roll(lab, 1, axis=ax), lab)
    if dyadic:
        pb = (rng.integers(0, 256, size=shape) / 256.0).astype(np.float32)
    else:
        pb = rng.random(shape, dtype=np.float32)
    return lab.astype(np.uint32), pb


def random_forest(ntree, mdim, max_depth=12, seed=0, nclass=2, orig_labels=(-1, 1), thr_pool=None,
                  leaf_prob=0.15, features=None):
    """Random forest in the rf_old / classRF array layout (ml/rf/ml_rf.h:97-157): per tree a
    padded block of nrnodes nodes; treemap 1-based child ids, nodestatus -1 = terminal,
    bestvar 1-based feature, nodeclass 1-based class.  thr_pool: optional [n, mdim] sample matrix
    whose values are used as thresholds (so that x <= t ties really occur)."""
    rng = np.random.default_rng(seed)
    trees = []
    for _ in range(ntree):
        tm, ns, xb, bv, nc = [[0, 0]], [0], [0.0], [0], [0]
        stack = [(0, 0)]
        while stack:
            k, depth = stack.pop()
            if depth >= max_depth or (depth > 1 and rng.random() < leaf_prob):
                ns[k] = -1
                nc[k] = int(rng.integers(1, nclass + 1))
                continue
            f = int(rng.integers(0, mdim)) if features is None else int(features[rng.integers(0, len(features))])
            t = float(thr_pool[rng.integers(0, len(thr_pool)), f]) if thr_pool is not None else float(rng.random())
            ns[k], bv[k], xb[k] = 1, f + 1, t
            for side in range(2):
                tm.append([0, 0]); ns.append(0); xb.append(0.0); bv.append(0); nc.append(0)
                tm[k][side] = len(ns)  # 1-based id of the new node
                stack.append((len(ns) - 1, depth + 1))
        trees.append((tm, ns, xb, bv, nc))
    nr = max(len(t[1]) for t in trees)
    f = dict(treemap=np.zeros((ntree, nr, 2), np.int32), nodestatus=np.zeros((ntree, nr), np.int32),
             xbestsplit=np.zeros((ntree, nr)), bestvar=np.zeros((ntree, nr), np.int32),
             nodeclass=np.zeros((ntree, nr), np.int32), ndbigtree=np.zeros(ntree, np.int32),
             orig_labels=np.asarray(orig_labels, np.int32))
    for i, (tm, ns, xb, bv, nc) in enumerate(trees):
        n = len(ns)
        f["treemap"][i, :n] = tm
        f["nodestatus"][i, :n] = ns
        f["xbestsplit"][i, :n] = xb
        f["bestvar"][i, :n] = bv
        f["nodeclass"][i, :n] = nc
        f["ndbigtree"][i] = n
    return f


# ---- bench forest: 255 trees trained with scikit-learn on feature rows of a small
# synthetic volume against a synthetic ground truth, exported to the rf_old array layout --------------
def gt_merge_labels(keys, shape, cell, block=3):
    """Synthetic ground truth for candidate pairs of a voronoi_volume: superpixels whose grid cells fall into
    the same block^D group of cells form one true segment.  Returns -1 (merge, BC_LABEL_MERGE) / +1 per pair."""
    sx = list(shape[::-1])
    nc = [(s + cell - 1) // cell for s in sx]

    def seg(lab):
        ci = lab.astype(np.int64) - 1
        s, mul = np.zeros_like(ci), 1
        for d in range(len(sx)):
            c = ci % nc[d]
            ci = ci // nc[d]
            s += (c // block) * mul
            mul *= (nc[d] + block - 1) // block
        return s
    return np.where(seg(keys[:, 0]) == seg(keys[:, 1]), -1, 1).astype(np.int32)


def sklearn_forest(rows, y, ntree=255, seed=0, max_depth=20):
    """Trains a scikit-learn forest (x <= t goes left, as classForest) and exports it as rf_old arrays."""
    from sklearn.ensemble import RandomForestClassifier
    X = np.ascontiguousarray(rows, dtype=np.float32)  # sklearn trains on float32; thresholds are doubles
    clf = RandomForestClassifier(n_estimators=ntree, max_depth=max_depth, random_state=seed, n_jobs=-1).fit(X, y)
    trees = [e.tree_ for e in clf.estimators_]
    classes = [int(c) for c in clf.classes_]
    nr = max(t.node_count for t in trees)
    f = dict(treemap=np.zeros((ntree, nr, 2), np.int32), nodestatus=np.zeros((ntree, nr), np.int32),
             xbestsplit=np.zeros((ntree, nr)), bestvar=np.zeros((ntree, nr), np.int32),
             nodeclass=np.zeros((ntree, nr), np.int32), ndbigtree=np.zeros(ntree, np.int32),
             orig_labels=np.asarray(classes, np.int32))
    for i, t in enumerate(trees):
        n = t.node_count
        leaf = t.children_left == -1
        f["treemap"][i, :n, 0] = np.where(leaf, 0, t.children_left + 1)
        f["treemap"][i, :n, 1] = np.where(leaf, 0, t.children_right + 1)
        f["nodestatus"][i, :n] = np.where(leaf, -1, 1)
        f["xbestsplit"][i, :n] = t.threshold
        f["bestvar"][i, :n] = np.where(leaf, 0, t.feature + 1)
        f["nodeclass"][i, :n] = np.where(leaf, t.value[:, 0, :].argmax(axis=1) + 1, 0)
        f["ndbigtree"][i] = n
    return f


def mean_path_length(f, X, max_rows=2000):
    """Mean number of internal nodes a row visits per tree (x <= t goes left): the work unit of RF scoring."""
    X = np.asarray(X, np.float64)[:max_rows]
    n, tot = len(X), 0
    ar = np.arange(n)
    for t in range(len(f["ndbigtree"])):
        k = np.zeros(n, np.int64)
        ns, bv, xb, tm = f["nodestatus"][t], f["bestvar"][t], f["xbestsplit"][t], f["treemap"][t]
        while True:
            act = ns[k] != -1
            if not act.any():
                break
            tot += int(act.sum())
            ka = k[act]
            left = X[ar[act], bv[ka] - 1] <= xb[ka]
            k[act] = np.where(left, tm[ka, 0], tm[ka, 1]) - 1
    return tot / max(n * len(f["ndbigtree"]), 1)
