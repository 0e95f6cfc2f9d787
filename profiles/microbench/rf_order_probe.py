"""Does the ORDER of the candidates matter to the RF kernel?  ("order the candidates by
their path through the top levels of a few anchor trees so that lanes share sectors deeper".)

Harness only: featurises the c4s volume (512^3) with the library, takes the rows, scores them with
glia_b200_rf_score (device rows) in different row orders and times the scoring with CUDA events.  No library change -
if an order wins by enough, the library would sort an index (cub radix sort on a 16-bit key) and walk through it.

    python profiles/microbench/rf_order_probe.py [--shape 512 512 512]
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import torch  # noqa: E402

import bench  # noqa: E402
from glia_b200 import api, synth  # noqa: E402


def top_path_key(f, X, trees, levels):
    """bits: for each anchor tree, the left/right decisions of its first `levels` levels (a leaf repeats its last bit)."""
    n = len(X)
    key = np.zeros(n, np.int64)
    ar = np.arange(n)
    for t in trees:
        ns, bv, xb, tm = f["nodestatus"][t], f["bestvar"][t], f["xbestsplit"][t], f["treemap"][t]
        k = np.zeros(n, np.int64)
        last = np.zeros(n, np.int64)
        for _ in range(levels):
            act = ns[k] != -1
            right = np.zeros(n, bool)
            ka = k[act]
            right[act] = ~(X[ar[act], bv[ka] - 1] <= xb[ka])
            k[act] = np.where(right[act], tm[ka, 1], tm[ka, 0]) - 1
            last = np.where(act, right.astype(np.int64), last)
            key = key * 2 + last
    return key


def leaf_key(f, X, t):
    n = len(X)
    ar = np.arange(n)
    ns, bv, xb, tm = f["nodestatus"][t], f["bestvar"][t], f["xbestsplit"][t], f["treemap"][t]
    k = np.zeros(n, np.int64)
    depth = np.zeros(n, np.int64)
    while True:
        act = ns[k] != -1
        if not act.any():
            return k, depth
        ka = k[act]
        left = X[ar[act], bv[ka] - 1] <= xb[ka]
        k[act] = np.where(left, tm[ka, 0], tm[ka, 1]) - 1
        depth[act] += 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", type=int, nargs=3, default=[512, 512, 512])
    ap.add_argument("--cell", type=int, default=10)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    shape = tuple(a.shape)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    L = api.lib()
    nvox = int(np.prod(shape))
    lab = torch.empty(nvox, dtype=torch.int32, device=dev)
    pb = torch.empty(nvox, dtype=torch.float32, device=dev)
    dims = (C.c_int * 3)(*shape[::-1])
    ncells = C.c_uint64(0)
    stream = torch.cuda.current_stream().cuda_stream
    assert L.glia_b200_synth_voronoi(3, dims, a.cell, 1234, lab.data_ptr(), pb.data_ptr(), C.c_void_p(stream), C.byref(ncells)) == 0
    torch.cuda.synchronize()
    dim = bench.row_dim(3)
    farr, forest = bench.make_forest(api, synth, torch, dim, 3, 0)
    spec = [(pb.data_ptr(), 16, 0.0, 1.0)]
    b = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec, thr=bench.THR), device=0, device_pointers=True,
                             label_bound=int(ncells.value) + 1, stream=stream)
    b.build(lab.data_ptr(), pb.data_ptr(), None)
    n = int(b.Cn)
    keys_d = torch.empty((n, 2), dtype=torch.int32, device=dev)  # a device-pointer handle writes to device buffers
    Xd = torch.empty((n, dim), dtype=torch.float64, device=dev)
    b._check(L.glia_b200_init_candidates(b.h, C.c_void_p(keys_d.data_ptr()), C.c_void_p(Xd.data_ptr())))
    torch.cuda.synchronize()
    rows = Xd.cpu().numpy()
    print("candidates", n, "dim", rows.shape[1], "forest", bench.FOREST_INFO, flush=True)
    prob = torch.empty(n, dtype=torch.float64, device=dev)

    def timed(order, name):
        Xp = Xd if order is None else Xd.index_select(0, torch.from_numpy(order).to(dev))
        ms = []
        for _ in range(a.reps + 2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            forest.predict_device(Xp.data_ptr(), n, dim, prob.data_ptr(), stream=stream)
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        p = prob.cpu().numpy()
        print("%-58s %.3f ms (min of %d)" % (name, min(ms[2:]), a.reps), flush=True)
        return p

    p0 = timed(None, "boundary order (as the library emits the candidates)")
    rng = np.random.default_rng(0)
    perm = rng.permutation(n)
    p = timed(perm, "random permutation")
    assert np.array_equal(p, p0[perm])
    for trees, levels in (((0,), 8), ((0, 1), 8), ((0, 1, 2, 3), 4), ((0, 1, 2, 3, 4, 5, 6, 7), 2)):
        k = top_path_key(farr, rows, trees, levels)
        o = np.argsort(k, kind="stable")
        p = timed(o, "top %d levels of trees %s (%d distinct keys)" % (levels, list(trees), len(np.unique(k))))
        assert np.array_equal(p, p0[o])
    leaf, depth = leaf_key(farr, rows, 0)
    o = np.argsort(leaf, kind="stable")
    timed(o, "leaf of tree 0 (%d leaves reached)" % len(np.unique(leaf)))
    o = np.argsort(p0, kind="stable")
    timed(o, "sorted by the forest's own output (not available in advance)")
    # total path length over a sample of trees: the best any ordering could do for idle lanes
    tot = np.zeros(n, np.int64)
    for t in range(0, 255, 16):
        tot += leaf_key(farr, rows, t)[1]
    o = np.argsort(tot, kind="stable")
    timed(o, "sorted by summed path length over 16 trees (not available)")
    b.close()


if __name__ == "__main__":
    main()
