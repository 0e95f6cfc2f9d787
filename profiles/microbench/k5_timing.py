"""Harness-only timing of the bc_feat stage (K5, glia_b200_bc_feat) for a spanning merge order of the whole graph:
C2 size on the GPU, and the compiled reference's own bc_feat call sequence (oracle/_ref, mode 3, all host threads) on
a C1-size slice beside it.    python profiles/microbench/k5_timing.py"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from glia_b200 import api  # noqa: E402


def spanning_order(keys, labels, seed=0):
    """A valid merge order (R - 1 merges for a connected graph): candidates in random order, merged when their ends lie
    in different current regions (union-find); merged ids continue after the largest label."""
    rng = np.random.default_rng(seed)
    idx = {int(l): i for i, l in enumerate(labels)}
    parent = list(range(len(labels)))
    cur = [int(l) for l in labels]
    nxt = int(labels.max()) + 1
    out = []

    def find(x):
        while parent[x] != x:
            parent[x] = parent[parent[x]]
            x = parent[x]
        return x
    for a, b in keys[rng.permutation(len(keys))]:
        ra, rb = find(idx[int(a)]), find(idx[int(b)])
        if ra == rb:
            continue
        out.append((cur[ra], cur[rb], nxt))
        parent[rb] = ra
        cur[ra] = nxt
        nxt += 1
    return np.array(out, np.uint32)


def synth(shape, cell, seed=1234):
    L = api.lib()
    nd, n = len(shape), int(np.prod(shape))
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    pb = torch.empty(n, dtype=torch.float32, device="cuda")
    dims = (C.c_int * 3)(*(list(shape[::-1]) + [1] * (3 - nd)))
    nc = C.c_uint64(0)
    assert L.glia_b200_synth_voronoi(nd, dims, cell, seed, lab.data_ptr(), pb.data_ptr(), None, C.byref(nc)) == 0
    torch.cuda.synchronize()
    return lab.cpu().numpy().view(np.uint32).reshape(shape), pb.cpu().numpy().reshape(shape), int(nc.value) + 1


def main():
    for shape, cell in (((4096, 4096), 13), ((256, 256, 256), 10)):
        lab, pb, lb = synth(shape, cell)
        spec = [(pb, 16, 0.0, 1.0)]
        g = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec), label_bound=lb).build(lab, pb)
        keys, _ = g.init_candidates(rows=False)
        order = spanning_order(keys, g.regions()["label"])
        g.bc_feat(order)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); rows = g.bc_feat(order); ts.append(time.perf_counter() - t0)
        print("GPU  bc_feat %s: R=%d merges=%d dim=%d  %.1f ms per call (host order in, host rows out, %.0f MB)  = %.2f M merges/s"
              % (shape, g.R, len(order), rows.shape[1], 1e3 * min(ts), rows.nbytes / 1e6, len(order) / min(ts) / 1e6), flush=True)
        g.close()
    try:
        import oracle_lib as ol
        shape, cell = (512, 512), 11
        lab, pb, lb = synth(shape, cell)
        spec = [(pb, 16, 0.0, 1.0)]
        cfg = ol.FeatureConfig(rimgs=spec, bimgs=spec)
        g = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec), label_bound=lb).build(lab, pb)
        keys, _ = g.init_candidates(rows=False)
        order = spanning_order(keys, g.regions()["label"])
        t0 = time.perf_counter(); rows = g.bc_feat(order); tg = time.perf_counter() - t0
        t0 = time.perf_counter(); ref = ol.ref_run(3, lab, pb, cfg, order=order, threads=0); tr = time.perf_counter() - t0
        same = np.allclose(rows, ref["merge_rows"], rtol=1e-5, atol=1e-12)
        print("REF  bc_feat %s (oracle/_ref, %d threads): merges=%d  %.2f s total (rmap %.2f s, features %.2f s) | GPU %.1f ms | rows agree: %s"
              % (shape, os.cpu_count(), len(order), tr, ref["t_rmap"], ref["t_feat"], 1e3 * tg, same), flush=True)
    except Exception as e:  # noqa: BLE001
        print("reference leg unavailable:", e)


if __name__ == "__main__":
    main()
