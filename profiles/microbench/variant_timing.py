"""Harness-only timing of the hist / median feature layouts (not a bench.py line): build + featurise the initial
candidates with device-resident inputs, default layout against use_hist + use_median.
    python profiles/microbench/variant_timing.py [c2|c4s]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from glia_b200 import api  # noqa: E402

WORK = {"c2": ((4096, 4096), 13), "c4s": ((512, 512, 512), 10)}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    shape, cell = WORK[name]
    nd, n = len(shape), int(np.prod(shape))
    L = api.lib()
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    pb = torch.empty(n, dtype=torch.float32, device="cuda")
    dims = (C.c_int * 3)(*(list(shape[::-1]) + [1] * (3 - nd)))
    nc = C.c_uint64(0)
    assert L.glia_b200_synth_voronoi(nd, dims, cell, 1234, lab.data_ptr(), pb.data_ptr(), None, C.byref(nc)) == 0
    torch.cuda.synchronize()
    spec = [(pb.data_ptr(), 16, 0.0, 1.0)]
    for kw in (dict(), dict(use_hist=True), dict(use_hist=True, use_median=True)):
        g = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec, **kw), device_pointers=True,
                                 label_bound=int(nc.value) + 1)
        g.build(lab.data_ptr(), pb.data_ptr())
        rows = torch.empty((g.Cn, g.row_dim()), dtype=torch.float64, device="cuda")
        keys = torch.empty((g.Cn, 2), dtype=torch.int32, device="cuda")
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        tb, tr = [], []
        for it in range(4):
            ev[0].record()
            g.build(lab.data_ptr(), pb.data_ptr())
            ev[1].record()
            g._check(L.glia_b200_init_candidates(g.h, C.c_void_p(keys.data_ptr()), C.c_void_p(rows.data_ptr())))
            ev[2].record()
            torch.cuda.synchronize()
            if it:
                tb.append(ev[0].elapsed_time(ev[1])); tr.append(ev[1].elapsed_time(ev[2]))
        print("%s %-34s R=%d C=%d dim=%3d  build %.2f ms  rows %.2f ms" % (name, kw or "default", g.R, g.Cn, g.row_dim(),
                                                                         np.mean(tb), np.mean(tr)))
        g.close()
        del rows, keys


if __name__ == "__main__":
    main()
