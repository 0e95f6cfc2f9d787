"""Secondary metric: wall time of the full greedy merge order (merge_order_bc stage) and
of the bc_feat stage through the C ABI with host buffers.  usage: python profiles/measure_merge_order.py"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from glia_b200 import api, synth  # noqa: E402

for shape, cell in (((512, 512), 11), ((1024, 1024), 11), ((2048, 2048), 13), ((4096, 4096), 13), ((128, 128, 128), 10)):
    lab, pb = synth.voronoi_volume(shape, cell, seed=1234)
    cfg = api.default_config(pb)
    g = api.MergeTreeBuilder(shape, cfg, label_bound=int(lab.max()) + 1)
    t0 = time.perf_counter(); g.build(lab, pb); t1 = time.perf_counter()
    _, rows = g.init_candidates()
    f = synth.random_forest(255, g.row_dim(), max_depth=20, seed=0, thr_pool=rows, leaf_prob=0.3)
    forest = api.Forest.from_arrays(f, g.row_dim())
    g.merge_order_bc(forest) if shape == (512, 512) else None  # warm-up
    t2 = time.perf_counter(); r = g.merge_order_bc(forest); t3 = time.perf_counter()
    t4 = time.perf_counter(); g.bc_feat(r["order"], r["sal"]); t5 = time.perf_counter()
    print("%-14s R=%7d C0=%8d merges=%7d featurised=%9d | build %.3f s, merge_order_bc %.3f s (%.0f candidates/s, %.1f us/merge), bc_feat %.3f s"
          % ("x".join(map(str, shape)), g.R, g.Cn, len(r["order"]), r["n_featurized"], t1 - t0, t3 - t2,
             r["n_featurized"] / (t3 - t2), 1e6 * (t3 - t2) / max(1, len(r["order"])), t5 - t4), flush=True)
    g.close()
