"""Generates tests/golden/seg/*.npz from the compiled, unmodified reference (oracle/_ref): the call sequence of
hmt/main_segment_greedy.cxx:30-95 (trees with node potentials, greedy resolution over one to three trees, final label
volume) on small seeded inputs.    python tests/golden/make_golden_seg.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from glia_b200 import synth  # noqa: E402

CASES = {"s2d_1tree": ((26, 30), 40, 71, False, 1), "s2d_mask_2trees": ((23, 27), 28, 72, True, 2),
         "s3d_3trees": ((8, 11, 12), 18, 73, False, 3)}


def inputs(shape, nl, seed, um, ntree):
    lab, pb = synth.speckle_volume(shape, nl, seed=seed)
    mask = (np.random.default_rng(seed + 100).random(shape) > 0.04).astype(np.uint32) if um else None
    rng = np.random.default_rng(seed)
    o = ol.Oracle(lab, pb, ol.default_config(pb), mask=mask)
    orders, probs, rprobs = [], [], []
    for t in range(ntree):
        if t == 0:
            res = ol.Oracle(lab, pb, ol.default_config(pb, contour_only=True), mask=mask).merge_order_pb(2)
        else:
            res = o.greedy(fn=lambda row, t=t: float(np.floor((np.sum(np.abs(row[:9])) * (3.1 + t)) % 1.0 * 16) / 16.0))
        orders.append(res["order"])
        probs.append(np.floor(rng.random(len(res["order"])) * 8) / 8.0)
        rprobs.append(np.floor(rng.random(2 * len(res["order"]) + 1) * 4) / 4.0)
    return lab, mask, orders, probs, rprobs


def main():
    ol.build_ref()
    os.makedirs(os.path.join(HERE, "seg"), exist_ok=True)
    for name, case in CASES.items():
        lab, mask, orders, probs, rprobs = inputs(*case)
        out = dict(labels=lab, mask=np.zeros(0, np.uint32) if mask is None else mask, ntree=np.int64(len(orders)))
        for i, (o, p, r) in enumerate(zip(orders, probs, rprobs)):
            out["order%d" % i], out["probs%d" % i], out["rprobs%d" % i] = o, p, r
        for tag, kw in (("probs", dict(merge_probs=probs)), ("noprobs", dict()), ("rp", dict(merge_probs=probs, region_probs=rprobs)),
                        ("exact", dict(merge_probs=probs, ignore=False))):
            l, pk, pot = ol.ref_segment_greedy(lab, orders, mask=mask, **kw)
            out["labels_" + tag], out["picks_" + tag] = l, pk
        np.savez_compressed(os.path.join(HERE, "seg", name + ".npz"), **out)
        print(name, case[0], "trees", len(orders), "picks", len(out["picks_probs"]))


if __name__ == "__main__":
    main()
