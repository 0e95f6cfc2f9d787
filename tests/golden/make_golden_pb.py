"""Generates tests/golden/pb/*.npz from the compiled, unmodified reference (oracle/_ref): the call sequences of
hmt/main_merge_order_pb.cxx (types 1 and 2) and gadget/main_pre_merge.cxx (without the ITK relabel) on small seeded
inputs.    python tests/golden/make_golden_pb.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from glia_b200 import synth  # noqa: E402

CASES = {
    "p2d": ((26, 30), 45, 51, False),
    "p2d_mask": ((23, 27), 30, 52, True),
    "p3d_mask": ((8, 11, 12), 18, 53, True),
}


def main():
    ol.build_ref()
    os.makedirs(os.path.join(HERE, "pb"), exist_ok=True)
    for name, (shape, nl, seed, um) in CASES.items():
        lab, pb = synth.speckle_volume(shape, nl, seed=seed)
        mask = (np.random.default_rng(seed + 100).random(shape) > 0.04).astype(np.uint32) if um else None
        med = ol.ref_run_pb(4, lab, pb, mask=mask, type=1)
        mean = ol.ref_run_pb(4, lab, pb, mask=mask, type=2)
        pm = ol.ref_run_pb(5, lab, pb, mask=mask, size_thr=(25, 60), rpb_thr=0.4)
        np.savez_compressed(os.path.join(HERE, "pb", name + ".npz"), labels=lab, pb=pb,
                            mask=np.zeros(0, np.uint32) if mask is None else mask,
                            median_order=med["order"], median_sal=med["sal"], mean_order=mean["order"], mean_sal=mean["sal"],
                            pm_order=pm["order"], pm_sal=pm["sal"], pm_labels=pm["labels"])
        print(name, shape, "merges: median %d mean %d pre_merge %d" % (len(med["order"]), len(mean["order"]), len(pm["order"])))


if __name__ == "__main__":
    main()
