"""Generates tests/golden/scfeat/*.npz from the compiled reference (oracle/_ref glia_ref_sc_feat: the reference's
section-classifier feature classes, hmt/sc_feat.hxx, in the call sequence of hmt/main_sc_feat.cxx) on small seeded pairs
of 2D sections.    python tests/golden/make_golden_scfeat.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from glia_b200 import synth  # noqa: E402

CASES = {"sc2d": ((26, 30), 36, 131, False), "sc2d_mask": ((23, 29), 24, 132, True)}
VARIANTS = {"default": dict(), "logs": dict(use_log=True), "raw": dict(norm_area=1.0, norm_len=1.0), "raw_logs": dict(norm_area=1.0, norm_len=1.0, use_log=True)}


def sections(shape, nl, seed, um):
    lab0, pb0 = synth.speckle_volume(shape, nl, seed=seed)
    lab1 = np.roll(lab0, (1, 2), axis=(0, 1)).copy()
    rng = np.random.default_rng(seed + 9)
    lab1[rng.random(shape) < 0.05] = rng.integers(1, nl + 1)
    pb1 = np.roll(pb0, (1, 2), axis=(0, 1)).copy()
    m0 = (rng.random(shape) > 0.04).astype(np.uint32) if um else None
    m1 = (rng.random(shape) > 0.04).astype(np.uint32) if um else None
    ok0 = np.ones(shape, bool) if m0 is None else m0 != 0
    ok1 = np.ones(shape, bool) if m1 is None else m1 != 0
    l0, l1 = np.unique(lab0[ok0]), np.unique(lab1[ok1])
    pairs = np.stack([rng.choice(l0, 40), rng.choice(l1, 40)], axis=1).astype(np.uint32)
    pairs[:10, 1] = [l if l in set(l1.tolist()) else l1[0] for l in pairs[:10, 0]]
    return lab0, pb0, m0, lab1, pb1, m1, pairs


def main():
    ol.build_ref()
    os.makedirs(os.path.join(HERE, "scfeat"), exist_ok=True)
    for name, case in CASES.items():
        lab0, pb0, m0, lab1, pb1, m1, pairs = sections(*case)
        e = np.zeros(0, np.uint32)
        out = dict(lab0=lab0, pb0=pb0, lab1=lab1, pb1=pb1, mask0=e if m0 is None else m0, mask1=e if m1 is None else m1, pairs=pairs)
        for tag, kw in VARIANTS.items():
            out["rows_" + tag] = ol.ref_sc_feat(lab0, pb0, lab1, pb1, pairs, mask0=m0, mask1=m1, **kw)
        np.savez_compressed(os.path.join(HERE, "scfeat", name + ".npz"), **out)
        print(name, case[0], "pairs", len(pairs), "dim", out["rows_default"].shape[1])


if __name__ == "__main__":
    main()
