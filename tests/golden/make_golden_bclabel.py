"""Generates tests/golden/bclabel/*.npz from the compiled reference (oracle/_ref, glia_ref_bc_label: the reference's
label templates of hmt/bc_label.hxx with the option logic of hmt/main_bc_label_ri.cxx) on small seeded inputs.
    python tests/golden/make_golden_bclabel.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from glia_b200 import synth  # noqa: E402

CASES = {"l2d": ((26, 30), 40, 101, False, 5), "l2d_mask": ((23, 27), 28, 102, True, 4), "l3d_mask": ((8, 11, 12), 18, 103, True, 3)}
VARIANTS = {"default": dict(), "tweak": dict(tweak=True), "mpd": dict(max_prec_drop=0.05), "optsplit": dict(opt_split=True),
            "ri": dict(use_f1=False), "g1": dict(global_opt=1), "g2": dict(global_opt=2)}


def quant_pred(row):
    s = float(np.sum(np.abs(row[:20]))) * 3.7
    return np.floor((s - np.floor(s)) * 8) / 8.0


def truth_of(lab, seed, ntruth):
    rng = np.random.default_rng(seed + 500)
    group = rng.integers(1, ntruth + 1, size=int(lab.max()) + 1).astype(np.uint32)
    t = group[lab]
    noise = rng.random(lab.shape)
    t = np.where(noise < 0.10, rng.integers(1, ntruth + 1, size=lab.shape).astype(np.uint32), t)
    t[noise > 0.94] = 0
    return t.astype(np.uint32)


def main():
    ol.build_ref()
    os.makedirs(os.path.join(HERE, "bclabel"), exist_ok=True)
    for name, (shape, nl, seed, um, ntruth) in CASES.items():
        lab, pb = synth.speckle_volume(shape, nl, seed=seed)
        mask = (np.random.default_rng(seed + 100).random(shape) > 0.04).astype(np.uint32) if um else None
        truth = truth_of(lab, seed, ntruth)
        order = ol.Oracle(lab, pb, ol.default_config(pb), mask=mask).greedy(fn=quant_pred)["order"]
        out = dict(labels=lab, truth=truth, mask=np.zeros(0, np.uint32) if mask is None else mask, order=order)
        for tag, kw in VARIANTS.items():
            l, ms, ss = ol.ref_bc_label(lab, truth, order, mask=mask, **kw)
            out["bc_" + tag], out["ms_" + tag], out["ss_" + tag] = l, ms, ss
        np.savez_compressed(os.path.join(HERE, "bclabel", name + ".npz"), **out)
        print(name, shape, "merges", len(order), "merge labels (default)", int((out["bc_default"] == -1).sum()))


if __name__ == "__main__":
    main()
