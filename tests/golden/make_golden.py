"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref, compiled from
/root/reference/code by oracle/Makefile) on small seeded inputs.  Run in the build container:
    python tests/golden/make_golden.py
The fixtures pin both the CPU oracle (tests/test_golden.py, CPU) and the CUDA path (GPU)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from glia_b200 import synth  # noqa: E402


def quant_pred(row):
    s = float(np.sum(np.abs(row[:20]))) * 3.7
    return np.floor((s - np.floor(s)) * 8) / 8.0


CASES = {
    # name: (shape, nlabels, seed, mask?, normalize, logs, simple)
    "g2d_plain": ((22, 26), 24, 21, False, False, False, False),
    "g2d_mask_ns_logs": ((19, 23), 14, 22, True, True, True, False),
    "g2d_simple": ((18, 20), 16, 23, False, False, False, True),
    "g3d_plain": ((8, 9, 10), 12, 24, False, False, False, False),
    "g3d_mask_ns_logs": ((7, 10, 9), 9, 25, True, True, True, False),
}


def main():
    ol.build_ref()
    for name, (shape, nl, seed, um, ns, lg, simp) in CASES.items():
        lab, pb = synth.speckle_volume(shape, nl, seed=seed)
        mask = None
        if um:
            mask = (np.random.default_rng(seed + 100).random(shape) > 0.05).astype(np.uint32)
        cfg = ol.default_config(pb, normalize=ns, use_log=lg, simple=simp)
        r0 = ol.ref_run(0, lab, pb, cfg, mask=mask)
        r2 = ol.ref_run(2, lab, pb, cfg, mask=mask, pred=quant_pred)
        n0 = len(ol.ref_run(1, lab, pb, cfg, mask=mask, pred=quant_pred)["cand_keys"])
        r3 = ol.ref_run(3, lab, pb, cfg, mask=mask, order=r2["order"], sal=r2["sal"], init_sal=1.0, sal_bias=0.25)
        r3n = ol.ref_run(3, lab, pb, cfg, mask=mask, order=r2["order"])
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"), labels=lab, pb=pb, mask=np.zeros(0, np.uint32) if mask is None else mask,
            flags=np.array([ns, lg, simp], np.int32),
            reg_label=r0["reg_label"], reg_count=r0["reg_count"], reg_border=r0["reg_border"],
            edge_a=r0["edge_a"], edge_b=r0["edge_b"], edge_count=r0["edge_count"],
            n_init=np.int64(n0), cand_keys=r2["cand_keys"], cand_rows=r2["cand_rows"], cand_scores=r2["cand_scores"],
            order=r2["order"], sal=r2["sal"], merge_rows=r2["merge_rows"],
            bcfeat_rows_sal=r3["merge_rows"], bcfeat_rows=r3n["merge_rows"])
        print(name, shape, "R=%d E=%d cands=%d merges=%d dim=%d" % (
            len(r0["reg_label"]), len(r0["edge_a"]), len(r2["cand_keys"]), len(r2["order"]), r2["dim"]))


if __name__ == "__main__":
    main()
