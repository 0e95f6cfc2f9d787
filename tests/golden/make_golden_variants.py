"""Generates tests/golden/variants/*.npz by running reference builds compiled with the two optional feature layouts
(CMakeLists.txt:53-64: -DGLIA_USE_HISTOGRAM_AS_FEATS / -DGLIA_USE_MEDIAN_AS_FEATS; oracle/Makefile builds them from
/root/reference/code as oracle/_ref/libglia_ref{2d,3d}_{hist,median,histmedian}.so) on small seeded inputs:
    python tests/golden/make_golden_variants.py
The fixtures pin the oracle's restatement of both layouts (CPU) and the CUDA path (GPU)."""
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
    # name: (shape, nlabels, seed, mask?, normalize, logs, simple, hist, median, second image?)
    "v2d_hist": ((21, 25), 20, 31, False, False, False, False, True, False, False),
    "v2d_median_mask_logs": ((19, 22), 14, 32, True, True, True, False, False, True, False),
    "v2d_median_simple": ((18, 21), 15, 33, False, False, False, True, False, True, False),
    "v2d_histmedian_two": ((20, 23), 16, 34, False, False, False, False, True, True, True),
    "v3d_histmedian": ((7, 9, 10), 10, 35, True, False, True, False, True, True, False),
}


def second_image(shape, seed):
    return (np.random.default_rng(seed + 7).integers(0, 64, size=shape) / 64.0).astype(np.float32)


def config(mod, pb, img2, ns, lg, simp, uh, um):
    kw = dict(normalize=ns, use_log=lg, simple=simp, use_hist=uh, use_median=um)
    if img2 is None:
        return mod.default_config(pb, **kw)
    return mod.FeatureConfig(rimgs=[(pb, 16, 0.0, 1.0), (img2, 8, 0.0, 1.0)], rlimgs=[(img2, 4, 0.0, 1.0)],
                             bimgs=[(pb, 16, 0.0, 1.0), (img2, 5, 0.0, 1.0)], **kw)


def main():
    ol.build_ref()
    os.makedirs(os.path.join(HERE, "variants"), exist_ok=True)
    for name, (shape, nl, seed, um_, ns, lg, simp, uh, umed, two) in CASES.items():
        lab, pb = synth.speckle_volume(shape, nl, seed=seed)
        mask = (np.random.default_rng(seed + 100).random(shape) > 0.05).astype(np.uint32) if um_ else None
        img2 = second_image(shape, seed) if two else None
        cfg = config(ol, pb, img2, ns, lg, simp, uh, umed)
        r1 = ol.ref_run(1, lab, pb, cfg, mask=mask, pred=quant_pred)
        r2 = ol.ref_run(2, lab, pb, cfg, mask=mask, pred=quant_pred)
        out = dict(labels=lab, pb=pb, mask=np.zeros(0, np.uint32) if mask is None else mask,
                   img2=np.zeros(0, np.float32) if img2 is None else img2,
                   flags=np.array([ns, lg, simp, uh, umed], np.int32), n_init=np.int64(len(r1["cand_keys"])),
                   init_keys=r1["cand_keys"], init_rows=r1["cand_rows"], cand_keys=r2["cand_keys"],
                   order=r2["order"], sal=r2["sal"], merge_rows=r2["merge_rows"])
        if not simp:
            out["bcfeat_rows_sal"] = ol.ref_run(3, lab, pb, cfg, mask=mask, order=r2["order"], sal=r2["sal"], init_sal=1.0,
                                                sal_bias=0.25)["merge_rows"]
            out["bcfeat_rows"] = ol.ref_run(3, lab, pb, cfg, mask=mask, order=r2["order"])["merge_rows"]
        np.savez_compressed(os.path.join(HERE, "variants", name + ".npz"), **out)
        print(name, shape, "cands=%d merges=%d dim=%d" % (len(r1["cand_keys"]), len(r2["order"]), r2["dim"]))


if __name__ == "__main__":
    main()
