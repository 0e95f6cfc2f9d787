"""Golden vectors produced by the unmodified reference (tests/golden/make_golden.py).
CPU: the oracle reproduces them bit-for-bit.  GPU: the CUDA path reproduces the RAG exactly and
the feature rows within 1e-5 relative (bit-identical except log/log2/pow terms)."""
import glob
import os

import numpy as np
import pytest

import oracle_lib as ol

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz")))
RTOL = 1e-5


def quant_pred(row):
    s = float(np.sum(np.abs(row[:20]))) * 3.7
    return np.floor((s - np.floor(s)) * 8) / 8.0


def _load(path):
    z = np.load(path)
    mask = z["mask"] if z["mask"].size else None
    ns, lg, simp = (bool(v) for v in z["flags"])
    return z, z["labels"], z["pb"], mask, dict(normalize=ns, use_log=lg, simple=simp)


def test_fixtures_present():
    assert len(GOLD) >= 5


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_oracle_reproduces_reference_golden(path):
    z, lab, pb, mask, kw = _load(path)
    o = ol.Oracle(lab, pb, ol.default_config(pb, **kw), mask=mask)
    r, e = o.regions(), o.edges()
    assert np.array_equal(r["label"], z["reg_label"]) and np.array_equal(r["count"], z["reg_count"])
    assert np.array_equal(r["border"], z["reg_border"])
    assert np.array_equal(e["a"], z["edge_a"]) and np.array_equal(e["b"], z["edge_b"])
    assert np.array_equal(e["count"], z["edge_count"])
    n0 = int(z["n_init"])
    rk = z["cand_keys"][:n0]
    ks = np.sort(rk[:, :2], axis=1)
    perm = np.lexsort((ks[:, 1], ks[:, 0]))
    swap = (rk[perm, 0] > rk[perm, 1]).astype(np.uint8)
    res = o.greedy(fn=quant_pred, want_cands=True, init_swap=swap)
    assert np.array_equal(res["order"], z["order"])
    assert np.array_equal(res["sal"], z["sal"])
    assert np.array_equal(res["rows"], z["merge_rows"])
    assert np.array_equal(res["cand_rows"][n0:], z["cand_rows"][n0:])
    assert np.array_equal(o.bc_feat(z["order"]), z["bcfeat_rows"])
    assert np.array_equal(o.bc_feat(z["order"], sal=z["sal"], init_sal=1.0, sal_bias=0.25), z["bcfeat_rows_sal"])


def _swap_x1_x2(row, ndim, simple, dim):
    """The reference row with the roles of x1 and x2 exchanged (hmt/bc_feat.hxx:232-238 layout)."""
    r = row.copy()
    if simple:  # selectFeatures: area1, area2, perim1, perim2, ... (hmt/bc_feat.hxx:247-279); the rest is symmetric
        r[[0, 1, 2, 3]] = row[[1, 0, 3, 2]]
        return r
    xi = ndim + 4 + 2 * 3 + 5 + 5
    x0 = dim - 3 * xi
    # x0: |a1-a2|, ./a1, ./a2, |p1-p2|, ./p1, ./p2, bl, bl/a1, bl/a2, bl/p1, bl/p2, vb[t], vb/bl, vb/p1, vb/p2 (3 each)
    r[[1, 2]] = row[[2, 1]]; r[[4, 5]] = row[[5, 4]]; r[[7, 8]] = row[[8, 7]]; r[[9, 10]] = row[[10, 9]]
    r[17:20], r[20:23] = row[20:23].copy(), row[17:20].copy()
    r[x0:x0 + xi], r[x0 + xi:x0 + 2 * xi] = row[x0 + xi:x0 + 2 * xi].copy(), row[x0:x0 + xi].copy()
    return r


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_cuda_reproduces_reference_golden(path):
    from glia_b200 import api
    z, lab, pb, mask, kw = _load(path)
    g = api.MergeTreeBuilder(lab.shape, api.default_config(pb, **kw)).build(lab, pb, mask)
    r, e = g.regions(), g.edges()
    assert np.array_equal(r["label"], z["reg_label"]) and np.array_equal(r["count"], z["reg_count"])
    assert np.array_equal(r["border"], z["reg_border"])
    assert np.array_equal(e["a"], z["edge_a"]) and np.array_equal(e["b"], z["edge_b"])
    assert np.array_equal(e["count"], z["edge_count"])
    n0 = int(z["n_init"])
    keys, rows = g.init_candidates()
    rk = z["cand_keys"][:n0]
    gold = {tuple(sorted(k[:2])): (z["cand_rows"][i], int(k[0] > k[1])) for i, k in enumerate(rk)}
    assert len(gold) == len(keys)
    cnt = dict(zip(r["label"].tolist(), r["count"].tolist()))
    tie, nties = g.candidate_ties()
    assert nties == int(tie.sum()) == sum(cnt[int(k[0])] == cnt[int(k[1])] for k in keys)
    # (1) default orientation (x1 = smaller label on equal areas): every row equals the reference row, or - for an
    #     enumerated equal-area pair that the reference happened to call the other way round - the reference row
    #     with its x1 / x2 parts exchanged.  Nothing is skipped.
    nswapped = 0
    for i, (k, row) in enumerate(zip(keys, rows)):
        ref_row, ref_second_first = gold[(int(k[0]), int(k[1]))]
        if tie[i] and ref_second_first:
            ref_row = _swap_x1_x2(ref_row, lab.ndim, kw["simple"], rows.shape[1]); nswapped += 1
        np.testing.assert_allclose(row, ref_row, rtol=RTOL, atol=1e-12)
    # (2) the replay of the reference's unordered_map order gives the call order of EVERY candidate of the compiled
    #     reference (oracle/ref_driver.cxx reserves 4 n + 16 buckets before the loop), ties or not
    emu = g.reference_orientation(4 * g.R + 16)
    ref_bits = np.array([gold[(int(k[0]), int(k[1]))][1] for k in keys], np.uint8)
    assert np.array_equal(emu, ref_bits)
    # (3) with those bits installed all rows match the reference as they stand
    g.set_tie_orientation(emu)
    _, rows2 = g.init_candidates()
    for k, row in zip(keys, rows2):
        np.testing.assert_allclose(row, gold[(int(k[0]), int(k[1]))][0], rtol=RTOL, atol=1e-12)
    assert nswapped == int((tie & emu).sum())
    g.set_tie_orientation(None)
    np.testing.assert_allclose(g.bc_feat(z["order"]), z["bcfeat_rows"], rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(g.bc_feat(z["order"], sal=z["sal"], init_sal=1.0, sal_bias=0.25),
                               z["bcfeat_rows_sal"], rtol=RTOL, atol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("path", [p for p in GOLD if "simple" not in p], ids=lambda p: os.path.basename(p)[:-4])
def test_cuda_greedy_with_reference_call_order(path):
    """Greedy merge order with the reference's equal-area call order installed: order and saliencies equal the
    oracle run that is fed the compiled reference's own swap bits (the oracle is pinned to the reference with them)."""
    from glia_b200 import api, synth
    z, lab, pb, mask, kw = _load(path)
    co, cg = ol.default_config(pb, **kw), api.default_config(pb, **kw)
    o = ol.Oracle(lab, pb, co, mask=mask)
    g = api.MergeTreeBuilder(lab.shape, cg).build(lab, pb, mask)
    n0 = int(z["n_init"])
    rk = z["cand_keys"][:n0]
    ks = np.sort(rk[:, :2], axis=1)
    perm = np.lexsort((ks[:, 1], ks[:, 0]))
    swap = (rk[perm, 0] > rk[perm, 1]).astype(np.uint8)
    assert np.array_equal(g.reference_orientation(4 * g.R + 16), swap)
    g.set_tie_orientation(swap)
    _, rowo = o.init_candidates(swap=swap)
    _, rowg = g.init_candidates()
    cols = np.flatnonzero((rowo == rowg).all(axis=0))
    # the forest looks at the x1 / x2 blocks, where the orientation shows
    f = synth.random_forest(15, o.row_dim(), max_depth=5, seed=3, thr_pool=rowg, features=cols)
    orf, grf = ol.OracleRF.from_arrays(f, o.row_dim()), api.Forest.from_arrays(f, o.row_dim())
    ro = o.greedy(rf=orf, init_swap=swap)
    rg = g.merge_order_bc(grf)
    assert np.array_equal(rg["order"], ro["order"])
    assert np.array_equal(rg["sal"], ro["sal"])


# ---------------------------------------------------------------- optional feature layouts (hist / median builds)
VGOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "variants", "*.npz")))


def _vload(path, mod):
    z = np.load(path)
    mask = z["mask"] if z["mask"].size else None
    ns, lg, simp, uh, um = (bool(v) for v in z["flags"])
    kw = dict(normalize=ns, use_log=lg, simple=simp, use_hist=uh, use_median=um)
    pb = z["pb"]
    if z["img2"].size:
        img2 = z["img2"]
        cfg = mod.FeatureConfig(rimgs=[(pb, 16, 0.0, 1.0), (img2, 8, 0.0, 1.0)], rlimgs=[(img2, 4, 0.0, 1.0)],
                                bimgs=[(pb, 16, 0.0, 1.0), (img2, 5, 0.0, 1.0)], **kw)
    else:
        cfg = mod.default_config(pb, **kw)
    return z, z["labels"], pb, mask, cfg, kw


def _init_swap(z):
    rk = z["init_keys"]
    ks = np.sort(rk[:, :2], axis=1)
    assert np.array_equal(ks, ks[np.lexsort((ks[:, 1], ks[:, 0]))])  # mode 1 emits in table (key) order
    return ks, (rk[:, 0] > rk[:, 1]).astype(np.uint8)


def test_variant_fixtures_present():
    assert len(VGOLD) >= 5


@pytest.mark.parametrize("path", VGOLD, ids=[os.path.basename(p)[:-4] for p in VGOLD])
def test_oracle_reproduces_variant_golden(path):
    """Fixtures from reference builds with GLIA_USE_HISTOGRAM_AS_FEATS / GLIA_USE_MEDIAN_AS_FEATS.  Median builds compute
    stddev two-pass over a pixel order the oracle does not keep (util/stats.hxx:62-71): 1e-12 there, bit-exact elsewhere."""
    z, lab, pb, mask, cfg, kw = _vload(path, ol)
    o = ol.Oracle(lab, pb, cfg, mask=mask)
    ks, swap = _init_swap(z)

    def same(a, b):
        assert a.shape == b.shape
        if kw["use_median"]:
            np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-14)
        else:
            assert np.array_equal(a, b)

    keys, rows = o.init_candidates(swap=swap)
    assert np.array_equal(keys, ks)
    same(rows, z["init_rows"])
    n0 = int(z["n_init"])
    rk = z["cand_keys"][:n0]
    k2 = np.sort(rk[:, :2], axis=1)
    perm = np.lexsort((k2[:, 1], k2[:, 0]))
    res = o.greedy(fn=quant_pred, init_swap=(rk[perm, 0] > rk[perm, 1]).astype(np.uint8))
    assert np.array_equal(res["order"], z["order"]) and np.array_equal(res["sal"], z["sal"])
    same(res["rows"], z["merge_rows"])
    if not kw["simple"]:
        same(o.bc_feat(z["order"]), z["bcfeat_rows"])
        same(o.bc_feat(z["order"], sal=z["sal"], init_sal=1.0, sal_bias=0.25), z["bcfeat_rows_sal"])


@pytest.mark.gpu
@pytest.mark.parametrize("path", VGOLD, ids=[os.path.basename(p)[:-4] for p in VGOLD])
def test_cuda_reproduces_variant_golden(path):
    """The CUDA path against the variant reference builds: initial rows with the reference's call order replayed
    (orient.cu), and the bc_feat stage.  Rows within 1e-5 relative."""
    from glia_b200 import api
    z, lab, pb, mask, cfg, kw = _vload(path, api)
    g = api.MergeTreeBuilder(lab.shape, cfg).build(lab, pb, mask)
    ks, swap = _init_swap(z)
    assert g.row_dim() == z["init_rows"].shape[1]
    emu = g.reference_orientation(4 * g.R + 16)
    assert np.array_equal(emu, swap)
    g.set_tie_orientation(emu)
    keys, rows = g.init_candidates()
    assert np.array_equal(keys, ks)
    np.testing.assert_allclose(rows, z["init_rows"], rtol=RTOL, atol=1e-9)
    g.set_tie_orientation(None)
    if not kw["simple"]:
        np.testing.assert_allclose(g.bc_feat(z["order"]), z["bcfeat_rows"], rtol=RTOL, atol=1e-9)
        np.testing.assert_allclose(g.bc_feat(z["order"], sal=z["sal"], init_sal=1.0, sal_bias=0.25),
                                   z["bcfeat_rows_sal"], rtol=RTOL, atol=1e-9)


# ---------------------------------------------------------------- merge_order_pb / pre_merge
PGOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pb", "*.npz")))


def _pb_case(path, mod, **kw):
    z = np.load(path)
    mask = z["mask"] if z["mask"].size else None
    spec = [(z["pb"], 1, 0.0, 1.0)]
    return z, z["labels"], z["pb"], mask, spec


def test_pb_fixtures_present():
    assert len(PGOLD) >= 3


@pytest.mark.parametrize("path", PGOLD, ids=[os.path.basename(p)[:-4] for p in PGOLD])
def test_oracle_reproduces_pb_golden(path):
    z, lab, pb, mask, spec = _pb_case(path, ol)
    for t, key in ((1, "median"), (2, "mean")):
        o = ol.Oracle(lab, pb, ol.FeatureConfig(bimgs=spec, thr=(), use_median=(t == 1), contour_only=True), mask=mask)
        r = o.merge_order_pb(t)
        assert np.array_equal(r["order"], z[key + "_order"]) and np.array_equal(r["sal"], z[key + "_sal"])
    r = ol.Oracle(lab, pb, ol.FeatureConfig(rimgs=spec, bimgs=spec, thr=()), mask=mask).pre_merge((25, 60), 0.4)
    assert np.array_equal(r["order"], z["pm_order"]) and np.array_equal(r["sal"], z["pm_sal"])
    assert np.array_equal(r["labels"], z["pm_labels"])


@pytest.mark.gpu
@pytest.mark.parametrize("path", PGOLD, ids=[os.path.basename(p)[:-4] for p in PGOLD])
def test_cuda_reproduces_pb_golden(path):
    from glia_b200 import api
    z, lab, pb, mask, spec = _pb_case(path, api)
    for t, key in ((1, "median"), (2, "mean")):
        g = api.MergeTreeBuilder(lab.shape, api.FeatureConfig(bimgs=spec, thr=(), use_median=(t == 1), contour_only=True))
        r = g.build(lab, pb, mask).merge_order_pb(t)
        assert np.array_equal(r["order"], z[key + "_order"]) and np.array_equal(r["sal"], z[key + "_sal"])
        g.close()
    g = api.MergeTreeBuilder(lab.shape, api.FeatureConfig(rimgs=spec, bimgs=spec, thr=())).build(lab, pb, mask)
    r = g.pre_merge((25, 60), 0.4)
    assert np.array_equal(r["order"], z["pm_order"]) and np.array_equal(r["sal"], z["pm_sal"])
    assert np.array_equal(r["labels"], z["pm_labels"])


# ---------------------------------------------------------------- segment_greedy
SGOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "seg", "*.npz")))
SEG_VARIANTS = (("probs", True, False, True), ("noprobs", False, False, True), ("rp", True, True, True), ("exact", True, False, False))


def _seg_case(path):
    z = np.load(path)
    n = int(z["ntree"])
    mask = z["mask"] if z["mask"].size else None
    return z, z["labels"], mask, [z["order%d" % i] for i in range(n)], [z["probs%d" % i] for i in range(n)], \
        [z["rprobs%d" % i] for i in range(n)]


def test_seg_fixtures_present():
    assert len(SGOLD) >= 3


@pytest.mark.parametrize("path", SGOLD, ids=[os.path.basename(p)[:-4] for p in SGOLD])
def test_python_oracle_reproduces_seg_golden(path):
    import sys
    sys.path.insert(0, ol.ORACLE_DIR)
    import segment_oracle as so
    z, lab, mask, orders, probs, rprobs = _seg_case(path)
    for tag, use_p, use_rp, ignore in SEG_VARIANTS:
        l, pk, _ = so.segment_greedy(lab, orders, merge_probs=probs if use_p else None, region_probs=rprobs if use_rp else None,
                                     mask=mask, ignore=ignore)
        assert np.array_equal(np.array(pk, np.int64).reshape(-1, 2), z["picks_" + tag]), tag
        assert np.array_equal(l, z["labels_" + tag]), tag


@pytest.mark.gpu
@pytest.mark.parametrize("path", SGOLD, ids=[os.path.basename(p)[:-4] for p in SGOLD])
def test_cuda_reproduces_seg_golden(path):
    from glia_b200 import api
    z, lab, mask, orders, probs, rprobs = _seg_case(path)
    zero = np.zeros(lab.shape, np.float32)
    g = api.MergeTreeBuilder(lab.shape, api.FeatureConfig(thr=())).build(lab, zero, mask)
    for tag, use_p, use_rp, ignore in SEG_VARIANTS:
        l, pk = g.segment_greedy(orders, merge_probs=probs if use_p else None, region_probs=rprobs if use_rp else None,
                                 ignore_missing=ignore)
        assert np.array_equal(pk.astype(np.int64), z["picks_" + tag]), tag
        assert np.array_equal(l, z["labels_" + tag]), tag


# ---------------------------------------------------------------- bc label generation
LGOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bclabel", "*.npz")))
BCL_VARIANTS = {"default": dict(), "tweak": dict(tweak=True), "mpd": dict(max_prec_drop=0.05), "optsplit": dict(opt_split=True),
                "ri": dict(use_f1=False), "g1": dict(global_opt=1), "g2": dict(global_opt=2)}


def test_bclabel_fixtures_present():
    assert len(LGOLD) >= 3


@pytest.mark.parametrize("path", LGOLD, ids=[os.path.basename(p)[:-4] for p in LGOLD])
def test_python_oracle_reproduces_bclabel_golden(path):
    import sys
    sys.path.insert(0, ol.ORACLE_DIR)
    import bc_label_oracle as bo
    z = np.load(path)
    mask = z["mask"] if z["mask"].size else None
    for tag, kw in BCL_VARIANTS.items():
        l, ms, ss = bo.bc_labels(z["labels"], z["truth"], z["order"], mask=mask, **kw)
        assert np.array_equal(l, z["bc_" + tag]), tag
        if kw.get("global_opt", 0) == 0:
            assert np.array_equal(ms, z["ms_" + tag], equal_nan=True) and np.array_equal(ss, z["ss_" + tag], equal_nan=True), tag


@pytest.mark.gpu
@pytest.mark.parametrize("path", LGOLD, ids=[os.path.basename(p)[:-4] for p in LGOLD])
def test_cuda_reproduces_bclabel_golden(path):
    from glia_b200 import api
    z = np.load(path)
    lab = z["labels"]
    mask = z["mask"] if z["mask"].size else None
    g = api.MergeTreeBuilder(lab.shape, api.FeatureConfig(thr=())).build(lab, np.zeros(lab.shape, np.float32), mask)
    for tag, kw in BCL_VARIANTS.items():
        l, ms, ss = g.bc_label(z["truth"], z["order"], **kw)
        assert np.array_equal(l, z["bc_" + tag]), tag
        if kw.get("global_opt", 0) == 0:   # integer pair counts -> the same doubles, bit for bit
            assert np.array_equal(ms, z["ms_" + tag], equal_nan=True) and np.array_equal(ss, z["ss_" + tag], equal_nan=True), tag


# ---------------------------------------------------------------- section-classifier features
CGOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scfeat", "*.npz")))
SC_VARIANTS = {"default": dict(), "logs": dict(use_log=True), "raw": dict(norm_area=1.0, norm_len=1.0),
               "raw_logs": dict(norm_area=1.0, norm_len=1.0, use_log=True)}


def _sc_case(path):
    z = np.load(path)
    return z, (z["mask0"] if z["mask0"].size else None), (z["mask1"] if z["mask1"].size else None)


def test_scfeat_fixtures_present():
    assert len(CGOLD) >= 2


@pytest.mark.parametrize("path", CGOLD, ids=[os.path.basename(p)[:-4] for p in CGOLD])
def test_oracle_reproduces_scfeat_golden(path):
    z, m0, m1 = _sc_case(path)
    o0 = ol.Oracle(z["lab0"], z["pb0"], ol.default_config(z["pb0"]), mask=m0)
    o1 = ol.Oracle(z["lab1"], z["pb1"], ol.default_config(z["pb1"]), mask=m1)
    for tag, kw in SC_VARIANTS.items():
        np.testing.assert_allclose(ol.sc_feat(o0, o1, z["pairs"], **kw), z["rows_" + tag], rtol=1e-12, atol=1e-300, err_msg=tag)


@pytest.mark.gpu
@pytest.mark.parametrize("path", CGOLD, ids=[os.path.basename(p)[:-4] for p in CGOLD])
def test_cuda_reproduces_scfeat_golden(path):
    """Rows within 1e-5 relative; absolute slack only where a third-order moment of a near-symmetric region cancels."""
    from glia_b200 import api
    z, m0, m1 = _sc_case(path)
    g0 = api.MergeTreeBuilder(z["lab0"].shape, api.default_config(z["pb0"])).build(z["lab0"], z["pb0"], m0)
    g1 = api.MergeTreeBuilder(z["lab1"].shape, api.default_config(z["pb1"])).build(z["lab1"], z["pb1"], m1)
    for tag, kw in SC_VARIANTS.items():
        got, want = g0.sc_feat(g1, z["pairs"], **kw), z["rows_" + tag]
        assert got.shape == want.shape
        if kw.get("use_log"):
            # the logarithm of a central-moment difference (columns 15..21) that is rounding noise of the reference's
            # pixel sums (regions of equal shape) carries no information: compared where the difference is real
            raw = z["rows_" + ("raw" if "norm_len" in kw else "default")][:, 15:22]
            noise = raw < 1e-7 * np.maximum(1.0, np.abs(raw).max(axis=0))
            got[:, 15:22][noise] = want[:, 15:22][noise]
        np.testing.assert_allclose(got, want, rtol=RTOL, atol=1e-9 * max(1.0, float(np.abs(want).max())), err_msg=tag)
