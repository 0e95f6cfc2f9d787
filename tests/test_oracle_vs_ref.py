"""Pins the CPU oracle (oracle/glia_oracle.c) against the UNMODIFIED reference headers compiled
into oracle/_ref (built from /root/reference/code by oracle/Makefile).  Skipped where oracle/_ref
is absent and cannot be built (the GPU box: /root/reference does not exist there, the prebuilt
.so files travel with the snapshot)."""
import numpy as np
import pytest

import oracle_lib as ol
from glia_b200 import synth


def _have_ref(nd):
    if ol.ref_available(nd):
        return True
    try:
        ol.build_ref()
    except Exception:
        return False
    return ol.ref_available(nd)


def quant_pred(row):
    """Deterministic predictor quantised to 1/8: heavy saliency ties."""
    s = float(np.sum(np.abs(row[:20]))) * 3.7
    return np.floor((s - np.floor(s)) * 8) / 8.0


CASES = [
    # (shape, nlabels, seed, mask?, normalize, logs, simple)
    ((24, 28), 30, 1, False, False, False, False),
    ((31, 17), 12, 2, True, True, True, False),
    ((20, 22), 25, 3, False, False, False, True),
    ((9, 10, 11), 14, 4, False, False, False, False),
    ((8, 12, 10), 10, 5, True, True, True, False),
]


def _make(shape, nlabels, seed, use_mask):
    lab, pb = synth.speckle_volume(shape, nlabels, seed=seed)
    mask = None
    if use_mask:
        rng = np.random.default_rng(seed + 100)
        mask = (rng.random(shape) > 0.04).astype(np.uint32)
    return lab, pb, mask


@pytest.mark.parametrize("case", CASES)
def test_rag_matches_reference(case):
    shape, nl, seed, um, ns, lg, simp = case
    if not _have_ref(len(shape)):
        pytest.skip("oracle/_ref not available")
    lab, pb, mask = _make(shape, nl, seed, um)
    cfg = ol.default_config(pb, normalize=ns, use_log=lg, simple=simp)
    ref = ol.ref_run(0, lab, pb, cfg, mask=mask)
    o = ol.Oracle(lab, pb, cfg, mask=mask)
    r, e = o.regions(), o.edges()
    assert np.array_equal(r["label"], ref["reg_label"])
    assert np.array_equal(r["count"], ref["reg_count"])
    assert np.array_equal(r["border"], ref["reg_border"])
    assert np.array_equal(e["a"], ref["edge_a"])
    assert np.array_equal(e["b"], ref["edge_b"])
    assert np.array_equal(e["count"], ref["edge_count"])


@pytest.mark.parametrize("case", CASES)
def test_init_candidates_match_reference(case):
    shape, nl, seed, um, ns, lg, simp = case
    if not _have_ref(len(shape)):
        pytest.skip("oracle/_ref not available")
    lab, pb, mask = _make(shape, nl, seed, um)
    cfg = ol.default_config(pb, normalize=ns, use_log=lg, simple=simp)
    ref = ol.ref_run(1, lab, pb, cfg, mask=mask, pred=quant_pred)
    o = ol.Oracle(lab, pb, cfg, mask=mask)
    rk = ref["cand_keys"]
    swap = (rk[:, 0] > rk[:, 1]).astype(np.uint8)
    keys, rows = o.init_candidates(swap=swap)
    assert np.array_equal(np.sort(rk[:, :2], axis=1), keys)
    assert rows.shape == ref["cand_rows"].shape
    # dyadic pb: every sum is exact, rows are bit-identical
    assert np.array_equal(rows, ref["cand_rows"])


@pytest.mark.parametrize("case", CASES)
def test_greedy_matches_reference(case):
    shape, nl, seed, um, ns, lg, simp = case
    if not _have_ref(len(shape)):
        pytest.skip("oracle/_ref not available")
    lab, pb, mask = _make(shape, nl, seed, um)
    cfg = ol.default_config(pb, normalize=ns, use_log=lg, simple=simp)
    ref = ol.ref_run(2, lab, pb, cfg, mask=mask, pred=quant_pred)
    o = ol.Oracle(lab, pb, cfg, mask=mask)
    # the initial featurisation direction of equal-area pairs follows the reference's
    # unordered_map iteration order: take it from the reference's own call log
    n0 = o.Cn
    rk = ref["cand_keys"][:n0]
    key_sorted = np.sort(rk[:, :2], axis=1)
    perm = np.lexsort((key_sorted[:, 1], key_sorted[:, 0]))
    swap = (rk[perm, 0] > rk[perm, 1]).astype(np.uint8)
    res = o.greedy(fn=quant_pred, want_cands=True, init_swap=swap)
    assert np.array_equal(res["order"], ref["order"])
    assert np.array_equal(res["sal"], ref["sal"])
    assert np.array_equal(res["rows"], ref["merge_rows"])
    assert res["n_cand"] == len(ref["cand_keys"])
    # candidates featurised during the loop (after the initial ones) come in the same sequence
    assert np.array_equal(res["cand_keys"][n0:, :2], ref["cand_keys"][n0:, :2])
    assert np.array_equal(res["cand_rows"][n0:], ref["cand_rows"][n0:])


@pytest.mark.parametrize("case", CASES[:2] + CASES[3:4])
def test_bc_feat_matches_reference(case):
    shape, nl, seed, um, ns, lg, simp = case
    if not _have_ref(len(shape)):
        pytest.skip("oracle/_ref not available")
    lab, pb, mask = _make(shape, nl, seed, um)
    cfg = ol.default_config(pb, normalize=ns, use_log=lg, simple=simp)
    o = ol.Oracle(lab, pb, cfg, mask=mask)
    res = o.greedy(fn=quant_pred)
    for sal in (None, res["sal"]):
        ref = ol.ref_run(3, lab, pb, cfg, mask=mask, order=res["order"], sal=sal, init_sal=1.0,
                         sal_bias=0.25)
        rows = o.bc_feat(res["order"], sal=sal, init_sal=1.0, sal_bias=0.25)
        assert rows.shape == ref["merge_rows"].shape
        assert np.array_equal(rows, ref["merge_rows"])


# ---- the reference's optional feature layouts: GLIA_USE_HISTOGRAM_AS_FEATS / GLIA_USE_MEDIAN_AS_FEATS builds
VARIANTS = [
    # (shape, nlabels, seed, mask?, normalize, logs, simple, hist, median)
    ((24, 28), 30, 11, False, False, False, False, True, False),
    ((27, 19), 16, 12, True, True, True, False, False, True),
    ((22, 25), 22, 13, False, False, False, True, False, True),
    ((23, 21), 18, 14, False, False, True, False, True, True),
    ((9, 10, 11), 14, 15, True, True, False, False, True, True),
]


def _have_variant(nd, cfg):
    if ol.ref_available(nd, cfg):
        return True
    try:
        ol.build_ref()
    except Exception:
        return False
    return ol.ref_available(nd, cfg)


def _variant_cfg(pb, case, two_images=False):
    shape, nl, seed, um, ns, lg, simp, uh, umed = case
    kw = dict(normalize=ns, use_log=lg, simple=simp, use_hist=uh, use_median=umed)
    if not two_images:
        return ol.default_config(pb, **kw)
    rng = np.random.default_rng(seed + 7)
    img2 = (np.floor(rng.random(pb.shape) * 64) / 64).astype(np.float32)
    return ol.FeatureConfig(rimgs=[(pb, 16, 0.0, 1.0), (img2, 8, 0.0, 1.0)], rlimgs=[(img2, 4, 0.0, 1.0)],
                            bimgs=[(pb, 16, 0.0, 1.0), (img2, 5, 0.0, 1.0)], **kw)


@pytest.mark.parametrize("two", [False, True])
@pytest.mark.parametrize("case", VARIANTS)
def test_variant_layouts_match_reference(case, two):
    """Initial candidates, greedy order and the bc_feat stage under the hist / median builds.  Dyadic pb: every
    column is bit-identical except the median builds' stddev, which the reference computes two-pass
    (util/stats.hxx:62-71) from a pixel order the oracle does not keep."""
    shape, nl, seed, um, ns, lg, simp, uh, umed = case
    lab, pb, mask = _make(shape, nl, seed, um)
    cfg = _variant_cfg(pb, case, two)
    if not _have_variant(len(shape), cfg):
        pytest.skip("oracle/_ref variant not available")
    o = ol.Oracle(lab, pb, cfg, mask=mask)

    def same(a, b):
        assert a.shape == b.shape
        if umed:
            np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-14)
        else:
            assert np.array_equal(a, b)

    ref = ol.ref_run(1, lab, pb, cfg, mask=mask, pred=quant_pred)
    rk = ref["cand_keys"]
    keys, rows = o.init_candidates(swap=(rk[:, 0] > rk[:, 1]).astype(np.uint8))
    assert np.array_equal(np.sort(rk[:, :2], axis=1), keys)
    assert rows.shape[1] == ref["dim"] == o.row_dim()
    same(rows, ref["cand_rows"])

    ref = ol.ref_run(2, lab, pb, cfg, mask=mask, pred=quant_pred)
    n0 = o.Cn
    rk = ref["cand_keys"][:n0]
    ks = np.sort(rk[:, :2], axis=1)
    perm = np.lexsort((ks[:, 1], ks[:, 0]))
    res = o.greedy(fn=quant_pred, want_cands=True, init_swap=(rk[perm, 0] > rk[perm, 1]).astype(np.uint8))
    assert np.array_equal(res["order"], ref["order"])
    same(res["rows"], ref["merge_rows"])
    if simp:
        return
    for sal in (None, res["sal"]):
        rb = ol.ref_run(3, lab, pb, cfg, mask=mask, order=res["order"], sal=sal, init_sal=1.0, sal_bias=0.25)
        rows = o.bc_feat(res["order"], sal=sal, init_sal=1.0, sal_bias=0.25)
        same(rows, rb["merge_rows"])


def test_nondyadic_within_tolerance():
    if not _have_ref(2):
        pytest.skip("oracle/_ref not available")
    lab, pb = synth.speckle_volume((26, 23), 20, seed=9, dyadic=False)
    cfg = ol.default_config(pb)
    ref = ol.ref_run(1, lab, pb, cfg, pred=quant_pred)
    o = ol.Oracle(lab, pb, cfg)
    rk = ref["cand_keys"]
    keys, rows = o.init_candidates(swap=(rk[:, 0] > rk[:, 1]).astype(np.uint8))
    np.testing.assert_allclose(rows, ref["cand_rows"], rtol=1e-5, atol=1e-12)


def test_rf_model_file_layout_pinned_by_reference_reader(tmp_path):
    """rf_old binary layout: a file written by the oracle is read back by the reference's own
    readModelFromBinaryFile (ml/rf/ml_rf_model.cxx:459-563), and a file re-written by the
    reference's writer is read back by the oracle, array for array."""
    import ctypes as C
    import os
    so = os.path.join(ol.ORACLE_DIR, "_ref", "libglia_refrf.so")
    if not os.path.exists(so):
        try:
            ol.build_ref()
        except Exception:
            pytest.skip("oracle/_ref not available")
    L = C.CDLL(so)
    assert L.glia_refrf_sizeof_model() == 520
    f = synth.random_forest(23, 11, max_depth=8, seed=5, nclass=3, orig_labels=(-1, 1, 2))
    orf = ol.OracleRF.from_arrays(f, 11)
    p1, p2 = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    assert orf.save(p1) == 0
    nt, nr, nc = C.c_int(), C.c_int(), C.c_int()
    L.glia_refrf_read.argtypes = [C.c_char_p] + [C.POINTER(C.c_int)] * 3 + [C.c_void_p] * 7
    L.glia_refrf_read(p1.encode(), C.byref(nt), C.byref(nr), C.byref(nc), *([None] * 7))
    assert (nt.value, nr.value, nc.value) == (23, f["nodestatus"].shape[1], 3)
    g = dict(treemap=np.zeros_like(f["treemap"]), nodestatus=np.zeros_like(f["nodestatus"]),
             xbestsplit=np.zeros_like(f["xbestsplit"]), bestvar=np.zeros_like(f["bestvar"]),
             nodeclass=np.zeros_like(f["nodeclass"]), ndbigtree=np.zeros_like(f["ndbigtree"]),
             orig_labels=np.zeros_like(f["orig_labels"]))
    L.glia_refrf_read(p1.encode(), C.byref(nt), C.byref(nr), C.byref(nc),
                      *[g[k].ctypes.data_as(C.c_void_p) for k in ("treemap", "nodestatus", "xbestsplit", "bestvar",
                                                                  "nodeclass", "ndbigtree", "orig_labels")])
    for k in g:
        assert np.array_equal(g[k], f[k]), k
    # the other direction: a file written by the reference's writer the way training fills the
    # Model (ml_rf_train.cxx:696-719) is read back by the oracle
    L.glia_refrf_write_native.argtypes = [C.c_char_p] + [C.c_int] * 4 + [C.c_void_p] * 7
    L.glia_refrf_write_native(p2.encode(), 23, nr.value, 3, 1,
                              *[f[k].ctypes.data_as(C.c_void_p) for k in ("treemap", "nodestatus", "xbestsplit",
                                                                          "bestvar", "nodeclass", "ndbigtree",
                                                                          "orig_labels")])
    back = ol.OracleRF.load(p2).arrays()
    for k in g:
        assert np.array_equal(back[k], f[k]), k


def test_rf_oracle_agrees_with_sklearn_rule():
    """Independent cross-check of the restated traversal (x <= t goes left, leaf class votes):
    sklearn trees use the same rule."""
    sk = pytest.importorskip("sklearn.ensemble")
    rng = np.random.default_rng(0)
    X = rng.random((400, 6)); y = np.where(X[:, 0] + 0.3 * X[:, 3] + 0.1 * rng.standard_normal(400) > 0.6, 1, -1)
    clf = sk.RandomForestClassifier(n_estimators=15, max_depth=7, random_state=0).fit(X, y)
    trees = [e.tree_ for e in clf.estimators_]
    nr = max(t.node_count for t in trees)
    f = dict(treemap=np.zeros((15, nr, 2), np.int32), nodestatus=np.zeros((15, nr), np.int32),
             xbestsplit=np.zeros((15, nr)), bestvar=np.zeros((15, nr), np.int32),
             nodeclass=np.zeros((15, nr), np.int32), ndbigtree=np.zeros(15, np.int32),
             orig_labels=np.array([-1, 1], np.int32))
    for i, t in enumerate(trees):
        n = t.node_count
        leaf = t.children_left == -1
        f["treemap"][i, :n, 0] = np.where(leaf, 0, t.children_left + 1)
        f["treemap"][i, :n, 1] = np.where(leaf, 0, t.children_right + 1)
        f["nodestatus"][i, :n] = np.where(leaf, -1, 1)
        f["xbestsplit"][i, :n] = t.threshold
        f["bestvar"][i, :n] = np.where(leaf, 0, t.feature + 1)
        f["nodeclass"][i, :n] = np.where(leaf, t.value[:, 0, :].argmax(axis=1) + 1, 0)
        f["ndbigtree"][i] = n
    orf = ol.OracleRF.from_arrays(f, 6)
    Xt = rng.random((300, 6))
    votes, _ = orf.predict(Xt)
    # sklearn compares float32(x) <= threshold: feed float32-representable inputs
    Xt = Xt.astype(np.float32).astype(np.float64)
    votes, _ = orf.predict(Xt)
    sk_votes = np.zeros_like(votes)
    for e in clf.estimators_:
        c = e.predict(Xt).astype(int)  # class index 0/1
        sk_votes[np.arange(len(Xt)), c] += 1
    assert np.array_equal(votes, sk_votes)


# ---- merge_order_pb and pre_merge
PB_CASES = [((24, 28), 40, 41, False), ((31, 22), 25, 42, True), ((9, 10, 11), 20, 43, False), ((8, 12, 10), 14, 44, True)]


@pytest.mark.parametrize("type_", [1, 2])
@pytest.mark.parametrize("case", PB_CASES)
def test_merge_order_pb_matches_reference(case, type_):
    """hmt/main_merge_order_pb.cxx:14-40 (type 1 = approximate median, 2 = mean): order and saliencies bit for bit
    (dyadic pb; tie-heavy: pb takes 256 levels, medians collide all the time)."""
    shape, nl, seed, um = case
    if not _have_ref(len(shape)):
        pytest.skip("oracle/_ref not available")
    lab, pb, mask = _make(shape, nl, seed, um)
    ref = ol.ref_run_pb(4, lab, pb, mask=mask, type=type_)
    o = ol.Oracle(lab, pb, ol.default_config(pb, use_median=(type_ == 1), contour_only=True), mask=mask)
    res = o.merge_order_pb(type_)
    assert len(ref["order"]) > 3
    assert np.array_equal(res["order"], ref["order"])
    assert np.array_equal(res["sal"], ref["sal"])


@pytest.mark.parametrize("thr", [(30,), (25, 60)])
@pytest.mark.parametrize("case", PB_CASES)
def test_pre_merge_matches_reference(case, thr):
    """gadget/main_pre_merge.cxx:7-87 (no relabel): the conditional greedy run and the renamed label image."""
    shape, nl, seed, um = case
    if not _have_ref(len(shape)):
        pytest.skip("oracle/_ref not available")
    lab, pb, mask = _make(shape, nl, seed, um)
    ref = ol.ref_run_pb(5, lab, pb, mask=mask, size_thr=thr, rpb_thr=0.4)
    o = ol.Oracle(lab, pb, ol.default_config(pb), mask=mask)
    res = o.pre_merge(thr, 0.4)
    assert np.array_equal(res["order"], ref["order"])
    assert np.array_equal(res["sal"], ref["sal"])
    assert np.array_equal(res["labels"], ref["labels"])


# ---- segment_greedy (tree resolution + final segmentation)
def _seg_inputs(shape, nl, seed, um, ntree):
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(ol.ORACLE_DIR), "oracle"))
    lab, pb, mask = _make(shape, nl, seed, um)
    rng = np.random.default_rng(seed)
    o = ol.Oracle(lab, pb, ol.default_config(pb), mask=mask)
    orders, probs, rprobs = [], [], []
    for t in range(ntree):
        # different trees over the same superpixels: mean-pb order, and greedy orders under random predictors
        if t == 0:
            o2 = ol.Oracle(lab, pb, ol.default_config(pb, contour_only=True), mask=mask)
            res = o2.merge_order_pb(2)
        else:
            res = o.greedy(fn=lambda row, t=t: float(np.floor((np.sum(np.abs(row[:9])) * (3.1 + t)) % 1.0 * 16) / 16.0))
        orders.append(res["order"])
        probs.append(np.floor(rng.random(len(res["order"])) * 8) / 8.0)       # tie-heavy probabilities
        rprobs.append(np.floor(rng.random(2 * len(res["order"]) + 1) * 4) / 4.0)
    return lab, mask, orders, probs, rprobs


@pytest.mark.parametrize("variant", ["probs", "noprobs", "regionprobs", "exact"])
@pytest.mark.parametrize("case", [((24, 28), 30, 61, False, 1), ((22, 25), 20, 62, True, 2), ((8, 10, 9), 14, 63, False, 3)])
def test_segment_greedy_matches_reference(case, variant):
    """hmt/main_segment_greedy.cxx:30-95: potentials, picks and the final label volume of the Python restatement
    (oracle/segment_oracle.py) against the compiled reference, one to three trees."""
    shape, nl, seed, um, ntree = case
    if not _have_ref(len(shape)):
        pytest.skip("oracle/_ref not available")
    lab, mask, orders, probs, rprobs = _seg_inputs(shape, nl, seed, um, ntree)
    import segment_oracle as so
    kw = dict(merge_probs=None if variant == "noprobs" else probs, region_probs=rprobs if variant == "regionprobs" else None,
              mask=mask, ignore=variant != "exact")
    want_lab, want_picks, want_pot = ol.ref_segment_greedy(lab, orders, **kw)
    got_lab, got_picks, got_pot = so.segment_greedy(lab, orders, **kw)
    assert np.array_equal(np.concatenate(got_pot), want_pot)
    assert np.array_equal(np.array(got_picks, np.int64).reshape(-1, 2), want_picks)
    assert np.array_equal(got_lab, want_lab)


# ---- bc label generation (hmt/main_bc_label_ri.cxx)
def _truth(lab, seed, ntruth):
    """A coarser segmentation: superpixels grouped at random into ntruth truth segments, 10 % of the voxels moved to
    another segment, 6 % background (0 = BG_VAL, excluded from the pair counts)."""
    rng = np.random.default_rng(seed + 500)
    group = rng.integers(1, ntruth + 1, size=int(lab.max()) + 1).astype(np.uint32)
    t = group[lab]
    noise = rng.random(lab.shape)
    t = np.where(noise < 0.10, rng.integers(1, ntruth + 1, size=lab.shape).astype(np.uint32), t)
    t[noise > 0.94] = 0
    return t.astype(np.uint32)


BCL_VARIANTS = [dict(), dict(tweak=True), dict(max_prec_drop=0.05), dict(opt_split=True), dict(use_f1=False),
                dict(global_opt=1), dict(global_opt=2), dict(tweak=True, max_prec_drop=0.2, opt_split=True)]


@pytest.mark.parametrize("kw", BCL_VARIANTS, ids=lambda k: "-".join("%s=%s" % i for i in k.items()) or "default")
@pytest.mark.parametrize("case", [((24, 28), 30, 91, False, 5), ((21, 26), 22, 92, True, 3), ((8, 10, 9), 14, 93, True, 4)])
def test_bc_label_matches_reference(case, kw):
    """Labels (and the F1 / RI scores behind them) of the Python restatement against the reference's own label templates
    (instantiated with __int128 for boost's int512_t) on a greedy merge order."""
    shape, nl, seed, um, ntruth = case
    if not _have_ref(len(shape)):
        pytest.skip("oracle/_ref not available")
    import sys
    sys.path.insert(0, ol.ORACLE_DIR)
    import bc_label_oracle as bo
    lab, pb, mask = _make(shape, nl, seed, um)
    truth = _truth(lab, seed, ntruth)
    order = ol.Oracle(lab, pb, ol.default_config(pb), mask=mask).greedy(fn=quant_pred)["order"]
    want, wm, ws = ol.ref_bc_label(lab, truth, order, mask=mask, **kw)
    got, gm, gs = bo.bc_labels(lab, truth, order, mask=mask, **kw)
    assert len(want) == len(order)
    assert np.array_equal(got, want)
    if kw.get("global_opt", 0) == 0:
        assert np.array_equal(gm, wm, equal_nan=True) and np.array_equal(gs, ws, equal_nan=True)
    if not kw:
        assert len(set(want.tolist())) == 2   # both labels occur: the test discriminates


# ---- section-classifier features (hmt/main_sc_feat.cxx)
def _sections(shape, nl, seed, um):
    """Two adjacent sections: the second is the first shifted by a voxel with some superpixels relabelled."""
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
    pairs[:10, 1] = [l if l in set(l1.tolist()) else l1[0] for l in pairs[:10, 0]]   # overlapping pairs
    return lab0, pb0, m0, lab1, pb1, m1, pairs


@pytest.mark.parametrize("kw", [dict(), dict(use_log=True), dict(norm_area=1.0, norm_len=1.0), dict(norm_area=1.0, norm_len=1.0, use_log=True)],
                         ids=["default", "logs", "unnormalised", "unnormalised-logs"])
@pytest.mark.parametrize("case", [((24, 28), 30, 121, False), ((27, 22), 18, 122, True)])
def test_sc_feat_matches_reference(case, kw):
    """hmt/main_sc_feat.cxx:27-117 through the reference's own feature classes (moments, Hu moments, eccentricity,
    overlaps, location) against the C restatement: dyadic pb, so everything but pow / log terms is bit-identical."""
    shape, nl, seed, um = case
    if not _have_ref(2):
        pytest.skip("oracle/_ref not available")
    lab0, pb0, m0, lab1, pb1, m1, pairs = _sections(shape, nl, seed, um)
    want = ol.ref_sc_feat(lab0, pb0, lab1, pb1, pairs, mask0=m0, mask1=m1, **kw)
    o0 = ol.Oracle(lab0, pb0, ol.default_config(pb0), mask=m0)
    o1 = ol.Oracle(lab1, pb1, ol.default_config(pb1), mask=m1)
    got = ol.sc_feat(o0, o1, pairs, **kw)
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-300)
    assert (got == want).mean() > 0.95
