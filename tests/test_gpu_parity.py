"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded
inputs.  Bit-exact for the RAG, integer statistics and RF votes; floating-point features within
1e-5 relative (north_star), bit-identical sums on dyadic pb."""
import numpy as np
import pytest

import oracle_lib as ol
from glia_b200 import api, synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5  # BASELINE.json north_star: floating-point features within 1e-5 relative


def _cfg_pair(pb, extra=None, **kw):
    """Same feature configuration for the oracle and the product."""
    r = [(pb, 16, 0.0, 1.0)]
    b = [(pb, 16, 0.0, 1.0)]
    rl = []
    if extra is not None:
        r.append((extra, 7, -0.5, 2.0))
        rl.append((extra, 5, 0.0, 3.0))
        b.append((extra, 9, 0.1, 0.9))
    return (ol.FeatureConfig(rimgs=r, rlimgs=rl, bimgs=b, **kw),
            api.FeatureConfig(rimgs=r, rlimgs=rl, bimgs=b, **kw))


def _compare_graph(o, g, cfg_o):
    assert (g.R, g.E, g.Cn) == (o.R, o.E, o.Cn)
    ro, rg = o.regions(), g.regions()
    nd = o.labels.ndim
    for k in ("label", "count", "border"):
        assert np.array_equal(ro[k], rg[k]), k
    assert np.array_equal(ro["bbmin"][:, :nd], rg["bbmin"][:, :nd])
    assert np.array_equal(ro["bbmax"][:, :nd], rg["bbmax"][:, :nd])
    eo, eg = o.edges(), g.edges()
    for k in ("a", "b", "count", "mutual", "thr"):
        assert np.array_equal(eo[k], eg[k]), k
    for kind, lst in ((0, cfg_o.rimgs), (1, cfg_o.rlimgs)):
        for i in range(len(lst)):
            so, sg = o.region_stats(i, kind), g.region_stats(i, kind)
            assert np.array_equal(so["hist"], sg["hist"])
            assert np.array_equal(so["sum"], sg["sum"]) and np.array_equal(so["sum2"], sg["sum2"])
            assert np.array_equal(so["min"], sg["min"]) and np.array_equal(so["max"], sg["max"])
    for i in range(len(cfg_o.bimgs)):
        so, sg = o.edge_stats(i), g.edge_stats(i)
        assert np.array_equal(so["hist"], sg["hist"])
        assert np.array_equal(so["sum"], sg["sum"]) and np.array_equal(so["sum2"], sg["sum2"])
        assert np.array_equal(so["min"], sg["min"]) and np.array_equal(so["max"], sg["max"])


CASES = [
    ((24, 28), 30, 1, False, {}),
    ((31, 17), 12, 2, True, dict(normalize=True, use_log=True)),
    ((20, 22), 25, 3, False, dict(simple=True)),
    ((9, 10, 11), 14, 4, False, {}),
    ((8, 12, 10), 10, 5, True, dict(normalize=True, use_log=True)),
    ((1, 40), 6, 6, False, {}),       # single row
    ((40, 1), 6, 7, True, {}),        # single column
    ((3, 1, 5), 4, 8, False, {}),     # 3D with a unit axis: every voxel is border or boundary
    ((70, 300), 400, 9, False, {}),   # wider than one tile, many labels per tile (table overflow)
    ((20, 40, 70), 3000, 10, True, {}),
]


@pytest.mark.parametrize("case", CASES)
def test_graph_and_rows_match_oracle(case):
    shape, nl, seed, use_mask, kw = case
    lab, pb = synth.speckle_volume(shape, nl, seed=seed)
    rng = np.random.default_rng(seed + 50)
    extra = (rng.integers(0, 512, size=shape) / 256.0).astype(np.float32) if seed % 2 == 0 else None
    mask = (rng.random(shape) > 0.05).astype(np.uint32) if use_mask else None
    co, cg = _cfg_pair(pb, extra, **kw)
    o = ol.Oracle(lab, pb, co, mask=mask)
    g = api.MergeTreeBuilder(shape, cg).build(lab, pb, mask)
    _compare_graph(o, g, co)
    ko, rowo = o.init_candidates()
    kg, rowg = g.init_candidates()
    assert np.array_equal(ko, kg)
    assert rowo.shape == rowg.shape
    np.testing.assert_allclose(rowg, rowo, rtol=RTOL, atol=1e-12)
    # everything except the transcendental terms (log2/log/pow) is bit-identical on dyadic input
    same = rowg == rowo
    assert same.mean() > 0.9


def test_voronoi_c1_matches_oracle():
    """configs[0]: 2D 512x512, ~2k superpixels."""
    lab, pb = synth.voronoi_volume((512, 512), 11, seed=1234)
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co)
    g = api.MergeTreeBuilder((512, 512), cg).build(lab, pb)
    _compare_graph(o, g, co)
    ko, rowo = o.init_candidates()
    kg, rowg = g.init_candidates()
    assert np.array_equal(ko, kg)
    np.testing.assert_allclose(rowg, rowo, rtol=RTOL, atol=1e-12)


def test_voronoi_3d_matches_oracle():
    lab, pb = synth.voronoi_volume((48, 64, 80), 9, seed=5)
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co)
    g = api.MergeTreeBuilder((48, 64, 80), cg, label_bound=int(lab.max()) + 1).build(lab, pb)
    _compare_graph(o, g, co)
    ko, rowo = o.init_candidates()
    kg, rowg = g.init_candidates()
    assert np.array_equal(ko, kg)
    np.testing.assert_allclose(rowg, rowo, rtol=RTOL, atol=1e-12)


def test_nondyadic_within_tolerance():
    lab, pb = synth.speckle_volume((40, 50), 60, seed=11, dyadic=False)
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co)
    g = api.MergeTreeBuilder((40, 50), cg).build(lab, pb)
    eo, eg = o.edges(), g.edges()
    for k in ("a", "b", "count", "mutual", "thr"):
        assert np.array_equal(eo[k], eg[k])
    assert np.array_equal(o.region_stats(0)["hist"], g.region_stats(0)["hist"])
    np.testing.assert_allclose(g.region_stats(0)["sum"], o.region_stats(0)["sum"], rtol=1e-12)
    _, rowo = o.init_candidates()
    _, rowg = g.init_candidates()
    np.testing.assert_allclose(rowg, rowo, rtol=RTOL, atol=1e-9)


def test_hist_bounds_quirk():
    """util/image_stats.hxx:17-36: bounds are cumulative from `interval`, not offset by lo; a value in
    (lo, hi) at or beyond the last bound is counted in n but in no bin."""
    shape = (16, 16)
    lab = np.ones(shape, np.uint32); lab[:, 8:] = 2
    pb = np.linspace(0.0, 1.0, 256, dtype=np.float32).reshape(shape)
    r = [(pb, 4, 0.5, 1.0)]  # interval .125 -> bounds .125 .25 .375 .5: values in (.5,1) drop out
    co = ol.FeatureConfig(rimgs=r, bimgs=r)
    cg = api.FeatureConfig(rimgs=r, bimgs=r)
    o = ol.Oracle(lab, pb, co)
    g = api.MergeTreeBuilder(shape, cg).build(lab, pb)
    ho, hg = o.region_stats(0)["hist"], g.region_stats(0)["hist"]
    assert np.array_equal(ho, hg)
    assert hg.sum() < 256  # some voxels fell in no bin
    assert np.array_equal(o.edge_stats(0)["hist"], g.edge_stats(0)["hist"])


def test_rf_votes_bit_exact():
    rng = np.random.default_rng(3)
    D = 50
    X = rng.random((20000, D))
    f = synth.random_forest(255, D, max_depth=14, seed=1, thr_pool=X[:500])
    X[:1000] = X[rng.integers(0, 500, 1000)]  # rows sitting exactly on thresholds (x <= t ties)
    orf = ol.OracleRF.from_arrays(f, D)
    grf = api.Forest.from_arrays(f, D)
    vo, po = orf.predict(X)
    vg, pg = grf.predict(X)
    assert np.array_equal(vo, vg)
    assert np.array_equal(po, pg)
    assert vg.sum(axis=1).min() == 255 and vg.sum(axis=1).max() == 255


def test_rf_probability_kernel_nan_inf_and_stump_trees():
    """The probability-only path (rows staged in shared memory, walk-encoded nodes; the one the handle's
    score_candidates and the C5 sweep use) against the oracle: rows with NaN / +-inf (NaN compares false, so
    it goes right: alg/rf.hxx:35-36 -> classForest), rows on thresholds, trees that are a single leaf, and
    row counts that do not fill the last block of 64."""
    rng = np.random.default_rng(11)
    D = 23
    for n, depth in ((1, 9), (63, 0), (64, 5), (4097, 16)):
        X = rng.random((n, D))
        f = synth.random_forest(37, D, max_depth=depth, seed=n, thr_pool=X[: max(1, n // 2)], leaf_prob=0.2)
        X[rng.random((n, D)) < 0.05] = np.nan
        X[rng.random((n, D)) < 0.02] = np.inf
        X[rng.random((n, D)) < 0.02] = -np.inf
        orf = ol.OracleRF.from_arrays(f, D)
        grf = api.Forest.from_arrays(f, D)
        vo, po = orf.predict(X)
        _, pg = grf.predict(X, want_votes=False)
        vg, pg2 = grf.predict(X)
        assert np.array_equal(po, pg), (n, depth)
        assert np.array_equal(vo, vg) and np.array_equal(po, pg2)


def test_rf_model_file_roundtrip(tmp_path):
    f = synth.random_forest(17, 9, max_depth=7, seed=2, nclass=3, orig_labels=(-1, 1, 4))
    orf = ol.OracleRF.from_arrays(f, 9)
    path = str(tmp_path / "model.bin")
    assert orf.save(path) == 0
    grf = api.Forest.load(path)
    assert (grf.ntree, grf.nclass) == (17, 3)
    X = np.random.default_rng(0).random((3000, 9))
    for label in (-1, 1, 4):
        vo, po = orf.predict(X, label)
        vg, pg = grf.predict(X, label)
        assert np.array_equal(vo, vg) and np.array_equal(po, pg)
    with pytest.raises(api.GliaError):
        grf.predict(X, label=7)  # unknown label: the reference perr()s (ml/rf/rf.hxx:370)


def test_score_candidates_matches_oracle():
    lab, pb = synth.voronoi_volume((96, 128), 8, seed=3)
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co)
    g = api.MergeTreeBuilder((96, 128), cg).build(lab, pb)
    _, rowo = o.init_candidates()
    _, rowg = g.init_candidates()
    # thresholds from ANOTHER image's rows: realistic splits without forcing x == t ties
    lab2, pb2 = synth.voronoi_volume((80, 80), 7, seed=8)
    o2 = ol.Oracle(lab2, pb2, _cfg_pair(pb2)[0])
    _, pool = o2.init_candidates()
    f = synth.random_forest(255, o.row_dim(), max_depth=12, seed=4, thr_pool=pool)
    orf, grf = ol.OracleRF.from_arrays(f, o.row_dim()), api.Forest.from_arrays(f, o.row_dim())
    sal = g.score_candidates(grf)
    # (1) votes are bit-exact on identical input rows
    _, p_same = orf.predict(rowg)
    assert np.array_equal(sal, p_same)
    # (2) against the oracle end to end, the only differences are RF decision flips where a
    # transcendental feature (entropy / compactness, last-bit differences) sits on a threshold:
    # enumerate them
    _, po = orf.predict(rowo)
    flips = np.flatnonzero(sal != po)
    assert len(flips) <= 0.02 * len(po), "flips: %s" % flips[:20]
    for c in flips:
        cols = np.flatnonzero(rowg[c] != rowo[c])
        assert len(cols) > 0
        np.testing.assert_allclose(rowg[c], rowo[c], rtol=RTOL, atol=1e-12)


# ---------------------------------------------------------------- bc_feat stage / merge order
def quant_pred(row):
    s = float(np.sum(np.abs(row[:20]))) * 3.7
    return np.floor((s - np.floor(s)) * 8) / 8.0


TREE_CASES = [
    ((24, 28), 30, 1, False, {}),
    ((31, 17), 12, 2, True, dict(normalize=True, use_log=True)),
    ((9, 10, 11), 14, 4, False, {}),
    ((8, 12, 10), 10, 5, True, dict(normalize=True, use_log=True)),
    ((40, 60), 150, 12, False, dict(simple=True)),
]


@pytest.mark.parametrize("case", TREE_CASES)
def test_bc_feat_stage_matches_oracle(case):
    """hmt/main_bc_feat.cxx: rows for every merge of a given order (with and without saliency)."""
    shape, nl, seed, use_mask, kw = case
    lab, pb = synth.speckle_volume(shape, nl, seed=seed)
    mask = (np.random.default_rng(seed + 50).random(shape) > 0.05).astype(np.uint32) if use_mask else None
    co, cg = _cfg_pair(pb, **kw)
    o = ol.Oracle(lab, pb, co, mask=mask)
    res = o.greedy(fn=quant_pred)  # a tie-heavy order: unbalanced trees, non-mutual boundaries inside
    g = api.MergeTreeBuilder(shape, cg).build(lab, pb, mask)
    for sal in (None, res["sal"]):
        ro = o.bc_feat(res["order"], sal=sal, init_sal=1.0, sal_bias=0.25)
        rg = g.bc_feat(res["order"], sal=sal, init_sal=1.0, sal_bias=0.25)
        assert ro.shape == rg.shape
        np.testing.assert_allclose(rg, ro, rtol=RTOL, atol=1e-12)
    # a partial order (prefix) is a forest, not a tree
    k = len(res["order"]) // 2
    np.testing.assert_allclose(g.bc_feat(res["order"][:k]), o.bc_feat(res["order"][:k]), rtol=RTOL, atol=1e-12)
    with pytest.raises(api.GliaError):
        bad = res["order"].copy(); bad[0, 0] = 4000000
        g.bc_feat(bad)


def _exact_columns(rowo, rowg):
    return np.flatnonzero((rowo == rowg).all(axis=0))


@pytest.mark.parametrize("case", TREE_CASES)
def test_merge_order_bc_matches_oracle(case):
    """hmt/main_merge_order_bc.cxx: full greedy chain; order, saliencies and -b rows."""
    shape, nl, seed, use_mask, kw = case
    lab, pb = synth.speckle_volume(shape, nl, seed=seed)
    mask = (np.random.default_rng(seed + 50).random(shape) > 0.05).astype(np.uint32) if use_mask else None
    co, cg = _cfg_pair(pb, **kw)
    o = ol.Oracle(lab, pb, co, mask=mask)
    g = api.MergeTreeBuilder(shape, cg).build(lab, pb, mask)
    _, rowo = o.init_candidates()
    _, rowg = g.init_candidates()
    # forest over the columns that are bit-identical (everything but log/log2/pow terms), few
    # trees and shallow => heavy saliency ties: the queue's tie-break order is exercised
    cols = _exact_columns(rowo, rowg)
    assert len(cols) >= 5
    f = synth.random_forest(9, o.row_dim(), max_depth=4, seed=seed, thr_pool=rowo, features=cols)
    orf, grf = ol.OracleRF.from_arrays(f, o.row_dim()), api.Forest.from_arrays(f, o.row_dim())
    ro = o.greedy(rf=orf)
    rg = g.merge_order_bc(grf, want_rows=True)
    assert np.array_equal(rg["order"], ro["order"])
    assert np.array_equal(rg["sal"], ro["sal"])
    assert rg["n_featurized"] == ro["n_cand"]
    np.testing.assert_allclose(rg["rows"], ro["rows"], rtol=RTOL, atol=1e-12)


@pytest.mark.parametrize("shape,cell", [((512, 512), 11), ((40, 64, 64), 8)])
def test_greedy_with_unrestricted_forest_first_divergence(shape, cell, record_property):
    """Greedy chain scored by a forest over ALL columns (entropy, compactness, log terms included), depth 12.  The
    transcendental columns agree to ~1e-15, not bit for bit, so a score may flip where a threshold falls between the two
    values.  The run must equal the oracle's up to the first such flip, and that flip must be exactly that: the GPU score
    is the oracle forest's score of the GPU row, the two rows agree within 1e-5, and a threshold of the forest separates
    them in a differing column.  The index of the first divergence is reported (len == no divergence)."""
    lab, pb = synth.voronoi_volume(shape, cell, seed=1234)
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co)
    g = api.MergeTreeBuilder(shape, cg).build(lab, pb)
    _, rowo = o.init_candidates()
    f = synth.random_forest(63, o.row_dim(), max_depth=12, seed=7, thr_pool=rowo, leaf_prob=0.1)
    orf, grf = ol.OracleRF.from_arrays(f, o.row_dim()), api.Forest.from_arrays(f, o.row_dim())
    ro = o.greedy(rf=orf, want_cands=True)
    rg = g.merge_order_bc(grf, want_rows=True)
    m = min(len(ro["order"]), len(rg["order"]))
    same = np.all(ro["order"][:m] == rg["order"][:m], axis=1) & (ro["sal"][:m] == rg["sal"][:m])
    n = m if same.all() else int(np.argmin(same))
    record_property("first_divergent_merge", n)
    record_property("merges", m)
    print("first divergent merge: %d of %d" % (n, m))
    np.testing.assert_allclose(rg["rows"][:n], ro["rows"][:n], rtol=RTOL, atol=1e-12)
    if n == m:
        assert len(ro["order"]) == len(rg["order"])
        return
    # Both chains hold the same candidates at step n (they were identical so far) but popped different ones: one of the
    # two popped candidates must carry a different score in the two runs.  Candidate rows are fixed at creation, so the
    # GPU's row of the oracle's choice is the row of that pair when the GPU merges it later.
    thr = np.unique(f["xbestsplit"][f["nodestatus"] != -1])
    ck = np.sort(ro["cand_keys"][:, :2].astype(np.int64), axis=1)

    def explained(key, row_g, sal_g):
        hit = np.flatnonzero((ck[:, 0] == key[0]) & (ck[:, 1] == key[1]))
        if not len(hit):
            return False
        row_o, sal_o = ro["cand_rows"][hit[-1]], ro["cand_scores"][hit[-1]]
        np.testing.assert_allclose(row_g, row_o, rtol=RTOL, atol=1e-12)
        _, p = orf.predict(row_g[None, :])
        assert p[0] == sal_g  # votes are bit-exact on identical rows
        if sal_g == sal_o:
            return False
        d = np.flatnonzero(row_g != row_o)
        lo, hi = np.minimum(row_g[d], row_o[d]), np.maximum(row_g[d], row_o[d])
        i = np.searchsorted(thr, lo, side="left")
        assert np.any((i < len(thr)) & (thr[np.minimum(i, len(thr) - 1)] <= hi))  # a threshold separates the two values
        return True

    key_g = tuple(sorted(int(x) for x in rg["order"][n][:2]))
    ok = explained(key_g, rg["rows"][n], rg["sal"][n])
    if not ok:
        key_o = tuple(sorted(int(x) for x in ro["order"][n][:2]))
        later = [j for j in range(n + 1, len(rg["order"])) if tuple(sorted(int(x) for x in rg["order"][j][:2])) == key_o]
        ok = bool(later) and explained(key_o, rg["rows"][later[0]], rg["sal"][later[0]])
    assert ok, "first divergence at merge %d is not explained by a decision flip of either popped candidate" % n


def test_merge_order_c1_voronoi():
    """configs[0]: 512x512, ~2.2k superpixels: the full greedy run the reference does on CPU."""
    lab, pb = synth.voronoi_volume((512, 512), 11, seed=1234)
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co)
    g = api.MergeTreeBuilder((512, 512), cg).build(lab, pb)
    _, rowo = o.init_candidates()
    _, rowg = g.init_candidates()
    cols = _exact_columns(rowo, rowg)
    f = synth.random_forest(31, o.row_dim(), max_depth=6, seed=1, thr_pool=rowo, features=cols)
    orf, grf = ol.OracleRF.from_arrays(f, o.row_dim()), api.Forest.from_arrays(f, o.row_dim())
    ro = o.greedy(rf=orf)
    rg = g.merge_order_bc(grf)
    assert len(rg["order"]) == g.R - 1
    assert np.array_equal(rg["order"], ro["order"])
    assert np.array_equal(rg["sal"], ro["sal"])


def test_generic_and_fast_voxel_kernels_agree():
    """The default configuration takes the fast kernel (voxel_fast.cu); forcing the generic one
    (voxel_pass.cu) must give identical records."""
    lab, pb = synth.voronoi_volume((40, 70, 90), 9, seed=21)
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co)
    g1 = api.MergeTreeBuilder(lab.shape, cg).build(lab, pb)
    g2 = api.MergeTreeBuilder(lab.shape, cg).tune("generic", 1).build(lab, pb)
    for g in (g1, g2):
        _compare_graph(o, g, co)


@pytest.mark.parametrize("shape,masked", [((24, 40, 64), True), ((24, 40, 64), False), ((70, 96), True), ((20, 30, 50), True)])
def test_generic_kernel_tma_staging_equals_plain_loads(shape, masked):
    """voxel_pass_generic stages its label (+ mask) tile with one TMA tensor copy per tile when the row length is a
    multiple of 4 voxels (16-byte strides); otherwise, or with tune("notma"), with plain loads.  Same records either way,
    and equal to the oracle."""
    lab, pb = synth.voronoi_volume(shape, 7, seed=31)
    mask = (np.random.default_rng(5).random(shape) > 0.05).astype(np.uint32) if masked else None
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co, mask=mask)
    g1 = api.MergeTreeBuilder(shape, cg).tune("generic", 1).build(lab, pb, mask)
    aligned = shape[-1] % 4 == 0
    if aligned:
        g1.tune("used_tma", 0)  # raises GLIA_E_STATE if the pass did not go through TMA
    else:
        with pytest.raises(api.GliaError):
            g1.tune("used_tma", 0)
    g2 = api.MergeTreeBuilder(shape, cg).tune("generic", 1).tune("notma", 1).build(lab, pb, mask)
    with pytest.raises(api.GliaError):
        g2.tune("used_tma", 0)
    for g in (g1, g2):
        _compare_graph(o, g, co)


def _same_graph(g1, g2):
    assert (g1.R, g1.E, g1.Cn) == (g2.R, g2.E, g2.Cn)
    r1, r2 = g1.regions(), g2.regions()
    for k in r1:
        assert np.array_equal(r1[k], r2[k]), k
    e1, e2 = g1.edges(), g2.edges()
    for k in e1:
        assert np.array_equal(e1[k], e2[k]), k
    for f in ("region_stats", "edge_stats"):
        s1, s2 = getattr(g1, f)(0), getattr(g2, f)(0)
        for k in s1:
            assert np.array_equal(s1[k], s2[k]), (f, k)


@pytest.mark.parametrize("shape,cell,masked", [((2080, 2048), 13, False), ((72, 250, 260), 9, False),
                                               ((2051, 2048), 17, True)])
def test_chunked_upload_equals_single_copy(shape, cell, masked):
    """Host inputs >= 16 MB with a known label bound are uploaded in z-chunks that overlap the voxel
    pass (api_build.cu); the result must equal the single-copy path (label_bound = 0 scans the labels
    first, which disables chunking) bit for bit, for the fast and the generic (mask) kernels."""
    lab, pb = synth.voronoi_volume(shape, cell, seed=21)
    mask = (np.random.RandomState(1).rand(*shape) > 0.02).astype(np.uint32) if masked else None
    cfg = api.FeatureConfig(rimgs=[(pb, 16, 0.0, 1.0)], bimgs=[(pb, 16, 0.0, 1.0)])
    g1 = api.MergeTreeBuilder(shape, cfg).build(lab, pb, mask)
    g2 = api.MergeTreeBuilder(shape, cfg, label_bound=int(lab.max()) + 1).build(lab, pb, mask)
    _same_graph(g1, g2)
    k1, rows1 = g1.init_candidates()
    k2, rows2 = g2.init_candidates()
    assert np.array_equal(k1, k2) and np.array_equal(rows1, rows2)


def test_set_image_rebinds_inputs():
    """One handle reused for a stream of slices (glia_b200_set_image) == a fresh handle per slice."""
    shape = (96, 128)
    la, pa = synth.voronoi_volume(shape, 9, seed=1)
    lb_, pb_ = synth.voronoi_volume(shape, 9, seed=2)
    cfg = lambda p: api.FeatureConfig(rimgs=[(p, 16, 0.0, 1.0)], bimgs=[(p, 16, 0.0, 1.0)])
    g = api.MergeTreeBuilder(shape, cfg(pa)).build(la, pa)
    g.set_image(0, 0, pb_); g.set_image(2, 0, pb_)
    g.build(lb_, pb_)
    _same_graph(g, api.MergeTreeBuilder(shape, cfg(pb_)).build(lb_, pb_))


def test_rf_block_kernel_is_exact_on_near_ties():
    """The prob-only RF kernel walks two tree levels per 32-byte block (rf.cu v5, exact double thresholds):
    rows that sit exactly on thresholds, one double ulp and 1e-9 relative beside them must score exactly
    as the oracle (and as the per-node kernel that also returns the votes)."""
    rng = np.random.default_rng(3)
    d = 37
    pool = rng.random((512, d)) * rng.choice([1e-3, 1.0, 1e6], size=(1, d))
    f = synth.random_forest(64, d, max_depth=14, seed=5, thr_pool=pool, leaf_prob=0.2)
    base = pool[rng.integers(0, len(pool), 6000)]
    X = np.concatenate([base, np.nextafter(base, np.inf), np.nextafter(base, -np.inf),
                        base * (1 + 1e-9), base * (1 - 1e-9), rng.random((3000, d))])
    orf, grf = ol.OracleRF.from_arrays(f, d), api.Forest.from_arrays(f, d)
    vo, po = orf.predict(X)
    _, pg3 = grf.predict(X, want_votes=False)  # default prob-only kernel (two-level blocks)
    assert np.array_equal(po, pg3)
    vg, pg2 = grf.predict(X, want_votes=True)  # one 16-byte node per level
    assert np.array_equal(po, pg2) and np.array_equal(vo, vg)


def _ensemble_pick(x, dim0, dim1, thr):
    """opt::ThresholdModelDistributor (type/function.hxx:71-85), restated for the checker."""
    return 0 if x[dim1] < thr else (1 if x[dim0] < thr else 2)


def test_rf_ensemble_matches_reference_rule():
    """alg::EnsembleRandomForest (alg/rf.hxx:63-98): per-row model choice, then that model's votes."""
    rng = np.random.default_rng(11)
    d, n = 23, 5000
    X = rng.random((n, d))
    X[:200, 4] = 0.5; X[200:400, 9] = 0.5  # rows exactly on the distributor threshold (strict <)
    fs = [synth.random_forest(nt, d, max_depth=9, seed=s, thr_pool=X[:300]) for nt, s in ((17, 1), (31, 2), (8, 3))]
    ofs = [ol.OracleRF.from_arrays(f, d) for f in fs]
    gfs = [api.Forest.from_arrays(f, d) for f in fs]
    ens = api.Forest.ensemble(gfs, 4, 9, 0.5)
    pick = np.array([_ensemble_pick(x, 4, 9, 0.5) for x in X])
    assert set(pick) == {0, 1, 2}
    votes = np.stack([o.predict(X)[0] for o in ofs]); probs = np.stack([o.predict(X)[1] for o in ofs])
    vo, po = votes[pick, np.arange(n)], probs[pick, np.arange(n)]
    vg, pg = ens.predict(X)
    _, pg2 = ens.predict(X, want_votes=False)
    assert np.array_equal(vo, vg) and np.array_equal(po, pg) and np.array_equal(po, pg2)
    two = api.Forest.ensemble(gfs[:2], 4, 9, 0.5)  # index 2 does not exist: clamped to the last model
    _, p2 = two.predict(X, want_votes=False)
    assert np.array_equal(p2, probs[np.minimum(pick, 1), np.arange(n)])


def test_merge_order_with_ensemble_matches_oracle():
    """--bcm m0 m1 m2 --bcmd dim0 dim1 thr (hmt/main_merge_order_bc.cxx:102-109) through the greedy chain."""
    shape = (26, 30)
    lab, pb = synth.speckle_volume(shape, 40, seed=8)
    co, cg = _cfg_pair(pb)
    o = ol.Oracle(lab, pb, co)
    g = api.MergeTreeBuilder(shape, cg).build(lab, pb)
    _, rowo = o.init_candidates()
    _, rowg = g.init_candidates()
    cols = _exact_columns(rowo, rowg)
    dim = o.row_dim()
    fs = [synth.random_forest(7, dim, max_depth=4, seed=s, thr_pool=rowo, features=cols) for s in (1, 2, 3)]
    ofs = [ol.OracleRF.from_arrays(f, dim) for f in fs]
    d0, d1 = int(cols[0]), int(cols[1])
    thr = float(np.median(rowo[:, d1]))
    ens = api.Forest.ensemble([api.Forest.from_arrays(f, dim) for f in fs], d0, d1, thr)
    fn = lambda x: float(ofs[_ensemble_pick(x, d0, d1, thr)].predict(x[None, :])[1][0])
    ro = o.greedy(fn=fn)
    rg = g.merge_order_bc(ens, want_rows=True)
    assert np.array_equal(rg["order"], ro["order"]) and np.array_equal(rg["sal"], ro["sal"])
    sal = g.score_candidates(ens)
    assert np.array_equal(sal, np.array([fn(r) for r in rowg]))
