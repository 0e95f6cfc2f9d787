"""GPU parity for the reference's two optional feature layouts (CMakeLists.txt:53-64), which are run-time parameters
here: use_hist (GLIA_USE_HISTOGRAM_AS_FEATS, type/feat.hxx:608-626) and use_median (GLIA_USE_MEDIAN_AS_FEATS,
type/feat.hxx:677-810, util/stats.hxx:83-91, hmt/bc_feat.hxx:252-268).  The oracle's restatement of both is pinned
against reference builds compiled with those definitions (tests/test_oracle_vs_ref.py)."""
import numpy as np
import pytest

import oracle_lib as ol
from glia_b200 import api, synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5  # north_star tolerance for floating-point features


def _cfgs(pb, extra, **kw):
    r, b, rl = [(pb, 16, 0.0, 1.0)], [(pb, 16, 0.0, 1.0)], []
    if extra is not None:
        r.append((extra, 7, -0.5, 2.0)); rl.append((extra, 5, 0.0, 3.0)); b.append((extra, 9, 0.1, 0.9))
    return (ol.FeatureConfig(rimgs=r, rlimgs=rl, bimgs=b, **kw), api.FeatureConfig(rimgs=r, rlimgs=rl, bimgs=b, **kw))


def _median_columns(o, cfg):
    """Indices of the median / medianDiff columns of a full row (they must be BIT-identical: order statistics)."""
    if not cfg.use_median or cfg.simple:
        return []
    D = o.labels.ndim
    T = len(cfg.thr)
    hist = cfg.use_hist
    cols, p = [], 11 + 4 * T
    for (_, bins, _, _) in cfg.rimgs:   # x0: l1, x2, entropyDiff, medianDiff, ...
        cols.append(p + 3); p += 8
    p += 3 * len(cfg.rlimgs)
    for (_, bins, _, _) in cfg.bimgs:   # x0 boundary ImageFeats: [hist] entropy median ...
        p += bins if hist else 0
        cols.append(p + 1); p += 6
    for _ in range(3):
        p += D + 4 + 2 * T
        for (_, bins, _, _) in cfg.rimgs:
            p += bins if hist else 0
            cols.append(p + 1); p += 6
        for (_, bins, _, _) in cfg.rlimgs:
            p += (bins if hist else 0) + 1
        for (_, bins, _, _) in cfg.bimgs:
            p += bins if hist else 0
            cols.append(p + 1); p += 6
    assert p == o.row_dim()
    return cols


CASES = [
    # shape, labels, seed, mask, two images, kwargs
    ((24, 28), 30, 21, False, False, dict(use_hist=True)),
    ((27, 19), 16, 22, True, True, dict(use_median=True, normalize=True, use_log=True)),
    ((22, 25), 22, 23, False, False, dict(use_median=True, simple=True)),
    ((33, 41), 40, 24, False, True, dict(use_hist=True, use_median=True)),
    ((9, 10, 11), 14, 25, True, False, dict(use_hist=True, use_median=True, use_log=True)),
    ((20, 24, 30), 200, 26, False, True, dict(use_median=True)),
]


@pytest.mark.parametrize("case", CASES)
def test_variant_rows_match_oracle(case):
    shape, nl, seed, use_mask, two, kw = case
    lab, pb = synth.speckle_volume(shape, nl, seed=seed)
    rng = np.random.default_rng(seed + 50)
    extra = (rng.integers(0, 512, size=shape) / 256.0).astype(np.float32) if two else None
    mask = (rng.random(shape) > 0.05).astype(np.uint32) if use_mask else None
    co, cg = _cfgs(pb, extra, **kw)
    o = ol.Oracle(lab, pb, co, mask=mask)
    g = api.MergeTreeBuilder(shape, cg).build(lab, pb, mask)
    assert g.row_dim() == o.row_dim()
    ko, rowo = o.init_candidates()
    kg, rowg = g.init_candidates()
    assert np.array_equal(ko, kg)
    np.testing.assert_allclose(rowg, rowo, rtol=RTOL, atol=1e-9)
    mc = _median_columns(o, co)
    if mc:
        assert np.array_equal(rowg[:, mc], rowo[:, mc])
    # the bc_feat stage on the oracle's greedy tree (quantised predictor: tie-heavy order), with saliency columns
    res = o.greedy(fn=lambda row: np.floor((np.sum(np.abs(row[:12])) * 3.7 % 1.0) * 8) / 8.0)
    if not co.simple:
        for sal in (None, res["sal"]):
            ro_ = o.bc_feat(res["order"], sal=sal, init_sal=1.0, sal_bias=0.25)
            rg_ = g.bc_feat(res["order"], sal=sal, init_sal=1.0, sal_bias=0.25)
            np.testing.assert_allclose(rg_, ro_, rtol=RTOL, atol=1e-9)
            if mc and sal is None:
                assert np.array_equal(rg_[:, mc], ro_[:, mc])


@pytest.mark.parametrize("case", CASES[1:5])
def test_variant_merge_order_matches_oracle(case):
    """Greedy chain under the variant layouts: order, saliencies and rows (forest over bit-identical columns)."""
    shape, nl, seed, use_mask, two, kw = case
    lab, pb = synth.speckle_volume(shape, nl, seed=seed)
    rng = np.random.default_rng(seed + 50)
    extra = (rng.integers(0, 512, size=shape) / 256.0).astype(np.float32) if two else None
    mask = (rng.random(shape) > 0.05).astype(np.uint32) if use_mask else None
    co, cg = _cfgs(pb, extra, **kw)
    o = ol.Oracle(lab, pb, co, mask=mask)
    g = api.MergeTreeBuilder(shape, cg).build(lab, pb, mask)
    _, rowo = o.init_candidates()
    _, rowg = g.init_candidates()
    cols = np.flatnonzero((rowo == rowg).all(axis=0))
    mc = _median_columns(o, co)
    assert set(mc) <= set(cols.tolist())
    f = synth.random_forest(9, o.row_dim(), max_depth=4, seed=seed, thr_pool=rowo, features=cols)
    orf, grf = ol.OracleRF.from_arrays(f, o.row_dim()), api.Forest.from_arrays(f, o.row_dim())
    ro = o.greedy(rf=orf)
    rg = g.merge_order_bc(grf, want_rows=True)
    assert np.array_equal(rg["order"], ro["order"])
    assert np.array_equal(rg["sal"], ro["sal"])
    np.testing.assert_allclose(rg["rows"], ro["rows"], rtol=RTOL, atol=1e-9)
    if mc:
        assert np.array_equal(rg["rows"][:, mc], ro["rows"][:, mc])


def test_median_needs_whole_volume():
    """A z-slab handle cannot hold the values of regions that cross its faces: creation is refused."""
    pb = np.zeros((8, 8, 8), np.float32)
    cfg = api.FeatureConfig(rimgs=[(pb, 16, 0.0, 1.0)], bimgs=[(pb, 16, 0.0, 1.0)], use_median=True)
    with pytest.raises(api.GliaError):
        api.MergeTreeBuilder((8, 8, 8), cfg, slab=(0, 4, 0, 1))


def test_median_variant_at_slice_size(record_property):
    """One C3-size slice (2048^2, ~25 k superpixels) and a 96^3 volume with hist + median features on: every row of the
    initial candidates against the oracle, the median columns bit for bit."""
    import ctypes as C
    import time
    import torch
    for shape, cell in (((2048, 2048), 13), ((96, 96, 96), 10)):
        L = api.lib()
        nd = len(shape)
        n = int(np.prod(shape))
        lab_d = torch.empty(n, dtype=torch.int32, device="cuda")
        pb_d = torch.empty(n, dtype=torch.float32, device="cuda")
        dims = (C.c_int * 3)(*(list(shape[::-1]) + [1] * (3 - nd)))
        nc = C.c_uint64(0)
        assert L.glia_b200_synth_voronoi(nd, dims, cell, 77, lab_d.data_ptr(), pb_d.data_ptr(), None, C.byref(nc)) == 0
        torch.cuda.synchronize()
        lab, pb = lab_d.cpu().numpy().view(np.uint32).reshape(shape), pb_d.cpu().numpy().reshape(shape)
        co, cg = _cfgs(pb, None, use_hist=True, use_median=True)
        o = ol.Oracle(lab, pb, co)
        g = api.MergeTreeBuilder(shape, cg, label_bound=int(nc.value) + 1)
        g.build(lab, pb)
        t0 = time.perf_counter()
        g.build(lab, pb)
        kg, rowg = g.init_candidates()
        dt = time.perf_counter() - t0
        ko, rowo = o.init_candidates()
        assert np.array_equal(ko, kg)
        np.testing.assert_allclose(rowg, rowo, rtol=RTOL, atol=1e-9)
        mc = _median_columns(o, co)
        assert np.array_equal(rowg[:, mc], rowo[:, mc])
        record_property("median_build_plus_rows_s_%dd" % nd, dt)
        print("median variant %s: R=%d C=%d dim=%d, build + rows (host buffers) %.3f s" % (shape, g.R, g.Cn, g.row_dim(), dt))
        g.close()
