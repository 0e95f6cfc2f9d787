"""GPU parity for SURVEY §8f rank 1: merge_order_pb (hmt/main_merge_order_pb.cxx) and pre_merge
(gadget/main_pre_merge.cxx) through the C ABI against the oracle, which is pinned to the compiled reference
(tests/test_oracle_vs_ref.py::test_merge_order_pb_matches_reference, ::test_pre_merge_matches_reference)."""
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
from glia_b200 import api, synth

pytestmark = pytest.mark.gpu
BIN = os.path.join(os.path.dirname(api.HERE), "glia_b200", "bin")

CASES = [((24, 28), 40, 41, False), ((31, 22), 25, 42, True), ((9, 10, 11), 20, 43, False), ((8, 12, 10), 14, 44, True),
         ((64, 80), 300, 45, False), ((20, 24, 28), 250, 46, True)]


def _make(shape, nl, seed, um):
    lab, pb = synth.speckle_volume(shape, nl, seed=seed)
    mask = (np.random.default_rng(seed + 100).random(shape) > 0.04).astype(np.uint32) if um else None
    return lab, pb, mask


@pytest.mark.parametrize("type_", [1, 2])
@pytest.mark.parametrize("case", CASES)
def test_merge_order_pb_matches_oracle(case, type_):
    """Order and saliencies bit for bit: dyadic pb makes the sums exact, medians are order statistics; pb has 256
    levels, so saliency ties (and with them the queue's tie-break rule) occur all the time."""
    shape, nl, seed, um = case
    lab, pb, mask = _make(shape, nl, seed, um)
    kw = dict(use_median=(type_ == 1), contour_only=True)
    spec = [(pb, 1, 0.0, 1.0)]
    o = ol.Oracle(lab, pb, ol.FeatureConfig(bimgs=spec, thr=(), **kw), mask=mask)
    g = api.MergeTreeBuilder(shape, api.FeatureConfig(bimgs=spec, thr=(), **kw)).build(lab, pb, mask)
    ro, rg = o.merge_order_pb(type_), g.merge_order_pb(type_)
    assert len(ro["order"]) > 3
    assert np.array_equal(rg["order"], ro["order"])
    assert np.array_equal(rg["sal"], ro["sal"])


@pytest.mark.parametrize("thr", [(30,), (25, 60)])
@pytest.mark.parametrize("case", CASES)
def test_pre_merge_matches_oracle(case, thr):
    shape, nl, seed, um = case
    lab, pb, mask = _make(shape, nl, seed, um)
    spec = [(pb, 1, 0.0, 1.0)]
    o = ol.Oracle(lab, pb, ol.FeatureConfig(rimgs=spec, bimgs=spec, thr=()), mask=mask)
    g = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec, thr=())).build(lab, pb, mask)
    ro, rg = o.pre_merge(thr, 0.4), g.pre_merge(thr, 0.4)
    assert np.array_equal(rg["order"], ro["order"])
    assert np.array_equal(rg["sal"], ro["sal"])
    assert np.array_equal(rg["labels"], ro["labels"])
    # every merged region got one key; untouched regions keep theirs
    merged = set(ro["order"][:, :2].ravel().tolist())
    keep = ~np.isin(lab, list(merged))
    assert np.array_equal(rg["labels"][keep], lab[keep])


def test_merge_order_pb_nondyadic_saliencies_within_tolerance():
    """Non-dyadic pb: the mean saliencies agree to 1e-12 (the reference sums floats into a double pixel by pixel, the
    voxel pass sums exactly); the median run stays bit-identical."""
    lab, pb = synth.speckle_volume((40, 50), 60, seed=11, dyadic=False)
    spec = [(pb, 1, 0.0, 1.0)]
    o = ol.Oracle(lab, pb, ol.FeatureConfig(bimgs=spec, thr=(), use_median=True, contour_only=True))
    g = api.MergeTreeBuilder((40, 50), api.FeatureConfig(bimgs=spec, thr=(), use_median=True, contour_only=True)).build(lab, pb)
    ro, rg = o.merge_order_pb(1), g.merge_order_pb(1)
    assert np.array_equal(rg["order"], ro["order"]) and np.array_equal(rg["sal"], ro["sal"])
    ro, rg = o.merge_order_pb(2), g.merge_order_pb(2)
    n = min(len(ro["sal"]), len(rg["sal"]))
    same = np.all(ro["order"][:n] == rg["order"][:n], axis=1)
    k = n if same.all() else int(np.argmin(same))  # a 1-ulp difference may reorder a near-tie: compare up to there
    assert k >= 1
    np.testing.assert_allclose(rg["sal"][:k], ro["sal"][:k], rtol=1e-12)


def _write_mha(path, arr, met_type):
    dims = " ".join(str(s) for s in arr.shape[::-1])
    hdr = ("ObjectType = Image\nNDims = %d\nBinaryData = True\nBinaryDataByteOrderMSB = False\n"
           "CompressedData = False\nDimSize = %s\nElementType = %s\nElementDataFile = LOCAL\n" % (arr.ndim, dims, met_type))
    with open(path, "wb") as f:
        f.write(hdr.encode())
        f.write(np.ascontiguousarray(arr).tobytes())


def test_cli_merge_order_pb_and_pre_merge(tmp_path):
    shape = (36, 48)
    lab, pb, mask = _make(shape, 70, 47, True)
    seg_f, pb_f, mask_f = str(tmp_path / "seg.npy"), str(tmp_path / "pb.mha"), str(tmp_path / "mask.npy")
    np.save(seg_f, lab); np.save(mask_f, mask.astype(np.uint8)); _write_mha(pb_f, pb, "MET_FLOAT")
    spec = [(pb, 1, 0.0, 1.0)]
    for t in (1, 2):
        o = ol.Oracle(lab, pb, ol.FeatureConfig(bimgs=spec, thr=(), use_median=(t == 1), contour_only=True), mask=mask)
        want = o.merge_order_pb(t)
        order_f, sal_f = str(tmp_path / ("order%d.txt" % t)), str(tmp_path / ("sal%d.txt" % t))
        r = subprocess.run([os.path.join(BIN, "merge_order_pb"), "-s", seg_f, "-p", pb_f, "-m", mask_f, "-t", str(t),
                            "-o", order_f, "-y", sal_f], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert np.array_equal(np.loadtxt(order_f, dtype=np.int64).reshape(-1, 3), want["order"])
        np.testing.assert_allclose(np.loadtxt(sal_f), want["sal"], rtol=1e-5, atol=1e-12)
    o = ol.Oracle(lab, pb, ol.FeatureConfig(rimgs=spec, bimgs=spec, thr=()), mask=mask)
    want = o.pre_merge((25, 60), 0.4)
    for out_f, dt in ((str(tmp_path / "out.npy"), np.uint32), (str(tmp_path / "out.mha"), np.uint16)):
        cmd = [os.path.join(BIN, "pre_merge"), "-s", seg_f, "-p", pb_f, "-m", mask_f, "-t", "25", "60", "-b", "0.4", "-o", out_f]
        if dt == np.uint16:
            cmd += ["-u", "1"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        if out_f.endswith(".npy"):
            got = np.load(out_f)
        else:
            raw = open(out_f, "rb").read()
            got = np.frombuffer(raw[raw.index(b"ElementDataFile = LOCAL\n") + 24:], dt).reshape(shape)
        assert got.dtype == dt and np.array_equal(got.astype(np.uint32), want["labels"])


# ---------------------------------------------------------------- segment_greedy
def _trees(lab, pb, mask, ntree, seed):
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
    return orders, probs, rprobs


@pytest.mark.parametrize("case", [((30, 36), 60, 81, False, 1), ((28, 31), 40, 82, True, 2), ((9, 12, 11), 30, 83, True, 3)])
def test_segment_greedy_matches_oracle(case):
    """hmt/main_segment_greedy.cxx: picks and the final label volume against the Python restatement (pinned to the
    compiled reference), with / without merge and region probabilities, ignore on / off, one to three trees."""
    import sys
    sys.path.insert(0, ol.ORACLE_DIR)
    import segment_oracle as so
    shape, nl, seed, um, ntree = case
    lab, pb, mask = _make(shape, nl, seed, um)
    orders, probs, rprobs = _trees(lab, pb, mask, ntree, seed)
    g = api.MergeTreeBuilder(shape, api.FeatureConfig(thr=())).build(lab, np.zeros(shape, np.float32), mask)
    for mp, rp, ign in ((probs, None, True), (None, None, True), (probs, rprobs, True), (probs, None, False)):
        want_l, want_pk, _ = so.segment_greedy(lab, orders, merge_probs=mp, region_probs=rp, mask=mask, ignore=ign)
        got_l, got_pk = g.segment_greedy(orders, merge_probs=mp, region_probs=rp, ignore_missing=ign)
        assert np.array_equal(got_pk.astype(np.int64), np.array(want_pk, np.int64).reshape(-1, 2))
        assert np.array_equal(got_l, want_l)


def test_cli_segment_greedy(tmp_path):
    import sys
    sys.path.insert(0, ol.ORACLE_DIR)
    import segment_oracle as so
    shape = (32, 40)
    lab, pb, mask = _make(shape, 50, 84, True)
    orders, probs, _ = _trees(lab, pb, mask, 2, 84)
    seg_f, mask_f, out_f = str(tmp_path / "seg.npy"), str(tmp_path / "mask.npy"), str(tmp_path / "final.npy")
    np.save(seg_f, lab); np.save(mask_f, mask.astype(np.uint8))
    of, pf = [], []
    for i, (o, p) in enumerate(zip(orders, probs)):
        of.append(str(tmp_path / ("order%d.txt" % i))); pf.append(str(tmp_path / ("prob%d.txt" % i)))
        np.savetxt(of[-1], o, fmt="%d"); np.savetxt(pf[-1], p, fmt="%.17g")
    r = subprocess.run([os.path.join(BIN, "segment_greedy"), "-s", seg_f, "-m", mask_f, "-o"] + of + ["-p"] + pf + ["-f", out_f],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    want, _, _ = so.segment_greedy(lab, orders, merge_probs=probs, mask=mask, ignore=True)
    assert np.array_equal(np.load(out_f), want)


# ---------------------------------------------------------------- bc label generation
def _truth_of(lab, seed, ntruth):
    rng = np.random.default_rng(seed + 500)
    group = rng.integers(1, ntruth + 1, size=int(lab.max()) + 1).astype(np.uint32)
    t = group[lab]
    noise = rng.random(lab.shape)
    t = np.where(noise < 0.10, rng.integers(1, ntruth + 1, size=lab.shape).astype(np.uint32), t)
    t[noise > 0.94] = 0
    return t.astype(np.uint32)


BCL_KW = [dict(), dict(tweak=True), dict(max_prec_drop=0.05), dict(opt_split=True), dict(use_f1=False), dict(global_opt=1),
          dict(global_opt=2), dict(tweak=True, max_prec_drop=0.2, opt_split=True)]


@pytest.mark.parametrize("case", [((30, 36), 60, 111, False, 6), ((28, 31), 40, 112, True, 4), ((9, 12, 11), 30, 113, True, 3)])
def test_bc_label_matches_oracle(case):
    """hmt/main_bc_label_ri.cxx: labels and the F1 / Rand-index values behind them against the voxel-set-by-voxel-set
    Python restatement (pinned to the reference's templates), every option; a prefix of the order (a forest) as well."""
    import sys
    sys.path.insert(0, ol.ORACLE_DIR)
    import bc_label_oracle as bo
    shape, nl, seed, um, ntruth = case
    lab, pb, mask = _make(shape, nl, seed, um)
    truth = _truth_of(lab, seed, ntruth)
    order = ol.Oracle(lab, pb, ol.default_config(pb), mask=mask).greedy(
        fn=lambda row: float(np.floor((np.sum(np.abs(row[:12])) * 3.7 % 1.0) * 8) / 8.0))["order"]
    g = api.MergeTreeBuilder(shape, api.FeatureConfig(thr=())).build(lab, np.zeros(shape, np.float32), mask)
    for kw in BCL_KW:
        for o in (order, order[:len(order) // 2]):
            want, wm, ws = bo.bc_labels(lab, truth, o, mask=mask, **kw)
            got, gm, gs = g.bc_label(truth, o, **kw)
            assert np.array_equal(got, want), kw
            if kw.get("global_opt", 0) == 0:
                assert np.array_equal(gm, wm, equal_nan=True) and np.array_equal(gs, ws, equal_nan=True), kw
    with pytest.raises(api.GliaError):
        bad = order.copy(); bad[0, 0] = 4000000
        g.bc_label(truth, bad)


def test_cli_bc_label_ri(tmp_path):
    import sys
    sys.path.insert(0, ol.ORACLE_DIR)
    import bc_label_oracle as bo
    shape = (32, 40)
    lab, pb, mask = _make(shape, 50, 114, True)
    truth = _truth_of(lab, 114, 5)
    order = ol.Oracle(lab, pb, ol.default_config(pb), mask=mask).merge_order_pb(2)["order"] if False else \
        ol.Oracle(lab, pb, ol.default_config(pb), mask=mask).greedy(fn=lambda row: float(np.sum(row[:5]) % 1.0))["order"]
    seg_f, mask_f, truth_f = str(tmp_path / "seg.npy"), str(tmp_path / "mask.npy"), str(tmp_path / "truth.npy")
    order_f, out_f = str(tmp_path / "order.txt"), str(tmp_path / "bclabel.txt")
    np.save(seg_f, lab); np.save(mask_f, mask.astype(np.uint8)); np.save(truth_f, truth.astype(np.uint16))
    np.savetxt(order_f, order, fmt="%d")
    for flags, kw in (([], dict()), (["-g", "2"], dict(global_opt=2)), (["-p", "1", "-w", "1", "-d", "0.2"], dict(opt_split=True, tweak=True, max_prec_drop=0.2)),
                      (["--f1", "0"], dict(use_f1=False))):
        r = subprocess.run([os.path.join(BIN, "bc_label_ri"), "-s", seg_f, "-o", order_f, "-t", truth_f, "-n", mask_f, "-l", out_f] + flags,
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        want, _, _ = bo.bc_labels(lab, truth, order, mask=mask, **kw)
        assert np.array_equal(np.loadtxt(out_f, dtype=np.int64), want), flags


# ---------------------------------------------------------------- section-classifier features
@pytest.mark.parametrize("case", [((40, 48), 60, 141, False), ((37, 45), 40, 142, True), ((256, 256), 0, 143, False)])
def test_sc_feat_matches_oracle(case):
    """hmt/main_sc_feat.cxx: every row against the C restatement (pinned to the reference's feature classes); the
    integer-derived columns are bit-identical, moments agree to ~1e-12 (closed form from integer power sums against a
    pixel-by-pixel double sum)."""
    shape, nl, seed, um = case
    rng = np.random.default_rng(seed)
    if nl:
        lab0, pb0 = synth.speckle_volume(shape, nl, seed=seed)
    else:
        lab0, pb0 = synth.voronoi_volume(shape, 11, seed=seed)
    lab1 = np.roll(lab0, (1, 2), axis=(0, 1)).copy()
    pb1 = np.roll(pb0, (1, 2), axis=(0, 1)).copy()
    m0 = (rng.random(shape) > 0.04).astype(np.uint32) if um else None
    m1 = (rng.random(shape) > 0.04).astype(np.uint32) if um else None
    l0 = np.unique(lab0[np.ones(shape, bool) if m0 is None else m0 != 0])
    l1 = np.unique(lab1[np.ones(shape, bool) if m1 is None else m1 != 0])
    pairs = np.stack([rng.choice(l0, 200), rng.choice(l1, 200)], axis=1).astype(np.uint32)
    both = np.intersect1d(l0, l1)
    pairs[:60, 0] = pairs[:60, 1] = rng.choice(both, 60)      # overlapping pairs
    o0 = ol.Oracle(lab0, pb0, ol.default_config(pb0), mask=m0)
    o1 = ol.Oracle(lab1, pb1, ol.default_config(pb1), mask=m1)
    g0 = api.MergeTreeBuilder(shape, api.default_config(pb0)).build(lab0, pb0, m0)
    g1 = api.MergeTreeBuilder(shape, api.default_config(pb1)).build(lab1, pb1, m1)
    for kw in (dict(), dict(use_log=True), dict(norm_area=1.0, norm_len=1.0), dict(norm_area=1.0, norm_len=1.0, use_log=True)):
        want = ol.sc_feat(o0, o1, pairs, **kw)
        got = g0.sc_feat(g1, pairs, **kw)
        assert got.shape == want.shape
        if kw.get("use_log"):
            # --logs takes the logarithm of the central-moment differences (columns 15..21): where two regions have the
            # same shape that difference is rounding noise of the pixel sums (1e-13) and its logarithm means nothing
            raw = ol.sc_feat(o0, o1, pairs, **dict(kw, use_log=False))[:, 15:22]
            noise = raw < 1e-7 * np.maximum(1.0, np.abs(raw).max(axis=0))
            got[:, 15:22][noise] = want[:, 15:22][noise]
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-9 * max(1.0, float(np.abs(want).max())), err_msg=str(kw))
        assert (got == want).mean() > 0.5
    assert (want[:60, 7] > 0).any()                          # the overlap column is exercised
    with pytest.raises(api.GliaError):
        bad = pairs.copy(); bad[0, 0] = 4000000
        g0.sc_feat(g1, bad)
