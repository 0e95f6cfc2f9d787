"""Multi-GPU path on ONE GPU: `world` ranks emulated as threads (parallel.ThreadComm), each with its
own handle over a z-slab + halo, must reproduce the single-handle result exactly: same candidates,
bit-identical rows on dyadic pb, identical RF scores.  The single-handle path itself is checked
against the oracle in test_gpu_parity.py, so this pins the slab/exchange logic (halo, global border
test, reduce keyed by label pair, ownership) — the NCCL transport is exercised by bench.py --gpus N."""
import threading

import numpy as np
import pytest

import oracle_lib as ol
from glia_b200 import api, parallel as par, synth

pytestmark = pytest.mark.gpu


def _run_sharded(world, shape, lab, pb, cfg_fn, mask=None, forest=None):
    comms = par.ThreadComm.group(world)
    res, errs = [None] * world, []
    lb = int(lab.max()) + 1

    def run(c):
        try:
            z0, z1, hl, hh = par.slab_with_halo(shape[0], c.world, c.rank)
            sl = slice(z0 - hl, z1 + hh)
            p = np.ascontiguousarray(pb[sl])
            b = par.slab_builder(shape, cfg_fn(p), c, label_bound=lb)
            t = {}
            par.build_sharded(c, b, np.ascontiguousarray(lab[sl]), p,
                              None if mask is None else np.ascontiguousarray(mask[sl]), timings=t)
            keys, rows = b.init_candidates()
            sal = b.score_candidates(forest) if forest is not None and b.Cn else np.zeros(b.Cn)
            res[c.rank] = (keys, rows, sal, b.R, t)
            b.close()
        except BaseException as e:  # noqa: BLE001
            errs.append(e)
            c.s.bar.abort()

    ts = [threading.Thread(target=run, args=(c,)) for c in comms]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    return res


def _cfg(p):
    return api.FeatureConfig(rimgs=[(p, 16, 0.0, 1.0)], bimgs=[(p, 16, 0.0, 1.0)])


@pytest.mark.parametrize("world,shape,cell,shuffle,masked", [
    (2, (24, 40, 56), 8, False, False),
    (3, (25, 33, 70), 7, False, False),   # ragged slabs, partial warps
    (2, (16, 24, 40), 6, True, False),    # labels not spatially ordered: most boundaries cross owners
    (4, (12, 20, 36), 5, True, True),     # mask -> generic voxel kernel; slabs of 3 planes
    (2, (2, 30, 33), 6, False, False),    # one plane per rank
])
def test_sharded_equals_single(world, shape, cell, shuffle, masked):
    lab, pb = synth.voronoi_volume(shape, cell, seed=11 + world)
    rng = np.random.RandomState(3)
    if shuffle:
        perm = rng.permutation(int(lab.max()) + 1).astype(np.uint32) + 1
        lab = perm[lab]
    mask = (rng.rand(*shape) > 0.04).astype(np.uint32) if masked else None
    g = api.MergeTreeBuilder(shape, _cfg(pb)).build(lab, pb, mask)
    k0, rows0 = g.init_candidates()
    f = synth.random_forest(31, g.row_dim(), max_depth=8, seed=2, thr_pool=rows0)
    forest = api.Forest.from_arrays(f, g.row_dim())
    sal0 = g.score_candidates(forest)
    res = _run_sharded(world, shape, lab, pb, _cfg, mask, forest)
    keys = np.concatenate([r[0] for r in res]); rows = np.concatenate([r[1] for r in res])
    sal = np.concatenate([r[2] for r in res])
    assert all(r[3] == g.R for r in res)
    assert len(keys) == g.Cn  # every candidate owned exactly once
    # owners hold ascending, disjoint region-index ranges -> concatenation is already in table order
    assert np.array_equal(keys, k0)
    assert np.array_equal(rows, rows0)  # dyadic pb: sums exact, so rows are bit-identical
    assert np.array_equal(sal, sal0)
    if world > 1 and not shuffle:
        # spatially ordered labels: only boundaries near slab faces travel
        assert sum(r[4]["sent_records"] for r in res) < 2 * g.E


def test_sharded_against_oracle():
    shape = (18, 26, 38)
    lab, pb = synth.voronoi_volume(shape, 7, seed=5)
    spec = [(pb, 16, 0.0, 1.0)]
    o = ol.Oracle(lab, pb, ol.FeatureConfig(rimgs=spec, bimgs=spec))
    ko, rowo = o.init_candidates()
    res = _run_sharded(2, shape, lab, pb, _cfg)
    keys = np.concatenate([r[0] for r in res]); rows = np.concatenate([r[1] for r in res])
    assert np.array_equal(keys, ko)
    np.testing.assert_allclose(rows, rowo, rtol=1e-5, atol=1e-12)


def test_sharded_rebuild_path(monkeypatch):
    """The fallback of build_sharded (own records exported too, tables rebuilt and grown from the received
    runs) gives the same candidates as the default path (own records kept in place)."""
    shape = (20, 30, 44)
    lab, pb = synth.voronoi_volume(shape, 6, seed=9)
    perm = np.random.RandomState(4).permutation(int(lab.max()) + 1).astype(np.uint32) + 1
    lab = perm[lab]
    a = _run_sharded(3, shape, lab, pb, _cfg)
    monkeypatch.setenv("GLIA_B200_FORCE_REBUILD", "1")
    b = _run_sharded(3, shape, lab, pb, _cfg)
    for ra, rb in zip(a, b):
        assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1])
