"""BASELINE.json's full-size configurations (C2 4096^2, C4 1024^3) through size-independent
properties: the oracle cannot run at these sizes, so the CUDA path is checked against invariants of
the reference's definitions computed independently with plain torch tensor ops on the same inputs
(type/neighbor.hxx:72-126, util/struct.hxx:95-108, type/boundary_table.hxx:99-102)."""
import ctypes as C

import numpy as np
import pytest
import torch

from glia_b200 import api, synth

pytestmark = pytest.mark.gpu


def _synth_device(shape, cell, seed):
    L = api.lib()
    nd = len(shape)
    n = int(np.prod(shape))
    dev = torch.device("cuda", 0)
    lab = torch.empty(n, dtype=torch.int32, device=dev)
    pb = torch.empty(n, dtype=torch.float32, device=dev)
    dims = (C.c_int * 3)(*(list(shape[::-1]) + [1] * (3 - nd)))
    nc = C.c_uint64(0)
    assert L.glia_b200_synth_voronoi(nd, dims, cell, seed, lab.data_ptr(), pb.data_ptr(), None, C.byref(nc)) == 0
    torch.cuda.synchronize()
    return lab.view(shape), pb.view(shape), int(nc.value) + 1


def _boundary_and_border(lab):
    """#voxels with a differing in-bounds neighbour; #other voxels lying on a face of the volume."""
    diff = torch.zeros_like(lab, dtype=torch.bool)
    face = torch.zeros_like(lab, dtype=torch.bool)
    for d in range(lab.dim()):
        a, b = lab.narrow(d, 0, lab.shape[d] - 1), lab.narrow(d, 1, lab.shape[d] - 1)
        ne = a != b
        diff.narrow(d, 0, lab.shape[d] - 1).logical_or_(ne)
        diff.narrow(d, 1, lab.shape[d] - 1).logical_or_(ne)
        del ne
        face.select(d, 0).fill_(True)
        face.select(d, lab.shape[d] - 1).fill_(True)
    nb = int(diff.sum().item())
    nborder = int((face & ~diff).sum().item())
    return nb, nborder


@pytest.mark.parametrize("shape,cell", [((4096, 4096), 13), ((1024, 1024, 1024), 10)])
def test_fullsize_invariants(shape, cell):
    free, _ = torch.cuda.mem_get_info()
    if free < 40 * 2 ** 30 and len(shape) == 3:
        pytest.skip("needs ~40 GB of free HBM")
    lab, pb, lb = _synth_device(shape, cell, 1234)
    n = lab.numel()
    spec = [(pb.data_ptr(), 16, 0.0, 1.0)]
    g = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec), device_pointers=True, label_bound=lb)
    g.build(lab.data_ptr(), pb.data_ptr())
    nb, nborder = _boundary_and_border(lab)
    r, e = g.regions(), g.edges()
    rs, es = g.region_stats(0), g.edge_stats(0)
    # every voxel in exactly one region; every boundary voxel in exactly one directed boundary
    assert int(r["count"].sum(dtype=np.uint64)) == n
    assert int(e["count"].sum(dtype=np.uint64)) == nb
    assert int(r["border"].sum(dtype=np.uint64)) == nborder
    assert g.R == int(torch.unique(lab).numel())
    assert np.all(np.diff(r["label"].astype(np.int64)) > 0)
    key = e["a"].astype(np.uint64) << np.uint64(32) | e["b"].astype(np.uint64)
    assert np.all(np.diff(key.astype(np.int64)) > 0)  # sorted, unique directed pairs
    # mutual flag == the reverse pair exists; candidates = mutual pairs counted once
    rev = e["b"].astype(np.uint64) << np.uint64(32) | e["a"].astype(np.uint64)
    assert np.array_equal(np.isin(rev, key, assume_unique=True), e["mutual"].astype(bool))
    assert int(e["mutual"].sum()) == 2 * g.Cn
    # histograms: pb is dyadic in [0, 1) so no value is dropped; thresholds are monotone
    assert np.array_equal(rs["hist"].sum(1), r["count"]) and np.array_equal(es["hist"].sum(1), e["count"])
    t = e["thr"].astype(np.int64)
    assert np.all(t[:, 0] >= t[:, 1]) and np.all(t[:, 1] >= t[:, 2]) and np.all(e["count"] >= t[:, 0])
    # sums are exact in double for dyadic pb: the grand totals must equal the volume's
    tot = float(pb.double().sum().item()); tot2 = float((pb.double() ** 2).sum().item())
    assert float(rs["sum"].sum()) == tot and float(rs["sum2"].sum()) == tot2
    assert float(rs["min"].min()) == float(pb.min().item()) and float(rs["max"].max()) == float(pb.max().item())
    # bbox: inside the volume, non-empty, and its box holds at least `count` voxels
    nd = len(shape)
    ext = (r["bbmax"][:, :nd] - r["bbmin"][:, :nd] + 1).astype(np.int64)
    assert np.all(r["bbmin"][:, :nd] >= 0) and np.all(r["bbmax"][:, :nd] < np.array(shape[::-1])) and np.all(ext.prod(1) >= r["count"])
    # scores: votes / 255, repeatable
    f = synth.random_forest(255, g.row_dim(), max_depth=12, seed=1)
    forest = api.Forest.from_arrays(f, g.row_dim())
    sal = torch.empty(g.Cn, dtype=torch.float64, device="cuda")
    g.score_candidates(forest, sal.data_ptr())
    s1 = sal.clone()
    k = torch.round(s1 * 255)
    assert float((k - s1 * 255).abs().max().item()) < 1e-9 and float(s1.min().item()) >= 0.0 and float(s1.max().item()) <= 1.0
    g.score_candidates(forest, sal.data_ptr())
    assert torch.equal(sal, s1)
    g.close()
