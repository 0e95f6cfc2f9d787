"""The drop-in command-line programs (glia_b200/bin): same flags and file formats as the
reference's merge_order_bc / bc_feat / pred_rf (SURVEY.md §8b "process level" seam)."""
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
from glia_b200 import api, synth

BIN = os.path.join(os.path.dirname(api.HERE), "glia_b200", "bin")


def _write_mha(path, arr, met_type):
    dims = " ".join(str(s) for s in arr.shape[::-1])
    hdr = ("ObjectType = Image\nNDims = %d\nBinaryData = True\nBinaryDataByteOrderMSB = False\n"
           "CompressedData = False\nDimSize = %s\nElementType = %s\nElementDataFile = LOCAL\n" % (arr.ndim, dims, met_type))
    with open(path, "wb") as f:
        f.write(hdr.encode())
        f.write(np.ascontiguousarray(arr).tobytes())


def test_binaries_exist_and_print_usage():
    for name in ("merge_order_bc", "bc_feat", "pred_rf", "merge_order_pb", "pre_merge", "segment_greedy", "bc_label_ri"):
        p = os.path.join(BIN, name)
        if not os.path.exists(p):
            api.build_library()
        r = subprocess.run([p, "--help"], capture_output=True, text=True)
        assert r.returncode == 0 and "Usage" in r.stderr


@pytest.mark.gpu
def test_cli_pipeline_matches_api_and_oracle(tmp_path):
    shape = (40, 56)
    lab, pb = synth.speckle_volume(shape, 60, seed=31)
    spec = [(pb, 16, 0.0, 1.0)]
    o = ol.Oracle(lab, pb, ol.FeatureConfig(rimgs=spec, bimgs=spec))
    g = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec)).build(lab, pb)
    _, rowo = o.init_candidates()
    _, rowg = g.init_candidates()
    cols = np.flatnonzero((rowo == rowg).all(axis=0))
    f = synth.random_forest(15, o.row_dim(), max_depth=5, seed=2, thr_pool=rowo, features=cols)
    orf = ol.OracleRF.from_arrays(f, o.row_dim())
    model = str(tmp_path / "model.bin")
    assert orf.save(model) == 0
    seg_f, pb_f = str(tmp_path / "seg.npy"), str(tmp_path / "pb.mha")
    np.save(seg_f, lab.astype(np.int32))          # a non-uint32 label file: converted on load
    _write_mha(pb_f, pb, "MET_FLOAT")
    order_f, sal_f, feat_f = (str(tmp_path / n) for n in ("order.txt", "sal.txt", "bcfeat.txt"))
    cmd = [os.path.join(BIN, "merge_order_bc"), "--bct", "1", "--bcm", model, "-s", seg_f, "--pb", pb_f,
           "--rbi", pb_f, "--rbb", "16", "--rbl", "0", "--rbu", "1", "--bt", "0.2", "0.5", "0.8",
           "-o", order_f, "--sal", sal_f, "-b", feat_f]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    ref = o.greedy(rf=orf)
    order = np.loadtxt(order_f, dtype=np.int64).reshape(-1, 3)
    assert np.array_equal(order, ref["order"])
    sal = np.loadtxt(sal_f)
    np.testing.assert_allclose(sal, ref["sal"], rtol=1e-5)        # default stream precision 6
    rows = np.loadtxt(feat_f).reshape(len(order), -1)
    np.testing.assert_allclose(rows, ref["rows"], rtol=1e-6, atol=1e-9)   # FLT_PREC = 8 digits
    # every row ends with "<value> \n" (util/text_io.hxx:120-133)
    assert open(feat_f).readline().endswith(" \n")
    # bc_feat on that order, with saliencies
    feat2_f = str(tmp_path / "bcfeat2.txt")
    cmd = [os.path.join(BIN, "bc_feat"), "-s", seg_f, "-o", order_f, "-y", sal_f, "--s0", "1.0", "--sb", "0.25",
           "--pb", pb_f, "--rbi", pb_f, "--rbb", "16", "--rbl", "0", "--rbu", "1", "--bt", "0.2", "0.5", "0.8",
           "-b", feat2_f]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rows2 = np.loadtxt(feat2_f).reshape(len(order), -1)
    want = o.bc_feat(ref["order"], sal=sal, init_sal=1.0, sal_bias=0.25)
    np.testing.assert_allclose(rows2, want, rtol=1e-6, atol=1e-9)
    # pred_rf on the feature file
    pred_f = str(tmp_path / "pred.txt")
    r = subprocess.run([os.path.join(BIN, "pred_rf"), "--m", model, "--f", feat_f, "--l", "-1", "--p", pred_f],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    _, want_p = orf.predict(rows)
    np.testing.assert_allclose(np.loadtxt(pred_f), want_p, rtol=1e-7)
    # errors go to stderr with exit code 1 (glia_base.hxx:66-69), never a crash
    r = subprocess.run([os.path.join(BIN, "pred_rf"), "--m", "/nonexistent", "--f", feat_f, "--l", "-1", "--p", pred_f],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "Error" in r.stderr


@pytest.mark.gpu
def test_cli_hist_and_median_flags(tmp_path):
    """--histf / --medf select the layouts the reference builds with GLIA_HMT_HIST_FEAT / GLIA_HMT_MEDIAN_FEAT."""
    shape = (30, 44)
    lab, pb = synth.speckle_volume(shape, 40, seed=41)
    spec = [(pb, 16, 0.0, 1.0)]
    kw = dict(use_hist=True, use_median=True)
    o = ol.Oracle(lab, pb, ol.FeatureConfig(rimgs=spec, bimgs=spec, **kw))
    res = o.greedy(fn=lambda row: np.floor((np.sum(np.abs(row[:12])) * 3.7 % 1.0) * 8) / 8.0)
    seg_f, pb_f = str(tmp_path / "seg.npy"), str(tmp_path / "pb.mha")
    np.save(seg_f, lab)
    _write_mha(pb_f, pb, "MET_FLOAT")
    order_f, feat_f = str(tmp_path / "order.txt"), str(tmp_path / "bcfeat.txt")
    np.savetxt(order_f, res["order"], fmt="%d")
    cmd = [os.path.join(BIN, "bc_feat"), "-s", seg_f, "-o", order_f, "--pb", pb_f, "--rbi", pb_f, "--rbb", "16",
           "--rbl", "0", "--rbu", "1", "--bt", "0.2", "0.5", "0.8", "--histf", "1", "--medf", "1", "-b", feat_f]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rows = np.loadtxt(feat_f).reshape(len(res["order"]), -1)
    want = o.bc_feat(res["order"])
    assert rows.shape == want.shape
    np.testing.assert_allclose(rows, want, rtol=1e-6, atol=1e-9)
