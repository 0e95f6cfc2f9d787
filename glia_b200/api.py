"""ctypes binding of libglia_b200.so (include/glia_b200.h) — the reference-side stub a maintainer
would write.  No torch types cross the C ABI; numpy arrays are host buffers,
integer addresses are device buffers.  There is no CPU fallback: a missing library or GPU raises.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libglia_b200.so")
MAX_IMG, MAX_THR = 4, 8
ERRORS = {1: "bad argument", 2: "CUDA error", 3: "edge table capacity", 4: "bad state",
          5: "model file I/O", 6: "bad merge order"}


class GliaError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__("glia_b200 error %d (%s): %s" % (code, ERRORS.get(code, "?"), msg))
        self.code = code


class ImageSpec(C.Structure):
    _fields_ = [("data", C.c_void_p), ("bins", C.c_int), ("lo", C.c_double), ("hi", C.c_double)]


class Params(C.Structure):
    _fields_ = [
        ("ndim", C.c_int), ("dims", C.c_int * 3),
        ("n_rimg", C.c_int), ("n_rlimg", C.c_int), ("n_bimg", C.c_int), ("n_thr", C.c_int),
        ("rimg", ImageSpec * MAX_IMG), ("rlimg", ImageSpec * MAX_IMG), ("bimg", ImageSpec * MAX_IMG),
        ("thr", C.c_double * MAX_THR),
        ("normalize", C.c_int), ("use_log", C.c_int), ("simple", C.c_int),
        ("device_pointers", C.c_int), ("device", C.c_int),
        ("label_bound", C.c_uint32), ("edge_capacity", C.c_uint64),
        ("z_begin", C.c_int), ("z_end", C.c_int), ("halo_lo", C.c_int), ("halo_hi", C.c_int),
        ("stream", C.c_void_p),
        ("use_hist_feat", C.c_int), ("use_median_feat", C.c_int), ("contour_only", C.c_int),
    ]


def build_library(verbose=False):
    """Compile libglia_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-s", "-j8", "-C", os.path.join(HERE, "csrc")],
                       capture_output=not verbose, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libglia_b200.so failed:\n" + (r.stdout or "") + (r.stderr or ""))


_lib = None
EXPORTS = ["glia_b200_create", "glia_b200_destroy", "glia_b200_last_error", "glia_b200_build",
           "glia_b200_counts", "glia_b200_row_dim", "glia_b200_get_regions",
           "glia_b200_get_region_stats", "glia_b200_get_edges", "glia_b200_get_edge_stats",
           "glia_b200_init_candidates", "glia_b200_score_candidates", "glia_b200_merge_order_bc",
           "glia_b200_bc_feat", "glia_b200_rf_load", "glia_b200_rf_from_arrays",
           "glia_b200_rf_destroy", "glia_b200_rf_info", "glia_b200_rf_score", "glia_b200_last_ms",
           "glia_b200_launch_count", "glia_b200_build_partial", "glia_b200_raw_region_tables",
           "glia_b200_finalize_regions", "glia_b200_edge_interest", "glia_b200_export_edges",
           "glia_b200_import_edges", "glia_b200_table_load", "glia_b200_finalize_edges", "glia_b200_synth_voronoi",
           "glia_b200_synth_voronoi_slab", "glia_b200_set_image", "glia_b200_rf_ensemble",
           "glia_b200_region_presence", "glia_b200_index_regions", "glia_b200_gather_regions",
           "glia_b200_tune", "glia_b200_candidate_ties", "glia_b200_set_tie_orientation",
           "glia_b200_reference_orientation", "glia_b200_comm_unique_id", "glia_b200_comm_create",
           "glia_b200_comm_destroy", "glia_b200_build_sharded", "glia_b200_comm_last_exchange_ms",
           "glia_b200_bc_feat_sharded", "glia_b200_merge_order_pb", "glia_b200_pre_merge",
           "glia_b200_segment_greedy", "glia_b200_rf_score_noalloc", "glia_b200_rf_score_noalloc_max_votes", "glia_b200_bc_label",
           "glia_b200_sc_row_dim", "glia_b200_sc_feat"]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libglia_b200.so is not built (run __graft_entry__.build()); "
                               "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.glia_b200_create.argtypes = [C.POINTER(Params), C.POINTER(vp)]
        L.glia_b200_destroy.argtypes = [vp]
        L.glia_b200_destroy.restype = None
        L.glia_b200_last_error.argtypes = [vp]
        L.glia_b200_last_error.restype = C.c_char_p
        L.glia_b200_build.argtypes = [vp, vp, vp, vp]
        L.glia_b200_tune.argtypes = [vp, C.c_char_p, C.c_int]
        L.glia_b200_comm_unique_id.argtypes = [vp]
        L.glia_b200_comm_create.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
        L.glia_b200_comm_destroy.argtypes = [vp]
        L.glia_b200_comm_destroy.restype = None
        L.glia_b200_build_sharded.argtypes = [vp, vp, vp, vp, vp]
        L.glia_b200_comm_last_exchange_ms.argtypes = [vp]
        L.glia_b200_comm_last_exchange_ms.restype = C.c_double
        L.glia_b200_candidate_ties.argtypes = [vp, vp, C.POINTER(C.c_int64)]
        L.glia_b200_set_tie_orientation.argtypes = [vp, vp]
        L.glia_b200_reference_orientation.argtypes = [vp, C.c_int64, vp]
        L.glia_b200_counts.argtypes = [vp] + [C.POINTER(C.c_int64)] * 3
        L.glia_b200_row_dim.argtypes = [vp, C.c_int]
        L.glia_b200_get_regions.argtypes = [vp] * 6
        L.glia_b200_get_region_stats.argtypes = [vp, C.c_int, C.c_int] + [vp] * 5
        L.glia_b200_get_edges.argtypes = [vp] * 6
        L.glia_b200_get_edge_stats.argtypes = [vp, C.c_int] + [vp] * 5
        L.glia_b200_init_candidates.argtypes = [vp, vp, vp]
        L.glia_b200_score_candidates.argtypes = [vp, vp, vp]
        L.glia_b200_merge_order_bc.argtypes = [vp, vp, vp, vp, vp, C.POINTER(C.c_int64),
                                               C.POINTER(C.c_int64)]
        L.glia_b200_bc_feat.argtypes = [vp, vp, C.c_int64, vp, C.c_double, C.c_double, vp]
        L.glia_b200_rf_score_noalloc.argtypes = [vp, vp, C.c_size_t, C.c_int, C.c_int, vp, vp, C.c_char_p, C.c_size_t]
        L.glia_b200_rf_score_noalloc_max_votes.argtypes = []
        L.glia_b200_rf_score_noalloc_max_votes.restype = C.c_size_t
        L.glia_b200_sc_row_dim.argtypes = [vp]
        L.glia_b200_sc_feat.argtypes = [vp, vp, vp, C.c_int64, C.c_double, C.c_double, C.c_int, vp]
        L.glia_b200_bc_label.argtypes = [vp, vp, vp, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, vp, vp, vp]
        L.glia_b200_segment_greedy.argtypes = [vp, C.c_int, vp, vp, vp, vp, C.c_int, vp, vp, C.POINTER(C.c_int64)]
        L.glia_b200_merge_order_pb.argtypes = [vp, C.c_int, vp, vp, C.POINTER(C.c_int64)]
        L.glia_b200_pre_merge.argtypes = [vp, vp, C.c_int, C.c_double, vp, vp, C.POINTER(C.c_int64), vp]
        L.glia_b200_bc_feat_sharded.argtypes = [vp, vp, vp, C.c_int64, vp, C.c_double, C.c_double, vp,
                                                C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.glia_b200_rf_load.argtypes = [C.c_char_p, C.c_int, C.POINTER(vp)]
        L.glia_b200_rf_from_arrays.argtypes = [C.c_int] * 4 + [vp] * 7 + [C.c_int, C.POINTER(vp)]
        L.glia_b200_rf_destroy.argtypes = [vp]
        L.glia_b200_rf_destroy.restype = None
        L.glia_b200_rf_info.argtypes = [vp] + [C.POINTER(C.c_int)] * 4
        L.glia_b200_rf_score.argtypes = [vp, vp, C.c_size_t, C.c_int, C.c_int, vp, vp, C.c_int, vp]
        L.glia_b200_last_ms.argtypes = [vp, C.c_int]
        L.glia_b200_last_ms.restype = C.c_double
        L.glia_b200_launch_count.argtypes = [vp]
        L.glia_b200_launch_count.restype = C.c_int64
        i64p = C.POINTER(C.c_int64)
        vpp = C.POINTER(vp)
        L.glia_b200_build_partial.argtypes = [vp, vp, vp, vp]
        L.glia_b200_raw_region_tables.argtypes = [vp, vpp, vpp, vpp, i64p, i64p, i64p]
        L.glia_b200_finalize_regions.argtypes = [vp, i64p]
        L.glia_b200_region_presence.argtypes = [vp, vpp, i64p]
        L.glia_b200_index_regions.argtypes = [vp, i64p]
        L.glia_b200_gather_regions.argtypes = [vp]
        L.glia_b200_edge_interest.argtypes = [vp, C.c_int, vpp, i64p]
        L.glia_b200_export_edges.argtypes = [vp, C.c_int, C.c_int, i64p, vpp, C.POINTER(C.c_int)]
        L.glia_b200_import_edges.argtypes = [vp, vp, C.c_int64, C.c_int]
        L.glia_b200_table_load.argtypes = [vp, i64p, i64p]
        L.glia_b200_finalize_edges.argtypes = [vp, C.c_int64, C.c_int64]
        L.glia_b200_synth_voronoi.argtypes = [C.c_int, C.POINTER(C.c_int), C.c_int, C.c_uint32, vp, vp, vp,
                                              C.POINTER(C.c_uint64)]
        L.glia_b200_synth_voronoi_slab.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_int, C.c_int, vp, vp,
                                                   vp, C.POINTER(C.c_uint64)]
        L.glia_b200_set_image.argtypes = [vp, C.c_int, C.c_int, vp]
        L.glia_b200_rf_ensemble.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.c_int, C.c_double, C.POINTER(vp)]
        _lib = L
    return _lib


def _addr(a):
    """numpy array -> host address; int -> device address; None -> NULL."""
    if a is None:
        return None
    if isinstance(a, (int, np.integer)):
        return C.c_void_p(int(a))
    return a.ctypes.data_as(C.c_void_p)


class FeatureConfig:
    """Mirrors the options of hmt/main_merge_order_bc.cxx:172-242 (--rbi/--ri/--rli/--bi, --bt,
    --ns, --logs, --simpf).  Each image entry: (array-or-device-address, bins, lo, hi)."""

    def __init__(self, rimgs=(), rlimgs=(), bimgs=(), thr=(0.2, 0.5, 0.8), normalize=False,
                 use_log=False, simple=False, use_hist=False, use_median=False, contour_only=False):
        self.contour_only = contour_only  # TRegionMap(..., onlyContour=true): masked-out voxels are classified too
        self.rimgs, self.rlimgs, self.bimgs = list(rimgs), list(rlimgs), list(bimgs)
        self.thr = list(thr)
        self.normalize, self.use_log, self.simple = normalize, use_log, simple
        # the reference's GLIA_USE_HISTOGRAM_AS_FEATS / GLIA_USE_MEDIAN_AS_FEATS builds (CMakeLists.txt:53-64)
        self.use_hist, self.use_median = use_hist, use_median


def default_config(pb, **kw):
    """--rbi pb --rbb 16 --rbl 0 --rbu 1 --bt 0.2 0.5 0.8"""
    return FeatureConfig(rimgs=[(pb, 16, 0.0, 1.0)], bimgs=[(pb, 16, 0.0, 1.0)], **kw)


class Forest:
    """alg::RandomForest (alg/rf.hxx:9-60) on the device."""

    def __init__(self, handle):
        self.h = handle
        v = [C.c_int() for _ in range(4)]
        lib().glia_b200_rf_info(self.h, *[C.byref(x) for x in v])
        self.ntree, self.nrnodes, self.nclass, self.mdim = (x.value for x in v)

    @classmethod
    def load(cls, path, device=0):
        h = C.c_void_p()
        st = lib().glia_b200_rf_load(path.encode(), device, C.byref(h))
        if st:
            raise GliaError(st, path)
        return cls(h)

    @classmethod
    def from_arrays(cls, f, mdim=0, device=0):
        nt, nr = f["nodestatus"].shape
        a = {k: np.ascontiguousarray(f[k], np.float64 if k == "xbestsplit" else np.int32) for k in
             ("treemap", "nodestatus", "xbestsplit", "bestvar", "nodeclass", "ndbigtree", "orig_labels")}
        h = C.c_void_p()
        st = lib().glia_b200_rf_from_arrays(nt, nr, len(a["orig_labels"]), mdim, _addr(a["treemap"]),
                                            _addr(a["nodestatus"]), _addr(a["xbestsplit"]),
                                            _addr(a["bestvar"]), _addr(a["nodeclass"]),
                                            _addr(a["ndbigtree"]), _addr(a["orig_labels"]), device,
                                            C.byref(h))
        if st:
            raise GliaError(st)
        return cls(h)

    @classmethod
    def ensemble(cls, forests, dim0, dim1, threshold):
        """alg::EnsembleRandomForest + ThresholdModelDistributor (alg/rf.hxx:63-98, type/function.hxx:71-85):
        model 0 if x[dim1] < threshold, else 1 if x[dim0] < threshold, else 2."""
        arr = (C.c_void_p * len(forests))(*[f.h for f in forests])
        h = C.c_void_p()
        st = lib().glia_b200_rf_ensemble(arr, len(forests), dim0, dim1, threshold, C.byref(h))
        if st:
            raise GliaError(st, "rf_ensemble")
        return cls(h)

    def predict(self, X, label=-1, want_votes=True):
        """Host rows [N][D] -> (votes [N][nclass] or None, prob [N])."""
        X = np.ascontiguousarray(X, np.float64)
        n, d = X.shape
        votes = np.zeros((n, self.nclass), np.int32) if want_votes else None
        prob = np.zeros(n)
        st = lib().glia_b200_rf_score(self.h, _addr(X), n, d, label, _addr(votes), _addr(prob), 0, None)
        if st:
            raise GliaError(st)
        return votes, prob

    def predict_device(self, x_ptr, n, d, prob_ptr, votes_ptr=None, label=-1, stream=None):
        st = lib().glia_b200_rf_score(self.h, _addr(int(x_ptr)), n, d, label,
                                      _addr(votes_ptr), _addr(int(prob_ptr)), 1,
                                      C.c_void_p(stream) if stream else None)
        if st:
            raise GliaError(st)

    def close(self):
        if self.h:
            lib().glia_b200_rf_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class NativeComm:
    """NCCL communicator pair owned by the library (include/glia_b200.h: glia_b200_comm_*), for
    MergeTreeBuilder.build_sharded.  `id256`: the 256 bytes of unique_id() of one rank, identical on all ranks."""

    def __init__(self, id256, world, rank, device=0):
        self.h = C.c_void_p()
        buf = (C.c_ubyte * 256).from_buffer_copy(bytes(id256))
        st = lib().glia_b200_comm_create(buf, world, rank, device, C.byref(self.h))
        if st:
            raise GliaError(st, "glia_b200_comm_create")
        self.world, self.rank = world, rank

    @staticmethod
    def unique_id():
        buf = (C.c_ubyte * 256)()
        st = lib().glia_b200_comm_unique_id(buf)
        if st:
            raise GliaError(st, "glia_b200_comm_unique_id (is libnccl.so.2 loadable?)")
        return bytes(buf)

    def last_exchange_ms(self):
        return float(lib().glia_b200_comm_last_exchange_ms(self.h))

    def close(self):
        if self.h:
            lib().glia_b200_comm_destroy(self.h)
            self.h = C.c_void_p()


class MergeTreeBuilder:
    """One handle = TRegionMap + TBoundaryTable of the reference for one volume on one GPU.

    shape is in numpy order ([z,] y, x).  With device_pointers=True every bulk argument is a
    device address (int) instead of a numpy array.
    """

    def __init__(self, shape, cfg, device=0, device_pointers=False, label_bound=0,
                 edge_capacity=0, slab=None, stream=None):
        self.shape = tuple(int(s) for s in shape)
        self.cfg = cfg
        self.device_pointers = device_pointers
        p = Params()
        p.ndim = len(self.shape)
        dims = list(self.shape[::-1]) + [1] * (3 - len(self.shape))
        for i in range(3):
            p.dims[i] = dims[i]
        self._keep = []

        def fill(dst, lst):
            for i, (img, bins, lo, hi) in enumerate(lst):
                if not device_pointers:
                    img = np.ascontiguousarray(img, np.float32)
                    self._keep.append(img)
                    dst[i].data = img.ctypes.data
                else:
                    dst[i].data = int(img)
                dst[i].bins, dst[i].lo, dst[i].hi = bins, lo, hi

        p.n_rimg, p.n_rlimg, p.n_bimg, p.n_thr = len(cfg.rimgs), len(cfg.rlimgs), len(cfg.bimgs), len(cfg.thr)
        fill(p.rimg, cfg.rimgs)
        fill(p.rlimg, cfg.rlimgs)
        fill(p.bimg, cfg.bimgs)
        for i, t in enumerate(cfg.thr):
            p.thr[i] = t
        p.normalize, p.use_log, p.simple = int(cfg.normalize), int(cfg.use_log), int(cfg.simple)
        p.use_hist_feat, p.use_median_feat = int(getattr(cfg, "use_hist", False)), int(getattr(cfg, "use_median", False))
        p.contour_only = int(getattr(cfg, "contour_only", False))
        p.device_pointers, p.device = int(device_pointers), device
        p.label_bound, p.edge_capacity = label_bound, edge_capacity
        if slab is not None:
            p.z_begin, p.z_end, p.halo_lo, p.halo_hi = slab
        p.stream = stream
        self.params = p
        self.h = C.c_void_p()
        st = lib().glia_b200_create(C.byref(p), C.byref(self.h))
        if st:
            self.h = None
            raise GliaError(st, "create")
        self.R = self.E = self.Cn = 0

    def _check(self, st):
        if st:
            raise GliaError(st, lib().glia_b200_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            lib().glia_b200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tune(self, key, value):
        """Per-handle knob (include/glia_b200.h: glia_b200_tune)."""
        self._check(lib().glia_b200_tune(self.h, key.encode(), int(value)))
        return self

    def candidate_ties(self):
        """(tie[C] uint8, count): initial candidates whose two regions have equal areas."""
        tie = np.zeros(max(self.Cn, 1), np.uint8)
        n = C.c_int64(0)
        self._check(lib().glia_b200_candidate_ties(self.h, _addr(tie), C.byref(n)))
        return tie[:self.Cn], int(n.value)

    def set_tie_orientation(self, second_first):
        if second_first is None:
            self._check(lib().glia_b200_set_tie_orientation(self.h, None))
            return self
        b = np.ascontiguousarray(second_first, np.uint8)
        assert len(b) == self.Cn
        self._keep_orient = b
        self._check(lib().glia_b200_set_tie_orientation(self.h, _addr(b)))
        return self

    def reference_orientation(self, reserve_hint=0):
        """Call order of a reference binary built with this libstdc++ (include/glia_b200.h)."""
        b = np.zeros(max(self.Cn, 1), np.uint8)
        self._check(lib().glia_b200_reference_orientation(self.h, int(reserve_hint), _addr(b)))
        return b[:self.Cn]

    def set_image(self, kind, i, data):
        """Re-point image i of kind 0 (rimg) / 1 (rlimg) / 2 (bimg) at another buffer of the same extent."""
        if not self.device_pointers:
            data = np.ascontiguousarray(data, np.float32)
            self._keep.append(data)
        self._check(lib().glia_b200_set_image(self.h, kind, i, _addr(data)))

    def build(self, labels, pb, mask=None):
        if not self.device_pointers:
            labels = np.ascontiguousarray(labels, np.uint32)
            pb = np.ascontiguousarray(pb, np.float32)
            mask = None if mask is None else np.ascontiguousarray(mask, np.uint32)
            self._keep += [labels, pb, mask]
        self._check(lib().glia_b200_build(self.h, _addr(labels), _addr(mask), _addr(pb)))
        v = [C.c_int64() for _ in range(3)]
        self._check(lib().glia_b200_counts(self.h, *[C.byref(x) for x in v]))
        self.R, self.E, self.Cn = (x.value for x in v)
        return self

    def row_dim(self, with_sal=False):
        return lib().glia_b200_row_dim(self.h, int(with_sal))

    def last_ms(self):
        return [lib().glia_b200_last_ms(self.h, i) for i in range(5)]

    def launch_count(self):
        return lib().glia_b200_launch_count(self.h)

    # ---- dumps (host) ----
    def regions(self):
        R = self.R
        lab, cnt, bor = (np.zeros(R, np.uint32) for _ in range(3))
        mn, mx = np.zeros((R, 3), np.int32), np.zeros((R, 3), np.int32)
        self._check(lib().glia_b200_get_regions(self.h, _addr(lab), _addr(cnt), _addr(bor), _addr(mn), _addr(mx)))
        return dict(label=lab, count=cnt, border=bor, bbmin=mn, bbmax=mx)

    def _stats(self, fn, n, bins, *pre):
        hist = np.zeros((n, bins), np.uint32)
        s, s2 = np.zeros(n), np.zeros(n)
        mn, mx = np.zeros(n, np.float32), np.zeros(n, np.float32)
        self._check(fn(self.h, *pre, _addr(hist), _addr(s), _addr(s2), _addr(mn), _addr(mx)))
        return dict(hist=hist, sum=s, sum2=s2, min=mn, max=mx)

    def region_stats(self, i, kind=0):
        bins = (self.cfg.rimgs if kind == 0 else self.cfg.rlimgs)[i][1]
        return self._stats(lib().glia_b200_get_region_stats, self.R, bins, kind, i)

    def edges(self):
        E, T = self.E, len(self.cfg.thr)
        a, b, c = (np.zeros(E, np.uint32) for _ in range(3))
        m = np.zeros(E, np.uint8)
        t = np.zeros((E, T), np.uint32)
        self._check(lib().glia_b200_get_edges(self.h, _addr(a), _addr(b), _addr(c), _addr(m), _addr(t)))
        return dict(a=a, b=b, count=c, mutual=m, thr=t)

    def edge_stats(self, i):
        return self._stats(lib().glia_b200_get_edge_stats, self.E, self.cfg.bimgs[i][1], i)

    # ---- candidates ----
    def init_candidates(self, rows=True):
        keys = np.zeros((self.Cn, 2), np.uint32)
        out = np.zeros((self.Cn, self.row_dim()), np.float64) if rows else None
        self._check(lib().glia_b200_init_candidates(self.h, _addr(keys), _addr(out)))
        return keys, out

    def score_candidates(self, forest, sal=None):
        """sal: numpy [C] (host mode) or device address; returns sal."""
        if sal is None:
            sal = np.zeros(self.Cn)
        self._check(lib().glia_b200_score_candidates(self.h, forest.h, _addr(sal)))
        return sal

    def merge_order_bc(self, forest, want_rows=False):
        M = max(self.R - 1, 0)
        order = np.zeros((M, 3), np.uint32)
        sal = np.zeros(M)
        rows = np.zeros((M, self.row_dim())) if want_rows else None
        nm, nf = C.c_int64(), C.c_int64()
        self._check(lib().glia_b200_merge_order_bc(self.h, forest.h, _addr(order), _addr(sal),
                                                   _addr(rows), C.byref(nm), C.byref(nf)))
        m = nm.value
        return dict(order=order[:m], sal=sal[:m], rows=None if rows is None else rows[:m],
                    n_featurized=nf.value)

    def merge_order_pb(self, type=1):
        """hmt/main_merge_order_pb.cxx: type 1 = approximate median (cfg.use_median), 2 = mean."""
        M = max(self.R - 1, 1)
        order = np.zeros((M, 3), np.uint32)
        sal = np.zeros(M)
        nm = C.c_int64()
        self._check(lib().glia_b200_merge_order_pb(self.h, type, _addr(order), _addr(sal), C.byref(nm)))
        return dict(order=order[:nm.value], sal=sal[:nm.value])

    def pre_merge(self, size_thr, rpb_thr=0.0, want_labels=True):
        """gadget/main_pre_merge.cxx (without the ITK relabel): order, saliencies and the renamed label volume."""
        M = max(self.R - 1, 1)
        order = np.zeros((M, 3), np.uint32)
        sal = np.zeros(M)
        thr = np.ascontiguousarray(size_thr, np.int32)
        out = np.zeros(self.shape, np.uint32) if want_labels else None
        nm = C.c_int64()
        self._check(lib().glia_b200_pre_merge(self.h, _addr(thr), len(thr), C.c_double(rpb_thr), _addr(order), _addr(sal),
                                              C.byref(nm), _addr(out)))
        return dict(order=order[:nm.value], sal=sal[:nm.value], labels=out)

    def sc_feat(self, other, pairs, norm_area=0.0, norm_len=0.0, use_log=False):
        """hmt/main_sc_feat.cxx: rows for (label in this section, label in `other`) pairs."""
        pairs = np.ascontiguousarray(pairs, np.uint32)
        rows = np.zeros((len(pairs), lib().glia_b200_sc_row_dim(self.h)))
        self._check(lib().glia_b200_sc_feat(self.h, other.h, _addr(pairs), len(pairs), C.c_double(norm_area), C.c_double(norm_len),
                                            int(use_log), _addr(rows)))
        return rows

    def bc_label(self, truth, order, use_f1=True, opt_split=False, global_opt=0, tweak=False, max_prec_drop=1.0):
        """hmt/main_bc_label_ri.cxx: (labels [n] in {-1 merge, +1 split}, merge scores, split scores)."""
        order = np.ascontiguousarray(order, np.uint32)
        if not self.device_pointers:
            truth = np.ascontiguousarray(truth, np.uint32)
        n = len(order)
        out = np.zeros(n, np.int32)
        ms, ss = np.zeros(n), np.zeros(n)
        self._check(lib().glia_b200_bc_label(self.h, _addr(truth), _addr(order), n, int(use_f1), int(opt_split), int(global_opt),
                                             int(tweak), C.c_double(max_prec_drop), _addr(out), _addr(ms), _addr(ss)))
        return out, ms, ss

    def segment_greedy(self, orders, merge_probs=None, region_probs=None, ignore_missing=True):
        """hmt/main_segment_greedy.cxx: resolve one or more merge trees; returns (labels, picks [(tree, node)])."""
        orders = [np.ascontiguousarray(o, np.uint32) for o in orders]
        n = len(orders)
        op = (C.c_void_p * n)(*[o.ctypes.data for o in orders])
        nm = (C.c_int64 * n)(*[len(o) for o in orders])

        def ptrs(lst):
            if lst is None:
                return None, None
            arrs = [np.ascontiguousarray(a, np.float64) for a in lst]
            return (C.c_void_p * n)(*[a.ctypes.data for a in arrs]), arrs
        mp, keep1 = ptrs(merge_probs)
        rp, keep2 = ptrs(region_probs)
        out = np.zeros(self.shape, np.uint32)
        picks = np.zeros((2 * max(self.R, 1) * n, 2), np.int32)
        npk = C.c_int64()
        self._check(lib().glia_b200_segment_greedy(self.h, n, op, nm, mp, rp, int(ignore_missing), _addr(out), _addr(picks),
                                                   C.byref(npk)))
        return out, picks[:npk.value]

    def bc_feat(self, order, sal=None, init_sal=1.0, sal_bias=0.0):
        order = np.ascontiguousarray(order, np.uint32)
        n = len(order)
        s = None if sal is None else np.ascontiguousarray(sal, np.float64)
        rows = np.zeros((n, self.row_dim(sal is not None)))
        self._check(lib().glia_b200_bc_feat(self.h, _addr(order), n, _addr(s), init_sal, sal_bias, _addr(rows)))
        return rows

    # ---- multi-GPU pieces (include/glia_b200.h "multi-GPU"; driven by glia_b200/parallel.py) ----
    def _counts(self):
        v = [C.c_int64() for _ in range(3)]
        self._check(lib().glia_b200_counts(self.h, *[C.byref(x) for x in v]))
        self.R, self.E, self.Cn = (x.value for x in v)

    def build_sharded(self, comm, labels, pb, mask=None):
        """The whole z-slab sequence (voxel pass, NCCL exchange, graph) in one stream-ordered library call."""
        if not self.device_pointers:
            labels = np.ascontiguousarray(labels, np.uint32)
            pb = np.ascontiguousarray(pb, np.float32)
            mask = None if mask is None else np.ascontiguousarray(mask, np.uint32)
            self._keep += [labels, pb, mask]
        self._check(lib().glia_b200_build_sharded(self.h, comm.h, _addr(labels), _addr(mask), _addr(pb)))
        self._counts()
        return self

    def bc_feat_sharded(self, comm, order, sal=None, init_sal=1.0, sal_bias=0.0):
        """bc_feat stage on a sharded build (collective): returns (first, rows) - this rank's share of the merges."""
        order = np.ascontiguousarray(order, np.uint32)
        n = len(order)
        s = None if sal is None else np.ascontiguousarray(sal, np.float64)
        share = (n + comm.world - 1) // comm.world
        rows = np.zeros((max(share, 1), self.row_dim(sal is not None)))
        first, count = C.c_int64(), C.c_int64()
        self._check(lib().glia_b200_bc_feat_sharded(self.h, comm.h, _addr(order), n, _addr(s), C.c_double(init_sal),
                                                    C.c_double(sal_bias), _addr(rows), C.byref(first), C.byref(count)))
        return first.value, rows[:count.value]

    def build_partial(self, labels, pb, mask=None):
        if not self.device_pointers:
            labels = np.ascontiguousarray(labels, np.uint32)
            pb = np.ascontiguousarray(pb, np.float32)
            mask = None if mask is None else np.ascontiguousarray(mask, np.uint32)
            self._keep += [labels, pb, mask]
        self._check(lib().glia_b200_build_partial(self.h, _addr(labels), _addr(mask), _addr(pb)))
        return self

    def raw_region_tables(self):
        """[(device address, word count)] for the add (u32), max (u32) and f64 parts."""
        p = [C.c_void_p() for _ in range(3)]
        n = [C.c_int64() for _ in range(3)]
        self._check(lib().glia_b200_raw_region_tables(self.h, *[C.byref(x) for x in p], *[C.byref(x) for x in n]))
        return [(x.value or 0, y.value) for x, y in zip(p, n)]

    def finalize_regions(self):
        r = C.c_int64()
        self._check(lib().glia_b200_finalize_regions(self.h, C.byref(r)))
        self.R = r.value
        return self.R

    def edge_interest(self, world):
        p, n = C.c_void_p(), C.c_int64()
        self._check(lib().glia_b200_edge_interest(self.h, world, C.byref(p), C.byref(n)))
        return p.value or 0, n.value

    def export_edges(self, world, self_rank=-1):
        counts = (C.c_int64 * world)()
        p, rb = C.c_void_p(), C.c_int()
        self._check(lib().glia_b200_export_edges(self.h, world, self_rank, counts, C.byref(p), C.byref(rb)))
        return list(counts), p.value or 0, rb.value

    def region_presence(self):
        p, n = C.c_void_p(), C.c_int64()
        self._check(lib().glia_b200_region_presence(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def index_regions(self):
        r = C.c_int64()
        self._check(lib().glia_b200_index_regions(self.h, C.byref(r)))
        self.R = r.value
        return r.value

    def gather_regions(self):
        self._check(lib().glia_b200_gather_regions(self.h))

    def table_load(self):
        o, c = C.c_int64(), C.c_int64()
        self._check(lib().glia_b200_table_load(self.h, C.byref(o), C.byref(c)))
        return o.value, c.value

    def import_edges(self, ptr, n, clear_first=True):
        self._check(lib().glia_b200_import_edges(self.h, C.c_void_p(int(ptr)) if ptr else None, n, int(clear_first)))

    def finalize_edges(self, lo, hi):
        self._check(lib().glia_b200_finalize_edges(self.h, lo, hi))
        self._counts()
        return self
