"""TEST INFRASTRUCTURE: ctypes bindings for the CPU oracle (oracle/libglia_oracle.so)
and, when present, the compiled unmodified reference (oracle/_ref/*.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs import this.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
MAX_IMG, MAX_THR = 4, 8

u32p = np.ctypeslib.ndpointer(np.uint32, flags="C")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C")
f64p = np.ctypeslib.ndpointer(np.float64, flags="C")


class GoCfg(C.Structure):
    _fields_ = [
        ("ndim", C.c_int), ("dims", C.c_int * 3),
        ("n_rimg", C.c_int), ("n_rlimg", C.c_int), ("n_bimg", C.c_int), ("n_thr", C.c_int),
        ("rimg", C.c_void_p * MAX_IMG), ("rbins", C.c_int * MAX_IMG),
        ("rlo", C.c_double * MAX_IMG), ("rhi", C.c_double * MAX_IMG),
        ("rlimg", C.c_void_p * MAX_IMG), ("rlbins", C.c_int * MAX_IMG),
        ("rllo", C.c_double * MAX_IMG), ("rlhi", C.c_double * MAX_IMG),
        ("bimg", C.c_void_p * MAX_IMG), ("bbins", C.c_int * MAX_IMG),
        ("blo", C.c_double * MAX_IMG), ("bhi", C.c_double * MAX_IMG),
        ("thr", C.c_double * MAX_THR),
        ("normalize", C.c_int), ("use_log", C.c_int), ("simple", C.c_int),
        ("use_hist", C.c_int), ("use_median", C.c_int), ("contour_only", C.c_int),
    ]


PRED_FN = C.CFUNCTYPE(C.c_double, C.POINTER(C.c_double), C.c_int, C.c_void_p)


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle"], check=True)


def build_ref():
    """Compile the unmodified reference headers into oracle/_ref (only where /root/reference exists)."""
    if not os.path.isdir("/root/reference/code"):
        raise RuntimeError("reference tree absent")
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "ref"], check=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(ORACLE_DIR, "libglia_oracle.so")
        if not os.path.exists(path):
            build_oracle()
        L = C.CDLL(path)
        L.go_build.restype = C.c_void_p
        L.go_build.argtypes = [C.POINTER(GoCfg), C.c_void_p, C.c_void_p, C.c_void_p]
        L.go_free.argtypes = [C.c_void_p]
        for n in ("go_num_regions", "go_num_edges", "go_num_candidates"):
            getattr(L, n).argtypes = [C.c_void_p]
        L.go_row_dim.argtypes = [C.c_void_p, C.c_int]
        L.go_get_regions.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.go_get_region_stats.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 5
        L.go_get_edges.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.go_get_edge_stats.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5
        L.go_init_candidates.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.go_greedy.restype = C.c_int
        L.go_greedy.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_size_t, C.POINTER(C.c_size_t), C.c_void_p]
        L.go_bc_feat.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_double,
                                 C.c_double, C.c_void_p]
        L.go_merge_order_pb.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.go_pre_merge.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p]
        L.go_sc_row_dim.argtypes = [C.c_void_p]
        L.go_sc_feat.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                 C.c_double, C.c_double, C.c_int, C.c_void_p]
        L.rf_oracle_load.restype = C.c_void_p
        L.rf_oracle_load.argtypes = [C.c_char_p]
        L.rf_oracle_from_arrays.restype = C.c_void_p
        L.rf_oracle_from_arrays.argtypes = [C.c_int] * 4 + [C.c_void_p] * 7
        L.rf_oracle_free.argtypes = [C.c_void_p]
        L.rf_oracle_info.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4
        L.rf_oracle_arrays.argtypes = [C.c_void_p] + [C.c_void_p] * 7
        L.rf_oracle_predict.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_int]
        L.rf_oracle_predict1.restype = C.c_double
        L.rf_oracle_predict1.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.rf_oracle_save.argtypes = [C.c_void_p, C.c_char_p]
        L.rf_oracle_pred_cb.restype = C.c_double
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class FeatureConfig:
    """Mirrors the CLI options of hmt/main_merge_order_bc.cxx:172-242."""

    def __init__(self, rimgs=(), rlimgs=(), bimgs=(), thr=(0.2, 0.5, 0.8), normalize=False,
                 use_log=False, simple=False, use_hist=False, use_median=False, contour_only=False):
        self.contour_only = contour_only  # TRegionMap(..., onlyContour=true)
        # each image entry: (array float32, bins, lo, hi)
        self.rimgs, self.rlimgs, self.bimgs = list(rimgs), list(rlimgs), list(bimgs)
        self.thr = list(thr)
        self.normalize, self.use_log, self.simple = normalize, use_log, simple
        # the reference's GLIA_USE_HISTOGRAM_AS_FEATS / GLIA_USE_MEDIAN_AS_FEATS builds (CMakeLists.txt:53-64)
        self.use_hist, self.use_median = use_hist, use_median


def default_config(pb, **kw):
    """--rbi pb --rbb 16 --rbl 0 --rbu 1 --bt 0.2 0.5 0.8."""
    return FeatureConfig(rimgs=[(pb, 16, 0.0, 1.0)], bimgs=[(pb, 16, 0.0, 1.0)], **kw)


class Oracle:
    def __init__(self, labels, pb, cfg, mask=None):
        self.labels = np.ascontiguousarray(labels, np.uint32)
        self.pb = np.ascontiguousarray(pb, np.float32)
        self.mask = None if mask is None else np.ascontiguousarray(mask, np.uint32)
        self.cfg = cfg
        shape = self.labels.shape  # numpy order: ([z,] y, x)
        g = GoCfg()
        g.ndim = len(shape)
        dims = list(shape[::-1]) + [1] * (3 - len(shape))
        for i in range(3):
            g.dims[i] = dims[i]
        self._keep = []
        for name, lst in (("r", cfg.rimgs), ("rl", cfg.rlimgs), ("b", cfg.bimgs)):
            setattr(g, "n_%simg" % name, len(lst))
            for i, (img, bins, lo, hi) in enumerate(lst):
                a = np.ascontiguousarray(img, np.float32)
                self._keep.append(a)
                getattr(g, name + "img")[i] = a.ctypes.data
                getattr(g, name + "bins")[i] = bins
                getattr(g, name + "lo")[i] = lo
                getattr(g, name + "hi")[i] = hi
        g.n_thr = len(cfg.thr)
        for i, t in enumerate(cfg.thr):
            g.thr[i] = t
        g.normalize, g.use_log, g.simple = int(cfg.normalize), int(cfg.use_log), int(cfg.simple)
        g.use_hist, g.use_median = int(getattr(cfg, "use_hist", False)), int(getattr(cfg, "use_median", False))
        g.contour_only = int(getattr(cfg, "contour_only", False))
        self.g = g
        self.h = lib().go_build(C.byref(g), _ptr(self.labels), _ptr(self.mask), _ptr(self.pb))
        self.R = lib().go_num_regions(self.h)
        self.E = lib().go_num_edges(self.h)
        self.Cn = lib().go_num_candidates(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().go_free(self.h)
            self.h = None

    def row_dim(self, with_sal=False):
        return lib().go_row_dim(self.h, int(with_sal))

    def regions(self):
        R = self.R
        lab, cnt, bor = (np.zeros(R, np.uint32) for _ in range(3))
        mn, mx = np.zeros((R, 3), np.int32), np.zeros((R, 3), np.int32)
        lib().go_get_regions(self.h, _ptr(lab), _ptr(cnt), _ptr(bor), _ptr(mn), _ptr(mx))
        return dict(label=lab, count=cnt, border=bor, bbmin=mn, bbmax=mx)

    def _stats(self, fn, n, bins, *pre):
        hist = np.zeros((n, bins), np.uint32)
        s, s2 = np.zeros(n), np.zeros(n)
        mn, mx = np.zeros(n, np.float32), np.zeros(n, np.float32)
        fn(self.h, *pre, _ptr(hist), _ptr(s), _ptr(s2), _ptr(mn), _ptr(mx))
        return dict(hist=hist, sum=s, sum2=s2, min=mn, max=mx)

    def region_stats(self, i, kind=0):
        bins = (self.cfg.rimgs if kind == 0 else self.cfg.rlimgs)[i][1]
        return self._stats(lib().go_get_region_stats, self.R, bins, kind, i)

    def edges(self):
        E, T = self.E, len(self.cfg.thr)
        a, b, c = (np.zeros(E, np.uint32) for _ in range(3))
        m = np.zeros(E, np.uint8)
        t = np.zeros((E, T), np.uint32)
        lib().go_get_edges(self.h, _ptr(a), _ptr(b), _ptr(c), _ptr(m), _ptr(t))
        return dict(a=a, b=b, count=c, mutual=m, thr=t)

    def edge_stats(self, i):
        return self._stats(lib().go_get_edge_stats, self.E, self.cfg.bimgs[i][1], i)

    def init_candidates(self, rows=True, swap=None):
        keys = np.zeros((self.Cn, 2), np.uint32)
        out = np.zeros((self.Cn, self.row_dim()), np.float64) if rows else None
        sw = None if swap is None else np.ascontiguousarray(swap, np.uint8)
        lib().go_init_candidates(self.h, _ptr(keys), _ptr(out), _ptr(sw))
        return keys, out

    def greedy(self, rf=None, fn=None, want_cands=False, init_swap=None):
        dim = self.row_dim()
        M = max(self.R - 1, 0)
        order = np.zeros((M, 3), np.uint32)
        sal = np.zeros(M)
        rows = np.zeros((M, dim))
        cb = PRED_FN(lambda p, d, ctx: fn(np.ctypeslib.as_array(p, (d,)))) if fn else None
        maxc = 40 * max(self.R, 1) + 64 if want_cands else 0
        ck = np.zeros((maxc, 3), np.uint32) if want_cands else None
        cr = np.zeros((maxc, dim)) if want_cands else None
        cs = np.zeros(maxc) if want_cands else None
        nc = C.c_size_t(0)
        sw = None if init_swap is None else np.ascontiguousarray(init_swap, np.uint8)
        m = lib().go_greedy(self.h, rf.h if rf else None, C.cast(cb, C.c_void_p) if cb else None,
                            None, _ptr(order), _ptr(sal), _ptr(rows), _ptr(ck), _ptr(cr), _ptr(cs),
                            maxc, C.byref(nc), _ptr(sw))
        res = dict(order=order[:m], sal=sal[:m], rows=rows[:m], n_cand=nc.value)
        if want_cands:
            n = min(nc.value, maxc)
            res.update(cand_keys=ck[:n], cand_rows=cr[:n], cand_scores=cs[:n])
        return res

    def merge_order_pb(self, type=1):
        """hmt/main_merge_order_pb.cxx: type 1 = approximate median (cfg.use_median), 2 = mean."""
        order = np.zeros((max(self.R, 1), 3), np.uint32)
        sal = np.zeros(max(self.R, 1))
        m = lib().go_merge_order_pb(self.h, type, _ptr(order), _ptr(sal))
        assert m >= 0
        return dict(order=order[:m], sal=sal[:m])

    def pre_merge(self, size_thr, rpb_thr=0.0):
        """gadget/main_pre_merge.cxx without the final ITK relabel."""
        order = np.zeros((max(self.R, 1), 3), np.uint32)
        sal = np.zeros(max(self.R, 1))
        thr = np.ascontiguousarray(size_thr, np.int32)
        out = np.zeros(self.labels.shape, np.uint32)
        m = lib().go_pre_merge(self.h, _ptr(thr), len(thr), rpb_thr, _ptr(order), _ptr(sal), _ptr(self.labels),
                               _ptr(self.mask), _ptr(out))
        assert m >= 0
        return dict(order=order[:m], sal=sal[:m], labels=out)

    def bc_feat(self, order, sal=None, init_sal=1.0, sal_bias=0.0):
        order = np.ascontiguousarray(order, np.uint32)
        n = len(order)
        s = None if sal is None else np.ascontiguousarray(sal, np.float64)
        rows = np.zeros((n, self.row_dim(sal is not None)))
        st = lib().go_bc_feat(self.h, _ptr(order), n, _ptr(s), init_sal, sal_bias, _ptr(rows))
        if st:
            raise ValueError("order references an unknown region")
        return rows


class OracleRF:
    def __init__(self, h):
        if not h:
            raise ValueError("rf_oracle: cannot load model")
        self.h = h
        v = [C.c_int() for _ in range(4)]
        lib().rf_oracle_info(h, *[C.byref(x) for x in v])
        self.ntree, self.nrnodes, self.nclass, self.mdim = (x.value for x in v)

    @classmethod
    def load(cls, path):
        return cls(lib().rf_oracle_load(path.encode()))

    @classmethod
    def from_arrays(cls, f, mdim):
        """f: dict with treemap [nt,nr,2], nodestatus, xbestsplit, bestvar, nodeclass [nt,nr],
        ndbigtree [nt], orig_labels [nclass]."""
        nt, nr = f["nodestatus"].shape
        a = {k: np.ascontiguousarray(f[k], np.float64 if k == "xbestsplit" else np.int32) for k in
             ("treemap", "nodestatus", "xbestsplit", "bestvar", "nodeclass", "ndbigtree", "orig_labels")}
        return cls(lib().rf_oracle_from_arrays(nt, nr, len(a["orig_labels"]), mdim, _ptr(a["treemap"]),
                                               _ptr(a["nodestatus"]), _ptr(a["xbestsplit"]),
                                               _ptr(a["bestvar"]), _ptr(a["nodeclass"]),
                                               _ptr(a["ndbigtree"]), _ptr(a["orig_labels"])))

    def arrays(self):
        nt, nr = self.ntree, self.nrnodes
        f = dict(treemap=np.zeros((nt, nr, 2), np.int32), nodestatus=np.zeros((nt, nr), np.int32),
                 xbestsplit=np.zeros((nt, nr)), bestvar=np.zeros((nt, nr), np.int32),
                 nodeclass=np.zeros((nt, nr), np.int32), ndbigtree=np.zeros(nt, np.int32),
                 orig_labels=np.zeros(self.nclass, np.int32))
        lib().rf_oracle_arrays(self.h, *[_ptr(f[k]) for k in ("treemap", "nodestatus", "xbestsplit",
                                                             "bestvar", "nodeclass", "ndbigtree",
                                                             "orig_labels")])
        return f

    def predict(self, X, label=-1, threads=0):
        X = np.ascontiguousarray(X, np.float64)
        n, d = X.shape
        votes = np.zeros((n, self.nclass), np.int32)
        prob = np.zeros(n)
        st = lib().rf_oracle_predict(self.h, _ptr(X), n, d, label, _ptr(votes), _ptr(prob), threads)
        if st:
            raise ValueError("rf_oracle_predict failed")
        return votes, prob

    def save(self, path):
        return lib().rf_oracle_save(self.h, path.encode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().rf_oracle_free(self.h)
            self.h = None


# ---------------------------------------------------------------- reference (oracle/_ref)
def _ref_variant(cfg):
    """Suffix of the reference build that has cfg's feature layout compiled in."""
    if cfg is None:
        return ""
    v = ("hist" if getattr(cfg, "use_hist", False) else "") + ("median" if getattr(cfg, "use_median", False) else "")
    return "_" + v if v else ""


def ref_available(ndim=2, cfg=None):
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libglia_ref%dd%s.so" % (ndim, _ref_variant(cfg))))


_ref = {}


def ref_lib(ndim, cfg=None):
    key = (ndim, _ref_variant(cfg))
    ndim_, ndim = ndim, key
    if ndim not in _ref:
        L = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libglia_ref%dd%s.so" % key))
        L.glia_ref_time.restype = C.c_double
        L.glia_ref_time_leaf_feats.restype = C.c_double
        for n in ("reg_label", "reg_count", "reg_border", "edge_a", "edge_b", "edge_count",
                  "cand_keys", "cand_rows", "cand_scores", "order", "sal", "merge_rows", "out_labels"):
            getattr(L, "glia_ref_n_" + n).restype = C.c_size_t
            getattr(L, "glia_ref_" + n).restype = C.c_void_p
        _ref[ndim] = L
    return _ref[ndim]


def _ref_get(L, name, dtype):
    n = getattr(L, "glia_ref_n_" + name)()
    p = getattr(L, "glia_ref_" + name)()
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(np.ctypeslib.as_ctypes_type(dtype))), (n,)).copy()


def rf_pred_callback(orf):
    """(function pointer, ctx) pair scoring a row with an OracleRF without going through Python."""
    fn = C.cast(lib().rf_oracle_pred_cb, C.c_void_p)
    return fn, C.c_void_p(orf.h)


def ref_run_pb(mode, labels, pb, mask=None, type=1, size_thr=(), rpb_thr=0.0):
    """The compiled reference's merge_order_pb (mode 4) / pre_merge (mode 5) call sequences (oracle/ref_driver.cxx)."""
    labels = np.ascontiguousarray(labels, np.uint32)
    pb = np.ascontiguousarray(pb, np.float32)
    ndim = labels.ndim
    L = ref_lib(ndim)
    dims = (C.c_int * 3)(*(list(labels.shape[::-1]) + [1] * (3 - ndim)))
    m = None if mask is None else np.ascontiguousarray(mask, np.uint32)
    thr = (C.c_int * max(len(size_thr), 1))(*size_thr)
    st = L.glia_ref_run_pb(C.c_int(mode), dims, _ptr(labels), _ptr(m), _ptr(pb), C.c_int(type), thr, C.c_int(len(size_thr)),
                           C.c_double(rpb_thr))
    assert st == 0
    out = dict(order=_ref_get(L, "order", np.uint32).reshape(-1, 3), sal=_ref_get(L, "sal", np.float64))
    if mode == 5:
        out["labels"] = _ref_get(L, "out_labels", np.uint32).reshape(labels.shape)
    return out


def sc_feat(o0, o1, pairs, norm_area=0.0, norm_len=0.0, use_log=False):
    """Section-classifier rows (hmt/main_sc_feat.cxx) from two Oracle objects built on the two sections."""
    pairs = np.ascontiguousarray(pairs, np.uint32)
    rows = np.zeros((len(pairs), lib().go_sc_row_dim(o0.h)))
    st = lib().go_sc_feat(o0.h, o1.h, _ptr(o0.labels), _ptr(o0.mask), _ptr(o1.labels), _ptr(o1.mask), _ptr(pairs), len(pairs),
                          norm_area, norm_len, int(use_log), _ptr(rows))
    assert st == 0
    return rows


def ref_sc_feat(lab0, pb0, lab1, pb1, pairs, mask0=None, mask1=None, bins=16, thr=(0.2, 0.5, 0.8), norm_area=0.0, norm_len=0.0,
                use_log=False):
    """The compiled reference's sc_feat call sequence (oracle/ref_driver.cxx glia_ref_sc_feat; 2D, --rbi pb)."""
    lab0, lab1 = np.ascontiguousarray(lab0, np.uint32), np.ascontiguousarray(lab1, np.uint32)
    pb0, pb1 = np.ascontiguousarray(pb0, np.float32), np.ascontiguousarray(pb1, np.float32)
    L = ref_lib(2)
    dims = (C.c_int * 3)(*(list(lab0.shape[::-1]) + [1]))
    m0 = None if mask0 is None else np.ascontiguousarray(mask0, np.uint32)
    m1 = None if mask1 is None else np.ascontiguousarray(mask1, np.uint32)
    pairs = np.ascontiguousarray(pairs, np.uint32)
    t = (C.c_double * max(len(thr), 1))(*thr)
    st = L.glia_ref_sc_feat(dims, _ptr(lab0), _ptr(m0), _ptr(pb0), _ptr(lab1), _ptr(m1), _ptr(pb1), C.c_int(bins), C.c_double(0.0),
                            C.c_double(1.0), C.c_int(len(thr)), t, C.c_double(norm_area), C.c_double(norm_len),
                            C.c_int(int(use_log)), _ptr(pairs), C.c_int(len(pairs)))
    assert st == 0
    return _ref_get(L, "merge_rows", np.float64).reshape(len(pairs), -1)


def ref_bc_label(labels, truth, order, mask=None, use_f1=True, opt_split=False, global_opt=0, tweak=False, max_prec_drop=1.0):
    """The compiled reference's bc_label_ri logic (oracle/ref_driver.cxx glia_ref_bc_label): (labels, merge scores, split
    scores)."""
    labels = np.ascontiguousarray(labels, np.uint32)
    truth = np.ascontiguousarray(truth, np.uint32)
    ndim = labels.ndim
    L = ref_lib(ndim)
    dims = (C.c_int * 3)(*(list(labels.shape[::-1]) + [1] * (3 - ndim)))
    m = None if mask is None else np.ascontiguousarray(mask, np.uint32)
    o = np.ascontiguousarray(order, np.uint32)
    st = L.glia_ref_bc_label(dims, _ptr(labels), _ptr(m), _ptr(truth), _ptr(o), C.c_int(len(o)), C.c_int(int(use_f1)),
                             C.c_int(int(opt_split)), C.c_int(global_opt), C.c_int(int(tweak)), C.c_double(max_prec_drop))
    assert st == 0
    return (_ref_get(L, "order", np.uint32).view(np.int32), _ref_get(L, "sal", np.float64),
            _ref_get(L, "cand_scores", np.float64))


def ref_segment_greedy(labels, orders, merge_probs=None, region_probs=None, mask=None, ignore=True):
    """The compiled reference's segment_greedy call sequence (oracle/ref_driver.cxx glia_ref_segment_greedy):
    returns (final labels, picks [(tree, node)], potentials of all trees' nodes concatenated)."""
    labels = np.ascontiguousarray(labels, np.uint32)
    ndim = labels.ndim
    L = ref_lib(ndim)
    dims = (C.c_int * 3)(*(list(labels.shape[::-1]) + [1] * (3 - ndim)))
    m = None if mask is None else np.ascontiguousarray(mask, np.uint32)
    n = len(orders)
    oa = [np.ascontiguousarray(o, np.uint32) for o in orders]
    op = (C.c_void_p * n)(*[a.ctypes.data for a in oa])
    nm = (C.c_int * n)(*[len(a) for a in oa])

    def ptrs(lst):
        if lst is None:
            return None, None
        arrs = [np.ascontiguousarray(a, np.float64) for a in lst]
        return (C.c_void_p * n)(*[a.ctypes.data for a in arrs]), arrs
    mp, k1 = ptrs(merge_probs)
    rp, k2 = ptrs(region_probs)
    st = L.glia_ref_segment_greedy(dims, _ptr(labels), _ptr(m), C.c_int(n), op, nm, mp, rp, C.c_int(int(ignore)))
    assert st == 0
    picks = _ref_get(L, "order", np.uint32).reshape(-1, 3)[:, :2].astype(np.int64)
    return _ref_get(L, "out_labels", np.uint32).reshape(labels.shape), picks, _ref_get(L, "sal", np.float64)


def ref_run(mode, labels, pb, cfg, mask=None, pred=None, order=None, sal=None, init_sal=1.0,
            sal_bias=0.0, threads=0, pred_c=None):
    """Run the compiled unmodified reference (oracle/ref_driver.cxx). mode: 0 RAG, 1 init
    candidates, 2 full greedy, 3 bc_feat stage."""
    labels = np.ascontiguousarray(labels, np.uint32)
    pb = np.ascontiguousarray(pb, np.float32)
    ndim = labels.ndim
    L = ref_lib(ndim, cfg)
    dims = (C.c_int * 3)(*(list(labels.shape[::-1]) + [1] * (3 - ndim)))
    keep = []

    def pack(lst):
        n = len(lst)
        arrs = [np.ascontiguousarray(x[0], np.float32) for x in lst]
        keep.extend(arrs)
        ptrs = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in arrs])
        bins = (C.c_int * max(n, 1))(*[x[1] for x in lst])
        lo = (C.c_double * max(n, 1))(*[x[2] for x in lst])
        hi = (C.c_double * max(n, 1))(*[x[3] for x in lst])
        return [C.c_int(n), ptrs, bins, lo, hi]

    thr = (C.c_double * max(len(cfg.thr), 1))(*cfg.thr)
    m = None if mask is None else np.ascontiguousarray(mask, np.uint32)
    cb = PRED_FN(lambda p, d, ctx: pred(np.ctypeslib.as_array(p, (d,)))) if pred else None
    o = None if order is None else np.ascontiguousarray(order, np.uint32)
    s = None if sal is None else np.ascontiguousarray(sal, np.float64)
    args = [C.c_int(mode), dims, _ptr(labels), _ptr(m), _ptr(pb)] + pack(cfg.rimgs) + \
        pack(cfg.rlimgs) + pack(cfg.bimgs) + [C.c_int(len(cfg.thr)), thr, C.c_int(int(cfg.normalize)),
                                              C.c_int(int(cfg.use_log)), C.c_int(int(cfg.simple)),
                                              (pred_c[0] if pred_c else (cb if cb else C.c_void_p(None))),
                                              (pred_c[1] if pred_c else C.c_void_p(None)), _ptr(o),
                                              C.c_int(0 if o is None else len(o)), _ptr(s),
                                              C.c_double(init_sal), C.c_double(sal_bias),
                                              C.c_int(threads)]
    st = L.glia_ref_run(*args)
    assert st == 0
    dim = L.glia_ref_dim()
    out = dict(dim=dim)
    for name, dt in (("reg_label", np.uint32), ("reg_count", np.uint32), ("reg_border", np.uint32),
                     ("edge_a", np.uint32), ("edge_b", np.uint32), ("edge_count", np.uint32),
                     ("cand_keys", np.uint32), ("cand_rows", np.float64), ("cand_scores", np.float64),
                     ("order", np.uint32), ("sal", np.float64), ("merge_rows", np.float64)):
        out[name] = _ref_get(L, name, dt)
    out["cand_keys"] = out["cand_keys"].reshape(-1, 3)
    out["order"] = out["order"].reshape(-1, 3)
    if dim:
        out["cand_rows"] = out["cand_rows"].reshape(-1, dim)
        out["merge_rows"] = out["merge_rows"].reshape(-1, dim)
    out["t_rmap"], out["t_feat"], out["t_pred"], out["t_total"] = (L.glia_ref_time(i) for i in range(4))
    out["nthreads"] = L.glia_ref_nthreads()
    return out
