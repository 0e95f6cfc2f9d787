#!/usr/bin/env python
"""bench.py — labels + boundary probabilities -> scored initial merge candidates.

A "step" is one pass of the hot path (voxel pass -> graph -> featurise -> RF score) over one
synthetic EM-like slice/volume per rank.  See DESIGN.md §6 for what each JSON key means.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c4|c1] [--impl reference]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# stdout carries exactly ONE JSON line.  Libraries write there too (NCCL prints its version banner on
# stdout when NCCL_DEBUG is VERSION or WARN), so file descriptor 1 points at stderr while the benchmark
# runs and is restored by emit() for the result line.
_STDOUT_FD = None


def quiet_stdout():
    global _STDOUT_FD
    if _STDOUT_FD is None:
        sys.stdout.flush()
        _STDOUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    global _STDOUT_FD
    sys.stdout.flush()
    if _STDOUT_FD is not None:
        try:  # C-level stdio of the libraries (NCCL's printf banner) is still buffered: push it to stderr now
            import ctypes
            ctypes.CDLL(None).fflush(None)
        except Exception:
            pass
        os.dup2(_STDOUT_FD, 1)
        os.close(_STDOUT_FD)
        _STDOUT_FD = None
    print(json.dumps(line), flush=True)

WORKLOADS = {
    # name: (shape in numpy order, cell, description)
    "c1": ((512, 512), 11, "2D 512x512, ~2.2k superpixels (BASELINE.json configs[0])"),
    "c2": ((4096, 4096), 13, "2D 4096x4096 synthetic EM slice, ~100k superpixels (BASELINE.json configs[1])"),
    "c3": ((2048, 2048), 13, "2D 2048x2048 slices, ~25k superpixels each (BASELINE.json configs[2])"),
    # the same slice / volume with a mask (1 % of the voxels masked out): the generic voxel kernel (TMA-staged tiles)
    "c2m": ((4096, 4096), 13, "2D 4096x4096 slice of configs[1] WITH a mask: generic voxel kernel"),
    "c4sm": ((512, 512, 512), 10, "3D 512^3 cut of configs[3] WITH a mask: generic voxel kernel"),
    "c4s": ((512, 512, 512), 10, "3D 512^3 synthetic EM volume, ~137k supervoxels (profiling-size cut of configs[3])"),
    "c4": ((1024, 1024, 1024), 10, "3D 1024^3 synthetic EM volume, ~1.09M supervoxels (BASELINE.json configs[3])"),
}
THR = (0.2, 0.5, 0.8)
RF_SWEEP = dict(n=10_000_000, d=50, ntree=255, desc="RF scoring sweep: 255-tree forest over 10M merge candidates "
                "with 50-dim bc features (BASELINE.json configs[4])")


def load_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of `kernel`, from the `ncu --set full` capture
    named in profiles/traffic.json (None = not captured for this workload)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f).get(workload, {}).get(kernel)
        return None if e is None else {"bytes": float(e["bytes"]), "capture": e["capture"]}
    except Exception:  # noqa: BLE001
        return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []
        self.t0 = self.t1 = None

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # samples inside the timed region; the sampler runs from before the warm-up steps (same load), so
        # when the region is shorter than nvidia-smi's sampling period the closest warm-up samples are used
        lines, window = self.lines, "timed region"
        if self.t0 is not None and self.t1 is not None:
            inside = [x for x in lines if self.t0 - 0.01 <= x[0] <= self.t1 + 0.03]
            if len(inside) >= 2:
                lines = inside
            else:
                lines = sorted(lines, key=lambda x: abs(x[0] - 0.5 * (self.t0 + self.t1)))[:5]
                window = "timed region shorter than the sampling period: nearest samples (warm-up, same load)"
        self.window = window
        for _, ln in lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons), "window": getattr(self, "window", "")}


FOREST_TRAIN = {2: ((384, 384), 9), 3: ((64, 96, 96), 8)}  # small synthetic image the bench forest is trained on


def train_forest(synth, keys, rows, shape, cell):
    """SURVEY §8d forest: 255 trees, 2 classes {-1, +1}, depth <= 20, trained with scikit-learn (seed 0) on the
    feature rows of a small synthetic image against a synthetic ground truth (blocks of 3^D cells are one true
    segment).  Falls back to the random-structure forest of round 1 when scikit-learn is unusable."""
    try:
        y = synth.gt_merge_labels(keys, shape, cell)
        f = synth.sklearn_forest(rows, y, ntree=255, seed=0, max_depth=20)
        kind = "sklearn-trained (seed 0, depth <= 20) on %d rows of a %s synthetic image" % (len(rows), "x".join(map(str, shape)))
    except Exception as e:  # noqa: BLE001
        f = synth.random_forest(255, rows.shape[1], max_depth=20, seed=0, thr_pool=rows, leaf_prob=0.3)
        kind = "random structure (scikit-learn unusable: %s)" % type(e).__name__
    info = {"kind": kind, "trees": int(len(f["ndbigtree"])), "nodes": int(f["ndbigtree"].sum()),
            "mean_path_len": round(synth.mean_path_length(f, rows), 2)}
    return f, info


def make_forest(api, synth, torch, dim, ndim, device):
    """The bench forest, trained on rows featurised by the GPU path itself (the product never touches oracle/)."""
    shape, cell = FOREST_TRAIN[ndim]
    lab, pb = synth.voronoi_volume(shape, cell, seed=99)
    spec = [(pb, 16, 0.0, 1.0)]
    g = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec, thr=THR), device=device).build(lab, pb)
    keys, rows = g.init_candidates()
    g.close()
    assert rows.shape[1] == dim
    f, info = train_forest(synth, keys, rows, shape, cell)
    FOREST_INFO.clear(); FOREST_INFO.update(info)
    return f, api.Forest.from_arrays(f, dim, device=device)


FOREST_INFO = {}


# ------------------------------------------------------------------ reference arm (CPU)
def _ref_worker(job):
    """One bounded sample through the UNMODIFIED reference (oracle/_ref, compiled from
    /root/reference/code) with the oracle's classForest restatement as the predictor: TRegionMap
    build + TBoundaryTable::init (fBcFeat + fBcPred for every initial candidate)."""
    shape, cell, seed, forest = job
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    from glia_b200 import synth
    lab, pb = synth.voronoi_volume(shape, cell, seed=seed)
    spec = [(pb, 16, 0.0, 1.0)]
    cfg = ol.FeatureConfig(rimgs=spec, bimgs=spec, thr=THR)
    use_ref = ol.ref_available(len(shape))
    t0 = time.perf_counter()
    if use_ref:
        orf = ol.OracleRF.from_arrays(forest, int(forest["bestvar"].max()))
        r = ol.ref_run(1, lab, pb, cfg, pred_c=ol.rf_pred_callback(orf), threads=1)
        ncand = len(r["cand_keys"])
    else:  # the oracle port (only where oracle/_ref is absent)
        orf = ol.OracleRF.from_arrays(forest, int(forest["bestvar"].max()))
        o = ol.Oracle(lab, pb, cfg)
        _, rows = o.init_candidates()
        orf.predict(rows)
        ncand = len(rows)
    return time.perf_counter() - t0, int(np.prod(shape)), ncand, "reference" if use_ref else "port"


def cpu_reference(wl, forest, steps, warmup, procs=None):
    """Times the reference on the host cores: `procs` independent processes (the reference's own
    data-parallel mechanism is one process per image, gadget/python/script_util.py:6-23), each
    step = one bounded sample per process."""
    import multiprocessing as mp
    shape, cell, _ = WORKLOADS[wl]
    nd = len(shape)
    sample = (1024, 1024) if nd == 2 else (96, 96, 96)
    procs = procs or os.cpu_count() or 1
    ctx = mp.get_context("spawn")
    times, vox, cand, kind = [], 0, 0, "reference"
    with ctx.Pool(procs) as pool:
        for s in range(warmup + steps):
            jobs = [(sample, cell, 5000 + 100 * s + i, forest) for i in range(procs)]
            res = pool.map(_ref_worker, jobs)
            dt = max(r[0] for r in res)  # the processes run side by side; input synthesis is not timed
            if s >= warmup:
                times.append(dt)
                vox += sum(r[1] for r in res); cand += sum(r[2] for r in res)
            kind = res[0][3]
    total = sum(times)
    return dict(mvox_s=vox / total / 1e6, cand_s=cand / total, ms_per_step=1e3 * total / max(len(times), 1),
                cores=procs, kind=kind,
                sample="%d processes x one %s synthetic sample (cell %d) per step: TRegionMap + TBoundaryTable::init "
                       "(fBcFeat+RF) per sample" % (procs, "x".join(map(str, sample)), cell))


def forest_arrays_cpu(synth, dim, ndim):
    """Same forest recipe as make_forest, trained on oracle rows (reference arm: no GPU code)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    shape, cell = FOREST_TRAIN[ndim]
    lab, pb = synth.voronoi_volume(shape, cell, seed=99)
    spec = [(pb, 16, 0.0, 1.0)]
    o = ol.Oracle(lab, pb, ol.FeatureConfig(rimgs=spec, bimgs=spec, thr=THR))
    keys, rows = o.init_candidates()
    assert rows.shape[1] == dim
    f, info = train_forest(synth, keys, rows, shape, cell)
    FOREST_INFO.clear(); FOREST_INFO.update(info)
    return f


def row_dim(ndim):
    return (11 + 4 * 3 + 7 + 5) + 3 * (ndim + 4 + 2 * 3 + 5 + 5)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    # default = the largest single-GPU configuration of BASELINE.json (configs[3], north_star's target): one
    # 1024^3 volume on one GPU; with N > 1 ranks the same volume z-slab sharded over NCCL (strong scaling)
    ap.add_argument("--workload", default=os.environ.get("GLIA_BENCH_WORKLOAD", "c4"),
                    choices=sorted(WORKLOADS) + ["c5"])  # c2m / c4sm: masked inputs -> voxel_pass_generic
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    quiet_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "c5":
        import bench_modes
        bench_modes.emit = emit
        return bench_modes.run_rf_sweep(args, rank, world, local_rank)
    shape, cell, desc = WORKLOADS[args.workload]
    nd = len(shape)
    nvox = int(np.prod(shape))
    dim = row_dim(nd)
    from glia_b200 import synth

    if args.impl == "reference":
        if rank != 0:
            return 0
        forest = forest_arrays_cpu(synth, dim, nd)
        steps, warmup = max(1, min(args.steps, 200)), max(0, min(args.warmup, 20))  # ~0.75 s of host work per step
        r = cpu_reference(args.workload, forest, steps, warmup)
        line = {"impl": "reference", "metric": "Mvoxels/s, labels+pb -> scored initial merge candidates",
                "value": r["mvox_s"], "unit": "Mvoxels/s", "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
                "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u32 labels / f32 pb / f64 features", "data": "synthetic",
                "config": {"workload": args.workload + ": " + desc, "sample": r["sample"], "forest": dict(FOREST_INFO),
                           "same_config": False,
                           "same_config_note": "the reference's point lists do not fit the full volume (SURVEY §5): each "
                                               "host process runs a bounded tile of the same synthetic recipe"},
                "candidates_per_s": r["cand_s"],
                "cpu_baseline": {"value": r["mvox_s"], "unit": "Mvoxels/s", "cores": r["cores"], "kind": r["kind"],
                                 "sample": r["sample"]},
                "e2e": {"value": r["mvox_s"], "unit": "Mvoxels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return 0

    import torch
    import torch.distributed as dist
    from glia_b200 import api
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if nd == 3 and world > 1:
        import bench_modes
        bench_modes.emit = emit
        return bench_modes.run_slab(args, rank, world, local_rank, dist)
    L = api.lib()
    import ctypes as C
    stream = torch.cuda.current_stream().cuda_stream

    # ---- synthetic inputs resident in HBM: NSETS distinct slices/volumes per rank, rotated so that
    # consecutive steps never see L2-resident inputs (2D) / one volume >> L2 (3D)
    nsets = 4 if nd == 2 else 1
    dims = (C.c_int * 3)(*(list(shape[::-1]) + [1] * (3 - nd)))
    sets, ncells = [], C.c_uint64(0)
    for i in range(nsets):
        lab = torch.empty(nvox, dtype=torch.int32, device=dev)
        pb = torch.empty(nvox, dtype=torch.float32, device=dev)
        st = L.glia_b200_synth_voronoi(nd, dims, cell, 1234 + 1000 * rank + i, lab.data_ptr(), pb.data_ptr(),
                                       C.c_void_p(stream), C.byref(ncells))
        assert st == 0
        sets.append((lab, pb))
    torch.cuda.synchronize()
    label_bound = int(ncells.value) + 1
    masked = args.workload.endswith("m")
    mask_dev = None
    if masked:
        g = torch.Generator(device=dev); g.manual_seed(5)
        mask_dev = (torch.rand(nvox, device=dev, generator=g) > 0.01).to(torch.int32)

    forest_arrays, forest = make_forest(api, synth, torch, dim, nd, local_rank)

    def new_builder(pb_ptr, device_pointers):
        spec = [(pb_ptr, 16, 0.0, 1.0)]
        b = api.MergeTreeBuilder(shape, api.FeatureConfig(rimgs=spec, bimgs=spec, thr=THR), device=local_rank,
                                 device_pointers=device_pointers, label_bound=label_bound,
                                 stream=stream if device_pointers else None)
        for key in ("tz", "slots"):  # kernel-shape experiments of the harness (the library reads no environment)
            if os.environ.get("GLIA_BENCH_" + key.upper()):
                b.tune(key, int(os.environ["GLIA_BENCH_" + key.upper()]))
        return b

    # ONE handle for the stream of inputs (its tables are scratch); only the inputs rotate
    builders = [new_builder(sets[0][1].data_ptr(), True)]
    sal_dev = torch.empty(64, dtype=torch.float64, device=dev)

    def step(i):
        nonlocal sal_dev
        b = builders[0]
        lab, pb = sets[i % nsets]
        b.set_image(0, 0, pb.data_ptr()); b.set_image(2, 0, pb.data_ptr())
        b.build(lab.data_ptr(), pb.data_ptr(), mask_dev.data_ptr() if mask_dev is not None else None)
        if sal_dev.numel() < b.Cn:
            sal_dev = torch.empty(int(b.Cn * 1.2) + 64, dtype=torch.float64, device=dev)
        b.score_candidates(forest, sal_dev.data_ptr())
        return b

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    sampler.start()
    for i in range(max(args.warmup, nsets)):  # every rotating input/handle is touched before timing
        step(i)
    launches0 = sum(b.launch_count() for b in builders)
    kern_ms = np.zeros(5)
    cands = 0
    barrier()
    sampler.mark_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        b = step(args.warmup + i)
        kern_ms += np.array(b.last_ms())
        cands += b.Cn
    e1.record()
    barrier()
    sampler.mark_end()
    clocks = sampler.stop()
    elapsed_ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([elapsed_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
        c = torch.tensor([cands], device=dev, dtype=torch.float64)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        cands_all = float(c.item())
    else:
        cands_all = float(cands)
    launches = sum(b.launch_count() for b in builders) - launches0
    kern_ms /= args.steps
    R, E, Cn = b.R, b.E, b.Cn
    value = world * nvox * args.steps / (elapsed_ms * 1e-3) / 1e6

    # ---- e2e: host (pinned) buffers through the C ABI, H2D + D2H inside the timed region
    e2e = None
    if not args.no_e2e and not masked:
        # two slices in flight per GPU (two handles, each with its own stream, driven by two host
        # threads): the upload of slice i+1 overlaps graph/featurise/RF of slice i.  Every step still
        # copies its inputs from pinned host memory and reads its saliencies back.
        # A 3D volume is uploaded from ONE pinned host copy by both in-flight handles (8.6 GB of pinned memory per
        # copy at 1024^3; the source of an upload is read-only); 2D slices rotate through distinct host buffers.
        nin = int(os.environ.get("GLIA_BENCH_INFLIGHT", "2"))
        host = []
        for j in range(nin):
            if nd == 3 and j > 0:
                hl, hp = host[0][4], host[0][5]
            else:
                hl = torch.empty(nvox, dtype=torch.int32, pin_memory=True)
                hp = torch.empty(nvox, dtype=torch.float32, pin_memory=True)
                hl.copy_(sets[j % nsets][0]); hp.copy_(sets[j % nsets][1])
            hln, hpn = hl.numpy().view(np.uint32).reshape(shape), hp.numpy().reshape(shape)
            host.append((hln, hpn, new_builder(hpn, False),
                         torch.empty(int(Cn * 1.5) + 64, dtype=torch.float64).pin_memory().numpy(), hl, hp))
        n_e2e = max(2, min(args.steps, 6))
        n_e2e = (n_e2e + nin - 1) // nin * nin
        for hln, hpn, hb, hsal, _, _ in host:
            for i in range(2 if nd == 2 else 1):
                hb.build(hln, hpn); hb.score_candidates(forest, hsal)
        errs = []
        # 3D: the two in-flight volumes take turns on the PCIe link (an upload runs at full rate while the other
        # volume is being featurised and scored) instead of sharing it
        turn = threading.Lock() if nd == 3 else None

        def worker(j):
            hln, hpn, hb, hsal, _, _ = host[j]
            try:
                torch.cuda.set_device(local_rank)
                for i in range(j, n_e2e, nin):
                    if turn is not None:
                        with turn:
                            hb.build(hln, hpn)
                    else:
                        hb.build(hln, hpn)
                    hb.score_candidates(forest, hsal)
            except BaseException as e:  # noqa: BLE001
                errs.append(e)

        barrier()
        t0 = time.perf_counter()
        ths = [threading.Thread(target=worker, args=(j,)) for j in range(nin)]
        [t.start() for t in ths]
        [t.join() for t in ths]
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if errs:
            raise errs[0]
        if world > 1:
            t = torch.tensor([dt], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": world * nvox * n_e2e / dt / 1e6, "unit": "Mvoxels/s", "h2d_bytes_per_step": 8 * nvox,
               "d2h_bytes_per_step": 8 * int(host[0][2].Cn), "steps": n_e2e, "ms_per_step": 1e3 * dt / n_e2e,
               "in_flight": nin}
        for x in host:
            x[2].close()
        # the ceiling of this leg: the same pinned buffers copied to the device with no kernel at all
        dl, dp = sets[0][0], sets[0][1]
        barrier()
        t0 = time.perf_counter()
        for i in range(n_e2e):
            dl.copy_(host[i % nin][4], non_blocking=True); dp.copy_(host[i % nin][5], non_blocking=True)
        torch.cuda.synchronize()
        dth = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dth], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dth = float(t.item())
        e2e["h2d_only_ms_per_step"] = 1e3 * dth / n_e2e
        e2e["h2d_only_gb_s_per_gpu"] = 8 * nvox * n_e2e / dth / 1e9
        e2e["frac_of_h2d_ceiling"] = dth / dt

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    peak, peak_src = load_peaks()
    # SURVEY §8d.  Voxel pass: label u32 + pb f32 read once.  RF scoring: 8*D + 8 bytes per candidate (row in,
    # probability out) + the forest once (16-byte nodes).  "roofline" is the kernel with the longer launch in
    # this run (RF scoring at C2, the voxel pass in 3D); "roofline_kernels" carries both.
    def roof(kernel, alg, ms):
        a = alg / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        tr = load_traffic(args.workload, kernel)
        return {"bound": "hbm", "kernel": kernel, "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak,
                "traffic": tr["bytes"] if tr else None, "traffic_capture": tr["capture"] if tr else None,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": alg, "launch_ms": ms}
    roofs = [roof("voxel_pass_generic" if masked else "voxel_pass_fast", (12.0 if masked else 8.0) * nvox, kern_ms[0]),
             roof("rf_score", float(Cn) * (8 * dim + 8) + 16.0 * float(forest_arrays["ndbigtree"].sum()),
                  kern_ms[3])]
    dominant = max(roofs, key=lambda r: r["launch_ms"])
    line = {"metric": "Mvoxels/s, labels+pb -> scored initial merge candidates", "value": value, "unit": "Mvoxels/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": elapsed_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32 labels / f32 pb / f64 features", "data": "synthetic",
            "config": {"workload": args.workload + ": " + desc, "per_gpu": "one %s per step" % "x".join(map(str, shape)),
                       "regions": int(R), "directed_boundaries": int(E), "candidates": int(Cn), "row_dim": dim,
                       "forest": dict(FOREST_INFO),
                       "l2": ("rotating %d distinct inputs (%d MB) through one handle; tables re-zeroed each step" % (nsets, nsets * 8 * nvox >> 20))
                       if nd == 2 else "input 8.6 GB >> 126 MB L2", "sharding": "independent inputs per rank, no collective"},
            "candidates_per_s": cands_all / (elapsed_ms * 1e-3),
            "rf_node_visits_per_s": (float(Cn) * FOREST_INFO.get("trees", 0) * FOREST_INFO.get("mean_path_len", 0.0) /
                                     (kern_ms[3] * 1e-3)) if kern_ms[3] > 0 else None,
            "stage_ms": {"table_init": kern_ms[4], "voxel_pass": kern_ms[0], "graph": kern_ms[1], "featurize": kern_ms[2],
                         "rf_score": kern_ms[3]},
            "roofline": dominant, "roofline_kernels": roofs,
            "clocks": clocks, "gpu_launches": int(launches), "e2e": e2e}
    if not args.no_cpu_baseline and world == 1:
        r = cpu_reference(args.workload, forest_arrays, 1, 0)
        line["cpu_baseline"] = {"value": r["mvox_s"], "unit": "Mvoxels/s", "cores": r["cores"], "kind": r["kind"],
                                "sample": r["sample"], "candidates_per_s": r["cand_s"]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
