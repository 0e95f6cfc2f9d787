"""bench.py modes beyond the per-rank slice loop:
  run_slab      --workload c4|c4s with N > 1 ranks: ONE volume, z-slab sharded with a one-plane halo,
                partial tables combined over NCCL (glia_b200/parallel.py) -> strong scaling
  run_rf_sweep  --workload c5: 255-tree forest over 10M x 50 candidate rows, rows sharded
                across ranks (independent candidates, forest replicated, no collective)
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def emit(line):  # replaced by bench.py's emit (which restores the real stdout first)
    print(json.dumps(line), flush=True)


def _max_over_ranks(torch, dist, world, dev, x):
    if world == 1:
        return float(x)
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _sum_over_ranks(torch, dist, world, dev, x):
    if world == 1:
        return float(x)
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def run_slab(args, rank, world, local_rank, dist):
    import torch
    import bench as B
    from glia_b200 import api, parallel as par, synth
    shape, cell, desc = B.WORKLOADS[args.workload]
    nz, ny, nx = shape
    dev = torch.device("cuda", local_rank)
    L = api.lib()
    stream = torch.cuda.current_stream().cuda_stream
    comm = par.TorchComm()
    z0, z1, hl, hh = par.slab_with_halo(nz, world, rank)
    nplanes = (z1 + hh) - (z0 - hl)
    nloc = nplanes * ny * nx
    lab = torch.empty(nloc, dtype=torch.int32, device=dev)
    pb = torch.empty(nloc, dtype=torch.float32, device=dev)
    dims = (C.c_int * 3)(nx, ny, nz)
    ncells = C.c_uint64(0)
    st = L.glia_b200_synth_voronoi_slab(dims, cell, 1234, z0 - hl, nplanes, lab.data_ptr(), pb.data_ptr(),
                                        C.c_void_p(stream), C.byref(ncells))
    assert st == 0
    torch.cuda.synchronize()
    label_bound = int(ncells.value) + 1
    dim = B.row_dim(3)
    forest_arrays, forest = B.make_forest(api, synth, torch, dim, 3, local_rank)
    spec = [(pb.data_ptr(), 16, 0.0, 1.0)]
    cfg = api.FeatureConfig(rimgs=spec, bimgs=spec, thr=B.THR)
    b = par.slab_builder(shape, cfg, comm, label_bound, device=local_rank, device_pointers=True, stream=stream)
    sal = torch.empty(64, dtype=torch.float64, device=dev)
    # exchange: the library's own stream-ordered NCCL sequence (glia_b200_build_sharded) unless GLIA_BENCH_EXCHANGE=torch
    # asks for the torch.distributed-driven one (glia_b200/parallel.py, also what the CPU / thread-emulated tests drive)
    native = None
    if os.environ.get("GLIA_BENCH_EXCHANGE", "native") == "native":
        native = par.native_comm(dist, world, rank, local_rank)

    def build(bb, l, p, tm=None):
        if native is None:
            return par.build_sharded(comm, bb, l, p, timings=tm)
        bb.build_sharded(native, l, p)
        if tm is not None:
            ms = bb.last_ms()
            tm.update(voxel_ms=ms[0] + ms[4], exchange_ms=native.last_exchange_ms(), graph_ms=ms[1],
                      exchange_detail_ms={"native (one stream-ordered call)": native.last_exchange_ms()},
                      sent_records=0, record_bytes=0)

    def step(tm=None):
        nonlocal sal
        build(b, lab.data_ptr(), pb.data_ptr(), tm)
        if sal.numel() < b.Cn:
            sal = torch.empty(int(b.Cn * 1.2) + 64, dtype=torch.float64, device=dev)
        if b.Cn:
            b.score_candidates(forest, sal.data_ptr())

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    sampler = B.ClockSampler(local_rank)
    sampler.start()
    for _ in range(args.warmup):
        step()
    launches0 = b.launch_count()
    acc = dict(voxel_ms=0.0, exchange_ms=0.0, graph_ms=0.0)
    detail = {}
    feat_ms = rf_ms = 0.0
    sent = 0
    barrier()
    sampler.mark_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        tm = {}
        step(tm)
        for k in acc:
            acc[k] += tm[k]
        for k, v in tm["exchange_detail_ms"].items():
            detail[k] = detail.get(k, 0.0) + v
        sent += tm["sent_records"] * tm["record_bytes"]
        ms = b.last_ms()
        feat_ms += ms[2]; rf_ms += ms[3]
    e1.record()
    barrier()
    sampler.mark_end()
    clocks = sampler.stop()
    elapsed_ms = _max_over_ranks(torch, dist, world, dev, e0.elapsed_time(e1))
    cands = _sum_over_ranks(torch, dist, world, dev, b.Cn)
    # per-rank picture (who waits for whom): candidates owned and mean stage times of every rank
    mine = torch.tensor([float(b.Cn), acc["voxel_ms"] / args.steps, acc["exchange_ms"] / args.steps, acc["graph_ms"] / args.steps,
                         feat_ms / args.steps, rf_ms / args.steps], device=dev, dtype=torch.float64)
    allr = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allr, mine)
    per_rank = [[round(float(x), 3) for x in t.tolist()] for t in allr]
    sent_all = _sum_over_ranks(torch, dist, world, dev, sent / args.steps)
    vox_kernel_ms = _max_over_ranks(torch, dist, world, dev, b.last_ms()[0])
    launches = b.launch_count() - launches0
    nvox = nz * ny * nx
    value = nvox * args.steps / (elapsed_ms * 1e-3) / 1e6
    # ---- e2e: this rank's slab from pinned host memory through the same path
    e2e = None
    if not args.no_e2e and 8 * nloc <= 5 * 2 ** 30:
        hl_t = torch.empty(nloc, dtype=torch.int32, pin_memory=True); hl_t.copy_(lab)
        hp_t = torch.empty(nloc, dtype=torch.float32, pin_memory=True); hp_t.copy_(pb)
        hln = hl_t.numpy().view(np.uint32).reshape(nplanes, ny, nx)
        hpn = hp_t.numpy().reshape(nplanes, ny, nx)
        hspec = [(hpn, 16, 0.0, 1.0)]
        hb = par.slab_builder(shape, api.FeatureConfig(rimgs=hspec, bimgs=hspec, thr=B.THR), comm, label_bound,
                              device=local_rank)
        hsal = torch.empty(int(b.Cn * 1.5) + 64, dtype=torch.float64).pin_memory().numpy()
        n_e2e = max(2, min(args.steps, 3))
        for i in range(1 + n_e2e):
            if i == 1:
                barrier()
                t0 = time.perf_counter()
            build(hb, hln, hpn)
            if hb.Cn:
                hb.score_candidates(forest, hsal)
        torch.cuda.synchronize()
        dt = _max_over_ranks(torch, dist, world, dev, time.perf_counter() - t0)
        e2e = {"value": nvox * n_e2e / dt / 1e6, "unit": "Mvoxels/s", "h2d_bytes_per_step": 8 * nloc * world,
               "d2h_bytes_per_step": int(8 * cands), "steps": n_e2e, "ms_per_step": 1e3 * dt / n_e2e}
        hb.close()
        # the ceiling of this leg: every rank copies the same pinned slab to its GPU, no kernel at all
        barrier()
        t0 = time.perf_counter()
        for i in range(n_e2e):
            lab.copy_(hl_t, non_blocking=True); pb.copy_(hp_t, non_blocking=True)
        torch.cuda.synchronize()
        dth = _max_over_ranks(torch, dist, world, dev, time.perf_counter() - t0)
        e2e["h2d_only_ms_per_step"] = 1e3 * dth / n_e2e
        e2e["h2d_only_gb_s_per_gpu"] = 8 * nloc * n_e2e / dth / 1e9
        e2e["frac_of_h2d_ceiling"] = dth / dt
    if rank == 0:
        peak, peak_src = B.load_peaks()
        alg = 8.0 * (z1 - z0) * ny * nx  # this rank's launch: owned voxels x (u32 label + f32 pb)
        achieved = alg / (vox_kernel_ms * 1e-3) / 1e9
        line = {"metric": "Mvoxels/s, labels+pb -> scored initial merge candidates", "value": value,
                "unit": "Mvoxels/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "u32 labels / f32 pb / f64 features", "data": "synthetic",
                "config": {"workload": args.workload + ": " + desc,
                           "sharding": "z-slabs of %d planes + 1-plane label halo; all-reduce of dense region tables, "
                                       "boundary records exchanged and reduced by label pair (NCCL)" % (z1 - z0),
                           "exchange": "native: glia_b200_build_sharded (C++, stream-ordered)" if native is not None
                                       else "torch.distributed collectives driven from glia_b200/parallel.py",
                           "row_dim": dim, "candidates": int(cands),
                           "forest": dict(B.FOREST_INFO),
                           "l2": "slab input %.2f GB >> 126 MB L2" % (8 * nloc / 1e9)},
                "candidates_per_s": cands * args.steps / (elapsed_ms * 1e-3),
                "stage_ms_rank0": {"voxel_pass+table_init": acc["voxel_ms"] / args.steps,
                                   "exchange": acc["exchange_ms"] / args.steps, "graph": acc["graph_ms"] / args.steps,
                                   "featurize": feat_ms / args.steps, "rf_score": rf_ms / args.steps},
                "exchange_detail_ms_rank0": {k: v / args.steps for k, v in detail.items()},
                "exchange_bytes_per_step": sent_all,
                "per_rank": {"columns": ["candidates", "voxel_ms", "exchange_ms", "graph_ms", "featurize_ms", "rf_ms"], "rows": per_rank},
                "roofline": {"bound": "hbm", "kernel": "voxel_pass_fast", "achieved": achieved, "peak": peak,
                             "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": alg},
                "rf_node_visits_per_s_rank0": (float(b.Cn) * B.FOREST_INFO.get("trees", 0) * B.FOREST_INFO.get("mean_path_len", 0.0) /
                                               (rf_ms / args.steps * 1e-3)) if rf_ms > 0 else None,
                "clocks": clocks, "gpu_launches": int(launches), "e2e": e2e}
        if not args.no_cpu_baseline:
            r = B.cpu_reference(args.workload, forest_arrays, 1, 0)
            line["cpu_baseline"] = {"value": r["mvox_s"], "unit": "Mvoxels/s", "cores": r["cores"], "kind": r["kind"],
                                    "sample": r["sample"], "candidates_per_s": r["cand_s"]}
        emit(line)
    b.close()
    if native is not None:
        native.close()
    dist.destroy_process_group()
    return 0


def _sweep_forest(synth, d, ntree):
    rng = np.random.default_rng(7)
    pool = rng.random((4096, d))
    return synth.random_forest(ntree, d, max_depth=20, seed=0, thr_pool=pool, leaf_prob=0.3), pool


def run_rf_sweep(args, rank, world, local_rank):
    import bench as B
    from glia_b200 import synth
    S = B.RF_SWEEP
    n_all, d = S["n"], S["d"]
    metric = "merge candidates RF-scored per second (255 trees, 50 features)"
    f, pool = _sweep_forest(synth, d, S["ntree"])
    if args.impl == "reference":
        if rank != 0:
            return 0
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        orf = ol.OracleRF.from_arrays(f, d)
        n = 200_000
        rng = np.random.default_rng(1)
        X = rng.random((n, d))
        X[: n // 20] = pool[rng.integers(0, len(pool), n // 20)]  # rows sitting exactly on thresholds
        steps, warmup = max(1, min(args.steps, 200)), max(0, min(args.warmup, 20))  # ~0.2 s of host work per step
        ts = []
        for s in range(warmup + steps):
            t0 = time.perf_counter()
            orf.predict(X, threads=os.cpu_count() or 1)  # OpenMP over rows: the fair parallel form of main_pred_rf.cxx:36's serial loop
            if s >= warmup:
                ts.append(time.perf_counter() - t0)
        v = n * len(ts) / sum(ts)
        cores = os.cpu_count() or 1
        sample = "%d rows x %d features per step, classForest restatement (oracle/rf_oracle.c), OpenMP over rows" % (n, d)
        emit({"impl": "reference", "metric": metric, "value": v, "unit": "candidates/s", "n_gpus": args.gpus,
                          "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * sum(ts) / len(ts),
                          "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                          "data": "synthetic", "config": {"workload": "c5: " + S["desc"], "sample": sample},
                          "cpu_baseline": {"value": v, "unit": "candidates/s", "cores": cores, "kind": "port", "sample": sample},
                          "e2e": {"value": v, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
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
    n = n_all // world + (1 if rank < n_all % world else 0)
    g = torch.Generator(device=dev); g.manual_seed(7 + rank)
    X = torch.rand((n, d), dtype=torch.float64, device=dev, generator=g)
    tp = torch.from_numpy(pool).to(dev)
    idx = torch.randint(0, len(pool), (n // 20,), device=dev, generator=g)
    X[: n // 20] = tp[idx]  # 5 % of the rows sit exactly on thresholds (x <= t ties)
    prob = torch.empty(n, dtype=torch.float64, device=dev)
    forest = api.Forest.from_arrays(f, d, device=local_rank)
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        forest.predict_device(X.data_ptr(), n, d, prob.data_ptr(), stream=stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = B.ClockSampler(local_rank)
    sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    sampler.mark_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    sampler.mark_end()
    clocks = sampler.stop()
    elapsed_ms = _max_over_ranks(torch, dist, world, dev, e0.elapsed_time(e1))
    value = n_all * args.steps / (elapsed_ms * 1e-3)
    # e2e: host rows through glia_b200_rf_score (chunked H2D + D2H inside)
    e2e = None
    if not args.no_e2e:
        ne = min(n, 2_000_000)
        hx = torch.empty((ne, d), dtype=torch.float64).pin_memory(); hx.copy_(X[:ne])
        hxn = hx.numpy()
        forest.predict(hxn[:1000], want_votes=False)
        barrier()
        t0 = time.perf_counter()
        forest.predict(hxn, want_votes=False)
        dt = _max_over_ranks(torch, dist, world, dev, time.perf_counter() - t0)
        e2e = {"value": ne * world / dt, "unit": "candidates/s", "h2d_bytes_per_step": 8 * d * ne * world,
               "d2h_bytes_per_step": 8 * ne * world, "steps": 1, "ms_per_step": 1e3 * dt,
               "rows": "first %d rows per rank from host memory" % ne}
    if rank == 0:
        peak, peak_src = B.load_peaks()
        alg = (8.0 * d + 8.0) * n
        achieved = alg / (elapsed_ms / args.steps * 1e-3) / 1e9
        nodes = int(f["ndbigtree"].sum())
        line = {"metric": metric, "value": value, "unit": "candidates/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "c5: " + S["desc"], "forest": "255 trees, %d nodes, depth <= 20" % nodes,
                           "rows": "U(0,1), 5% of rows exactly on thresholds", "l2": "rows %.1f GB >> 126 MB L2" % (8 * d * n / 1e9),
                           "sharding": "row ranges per rank, forest replicated, no collective"},
                "roofline": {"bound": "hbm", "kernel": "rf_score", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": alg,
                             "note": "tree traversal is shared-memory/L2 gather bound, not HBM bound (SURVEY §8d)"},
                "clocks": clocks, "gpu_launches": args.steps, "e2e": e2e}
        if not args.no_cpu_baseline and world == 1:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_lib as ol
            orf = ol.OracleRF.from_arrays(f, d)
            ns = 200_000
            Xs = X[:ns].cpu().numpy()
            t0 = time.perf_counter(); orf.predict(Xs, threads=os.cpu_count() or 1); dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": ns / dt, "unit": "candidates/s", "cores": os.cpu_count() or 1, "kind": "port",
                                    "sample": "%d rows, classForest restatement (oracle/rf_oracle.c), OpenMP over rows" % ns}
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0
