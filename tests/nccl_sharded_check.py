"""Run under torchrun on >= 2 GPUs: the library's native NCCL exchange (glia_b200_build_sharded) against the
single-handle result on rank 0 - same candidates, bit-identical rows and scores (dyadic pb).
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/nccl_sharded_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from glia_b200 import api, parallel as par, synth  # noqa: E402


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    native = par.native_comm(dist, world, rank, lr)
    ok = True
    for shape, cell, shuffle in (((48, 64, 96), 8, False), ((33, 40, 52), 6, True), ((2 * world, 30, 36), 5, False)):
        lab, pb = synth.voronoi_volume(shape, cell, seed=17)
        if shuffle:  # labels far from spatial order: most boundaries cross owners (and the rebuild path may trigger)
            perm = np.random.RandomState(3).permutation(int(lab.max()) + 1).astype(np.uint32) + 1
            lab = perm[lab]
        lb = int(lab.max()) + 1
        cfg = lambda p: api.FeatureConfig(rimgs=[(p, 16, 0.0, 1.0)], bimgs=[(p, 16, 0.0, 1.0)])
        z0, z1, hl, hh = par.slab_with_halo(shape[0], world, rank)
        sl = slice(z0 - hl, z1 + hh)
        p = np.ascontiguousarray(pb[sl])
        b = api.MergeTreeBuilder(shape, cfg(p), device=lr, label_bound=lb, slab=(z0, z1, hl, hh))
        for rep in range(2):  # twice: handle and communicator are reused per step
            b.build_sharded(native, np.ascontiguousarray(lab[sl]), p)
        keys, rows = b.init_candidates()
        g = api.MergeTreeBuilder(shape, cfg(pb), device=lr).build(lab, pb)
        k0, rows0 = g.init_candidates()
        f = synth.random_forest(31, g.row_dim(), max_depth=8, seed=2, thr_pool=rows0)
        forest = api.Forest.from_arrays(f, g.row_dim(), device=lr)
        sal0 = g.score_candidates(forest)
        sal = b.score_candidates(forest) if b.Cn else np.zeros(0)
        # gather every rank's part on rank 0 through torch (sizes differ)
        parts = [None] * world
        dist.all_gather_object(parts, (keys, rows, sal, b.R))
        if rank == 0:
            K = np.concatenate([x[0] for x in parts]); Rw = np.concatenate([x[1] for x in parts]); S = np.concatenate([x[2] for x in parts])
            good = (all(x[3] == g.R for x in parts) and np.array_equal(K, k0) and np.array_equal(Rw, rows0) and np.array_equal(S, sal0))
            print("shape", shape, "shuffle", shuffle, "R", g.R, "C", g.Cn, "native == single:", good, flush=True)
            ok = ok and good
        # the bc_feat stage on the sharded build (K5): a greedy order from the single handle, every rank featurises its
        # share of the merges, gathered rows == the single-handle stage (with and without saliency columns)
        res = g.merge_order_bc(forest)
        for sal_in in (None, res["sal"]):
            want = g.bc_feat(res["order"], sal=sal_in, init_sal=1.0, sal_bias=0.25)
            first, mine = b.bc_feat_sharded(native, res["order"], sal=sal_in, init_sal=1.0, sal_bias=0.25)
            parts = [None] * world
            dist.all_gather_object(parts, (first, mine))
            if rank == 0:
                parts.sort(key=lambda x: x[0])
                got = np.concatenate([x[1] for x in parts])
                good = got.shape == want.shape and np.array_equal(got, want)
                if not good and got.shape == want.shape:
                    bad = np.argwhere(got != want)
                    print("  first mismatches (row, col):", bad[:5].tolist(), "max abs diff", float(np.abs(got - want).max()), flush=True)
                print("  bc_feat sharded (%d merges, sal=%s) == single: %s" % (len(want), sal_in is not None, good), flush=True)
                ok = ok and good
        b.close(); g.close()
    native.close()
    flag = torch.tensor([1 if ok else 0], device=torch.device("cuda", lr))
    dist.broadcast(flag, src=0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
