"""Host-side logic of the multi-GPU path on CPU: slab / ownership partitions and the collectives
of glia_b200/parallel.py under torch.distributed (gloo, world_size 2) and under the in-process
ThreadComm used by the single-GPU emulation tests."""
import os
import socket
import threading

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from glia_b200 import parallel as par


def test_slab_partition_covers_volume():
    for nz in (1, 2, 7, 64, 1024):
        for world in (1, 2, 3, 4, 8):
            if world > nz:
                continue
            planes = []
            for r in range(world):
                z0, z1, hl, hh = par.slab_with_halo(nz, world, r)
                assert z1 > z0 and hl == (1 if z0 > 0 else 0) and hh == (1 if z1 < nz else 0)
                planes += list(range(z0, z1))
            assert planes == list(range(nz))  # every plane owned exactly once


def test_owner_ranges_match_device_rule():
    for R in (0, 1, 5, 100, 1091976):
        for world in (1, 2, 3, 8):
            chunk = max(1, -(-R // world))
            seen = 0
            for r in range(world):
                lo, hi = par.owner_range(R, world, r)
                assert lo == seen or (lo == R and hi == R)
                seen = max(seen, hi)
                if hi > lo:  # shard.cu owner_of: min(world-1, idx / chunk)
                    assert min(world - 1, lo // chunk) == r and min(world - 1, (hi - 1) // chunk) == r
            assert seen == R


def _expected(world):
    """Deterministic per-rank inputs and the expected results of each collective."""
    rng = np.random.RandomState(5)
    add = rng.randint(0, 1000, (world, 64)).astype(np.int32)
    mx = rng.randint(0, 2 ** 32, (world, 64), dtype=np.uint64).astype(np.uint32)  # top bit often set
    mx[:, :8] = 0
    f = rng.rand(world, 16)
    counts = rng.randint(0, 5, (world, world))
    return add, mx, f, counts


def _exercise(comm, out):
    world, r = comm.world, comm.rank
    add, mx, f, counts = _expected(world)
    t = torch.from_numpy(add[r].copy()); comm.all_reduce(t, "sum")
    assert np.array_equal(t.numpy(), add.sum(0))
    t = torch.from_numpy(mx[r].copy().view(np.int32)); comm.all_reduce(t, "max_u32")
    assert np.array_equal(t.numpy().view(np.uint32), mx.max(0))
    t = torch.from_numpy(f[r].copy()); comm.all_reduce(t, "sum")
    np.testing.assert_allclose(t.numpy(), f.sum(0), rtol=1e-15)
    t = torch.from_numpy((mx[r] & 1).astype(np.uint8)); comm.all_reduce(t, "max")
    assert np.array_equal(t.numpy(), (mx & 1).max(0).astype(np.uint8))
    # the same reductions started on the side communicator and finished later (build_sharded overlaps them)
    ta = torch.from_numpy(add[r].copy()); tm = torch.from_numpy(mx[r].copy().view(np.int32)); tf = torch.from_numpy(f[r].copy())
    pend = [comm.all_reduce_start(ta, "sum"), comm.all_reduce_start(tm, "max_u32"), comm.all_reduce_start(tf, "sum")]
    t = torch.from_numpy((mx[r] & 1).astype(np.uint8)); comm.all_reduce(t, "max")  # other traffic in between
    for hnd in pend:
        comm.all_reduce_finish(hnd)
    assert np.array_equal(ta.numpy(), add.sum(0)) and np.array_equal(tm.numpy().view(np.uint32), mx.max(0))
    np.testing.assert_allclose(tf.numpy(), f.sum(0), rtol=1e-15)
    # all_to_all_v: record j sent by rank s to rank d carries (s, d, j)
    rec = 12
    rows = [np.array([r, d, j], np.int32).view(np.uint8) for d in range(world) for j in range(counts[r, d])]
    send = torch.from_numpy(np.stack(rows) if rows else np.zeros((0, rec), np.uint8))
    recv = comm.all_to_all_v(send, [int(c) for c in counts[r]]).numpy().view(np.int32).reshape(-1, 3)
    exp = [(s, r, j) for s in range(world) for j in range(counts[s, r])]
    assert [tuple(x) for x in recv.tolist()] == exp
    # the same exchange when the receive counts are already known (build_sharded gathers them with the table load)
    info = comm.all_gather_fixed(torch.tensor([int(c) for c in counts[r]] + [r], dtype=torch.int64))
    info = info.view(world, world + 1).tolist()
    assert [row[:world] for row in info] == counts.tolist() and [row[world] for row in info] == list(range(world))
    recv2 = comm.all_to_all_v(send, [int(c) for c in counts[r]], [int(info[s][r]) for s in range(world)])
    assert [tuple(x) for x in recv2.numpy().view(np.int32).reshape(-1, 3).tolist()] == exp
    g = comm.all_gather_v(torch.arange(r + 1, dtype=torch.int64) + 10 * r)
    assert g.tolist() == [10 * s + i for s in range(world) for i in range(s + 1)]
    comm.barrier()
    out[r] = True


def _gloo_worker(rank, world, port):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    out = {}
    _exercise(par.TorchComm(), out)
    assert out[rank]
    dist.destroy_process_group()


def test_torch_comm_gloo_world2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_gloo_worker, args=(2, port), nprocs=2, join=True)


@pytest.mark.parametrize("world", [2, 3])
def test_thread_comm(world):
    comms = par.ThreadComm.group(world)
    out, errs = {}, []

    def run(c):
        try:
            _exercise(c, out)
        except BaseException as e:  # noqa: BLE001
            errs.append(e)
            c.s.bar.abort()

    ts = [threading.Thread(target=run, args=(c,)) for c in comms]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    assert len(out) == world
