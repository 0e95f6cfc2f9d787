"""Multi-GPU driver of the z-slab path: one process per GPU, collectives
over torch.distributed (NCCL on GPUs, gloo in the CPU tests).  The compute is libglia_b200's; this
file only moves the partial tables between ranks.

  comm = TorchComm()                       # or ThreadComm.group(n) to emulate n ranks in one process
  b = slab_builder(shape, cfg, comm, label_bound=..., device=...)
  build_sharded(comm, b, labels_slab, pb_slab)      # b now holds the candidates this rank owns
  sal = b.score_candidates(forest)
"""
import os
import threading

import numpy as np
import torch

from . import api


# --------------------------------------------------------------------------- helpers
def slab_range(nz, world, rank):
    """Owned planes [z0, z1) of rank `rank`: contiguous, sizes differ by at most one."""
    return nz * rank // world, nz * (rank + 1) // world


def slab_with_halo(nz, world, rank):
    """(z_begin, z_end, halo_lo, halo_hi): one label plane of halo on each inner face, because the
    z-1 / z+1 neighbours are inspected by getContourTraits (type/neighbor.hxx:78-88)."""
    z0, z1 = slab_range(nz, world, rank)
    return z0, z1, (1 if z0 > 0 else 0), (1 if z1 < nz else 0)


def owner_range(R, world, rank):
    """Region indices [lo, hi) owned by `rank` (same rule as shard.cu: r / ceil(R/world))."""
    chunk = max(1, -(-R // world))
    lo = min(R, rank * chunk)
    hi = R if rank == world - 1 else min(R, lo + chunk)
    return lo, hi


class _DevMem:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 2}


def dev_tensor(ptr, count, dtype, device):
    """A torch view (no copy) of `count` elements of `dtype` at device address `ptr`."""
    nbytes = count * torch.empty((), dtype=dtype).element_size()
    if nbytes == 0:
        return torch.empty(0, dtype=dtype, device=device)
    return torch.as_tensor(_DevMem(ptr, nbytes), device=device).view(dtype)


# --------------------------------------------------------------------------- communicators
class TorchComm:
    """torch.distributed process group (NCCL for CUDA tensors, gloo for CPU tensors)."""

    def __init__(self, group=None, side_group=True):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        # a second communicator (its own NCCL stream): the all-reduce of the region record tables runs on it
        # while the boundary records are exchanged on the first one.  new_group is collective: every rank
        # constructs its TorchComm at the same point.
        self.side = dist.new_group(ranks=dist.get_process_group_ranks(group) if group is not None else None) \
            if side_group and self.world > 1 else group

    def all_reduce(self, t, op):
        if t.numel() == 0:
            return
        d = self.dist
        if op == "max_u32":  # unsigned max through a signed reduce: flip the sign bit around it
            t.bitwise_xor_(-2 ** 31)
            d.all_reduce(t, op=d.ReduceOp.MAX, group=self.group)
            t.bitwise_xor_(-2 ** 31)
        else:
            d.all_reduce(t, op=d.ReduceOp.SUM if op == "sum" else d.ReduceOp.MAX, group=self.group)

    def all_reduce_start(self, t, op):
        """all_reduce on the side communicator; returns a handle for all_reduce_finish."""
        if t.numel() == 0:
            return None
        d = self.dist
        if op == "max_u32":
            t.bitwise_xor_(-2 ** 31)
            return (d.all_reduce(t, op=d.ReduceOp.MAX, group=self.side, async_op=True), t, True)
        return (d.all_reduce(t, op=d.ReduceOp.SUM if op == "sum" else d.ReduceOp.MAX, group=self.side, async_op=True),
                t, False)

    def all_reduce_finish(self, handle):
        if handle is None:
            return
        work, t, flip = handle
        work.wait()  # the current stream waits for the collective; the host does not block
        if flip:
            t.bitwise_xor_(-2 ** 31)

    def all_gather_fixed(self, t):
        """Concatenation over ranks of equally sized 1-D tensors."""
        out = torch.empty((self.world * t.numel(),), dtype=t.dtype, device=t.device)
        self.dist.all_gather_into_tensor(out, t, group=self.group)
        return out

    def all_to_all_v(self, send, counts, cout=None):
        """send: [sum(counts), rec] uint8 rows grouped by destination; returns the rows received.
        cout: rows to expect from each rank when the caller already knows them."""
        d = self.dist
        if cout is None:
            cin = torch.tensor(counts, dtype=torch.int64, device=send.device)
            cout = torch.empty_like(cin)
            d.all_to_all_single(cout, cin, group=self.group)
            cout = [int(x) for x in cout.tolist()]
        recv = torch.empty((sum(cout), send.shape[1]), dtype=send.dtype, device=send.device)
        d.all_to_all_single(recv, send, output_split_sizes=cout, input_split_sizes=list(counts), group=self.group)
        return recv

    def all_gather_v(self, t):
        """Concatenation over ranks of 1-D/2-D tensors whose first dimension differs per rank."""
        d = self.dist
        n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
        ns = [torch.empty_like(n) for _ in range(self.world)]
        d.all_gather(ns, n, group=self.group)
        ns = [int(x.item()) for x in ns]
        m = max(ns + [1])
        pad = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[:t.shape[0]] = t
        outs = [torch.empty_like(pad) for _ in range(self.world)]
        d.all_gather(outs, pad, group=self.group)
        return torch.cat([o[:k] for o, k in zip(outs, ns)])

    def barrier(self):
        self.dist.barrier(group=self.group)


class ThreadComm:
    """`world` ranks emulated as threads of one process (all handles on one GPU, or CPU tensors):
    the single-GPU test double of TorchComm.  Same collectives, same results."""

    class _Shared:
        def __init__(self, world):
            self.world = world
            self.bar = threading.Barrier(world)
            self.slots = [None] * world
            self.result = None

    def __init__(self, shared, rank):
        self.s, self.rank, self.world = shared, rank, shared.world

    @classmethod
    def group(cls, world):
        s = cls._Shared(world)
        return [cls(s, r) for r in range(world)]

    def _sync(self, t):
        if t.is_cuda:
            torch.cuda.synchronize(t.device)

    def all_reduce(self, t, op):
        s = self.s
        self._sync(t)
        s.slots[self.rank] = t
        s.bar.wait()
        if self.rank == 0:
            if op == "max_u32":
                st = torch.stack([x.bitwise_xor(-2 ** 31) for x in s.slots])
                s.result = st.max(dim=0).values.bitwise_xor(-2 ** 31)
            elif op == "sum":
                s.result = torch.stack(s.slots).sum(dim=0)
            else:
                s.result = torch.stack(s.slots).max(dim=0).values
            self._sync(t)
        s.bar.wait()
        t.copy_(s.result)
        self._sync(t)
        s.bar.wait()

    def all_gather_fixed(self, t):
        return self.all_gather_v(t)

    def all_reduce_start(self, t, op):
        self.all_reduce(t, op)
        return None

    def all_reduce_finish(self, handle):
        pass

    def all_to_all_v(self, send, counts, cout=None):
        s = self.s
        self._sync(send)
        s.slots[self.rank] = (send, list(counts))
        s.bar.wait()
        parts = []
        for src in range(self.world):
            buf, c = s.slots[src]
            off = sum(c[:self.rank])
            parts.append(buf[off:off + c[self.rank]])
        recv = torch.cat(parts) if parts else send[:0]
        self._sync(recv)
        s.bar.wait()
        return recv

    def all_gather_v(self, t):
        s = self.s
        self._sync(t)
        s.slots[self.rank] = t
        s.bar.wait()
        out = torch.cat(list(s.slots))
        self._sync(out)
        s.bar.wait()
        return out

    def barrier(self):
        self.s.bar.wait()


# --------------------------------------------------------------------------- the sharded build
def slab_builder(shape, cfg, comm, label_bound, device=0, device_pointers=False, edge_capacity=0, stream=None):
    """MergeTreeBuilder for this rank's z-slab of a 3D volume of full extent `shape` (z, y, x)."""
    if len(shape) != 3:
        raise ValueError("z-slab sharding is the 3D path; 2D slices are independent problems (one handle per slice)")
    if not label_bound:
        raise ValueError("label_bound (max label + 1 over the whole volume) must be the same on every rank")
    return api.MergeTreeBuilder(shape, cfg, device=device, device_pointers=device_pointers, label_bound=label_bound,
                                edge_capacity=edge_capacity, slab=slab_with_halo(shape[0], comm.world, comm.rank),
                                stream=stream)


def native_comm(dist, world, rank, device):
    """The library's own NCCL communicator pair (api.NativeComm): rank 0 draws the unique ids, torch.distributed only
    carries the 256 bytes to the other ranks."""
    dev = torch.device("cuda", device)
    t = torch.zeros(256, dtype=torch.uint8, device=dev)
    if rank == 0:
        t.copy_(torch.frombuffer(bytearray(api.NativeComm.unique_id()), dtype=torch.uint8))
    dist.broadcast(t, src=0)
    return api.NativeComm(bytes(t.cpu().numpy().tobytes()), world, rank, device)


def build_sharded(comm, b, labels, pb, mask=None, timings=None):
    """Voxel pass on the owned planes, then the exchange steps listed in include/glia_b200.h.
    labels/pb/mask hold planes [z_begin - halo_lo, z_end + halo_hi).  On return `b` is finalised and
    holds the candidates owned by this rank (b.Cn, b.init_candidates(), b.score_candidates())."""
    dev = torch.device("cuda", b.params.device)
    ev = []

    def mark():
        if timings is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            ev.append(e)

    # the library's launches and the collectives are ordered by the stream when the handle runs on torch's
    # current stream (bench.py does); otherwise fence each hand-over
    own_stream = not b.params.stream or b.params.stream != torch.cuda.current_stream(dev).cuda_stream

    def fence():
        if own_stream:
            torch.cuda.synchronize(dev)

    mark()
    b.build_partial(labels, pb, mask)
    mark()
    fence()
    # The record tables (131 MB at C4) start their all-reduce on the side communicator; only the 4-byte presence
    # flags are needed to index the regions, and the boundary exchange below needs the index, not the records.
    (pa, na), (pm, nm), (pf, nf) = b.raw_region_tables()
    ppres, npres = b.region_presence()
    fence()
    comm.all_reduce(dev_tensor(ppres, npres, torch.int32, dev), "max")  # first: the small one is on the critical path
    pending = [comm.all_reduce_start(dev_tensor(pa, na, torch.int32, dev), "sum"),      # counts, histogram bins
               comm.all_reduce_start(dev_tensor(pm, nm, torch.int32, dev), "max_u32"),  # bbox, min/max codes
               comm.all_reduce_start(dev_tensor(pf, nf, torch.float64, dev), "sum")]    # sum v, sum v^2
    fence()
    mark()
    R = b.index_regions()
    pi, ni = b.edge_interest(comm.world)
    fence()
    comm.all_reduce(dev_tensor(pi, ni, torch.uint8, dev), "max")
    fence()
    mark()
    # reduce keyed by label pair: every rank keeps its own records and merges what the others send; if the
    # merged set would not fit some rank's table (labels far from spatial order) all ranks repeat the exchange
    # with their own records included and rebuild (and grow) the tables from the received runs.  One small
    # all-gather carries every rank's send counts and table load, so all ranks size their receive buffers
    # and take the same decision without a second collective.
    rebuild = bool(os.environ.get("GLIA_B200_FORCE_REBUILD"))  # test hook: take the fallback path
    for attempt in range(2):
        counts, pe, rec = b.export_edges(comm.world, -1 if rebuild else comm.rank)
        occ, cap = b.table_load()
        info = comm.all_gather_fixed(torch.tensor(list(counts) + [occ, cap], dtype=torch.int64, device=dev))
        info = info.view(comm.world, comm.world + 2).tolist()
        cout = [int(info[r][comm.rank]) for r in range(comm.world)]
        send = dev_tensor(pe, sum(counts) * rec, torch.uint8, dev).view(-1, rec)
        if rebuild or not any((row[comm.world] + sum(info[r][q] for r in range(comm.world))) * 10 > row[comm.world + 1] * 8
                              for q, row in enumerate(info)):
            recv = comm.all_to_all_v(send, counts, cout)
            break
        rebuild = True
    fence()
    mark()
    b.import_edges(recv.data_ptr() if recv.numel() else 0, recv.shape[0], rebuild)
    lo, hi = owner_range(R, comm.world, comm.rank)
    b.finalize_edges(lo, hi)
    for hnd in pending:
        comm.all_reduce_finish(hnd)
    fence()
    b.gather_regions()
    mark()
    if timings is not None:
        torch.cuda.synchronize(dev)
        dt = [ev[i].elapsed_time(ev[i + 1]) for i in range(len(ev) - 1)]
        timings.update(voxel_ms=dt[0], exchange_ms=dt[1] + dt[2] + dt[3], graph_ms=dt[4],
                       exchange_detail_ms={"presence_all_reduce (record tables start on the side communicator)": dt[1],
                                           "region_index+interest": dt[2], "boundary_records_all_to_all": dt[3]},
                       sent_records=int(sum(counts)), record_bytes=rec)
    return b
