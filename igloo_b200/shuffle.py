"""Hash-repartitioned join across the GPUs of one node: the data plane of the
"worker fragment exchange" the reference only declares (FragmentType::Shuffle,
crates/coordinator/src/fragment.rs:8-13; DistributedPlanner pins the join fragment
to the coordinator, distributed_planner.rs:65-92; GetDataForTask returns empty
bytes, crates/worker/src/service.rs:26-32).

One process per GPU.  Per join:
  1. igl_partition_histogram on both sides (dest = hash of the join key)
  2. one all_gather of the 2 x n count vectors (torch.distributed; the only
     collective on the control path)
  3. rows travel ONCE: either
       mode="p2p"  igl_partition_scatter writes every row straight into the
                   destination GPU's receive slab through CUDA-IPC-mapped peer
                   pointers (NVLink stores issued by the partition kernel itself), or
       mode="nccl" scatter into local per-destination slabs, then
                   all_to_all_single per column (the NCCL baseline)
  4. purely local build + probe; the output stays sharded.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

from ._lib import DevCol, check, lib
from .expr import Field, col

_W = {2: 4, 3: 8, 4: 8}


@dataclass
class ExchangePlan:
    """Pure bookkeeping derived from the n x n count matrix (counts[src][dst])."""
    rank: int
    world: int
    send_counts: List[int]      # rows this rank sends to each destination
    recv_counts: List[int]      # rows this rank receives from each source
    recv_offsets: List[int]     # where source s lands inside this rank's receive slab
    dst_row_offset: List[int]   # where THIS rank's rows land inside destination p's slab
    total_recv: int

    @staticmethod
    def from_counts(counts: Sequence[Sequence[int]], rank: int) -> "ExchangePlan":
        world = len(counts)
        assert all(len(r) == world for r in counts), "count matrix must be square"
        send = [int(counts[rank][p]) for p in range(world)]
        recv = [int(counts[s][rank]) for s in range(world)]
        roff = [0] * world
        for s in range(1, world):
            roff[s] = roff[s - 1] + recv[s - 1]
        dst_off = [int(sum(counts[s][p] for s in range(rank))) for p in range(world)]
        return ExchangePlan(rank, world, send, recv, roff, dst_off, int(sum(recv)))


@dataclass
class RadixExchangePlan:
    """Bookkeeping of the exchange fused with the join's radix partitioning.

    counts[src][dest * parts + part] = rows of source `src` that belong to partition
    `part` of destination `dest`.  A destination's slab is laid out partition by
    partition and, inside a partition, source by source — so the rows a GPU receives are
    already ordered by the top bits of the join hash, the order its local build and
    probe want (igl_join_set_prepartitioned)."""
    rank: int
    world: int
    parts: int
    bin_row_offset: List[int]   # [world * parts]: where THIS rank's rows of each bin start in the destination's slab
    send_counts: List[int]      # rows this rank sends to each destination
    total_recv: int             # rows this rank receives

    @staticmethod
    def from_counts(counts, rank: int, world: int, parts: int) -> "RadixExchangePlan":
        c = np.asarray(counts, dtype=np.int64).reshape(world, world, parts)      # [src][dest][part]
        per_part = c.sum(axis=0)                                                  # [dest][part]
        part_start = np.cumsum(per_part, axis=1) - per_part                       # first row of a partition in its slab
        before_me = c[:rank].sum(axis=0)                                          # rows of lower-ranked sources, per bin
        off = (part_start + before_me).reshape(-1)
        return RadixExchangePlan(rank, world, parts, [int(x) for x in off], [int(x) for x in c[rank].sum(axis=1)],
                                 int(per_part[rank].sum()))


class RawBuffer:
    """A pool block from igl_dev_alloc (its own cudaMalloc, so IPC-exportable)."""

    def __init__(self, ctx, nbytes: int):
        self.ctx, self.nbytes = ctx, nbytes
        p = C.c_void_p()
        check(lib.igl_dev_alloc(ctx._h, max(nbytes, 256), C.byref(p)))
        self.ptr = p.value

    def data_ptr(self) -> int:
        return self.ptr

    def free(self):
        if self.ptr and getattr(self.ctx, "_h", None):
            lib.igl_dev_free(self.ctx._h, self.ptr)
        self.ptr = None


class _Side:
    def __init__(self, fields: Sequence[Field], key_idx: int):
        self.fields, self.key_idx = list(fields), key_idx
        self.cols = None
        self.nrows = 0
        self.slabs: List[RawBuffer] = []     # receive slab per column
        self.send: List[RawBuffer] = []      # nccl mode: per-destination-contiguous send slab per column
        self.capacity = 0
        self.peer_ptrs: List[List[int]] = [] # [peer][col] mapped pointers (p2p mode)


class PartitionedJoin:
    def __init__(self, ctx, rank: int, world: int, left_fields: Sequence[Field], right_fields: Sequence[Field],
                 left_key: int, right_key: int, join_type: int = 0, mode: str = "p2p"):
        from .device import HashJoin
        assert 1 <= world <= 16
        self.ctx, self.rank, self.world, self.mode = ctx, rank, world, mode
        self.L, self.R = _Side(left_fields, left_key), _Side(right_fields, right_key)
        self.join = HashJoin(ctx, join_type, left_fields, right_fields, col(left_fields[left_key].name),
                             col(right_fields[right_key].name))
        self.out = None
        self.out_capacity = 0
        import torch
        self._hist = torch.zeros(2, world, dtype=torch.int64, device=f"cuda:{ctx.device}")
        # p2p: the exchange also does the join's radix partitioning (set in prepare())
        self.log2_parts = 0
        self.fused = False

    # ---- set-up --------------------------------------------------------------
    def prepare(self, right_cols, n_right, left_cols, n_left, slack: float = 1.10):
        """Registers the local shards and sizes the receive slabs once (expected
        share * slack); the slabs' IPC handles are exchanged here, outside the step."""
        import torch
        import torch.distributed as dist
        self.R.cols, self.R.nrows = list(right_cols), n_right
        self.L.cols, self.L.nrows = list(left_cols), n_left
        tot = torch.tensor([n_right, n_left], dtype=torch.int64, device=f"cuda:{self.ctx.device}")
        dist.all_reduce(tot)
        if self.mode == "p2p":
            # the fan-out the local join would use for its share of the build side
            self.log2_parts = int(lib.igl_join_partition_bits(self.ctx._h, int(tot[0]) // self.world))
            # Runs of one bin inside a 4096-row tile are what crosses NVLink: below ~32 rows per run the
            # link efficiency collapses (8 GPUs, C3: 1024 bins 12.0 ms, 256 bins 6.1 ms, 128 bins 5.1 ms
            # per exchange), so the fan-out is capped at 128 bins; regions of up to 32 MB still sit in L2.
            while self.log2_parts > 0 and (self.world << self.log2_parts) > 128:
                self.log2_parts -= 1
            self.fused = ((self.world << self.log2_parts) <= 1024 and all(c.validity is None for c in self.R.cols + self.L.cols)
                          and all(lib.igl_partition_radix_supported(self._devcols(side), len(side.fields)) == 1 for side in (self.R, self.L)))
            if self.fused:
                self._hist = torch.zeros(2, self.world << self.log2_parts, dtype=torch.int64, device=f"cuda:{self.ctx.device}")
            check(lib.igl_join_set_prepartitioned(self.join._h, 1 if self.fused else 0))
        for side, total in ((self.R, int(tot[0])), (self.L, int(tot[1]))):
            self._alloc_slabs(side, int(total / self.world * slack) + 4096)
        dist.barrier()

    def _alloc_slabs(self, side, capacity: int):
        """(Re)allocates one side's receive slabs for `capacity` rows on every rank — a collective
        call — and, in p2p mode, maps the peers' slabs (CUDA IPC)."""
        import torch.distributed as dist
        if side.slabs:       # growing: drop the old mappings and blocks first
            for p, ptrs in enumerate(getattr(side, "peer_ptrs", []) or []):
                if p != self.rank:
                    for q in ptrs:
                        lib.igl_ipc_close(self.ctx._h, q)
            dist.barrier()   # nobody still maps a block that is about to be freed
            for b in side.slabs + (getattr(side, "send", None) or []):
                b.free()
        side.capacity = capacity
        side.slabs = [RawBuffer(self.ctx, side.capacity * _W[f.type] + 256) for f in side.fields]
        if self.mode == "nccl":
            side.send = [RawBuffer(self.ctx, max(side.nrows, 1) * _W[f.type] + 256) for f in side.fields]
            return
        handles = []
        for b in side.slabs:
            h = C.create_string_buffer(64)
            check(lib.igl_ipc_export(self.ctx._h, b.ptr, h))
            handles.append(h.raw)
        gathered = [None] * self.world
        dist.all_gather_object(gathered, handles)
        side.peer_ptrs = []
        for p in range(self.world):
            if p == self.rank:
                side.peer_ptrs.append([b.ptr for b in side.slabs])
            else:
                ptrs = []
                for hraw in gathered[p]:
                    out = C.c_void_p()
                    check(lib.igl_ipc_open(self.ctx._h, hraw, C.byref(out)))
                    ptrs.append(out.value)
                side.peer_ptrs.append(ptrs)

    def _fit_slabs(self, needed_rows):
        """Skewed keys (C5: one GPU can receive a fifth of all probe rows) outgrow slabs sized for an
        even share.  `needed_rows[side]` = the largest number of rows any rank receives, derived from
        the count matrix every rank holds, so all ranks take the same decision and grow together."""
        for side, need in zip((self.R, self.L), needed_rows):
            if need > side.capacity:
                self._alloc_slabs(side, int(need * 1.05) + 4096)

    # ---- one join --------------------------------------------------------------
    def _devcols(self, side, ptrs=None, validity=None):
        arr = (DevCol * len(side.fields))()
        for i, f in enumerate(side.fields):
            p = ptrs[i] if ptrs is not None else side.cols[i].data.data_ptr()
            arr[i] = DevCol(f.type, 0, p, None, None)
        return arr

    def run(self) -> int:
        import torch
        import torch.distributed as dist
        from .device import DeviceColumn
        ctx, W = self.ctx, self.world
        st = ctx.stream()
        marks = []

        def mark(name):   # phase boundaries on the device timeline (read back by phase_ms())
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            marks.append((name, ev))
        mark("start")
        if self.fused:
            return self._run_fused(mark, marks)
        # 1. histograms
        for i, side in enumerate((self.R, self.L)):
            key = side.cols[side.key_idx].c_struct()
            check(lib.igl_partition_histogram(ctx._h, C.byref(key), side.nrows, W, self._hist[i].data_ptr(), st))
        # 2. count exchange
        allh = torch.empty(W, 2, W, dtype=torch.int64, device=self._hist.device)
        dist.all_gather_into_tensor(allh.view(-1), self._hist.view(-1))
        counts = allh.cpu().numpy()
        mark("histogram+counts")
        plans = [ExchangePlan.from_counts(counts[:, i, :].tolist(), self.rank) for i in range(2)]
        self._fit_slabs([int(counts[:, i, :].sum(axis=0).max()) for i in range(2)])
        # 3. rows travel once
        if self.mode == "p2p":
            dist.barrier()                       # peers are done reading their slabs from the previous join
            for side, plan in zip((self.R, self.L), plans):
                nc = len(side.fields)
                dst = (C.c_void_p * (W * nc))(*[side.peer_ptrs[p][c] for p in range(W) for c in range(nc)])
                off = (C.c_int64 * W)(*plan.dst_row_offset)
                check(lib.igl_partition_scatter(ctx._h, self._devcols(side), nc, side.key_idx, side.nrows, W, dst, off, st))
            torch.cuda.synchronize()
            dist.barrier()                       # every rank's stores have landed
        else:
            for side, plan in zip((self.R, self.L), plans):
                nc = len(side.fields)
                send_off = np.concatenate([[0], np.cumsum(plan.send_counts)[:-1]]).astype(np.int64)
                dst = (C.c_void_p * (W * nc))(*[side.send[c].ptr for p in range(W) for c in range(nc)])
                off = (C.c_int64 * W)(*send_off.tolist())
                check(lib.igl_partition_scatter(ctx._h, self._devcols(side), nc, side.key_idx, side.nrows, W, dst, off, st))
                for c, f in enumerate(side.fields):
                    dt = {2: torch.int32, 3: torch.int64, 4: torch.float64}[f.type]
                    src = _as_tensor(side.send[c].ptr, sum(plan.send_counts), dt, ctx.device)
                    dstt = _as_tensor(side.slabs[c].ptr, plan.total_recv, dt, ctx.device)
                    dist.all_to_all_single(dstt, src, plan.recv_counts, plan.send_counts)
        mark("exchange")
        # 4. local join on what arrived
        nr, nl = plans[0].total_recv, plans[1].total_recv
        rcols = [DeviceColumn(f.type, nr, self.R.slabs[i]) for i, f in enumerate(self.R.fields)]
        lcols = [DeviceColumn(f.type, nl, self.L.slabs[i]) for i, f in enumerate(self.L.fields)]
        self.join.build(rcols, nr)
        mark("build")
        cap = self.join.probe_count(lcols, nl) if self.join.has_duplicates else nl
        if self.out is None or cap > self.out_capacity:
            self.out_capacity = int(cap * 1.05) + 1024
            self.out = self.join.alloc_outputs(self.out_capacity, lcols)
        _, n = self.join.probe(lcols, nl, out=self.out, capacity=self.out_capacity)
        mark("probe")
        self.last_plans = plans
        self._marks = marks
        return n

    def _run_fused(self, mark, marks) -> int:
        """p2p with the radix partitioning folded into the exchange: rows are sorted by
        (destination, partition) in shared memory and stored to the peers as runs."""
        import torch
        import torch.distributed as dist
        from .device import DeviceColumn
        ctx, W, lp = self.ctx, self.world, self.log2_parts
        st = ctx.stream()
        bins = W << lp
        for i, side in enumerate((self.R, self.L)):
            key = side.cols[side.key_idx].c_struct()
            check(lib.igl_partition_histogram_radix(ctx._h, C.byref(key), side.nrows, W, lp, self._hist[i].data_ptr(), st))
        allh = torch.empty(W, 2, bins, dtype=torch.int64, device=self._hist.device)
        dist.all_gather_into_tensor(allh.view(-1), self._hist.view(-1))
        counts = allh.cpu().numpy()
        mark("histogram+counts")
        plans = [RadixExchangePlan.from_counts(counts[:, i, :], self.rank, W, 1 << lp) for i in range(2)]
        self._fit_slabs([int(counts[:, i, :].reshape(W, W, 1 << lp).sum(axis=(0, 2)).max()) for i in range(2)])
        dist.barrier()                       # peers are done reading their slabs from the previous join
        mark("exchange:barrier")
        for side, plan in zip((self.R, self.L), plans):
            nc = len(side.fields)
            dst = (C.c_void_p * (W * nc))(*[side.peer_ptrs[p][c] for p in range(W) for c in range(nc)])
            off = (C.c_int64 * bins)(*plan.bin_row_offset)
            check(lib.igl_partition_scatter_radix(ctx._h, self._devcols(side), nc, side.key_idx, side.nrows, W, lp, dst, off, st))
        mark("exchange:scatter")
        torch.cuda.synchronize()
        dist.barrier()                       # every rank's stores have landed
        mark("exchange:landed")
        nr, nl = plans[0].total_recv, plans[1].total_recv
        rcols = [DeviceColumn(f.type, nr, self.R.slabs[i]) for i, f in enumerate(self.R.fields)]
        lcols = [DeviceColumn(f.type, nl, self.L.slabs[i]) for i, f in enumerate(self.L.fields)]
        self.join.build(rcols, nr)
        mark("build")
        cap = self.join.probe_count(lcols, nl) if self.join.has_duplicates else nl
        if self.out is None or cap > self.out_capacity:
            self.out_capacity = int(cap * 1.05) + 1024
            self.out = self.join.alloc_outputs(self.out_capacity, lcols)
        _, n = self.join.probe(lcols, nl, out=self.out, capacity=self.out_capacity)
        mark("probe")
        self.last_plans = plans
        self._marks = marks
        return n

    def phase_ms(self):
        """Device time of each phase of the last run() on this rank."""
        import torch
        torch.cuda.synchronize()
        return {b[0]: round(a[1].elapsed_time(b[1]), 3) for a, b in zip(self._marks, self._marks[1:])}


def _as_tensor(ptr: int, n: int, dtype, device: int):
    """A torch view over a raw device pointer (for NCCL only)."""
    import torch

    class _Iface:
        pass
    itemsize = torch.empty(0, dtype=dtype).element_size()
    typestr = {torch.int32: "<i4", torch.int64: "<i8", torch.float64: "<f8"}[dtype]
    obj = _Iface()
    obj.__cuda_array_interface__ = {"shape": (max(n, 0),), "typestr": typestr, "data": (ptr, False), "version": 3,
                                    "strides": None}
    if n == 0:
        return torch.empty(0, dtype=dtype, device=f"cuda:{device}")
    return torch.as_tensor(obj, device=f"cuda:{device}")
