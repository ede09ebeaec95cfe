"""Hash join across the GPUs of one node — a thin ctypes front end of
igl_comm_init / igl_join_shuffle_* (csrc/shuffle.cu).  Everything that moves or
joins rows lives behind the C ABI, so a Rust worker (crates/worker/src/main.rs:14-52)
binds the same calls; this module only distributes the NCCL id and keeps handles.

It is the data plane of the "worker fragment exchange" the reference only declares
(FragmentType::Shuffle, crates/coordinator/src/fragment.rs:8-13; the join fragment fed
by scan fragments, distributed_planner.rs:65-92; GetDataForTask returning empty
bytes, crates/worker/src/service.rs:26-32).

One process per GPU.  `init_comm` is called once per context; `PartitionedJoin.run`
is collective: every rank passes its row-range shard of both sides and gets its
share of the join output (library-owned device columns).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import numpy as np

from ._lib import DevCol, check, lib
from .expr import Field, col

EXCHANGE_P2P, EXCHANGE_NCCL = 0, 1
_MODES = {"p2p": EXCHANGE_P2P, "nccl": EXCHANGE_NCCL}
PHASES = ("counts", "partition", "exchange", "build", "probe", "closing_barrier", "total", "probe_kernel")
STATS = ("build_rows_received", "probe_rows_received", "bytes_read_from_peers", "log2_parts", "fused_probe", "output_rows",
         "probe_partition_us", "probe_wait_us", "prefetched_bytes", "prefetched_parts", "hot_keys")


def init_comm(ctx, rank: int, world: int, broadcast=None) -> None:
    """igl_comm_init for `ctx`.  The 128-byte id made by rank 0 (igl_comm_unique_id) has to
    reach every rank through the host: `broadcast(bytes_or_None) -> bytes` does that; by default
    torch.distributed's default group carries it (the tests and bench.py run under torchrun)."""
    ident = C.create_string_buffer(128)
    if rank == 0:
        check(lib.igl_comm_unique_id(ident))
    if broadcast is None:
        import torch.distributed as dist

        def broadcast(b):
            box = [b]
            dist.broadcast_object_list(box, src=0)
            return box[0]
    raw = broadcast(ident.raw if rank == 0 else None)
    check(lib.igl_comm_init(ctx._h, world, rank, raw))


def shuffle_plan(counts, rank: int, world: int, log2_parts: int, broadcast: int = 0):
    """igl_shuffle_plan[_ex] (pure host code): `counts[src][dest * parts + part]` -> the padded bin
    offsets of every source's staging buffer, and for `rank` the runs it reads, in walk order,
    as (source, first row in the source's staging buffer, rows); plus their total.
    broadcast=1: one more destination (index `world`) whose bins every rank reads."""
    parts, bins = 1 << log2_parts, (world + broadcast) << log2_parts
    c = np.ascontiguousarray(np.asarray(counts, dtype=np.int64).reshape(world, bins))
    offs = np.zeros((world, bins), np.int64)
    nseg = world * parts * (1 + broadcast)
    src = np.zeros(nseg, np.int32); row0 = np.zeros(nseg, np.int64); rows = np.zeros(nseg, np.int64)
    total = C.c_int64(0)
    check(lib.igl_shuffle_plan_ex(world, rank, log2_parts, broadcast, c.ctypes.data_as(C.POINTER(C.c_int64)),
                                  offs.ctypes.data_as(C.POINTER(C.c_int64)), src.ctypes.data_as(C.POINTER(C.c_int32)),
                                  row0.ctypes.data_as(C.POINTER(C.c_int64)), rows.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(total)))
    return offs, list(zip(src.tolist(), row0.tolist(), rows.tolist())), total.value


class PartitionedJoin:
    """igl_join_shuffle: HashJoinExec (hash_join.rs:35-213) over row-range shards held by
    `world` GPUs.  mode "p2p": the probe kernel reads its rows out of the peers' staging
    buffers over NVLink while it probes; "nccl": grouped ncclSend/ncclRecv first."""

    def __init__(self, ctx, rank: int, world: int, left_fields: Sequence[Field], right_fields: Sequence[Field],
                 left_key: int, right_key: int, join_type: int = 0, mode: str = "p2p"):
        from .device import join_output_fields
        self.ctx, self.rank, self.world, self.mode = ctx, rank, world, mode
        self.left, self.right = list(left_fields), list(right_fields)
        self._lk = ctx.compile(col(self.left[left_key].name), self.left)
        self._rk = ctx.compile(col(self.right[right_key].name), self.right)
        lt = (C.c_int32 * len(self.left))(*[f.type for f in self.left])
        rt = (C.c_int32 * len(self.right))(*[f.type for f in self.right])
        ln = bytes(1 if f.nullable else 0 for f in self.left)
        rn = bytes(1 if f.nullable else 0 for f in self.right)
        h = C.c_void_p()
        check(lib.igl_join_shuffle_create(ctx._h, join_type, lt, ln, len(self.left), rt, rn, len(self.right),
                                          self._lk._h, self._rk._h, _MODES[mode], C.byref(h)))
        self._h = h
        self.out_fields = join_output_fields(self.left, self.right, join_type)
        self.left_join = join_type in (1, 3)
        self._inputs = None
        self.out: List = []
        self.out_rows = 0

    def __del__(self):
        if getattr(self, "_h", None) and getattr(self.ctx, "_h", None):
            lib.igl_join_shuffle_destroy(self._h)
        self._h = None

    close = __del__

    def prepare(self, right_cols, n_right, left_cols, n_left):
        """Registers this rank's shard (device columns) for run()."""
        self._inputs = (list(right_cols), int(n_right), list(left_cols), int(n_left))

    def run(self) -> int:
        from .device import DeviceColumn
        rc, nr, lc, nl = self._inputs
        ra = (DevCol * len(rc))(*[c.c_struct() for c in rc])
        la = (DevCol * len(lc))(*[c.c_struct() for c in lc])
        n = C.c_int64(0)
        check(lib.igl_join_shuffle_run(self._h, ra, nr, la, nl, C.byref(n), self.ctx.stream()))
        nout = len(self.left) + len(self.right)
        arr = (DevCol * nout)()
        rows = C.c_int64(0)
        check(lib.igl_join_shuffle_output(self._h, arr, C.byref(rows)))
        self.out = [_BorrowedColumn(arr[i].type, rows.value, arr[i].data, arr[i].validity) for i in range(nout)]
        self.out_rows = rows.value
        return n.value

    def phase_ms(self):
        out = (C.c_double * 8)()
        check(lib.igl_join_shuffle_phase_ms(self._h, out))
        return {k: round(float(v), 3) for k, v in zip(PHASES, out)}

    def stats(self):
        out = (C.c_int64 * 12)()
        check(lib.igl_join_shuffle_stats(self._h, out))
        return {k: int(v) for k, v in zip(STATS, out)}


class _BorrowedColumn:
    """A device column owned by the library (igl_join_shuffle_output): pointers only."""

    def __init__(self, type_: int, length: int, data: int, validity):
        self.type, self.length, self._data, self._validity = type_, length, data, validity
        self.validity = None if not validity else self

    def c_struct(self) -> DevCol:
        return DevCol(self.type, 0, self._data, self._validity, None)
