"""Fragment-exchange wire format.

The reference declares the message that carries a batch between fragments,
`RecordBatchMessage { bytes schema = 1; bytes batch_data = 2; int64 num_rows = 3; }`
(crates/api/proto/distributed.proto:53-57), but leaves both ends as stubs:
`serialize_plan` returns empty bytes and `deserialize_batch` fabricates a dummy batch
(crates/coordinator/src/distributed_executor.rs:203-222).  Here the two byte fields are
Arrow IPC (the encapsulated schema message and the encapsulated record-batch message, the
format arrow-rs `arrow::ipc` reads and writes), and the message itself is encoded as the
protobuf wire bytes a tonic/prost peer produces for that definition — so a batch leaving a
GPU worker on the control path (small results, or a peer outside the NVLink domain) is
readable by the reference's Rust side once its stub is filled in.

Inside one NVLink box the join's rows do not take this path: they move GPU to GPU through
`igloo_b200.shuffle` (FragmentType.Shuffle below names that step in a fragment plan).
"""
from __future__ import annotations

import enum
from dataclasses import dataclass, field
from typing import List

import pyarrow as pa


class FragmentType(enum.Enum):
    """crates/coordinator/src/fragment.rs:8-13."""
    Scan = "Scan"
    Join = "Join"
    Compute = "Compute"
    Shuffle = "Shuffle"


@dataclass
class QueryFragment:
    """crates/coordinator/src/fragment.rs:16-23, with the plan held as this package's operator tree."""
    id: str
    fragment_type: FragmentType
    physical_plan: object
    worker_address: str = ""
    dependencies: List[str] = field(default_factory=list)


@dataclass
class RecordBatchMessage:
    schema: bytes
    batch_data: bytes
    num_rows: int

    # ---- protobuf wire format (proto3: field 1, 2 length-delimited; field 3 varint) ----
    def to_wire(self) -> bytes:
        out = bytearray()
        for tag, payload in ((0x0A, self.schema), (0x12, self.batch_data)):
            if payload:
                out.append(tag)
                out += _varint(len(payload))
                out += payload
        if self.num_rows:
            out.append(0x18)
            out += _varint(self.num_rows & 0xFFFFFFFFFFFFFFFF)
        return bytes(out)

    @staticmethod
    def from_wire(buf: bytes) -> "RecordBatchMessage":
        schema = batch = b""
        rows = 0
        i, n = 0, len(buf)
        while i < n:
            key, i = _read_varint(buf, i)
            fno, wt = key >> 3, key & 7
            if wt == 2:
                ln, i = _read_varint(buf, i)
                if i + ln > n:
                    raise ValueError("truncated RecordBatchMessage")
                payload = bytes(buf[i:i + ln]); i += ln
                if fno == 1:
                    schema = payload
                elif fno == 2:
                    batch = payload
            elif wt == 0:
                v, i = _read_varint(buf, i)
                if fno == 3:
                    rows = v - (1 << 64) if v >> 63 else v
            elif wt == 1:
                i += 8
            elif wt == 5:
                i += 4
            else:
                raise ValueError(f"unsupported wire type {wt}")
        return RecordBatchMessage(schema, batch, rows)


def _varint(v: int) -> bytes:
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: bytes, i: int):
    shift = v = 0
    while True:
        if i >= len(buf):
            raise ValueError("truncated varint")
        b = buf[i]; i += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, i
        shift += 7


def serialize_batch(batch: pa.RecordBatch) -> RecordBatchMessage:
    """The sending half the reference never wrote (a worker's ExecuteFragment stream item)."""
    return RecordBatchMessage(batch.schema.serialize().to_pybytes(), batch.serialize().to_pybytes(), batch.num_rows)


def deserialize_batch(message: RecordBatchMessage) -> pa.RecordBatch:
    """What `deserialize_batch` (distributed_executor.rs:209-222) stands in for."""
    schema = pa.ipc.read_schema(pa.py_buffer(message.schema))
    batch = pa.ipc.read_record_batch(pa.py_buffer(message.batch_data), schema)
    if batch.num_rows != message.num_rows:
        raise ValueError(f"RecordBatchMessage says {message.num_rows} rows, the batch holds {batch.num_rows}")
    return batch


# ---- fragments on the receiving side -------------------------------------------------------------
class WireScanExec:
    """ExecutionPlan over RecordBatchMessage wire bytes: what a fragment that depends on a remote
    fragment reads (distributed_executor.rs:105-166 collects the replies; :209-222 is the stub this
    fills).  Feeds any operator of igloo_b200.operators, e.g. the probe side of a HashJoinExec."""

    def __init__(self, messages, schema: pa.Schema):
        self._messages, self._schema = messages, schema

    def execute(self):
        for raw in self._messages:
            yield deserialize_batch(RecordBatchMessage.from_wire(raw))

    def schema(self):
        return self._schema


@dataclass
class ShuffleJoinSpec:
    """The plan a FragmentType.Shuffle fragment carries: a hash join whose two inputs are sharded over
    the worker's GPUs and must be repartitioned by key (the step the reference declares in
    fragment.rs:8-13 and never builds).  Executed by igl_join_shuffle_* (csrc/shuffle.cu)."""
    left_fields: list
    right_fields: list
    left_key: int
    right_key: int
    join_type: int = 0
    mode: str = "p2p"


def execute_fragment(fragment: QueryFragment, ctx=None, rank: int = 0, world: int = 1, shards=None):
    """The worker-side half of DistributedExecutor::execute_fragments (distributed_executor.rs:36-193)
    for one fragment.  Scan / Compute / Join fragments run their operator tree and reply with
    RecordBatchMessage wire bytes (one per batch); a Shuffle fragment runs the multi-GPU join on this
    rank's device-resident shards (`shards` = (right cols, n_right, left cols, n_left)) and returns
    (output columns, rows) — its rows never leave the GPUs."""
    if fragment.fragment_type is FragmentType.Shuffle:
        from .shuffle import PartitionedJoin
        spec = fragment.physical_plan
        if not isinstance(spec, ShuffleJoinSpec):
            raise TypeError("a Shuffle fragment carries a ShuffleJoinSpec")
        pj = PartitionedJoin(ctx, rank, world, spec.left_fields, spec.right_fields, spec.left_key, spec.right_key,
                             join_type=spec.join_type, mode=spec.mode)
        pj.prepare(*shards)
        n = pj.run()
        return pj, n
    return [serialize_batch(b).to_wire() for b in fragment.physical_plan.execute()]
