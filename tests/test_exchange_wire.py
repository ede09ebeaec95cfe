"""Fragment-exchange wire format (igloo_b200/exchange.py): Arrow IPC inside the reference's
RecordBatchMessage (distributed.proto:53-57), protobuf wire bytes outside."""
import numpy as np
import pyarrow as pa
import pytest

from igloo_b200.exchange import FragmentType, QueryFragment, RecordBatchMessage, deserialize_batch, serialize_batch


def _batch(n):
    rng = np.random.default_rng(n)
    return pa.record_batch({
        "id": pa.array(rng.integers(-2**40, 2**40, n), pa.int64()),
        "age": pa.array(rng.integers(0, 99, n), pa.int32(), mask=rng.random(n) < 0.2) if n else pa.array([], pa.int32()),
        "name": pa.array([None if i % 7 == 3 else f"user{i}" for i in range(n)], pa.string()),
        "x": pa.array(rng.random(n)),
        "ok": pa.array(rng.random(n) < 0.5),
    })


@pytest.mark.parametrize("n", [0, 1, 5, 1024, 70_001])
def test_round_trip_through_wire_bytes(n):
    b = _batch(n)
    wire = serialize_batch(b).to_wire()
    msg = RecordBatchMessage.from_wire(wire)
    assert msg.num_rows == n
    got = deserialize_batch(msg)
    assert got.schema.equals(b.schema) and got.equals(b)


def test_wire_layout_is_the_proto3_encoding_of_the_reference_message():
    msg = RecordBatchMessage(b"\x01\x02\x03", b"\xff" * 200, 300)
    w = msg.to_wire()
    # field 1 (schema): tag 0x0A, length 3; field 2 (batch_data): tag 0x12, length 200 = c8 01; field 3: tag 0x18, 300 = ac 02
    assert w[:5] == b"\x0a\x03\x01\x02\x03"
    assert w[5:8] == b"\x12\xc8\x01" and w[8:208] == b"\xff" * 200
    assert w[208:] == b"\x18\xac\x02"
    # unknown fields are skipped, defaults may be absent, a negative int64 is a 10-byte varint
    extra = w + b"\x20\x07" + b"\x2a\x02hi"
    assert RecordBatchMessage.from_wire(extra) == msg
    assert RecordBatchMessage.from_wire(b"") == RecordBatchMessage(b"", b"", 0)
    neg = RecordBatchMessage(b"", b"", -1)
    assert len(neg.to_wire()) == 11 and RecordBatchMessage.from_wire(neg.to_wire()).num_rows == -1
    with pytest.raises(ValueError):
        RecordBatchMessage.from_wire(w[:100])


def test_row_count_mismatch_is_an_error_and_fragment_types_match_the_reference():
    m = serialize_batch(_batch(10))
    m.num_rows = 11
    with pytest.raises(ValueError):
        deserialize_batch(m)
    assert [t.name for t in FragmentType] == ["Scan", "Join", "Compute", "Shuffle"]      # fragment.rs:8-13
    f = QueryFragment("f1", FragmentType.Shuffle, None, "gpu0", ["scan_l", "scan_r"])
    assert f.fragment_type is FragmentType.Shuffle and f.dependencies == ["scan_l", "scan_r"]
