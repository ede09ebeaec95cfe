"""Streaming operator API (igl_op_*) end to end from host Arrow batches: C2-shaped
FusedFilterProjectExec and a C3-shaped HashJoinExec, rows/s including host staging."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pyarrow as pa
from igloo_b200.device import Context
from igloo_b200.expr import col
from igloo_b200.operators import FusedFilterProjectExec, HashJoinExec, MemoryScanExec
from oracle import oracle as orc

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64_000_000
batch_rows = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
ctx = Context(0)
seed = 0x16100002
t = pa.table({"a": orc.gen_i32_uniform(seed, 0, 0, n, 1000), "b": orc.gen_i32_uniform(seed, 1, 0, n, 1 << 31), "x": orc.gen_f64_unit(seed, 2, 0, n)},
             schema=pa.schema([pa.field("a", pa.int32(), False), pa.field("b", pa.int32(), False), pa.field("x", pa.float64(), False)]))
exprs = [((col("b") + col("a")).alias("s"), "s"), ((col("x") * 1.5 + 0.25).alias("y"), "y")]
sch = pa.schema([pa.field("s", pa.int32(), False), pa.field("y", pa.float64(), False)])
for rep in range(2):
    plan = FusedFilterProjectExec(MemoryScanExec(t, batch_rows), col("a") < 500, exprs, sch, ctx)
    t0 = time.perf_counter(); rows = sum(b.num_rows for b in plan.execute()); dt = time.perf_counter() - t0
    print(f"filter+project via igl_op_*: {n} rows in {batch_rows}-row batches -> {rows} rows, {dt * 1e3:.0f} ms = {n / dt / 1e6:.0f} M rows/s "
          f"({n * 16 / dt / 1e9:.1f} GB/s of input)", flush=True)
bits = 22
nb, npr = 1 << bits, n // 2
right = pa.table({"k": orc.gen_perm_pow2(0, nb, bits), "v": orc.gen_i64_uniform(3, 1, 0, nb)},
                 schema=pa.schema([pa.field("k", pa.int64(), False), pa.field("v", pa.int64(), False)]))
left = pa.table({"k": orc.gen_i64_uniform(3, 2, 0, npr, nb), "w": orc.gen_i64_uniform(3, 3, 0, npr)},
                schema=pa.schema([pa.field("k", pa.int64(), False), pa.field("w", pa.int64(), False)]))
for rep in range(2):
    j = HashJoinExec(MemoryScanExec(left, batch_rows), MemoryScanExec(right, batch_rows), [col("k")], [col("k")], 0, ctx)
    t0 = time.perf_counter(); rows = sum(b.num_rows for b in j.execute()); dt = time.perf_counter() - t0
    print(f"hash join via igl_op_*: build {nb} x probe {npr} -> {rows} rows, {dt * 1e3:.0f} ms = {(nb + npr) / dt / 1e6:.0f} M input rows/s", flush=True)
