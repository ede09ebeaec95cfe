"""C4 (TPC-H-shaped lineitem JOIN orders, SF100 row counts / scale) on one GPU: timing and a
digest check against the answer recomputed from the key pattern (no hash table involved).
orders(o_orderkey:i64 sparse, o_custkey:i64, o_totalprice:f64, o_orderdate:i32) is the build side,
lineitem(l_orderkey:i64, l_partkey:i64, l_quantity:f64, l_extendedprice:f64, l_shipdate:i32) probes."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from igloo_b200.device import Context, DeviceColumn, HashJoin
from igloo_b200.expr import Field, col, T_F64, T_I32, T_I64

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 1
n_ord = 150_000_000 // scale
n_li = 600_037_902 // scale
ctx = Context(0)
dev = "cuda:0"
seed = 0x16100004
i = torch.arange(n_ord, device=dev, dtype=torch.int64)
o_key = (((i >> 3) << 5) | (i & 7)) + 1
o_cust = ctx.gen(2, seed, 1, 0, n_ord, m=15_000_000)
o_price = ctx.gen(3, seed, 2, 0, n_ord)
o_date = ctx.gen(1, seed, 3, 0, n_ord, m=2500)
# 1-7 lines per order, trimmed / padded to the exact lineitem count
cnt = 1 + (ctx.gen(2, seed, 4, 0, n_ord, m=7).data[:n_ord])
idx = torch.repeat_interleave(i, cnt)
if idx.numel() < n_li:
    idx = torch.cat([idx, idx[: n_li - idx.numel()]])
idx = idx[:n_li].contiguous()
del cnt
pad = lambda t: torch.cat([t, torch.zeros(16, dtype=t.dtype, device=dev)])   # noqa: E731  (Arrow-style slack behind the values)
l_key = DeviceColumn(T_I64, n_li, pad(o_key[idx]))
l_part = ctx.gen(2, seed, 5, 0, n_li, m=20_000_000)
l_qty = ctx.gen(3, seed, 6, 0, n_li)
l_price = ctx.gen(3, seed, 7, 0, n_li)
l_ship = ctx.gen(1, seed, 8, 0, n_li, m=2500)
right = [DeviceColumn(T_I64, n_ord, pad(o_key)), o_cust, o_price, o_date]
left = [l_key, l_part, l_qty, l_price, l_ship]
lf = [Field("l_orderkey", T_I64), Field("l_partkey", T_I64), Field("l_quantity", T_F64), Field("l_extendedprice", T_F64), Field("l_shipdate", T_I32)]
rf = [Field("o_orderkey", T_I64), Field("o_custkey", T_I64), Field("o_totalprice", T_F64), Field("o_orderdate", T_I32)]
j = HashJoin(ctx, 0, lf, rf, col("l_orderkey"), col("o_orderkey"))
j._right_cols = right
out = j.alloc_outputs(n_li, left)
times = []
for it in range(4):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    j.build(right, n_ord)
    _, n = j.probe(left, n_li, out=out, capacity=n_li)
    e1.record(); torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
    ph = j.phase_ms()
assert n == n_li, (n, n_li)
ms = statistics.median(times[1:])
alg = n_ord * 28 + n_li * 36 + n * 64
print(f"C4/{scale}: build {n_ord} x probe {n_li} -> {n} rows; {ms:.2f} ms/step = {(n_ord + n_li) / ms / 1e6:.2f} G input rows/s; "
      f"algorithmic {alg / 1e9:.1f} GB -> {alg / ms / 1e6:.0f} GB/s; phases(ms) part/build/part/probe = {[round(x, 2) for x in ph]}; "
      f"duplicates={j.has_duplicates}", flush=True)
# expected right columns straight from the key pattern
exp = [DeviceColumn(c.type, n_li, pad(c.data[:n_ord][idx])) for c in right]
got = ctx.digest(out, n)
want = ctx.digest(left + exp, n_li)
print("digest equal:", got == want, flush=True)
assert got == want
