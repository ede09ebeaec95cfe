"""C5-shaped skew on one GPU at 1/scale: (5a) unique build, Zipf-like probe keys; (5b) Zipf-like
build keys (heavy duplicate buckets), uniform probe.  Checks row counts against a torch count."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from igloo_b200.device import Context, DeviceColumn, HashJoin
from igloo_b200.expr import Field, col, T_I64

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 8
bits = 28 - (scale.bit_length() - 1)
nb, npr = 1 << bits, 2_000_000_000 // scale
ctx = Context(0); dev = "cuda:0"; seed = 0x16100005
pad = lambda t: torch.cat([t, torch.zeros(16, dtype=t.dtype, device=dev)])   # noqa: E731


def zipf_ranks(col_id, n):   # inverse CDF of a Pareto tail with exponent 1.1, clamped to the domain
    u = ctx.gen(3, seed, col_id, 0, n).data[:n]
    r = torch.clamp((1.0 - u).clamp_min(1e-18).pow(-10.0) - 1.0, max=float(nb - 1)).to(torch.int64)
    return r


def run(name, rk, lk):
    rv = ctx.gen(2, seed, 11, 0, rk.length); lw = ctx.gen(2, seed, 12, 0, lk.length)
    f = [Field("k", T_I64), Field("w", T_I64)]; g = [Field("k", T_I64), Field("v", T_I64)]
    j = HashJoin(ctx, 0, f, g, col("k"), col("k"))
    times = []
    for it in range(3):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        j.build([rk, rv], rk.length)
        n = j.probe_count([lk, lw], lk.length)
        out, n2 = j.probe([lk, lw], lk.length, capacity=n)
        e1.record(); torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
        del out
    assert n == n2
    # expected count: sum over probe keys of the key's multiplicity on the build side
    mult = torch.bincount(rk.data[:rk.length], minlength=nb)
    want = int(mult[lk.data[:lk.length]].sum().item())
    ms = statistics.median(times[1:])
    print(f"{name}: build {rk.length} x probe {lk.length} -> {n2} rows (expected {want}); {ms:.1f} ms = "
          f"{(rk.length + lk.length) / ms / 1e6:.2f} G input rows/s; duplicates={j.has_duplicates}; top build multiplicity {int(mult.max())}", flush=True)
    assert n2 == want


perm = ctx.gen(4, 0, 0, 0, nb, bits=bits)                              # unique keys, scattered
hot = perm.data[:nb]
lk_a = DeviceColumn(T_I64, npr, pad(hot[zipf_ranks(21, npr)]))         # rank -> key through the permutation
run(f"C5a/{scale}", perm, lk_a)
rk_b = DeviceColumn(T_I64, nb, pad(hot[zipf_ranks(22, nb)]))
lk_b = ctx.gen(2, seed, 23, 0, npr // 8, m=nb)                         # uniform probe over the domain (1/8 of the rows: the output explodes otherwise)
lk_b = DeviceColumn(T_I64, npr // 8, pad(hot[lk_b.data[:npr // 8]]))
run(f"C5b/{scale}", rk_b, lk_b)
