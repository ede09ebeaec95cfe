import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from igloo_b200.device import Context, FilterProject
from igloo_b200.expr import Field, col, T_F64, T_I32
n = int(sys.argv[1]) if len(sys.argv) > 1 else 250_000_000
ctx = Context(0)
a = ctx.gen(1, 0x16100002, 0, 0, n, 1000); b = ctx.gen(1, 0x16100002, 1, 0, n, 1 << 31); x = ctx.gen(3, 0x16100002, 2, 0, n)
fields = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
proj = [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")]
def timeit(name, pred, pr, bytes_fn):
    fp = FilterProject(ctx, fields, pred, pr)
    out = fp.alloc_outputs(n)
    for _ in range(3):
        fp.launch([a, b, x], n, out); kept = fp.finish()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    torch.cuda.synchronize(); ev[0].record()
    for i in range(5):
        fp.launch([a, b, x], n, out); ev[i + 1].record()
    torch.cuda.synchronize(); kept = fp.finish()
    ms = statistics.median(ev[i].elapsed_time(ev[i + 1]) for i in range(5))
    gb = bytes_fn(kept / n) * n / 1e9
    print(f"{name:28s} kept={kept/n:.3f} ms={ms:.3f} GB/s={gb/ms*1e3:.0f} rows/s={n/ms/1e6:.1f}G", flush=True)
timeit("projection only", None, proj, lambda s: 16 + 12)
timeit("copy a only (proj)", None, [col("a")], lambda s: 8)
for tp in ("0", "1"):
    os.environ["IGL_FP_TWO_PASS"] = tp
    print("IGL_FP_TWO_PASS=" + tp)
    timeit("filter+project T=500", col("a") < 500, proj, lambda s: 16 + 12 * s)
    timeit("filter+project T=10", col("a") < 10, proj, lambda s: 16 + 12 * s)
    timeit("filter+project T=1000", col("a") < 1000, proj, lambda s: 16 + 12 * s)
    timeit("filter only T=500", col("a") < 500, [], lambda s: 16 + 16 * s)
    timeit("filter a<500 -> a,b", col("a") < 500, [col("a"), col("b")], lambda s: 8 + 8 * s)
if len(sys.argv) > 2:
    for m in (1 << 20, 1 << 22, 1 << 24, 1 << 26):
        n = m
        for tp in ("0", "1"):
            os.environ["IGL_FP_TWO_PASS"] = tp
            timeit(f"n={m} two_pass={tp} T=500", col("a") < 500, proj, lambda s: 16 + 12 * s)
