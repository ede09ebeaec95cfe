import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from igloo_b200.device import Context, FilterProject
from igloo_b200.expr import Field, col, T_F64, T_I32
n = 125_000_000
ctx = Context(0)
a = ctx.gen(1, 0x16100002, 0, 0, n, 1000); b = ctx.gen(1, 0x16100002, 1, 0, n, 1 << 31); x = ctx.gen(3, 0x16100002, 2, 0, n)
fields = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
proj = [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")]
for pred in (None, col("a") < 500):
    fp = FilterProject(ctx, fields, pred, proj)
    out = fp.alloc_outputs(n)
    for _ in range(2):
        _, kept = fp.run([a, b, x], n, out)
    print(kept)
torch.cuda.synchronize()
