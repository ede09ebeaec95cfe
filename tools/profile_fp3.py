"""One two-pass filter+project launch at C2's shape, for an ncu capture of the three kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["IGL_FP_TWO_PASS"] = "1"
import torch
from igloo_b200.device import Context, FilterProject
from igloo_b200.expr import Field, col, T_F64, T_I32
n = 125_000_000
ctx = Context(0)
a = ctx.gen(1, 0x16100002, 0, 0, n, 1000); b = ctx.gen(1, 0x16100002, 1, 0, n, 1 << 31); x = ctx.gen(3, 0x16100002, 2, 0, n)
fields = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
proj = [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")]
fp = FilterProject(ctx, fields, col("a") < 500, proj)
out = fp.alloc_outputs(n)
for _ in range(2):
    _, kept = fp.run([a, b, x], n, out)
print(kept)
torch.cuda.synchronize()
