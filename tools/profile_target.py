"""One filter+project launch and one join (build + probe) for ncu captures."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from igloo_b200.device import Context, FilterProject, HashJoin
from igloo_b200.expr import Field, col, T_F64, T_I32, T_I64

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 16
what = sys.argv[2] if len(sys.argv) > 2 else "both"
ctx = Context(0)
if what in ("both", "fp"):
    n = 1_000_000_000 // scale
    a = ctx.gen(1, 0x16100002, 0, 0, n, 1000); b = ctx.gen(1, 0x16100002, 1, 0, n, 1 << 31); x = ctx.gen(3, 0x16100002, 2, 0, n)
    fields = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
    fp = FilterProject(ctx, fields, col("a") < 500, [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")])
    out = fp.alloc_outputs(n)
    for _ in range(3):
        _, kept = fp.run([a, b, x], n, out)
    print("fp kept", kept, "of", n)
    del a, b, x, out, fp
    torch.cuda.empty_cache()
if what in ("both", "join"):
    bits = 27 - (scale.bit_length() - 1)
    nb, npr = 1 << bits, 1_000_000_000 // scale
    rk = ctx.gen(4, 0, 0, 0, nb, bits=bits); rv = ctx.gen(2, 0x16100003, 1, 0, nb)
    lk = ctx.gen(2, 0x16100003, 2, 0, npr, m=nb); lw = ctx.gen(2, 0x16100003, 3, 0, npr)
    j = HashJoin(ctx, 0, [Field("k", T_I64), Field("w", T_I64)], [Field("k", T_I64), Field("v", T_I64)], col("k"), col("k"))
    j._right_cols = [rk, rv]
    out = j.alloc_outputs(npr, [lk, lw])
    for _ in range(3):
        j.build([rk, rv], nb)
        _, m = j.probe([lk, lw], npr, out=out, capacity=npr)
    print("join rows", m)
torch.cuda.synchronize()
