import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from igloo_b200.device import Context, HashJoin
from igloo_b200.expr import Field, col, T_I64
ctx = Context(0)
bits, npr = 23, 4_000_000
nb = 1 << bits
rk = ctx.gen(4, 0, 0, 0, nb, bits=bits); rv = ctx.gen(2, 3, 1, 0, nb)
lk = ctx.gen(2, 3, 2, 0, npr, m=nb); lw = ctx.gen(2, 3, 3, 0, npr)
j = HashJoin(ctx, 0, [Field("k", T_I64), Field("w", T_I64)], [Field("k", T_I64), Field("v", T_I64)], col("k"), col("k"))
j.build([rk, rv], nb)
out, n = j.probe([lk, lw], npr, capacity=npr)
k = out[0].data[:n].cpu().numpy().astype(np.uint64)
C = np.uint64(0xD6E8FEB86659FD93)
h = k.copy(); h ^= h >> np.uint64(32); h *= C; h ^= h >> np.uint64(32); h *= C; h ^= h >> np.uint64(32)
for lp in (4, 8):
    part = (h >> np.uint64(64 - lp)).astype(np.int64)
    print("lp", lp, "monotone:", bool((np.diff(part) >= 0).all()), "first parts", part[:5], "last", part[-5:], "changes", int((np.diff(part) != 0).sum()))
