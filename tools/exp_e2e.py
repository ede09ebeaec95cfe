"""e2e of the C3 join through the operator API only (bench.py's run_e2e), at 1/scale; IGL_OP_TRACE=1 prints where the host time goes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from igloo_b200.device import Context
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 4
ctx = Context(0)
wl = bench.C3(ctx, scale)
for pinned in (True, False):
    r = bench.run_e2e(ctx, wl, pinned=pinned, steps=2)
    print(pinned, f"{r['value'] / 1e9:.3f} G rows/s", f"{r['ms_per_step']:.1f} ms", flush=True)
