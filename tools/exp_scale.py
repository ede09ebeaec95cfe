import json, subprocess, sys, os
for sc in (256, 64, 16, 4):
    env = dict(os.environ, IGL_JOIN_PARTITION=sys.argv[1] if len(sys.argv) > 1 else "0")
    out = subprocess.run([sys.executable, "bench.py", "--scale", str(sc), "--steps", "5", "--warmup", "3", "--skip-fp", "--skip-cpu"],
                         capture_output=True, text=True, env=env).stdout.strip().splitlines()[-1]
    d = json.loads(out); w = d["config"]["workload"]
    npr = int(w.split("probe ")[1].split()[0]); nb = 1 << int(w.split("2^")[1].split()[0])
    pm = d["roofline"]["kernel_ms"]; bm = d["ms_per_step"] - pm
    print("scale", sc, "table MB", nb * 2 * 16 >> 20, "probe ms", round(pm, 3), "ps/probe", round(pm * 1e9 / npr, 1),
          "build+init ms", round(bm, 3), "ps/build row", round(bm * 1e9 / nb, 1), flush=True)
