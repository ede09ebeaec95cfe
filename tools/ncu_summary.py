"""Summarise an .ncu-rep (raw page) into the handful of numbers we track."""
import csv
import subprocess
import sys

rep = sys.argv[1]
json_out = sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None   # also write {kernel: dram bytes per launch}
traffic = {}
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__inst_executed.sum", "smsp__inst_executed.sum"]
stall = [h for h in hdr if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("=" * 100)
    for k in keys:
        if k in d:
            print(f"{k:70s} {d[k]} {units[hdr.index(k)]}")
    st = sorted(((float(d[h] or 0), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")) for h in stall), reverse=True)
    print("stalls per issue:", ", ".join(f"{n}={v:.2f}" for v, n in st[:7]))

    def _bytes(key):
        v, u = float(d[key].replace(",", "")), units[hdr.index(key)].lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}.get(u, 1)
    b = _bytes("dram__bytes_read.sum") + _bytes("dram__bytes_write.sum")
    name = d["Kernel Name"]
    if name not in traffic or b > traffic[name]["dram_bytes"]:   # the largest launch of a kernel = the full-size one
        traffic[name] = {"dram_bytes": int(b), "ms": float(d["gpu__time_duration.sum"].replace(",", "")) * ({"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(units[hdr.index("gpu__time_duration.sum")], 1.0))}
if json_out:
    import json
    import os
    json.dump({"source": os.path.basename(rep) + " (ncu --set full, one launch per kernel at the size the bench times)", "kernels": traffic},
              open(json_out, "w"), indent=1)
