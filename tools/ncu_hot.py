"""Hot spots of one kernel in an .ncu-rep: SASS instructions holding the most warp-stall samples.
usage: ncu_hot.py report.ncu-rep kernel-regex [min_pct]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
min_pct = float(sys.argv[3]) if len(sys.argv) > 3 else 1.2
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
print(rows[0][1][:100])
hdr = rows[1]; si = hdr.index('Source'); ii = hdr.index('Instructions Executed'); wi = hdr.index('Warp Stall Sampling (All Samples)')
data = []
for r in rows[2:]:
    if len(r) <= ii: continue
    try: data.append((r[si], int(r[ii] or 0), int(r[wi] or 0)))
    except ValueError: break      # the second kernel's header
tot = sum(d[2] for d in data); ex = sum(d[1] for d in data)
print(len(data), 'SASS instructions; warp instructions executed', ex, '; stall samples', tot)
for i, (s, c, w) in enumerate(data):
    if w > tot * min_pct / 100: print(f"{i:5d} {w / tot * 100:5.1f}% exec={c:9d} {s[:110]}")
# opcode histogram of executed warp instructions
import collections
h = collections.Counter()
for s_, c, w in data:
    op = s_.split()
    if not op: continue
    o = op[1] if op[0].startswith('@') and len(op) > 1 else op[0]
    h[o.split('.')[0]] += c
print('opcode mix:', ', '.join(f"{k}={v / ex * 100:.1f}%" for k, v in h.most_common(18)))
