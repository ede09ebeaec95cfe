#!/bin/bash
# usage: tools/breakdown.sh <scale> ; per-kernel durations / DRAM bytes of one partitioned join (ncu, serialised)
sc=${1:-1}
timeout 200 python tools/profile_target.py $sc join > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active -k regex:"radix|join_" -s 6 -c 7 --csv --log-file gpurun_out/l3.csv python tools/profile_target.py $sc join > /dev/null 2>&1
python - <<PY
import csv, collections
rows=list(csv.reader(open("gpurun_out/l3.csv")))
hdr=None; acc=collections.OrderedDict()
for r in rows:
    if "Kernel Name" in r: hdr=r; continue
    if hdr and len(r)==len(hdr):
        d=dict(zip(hdr,r)); acc.setdefault((d["ID"],d["Kernel Name"][:30]),{})[d["Metric Name"].split("__")[1][:14]]=d["Metric Value"]
for k,v in acc.items(): print(k[1], v)
PY
