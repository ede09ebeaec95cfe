#!/bin/bash
# usage: tools/ab.sh <scale> : join bench with partitioning off / on
sc=${1:-1}
for p in 0 1; do
  IGL_JOIN_PARTITION=$p timeout 300 python bench.py --scale $sc --steps 3 --warmup 2 --skip-fp --skip-cpu 2>/dev/null | tail -1 > gpurun_out/ab_$p.json
  python -c "
import json; d=json.load(open('gpurun_out/ab_$p.json'))
print('PART=$p  G rows/s', round(d['value']/1e9,2), ' ms/step', round(d['ms_per_step'],2), ' probe ms', round(d['roofline']['kernel_ms'],2), ' e2e G rows/s', round(d['e2e']['value']/1e9,2))"
done
