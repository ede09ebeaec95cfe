#!/bin/bash
# usage: tools/abe.sh <scale> VAR v1 v2 ... : join bench (device-resident only) with env VAR set to each value
sc=$1; var=$2; shift 2
for v in "$@"; do
  env $var=$v timeout 300 python bench.py --scale $sc --steps 5 --warmup 3 --skip-fp --skip-cpu --skip-e2e 2>/dev/null | tail -1 > gpurun_out/abv.json
  python -c "
import json; d=json.load(open('gpurun_out/abv.json'))
print('$var=$v', ' G rows/s', round(d['value']/1e9,2), ' ms/step', round(d['ms_per_step'],2), d['roofline']['phases_ms'])"
done
