#!/bin/bash
# usage: tools/abv.sh <scale> lib1.so lib2.so ... : join bench (device-resident only) with each library build
sc=$1; shift
for l in "$@"; do
  IGLOO_CUDA_LIB=$PWD/$l timeout 300 python bench.py --scale $sc --steps 5 --warmup 3 --skip-fp --skip-cpu --skip-e2e 2>/dev/null | tail -1 > gpurun_out/abv.json
  python -c "
import json; d=json.load(open('gpurun_out/abv.json'))
print('$l', ' G rows/s', round(d['value']/1e9,2), ' ms/step', round(d['ms_per_step'],2), ' probe ms', round(d['roofline']['kernel_ms'],2))"
done
