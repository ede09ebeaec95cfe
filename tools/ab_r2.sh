#!/bin/bash
# A/B on one GPU: C3 join phases (ms) per configuration (IGLOO_CUDA_LIB selects a variant build).
run() { echo "== $1"; shift; env "$@" python bench.py --skip-fp --skip-cpu --skip-e2e --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(round(d['ms_per_step'],3), 'ms/step', d['parity']['digest_equal'], {k: round(v,3) for k,v in d['roofline']['phases_ms'].items()})"; }
run "default" X=1
run "variant library" IGLOO_CUDA_LIB=$PWD/igloo_b200/libigloo_cuda_nc.so
