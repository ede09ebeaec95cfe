#!/bin/bash
# A/B of the round-2 kernel changes on one GPU: C3 join phases (ms) per configuration.
run() { echo "== $1"; shift; env "$@" python bench.py --skip-fp --skip-cpu --skip-e2e --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(round(d['ms_per_step'],3), 'ms/step', d['parity']['digest_equal'], {k: round(v,3) for k,v in d['roofline']['phases_ms'].items()})"; }
run "default (compact walk, 2x512 scatter)" X=1
run "1x1024 scatter" IGL_RADIX_CTAS2=0
run "old walk" IGLOO_CUDA_LIB=$PWD/igloo_b200/libigloo_cuda_nc.so
run "old walk, 1x1024 scatter" IGLOO_CUDA_LIB=$PWD/igloo_b200/libigloo_cuda_nc.so IGL_RADIX_CTAS2=0
