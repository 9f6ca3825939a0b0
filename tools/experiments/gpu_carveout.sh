#!/bin/bash
# Sweep of the compositors' shared-memory carve-out (the rest of the 256 KB is L1) and resident CTAs per SM.
set -u
TAG=${1:-carve}; O=gpurun_out/$TAG; mkdir -p $O
run() { # name, env...
  n=$1; shift
  env "$@" timeout 150 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > $O/bench_$n.json 2> $O/bench_$n.err
  echo "== $n exit $? $(python - <<P
import json
try:
    d=json.load(open('$O/bench_$n.json')); print(d['value'], d['e2e']['value'], {k:v for k,v in d['roofline']['per_kernel_ms'].items() if 'render' in k})
except Exception as e: print('unreadable', e)
P
)"
}
run base X=1
run bwd_c100 GUAVA_BWD_CARVEOUT=100
run bwd_c72_n10 GUAVA_BWD_CARVEOUT=72 GUAVA_BWD_CTAS=10
run bwd_c58_n8 GUAVA_BWD_CARVEOUT=58 GUAVA_BWD_CTAS=8
run bwd_n10 GUAVA_BWD_CTAS=10
run fwd_c100 GUAVA_FWD_CARVEOUT=100
run fwd_c72_n14 GUAVA_FWD_CARVEOUT=72 GUAVA_FWD_CTAS=14
run fwd_c58_n11 GUAVA_FWD_CARVEOUT=58 GUAVA_FWD_CTAS=11
run fwd_n14 GUAVA_FWD_CTAS=14
