#!/bin/bash
# A/B of one environment knob on the box: [the GPU suite,] then the bench line with the knob at each value, interleaved twice.
# usage: VAR=GUAVA_SCAN_SINGLE VALS="0 1" [TESTS=1] [EXTRA_ARGS="--P 202000"] bash tools/gpu_ab_env.sh <tag>
O=gpurun_out/${1:-ab}; mkdir -p $O
VAR=${VAR:-GUAVA_SCAN_SINGLE}; VALS=${VALS:-"0 1"}
if [ -n "$TESTS" ]; then timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; tail -3 $O/pytest_gpu.log; fi
for rep in 1 2; do for v in $VALS; do
  env $VAR=$v timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline $EXTRA_ARGS > $O/bench_${v}_$rep.json 2> $O/bench_${v}_$rep.err
  python - <<P
import json
d=json.loads(open('$O/bench_${v}_$rep.json').read().strip().splitlines()[-1]); r=d['roofline']['per_kernel_ms']
print('$VAR=$v rep $rep', d['value'], d['e2e']['value'], {k: r[k] for k in ('preprocess','tile_scan','scatter','tile_sort')})
P
done; done
