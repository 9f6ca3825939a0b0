#!/bin/bash
# Generic A/B: bench every guava_renderer_b200/lib/variants/lib_bwd<name>.so named in $VARIANTS (in-place swap of the
# library), GPU parity suite on those in $TEST, default restored at the end.  Output -> gpurun_out/<tag>/
set -u
TAG=${1:-ab}; O=gpurun_out/$TAG; mkdir -p $O
L=guava_renderer_b200/lib; V=$L/variants
for v in $VARIANTS; do
  cp $V/lib_bwd$v.so $L/libguava_rasterizer_b200.so
  timeout 150 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > $O/bench_$v.json 2> $O/bench_$v.err
  echo "== $v exit $? $(python - <<P
import json
try:
    d=json.load(open('$O/bench_$v.json')); print(d['value'], d['e2e']['value'], {k:v for k,v in d['roofline']['per_kernel_ms'].items() if 'render' in k})
except Exception as e: print('unreadable', e)
P
)"
done
for v in ${TEST:-}; do
  cp $V/lib_bwd$v.so $L/libguava_rasterizer_b200.so
  timeout 200 python -m pytest tests -m gpu -x -q > $O/pytest_$v.log 2>&1; echo "== pytest $v exit $?"; tail -3 $O/pytest_$v.log
done
cp $V/lib_bwd12.so $L/libguava_rasterizer_b200.so
