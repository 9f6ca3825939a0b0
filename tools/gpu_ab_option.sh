#!/bin/bash
# A/B of one library option on the box: the compositor-option parity tests, then the bench line with the option at
# each value, interleaved twice.  usage: OPT=TMA_STAGING VALS="0 2" [EXTRA_ARGS="--P 202000"] bash tools/gpu_ab_option.sh <tag>
O=gpurun_out/${1:-ab}; mkdir -p $O
OPT=${OPT:-TMA_STAGING}; VALS=${VALS:-"0 2"}
timeout 600 python -m pytest tests/test_gpu_parity.py -q -k "compositor_options or taken_masks_option" > $O/pytest.log 2>&1; tail -3 $O/pytest.log
for rep in 1 2; do for v in $VALS; do
  env GUAVA_OPT_$OPT=$v timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline $EXTRA_ARGS > $O/bench_${v}_$rep.json 2> $O/bench_${v}_$rep.err
  python - <<P
import json
d=json.loads(open('$O/bench_${v}_$rep.json').read().strip().splitlines()[-1]); r=d['roofline']['per_kernel_ms']
print('$OPT=$v rep $rep', d['value'], d['e2e']['value'], r['render_forward'], r['render_backward'])
P
done; done
