#!/bin/bash
# A/B of one bench.py flag on the box, interleaved twice.  usage: FLAG=--per-view-grads [EXTRA_ARGS=...] bash tools/gpu_ab_flag.sh <tag>
O=gpurun_out/${1:-ab}; mkdir -p $O
for rep in 1 2; do for v in with without; do
  f=""; [ $v = with ] && f="$FLAG"
  timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline $f $EXTRA_ARGS > $O/bench_${v}_$rep.json 2> $O/bench_${v}_$rep.err
  python - <<P
import json
d=json.loads(open('$O/bench_${v}_$rep.json').read().strip().splitlines()[-1]); r=d['roofline']['per_kernel_ms']
print('$v $FLAG rep $rep', d['value'], d['ms_per_step'], d['e2e']['value'], d.get('gpu_launches'), r['render_backward'], r['preprocess_backward'])
P
done; done
