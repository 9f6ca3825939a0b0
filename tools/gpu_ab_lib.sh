#!/bin/bash
# A/B of two builds of the libraries on the box (GUAVA_LIB_DIR): guava_renderer_b200/lib (A) against guava_renderer_b200/lib_variant (B),
# the forward parity tests on B first, then the bench line interleaved.  usage: bash tools/gpu_ab_lib.sh <tag>
O=gpurun_out/${1:-ablib}; mkdir -p $O
B=$PWD/guava_renderer_b200/lib_variant
GUAVA_LIB_DIR=$B timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "forward_vs_reference or compositor_options or golden or full_size_configs" > $O/pytest_variant.log 2>&1; tail -2 $O/pytest_variant.log
for rep in 1 2 3; do for v in A B; do
  d=""; [ $v = B ] && d=$B
  GUAVA_LIB_DIR=$d timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline $EXTRA_ARGS > $O/bench_${v}_$rep.json 2> $O/bench_${v}_$rep.err
  python - <<P
import json
d=json.loads(open('$O/bench_${v}_$rep.json').read().strip().splitlines()[-1]); r=d['roofline']['per_kernel_ms']
print('$v rep $rep', d['value'], d['ms_per_step'], r['render_forward'], r['render_backward'])
P
done; done
