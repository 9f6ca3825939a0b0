# Short GPU-box session: full GPU test suite and the bench line (default workload, twice; P = 202 k once).
set -x
TAG=${TAG:-r02_light}
O=gpurun_out/$TAG; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1
B="timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline"
$B > $O/bench_base.json 2> $O/bench_base.err
$B > $O/bench_base2.json 2> $O/bench_base2.err
$B --P 202000 > $O/bench_202k.json 2> $O/bench_202k.err
$EXTRA
tail -5 $O/pytest_gpu.log
for f in $O/bench_*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); r=d.get('roofline',{})
print(d.get('value'), d.get('e2e',{}).get('value'), r.get('per_kernel_ms'))"; done
