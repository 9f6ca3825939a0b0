# A/B of launch knobs with the bench command (one box): name value render_forward render_backward
O=gpurun_out/r02_sweep; mkdir -p $O
B="timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline"
run() { name=$1; shift; env "$@" $B > $O/$name.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/$name.json').read().strip().splitlines()[-1]); r=d['roofline']['per_kernel_ms']; print('$name', d['value'], r['render_forward'], r['render_backward'])"; }
run base X=1
run minb14 GUAVA_BWD_MINB=14
run minb16 GUAVA_BWD_MINB=16
run minb16_c14 GUAVA_BWD_MINB=16 GUAVA_BWD_CTAS=14
run minb16_c15 GUAVA_BWD_MINB=16 GUAVA_BWD_CTAS=15
run base2 X=1
B="$B --P 202000"; run base_202k X=1; run minb16_202k GUAVA_BWD_MINB=16
