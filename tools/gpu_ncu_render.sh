#!/bin/bash
# ncu --set full of the two compositing kernels only (after the plain run exited 0)
O=gpurun_out/${1:-ncu}
mkdir -p $O
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"render_" -s 8 -c 2 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
echo "ncu exit $?"; tail -2 $O/ncu_full.log
