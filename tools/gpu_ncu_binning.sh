#!/bin/bash
# ncu --set full of the binning chain's kernels (after the plain run exited 0)
O=gpurun_out/${1:-ncu_bin}
mkdir -p $O
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline $EXTRA_ARGS"
$CMD > $O/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"scatter_kernel|preprocess_forward|tile_sort|tile_scan|preprocess_backward" -s 10 -c 5 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
echo "ncu exit $?"; tail -2 $O/ncu_full.log
