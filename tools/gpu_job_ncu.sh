# One GPU-box session for the profiler evidence of the benchmarked step (B200_PROFILING.md recipe):
# plain run -> launch list (gpu__time_duration per launch) -> plain run -> `--set full` capture of the library's
# kernels.  Outputs under gpurun_out/$TAG; summarise here with tools/ncu_extract.py into profiles/.
set -x
TAG=${TAG:-r02_ncu}
O=gpurun_out/$TAG; mkdir -p $O
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$CMD > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:"render_|preprocess_|tile_|scatter_|pack_|image_dot" -s 30 -c 20 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $O
