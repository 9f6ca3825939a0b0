#!/bin/bash
# One GPU-box visit: parity tests, both bench arms, then (only after the plain run exited 0)
# the ncu launch list and one `--set full` capture of the library's kernels.  Output -> gpurun_out/
set -u
TAG=${1:-run}
O=gpurun_out/$TAG
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_reference.json 2> $O/bench_reference.err; echo "benchref exit $?"
cat $O/bench_reference.json
python bench.py --steps 50 --warmup 5 > $O/bench_own.json 2> $O/bench_own.err; echo "bench exit $?"
cat $O/bench_own.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$CMD > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:"render_|preprocess_|tile_|scatter_|pack_|image_dot" -s 30 -c 20 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $O
