#!/bin/bash
# One GPU-box visit: parity tests, both bench arms, optional variants.  Output -> gpurun_out/
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_own.json 2> gpurun_out/bench_own.err; echo "bench exit $?"
cat gpurun_out/bench_own.json
GUAVA_OPT_SCALAR_RED=1 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_own_scalar_red.json 2>&1; echo "bench(scalar red) exit $?"
cat gpurun_out/bench_own_scalar_red.json
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "benchref exit $?"
cat gpurun_out/bench_ref.json
