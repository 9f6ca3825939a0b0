#!/bin/bash
# tests + own bench (+ optional extra args) -> gpurun_out/<tag>/
set -u
TAG=${1:-qb}; shift || true
O=gpurun_out/$TAG; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $O/pytest_gpu.log
python bench.py --steps 50 --warmup 5 --no-cpu-baseline "$@" > $O/bench_own.json 2> $O/bench_own.err; echo "bench exit $?"
cat $O/bench_own.json; tail -3 $O/bench_own.err
