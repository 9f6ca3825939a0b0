#!/bin/bash
# quick GPU iteration: forward/backward parity subset + bench stage times
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "(reference_build and tile_buckets) or golden or backward_vs_cpu" 2>&1 | tail -3
python bench.py --steps 50 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['per_kernel_ms'])"
