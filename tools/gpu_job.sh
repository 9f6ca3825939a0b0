# One GPU-box session: full GPU test suite, A/B of the compositor options, both bench arms, other workloads,
# sort micro-benchmark.  Outputs under gpurun_out/$TAG.
set -x
TAG=${TAG:-r02_v2}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > $O/gpu.txt
rm -f gpurun_out/grad_parity_stats.jsonl
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1
cp gpurun_out/grad_parity_stats.jsonl $O/ 2>/dev/null
B="timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline"
$B > $O/bench_base.json 2> $O/bench_base.err
GUAVA_OPT_BLOCK_MASKS=1 $B > $O/bench_masks.json 2> $O/bench_masks.err
GUAVA_OPT_TMA_STAGING=1 $B > $O/bench_tma.json 2> $O/bench_tma.err
GUAVA_OPT_BLOCK_MASKS=1 GUAVA_OPT_TMA_STAGING=1 $B > $O/bench_masks_tma.json 2> $O/bench_masks_tma.err
$B --eager > $O/bench_eager.json 2> $O/bench_eager.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 200 python bench.py --steps 20 --warmup 5 > $O/bench_own_driverlike.json 2> $O/bench_own_driverlike.err
for part in 2048 4096 6144; do GUAVA_OS_PART=$part $B --P 202000 > $O/bench_202k_part$part.json 2> $O/bench_202k_part$part.err; done
timeout 200 python bench.py --workload C2 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err
timeout 300 python bench.py --workload C4 --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err
timeout 300 python bench.py --workload C4 --impl reference --steps 3 --warmup 3 > $O/bench_c4_ref.json 2> $O/bench_c4_ref.err
timeout 300 python tools/sort_vs_cub.py > $O/sort_vs_cub.jsonl 2> $O/sort_vs_cub.err
tail -40 $O/pytest_gpu.log
for f in $O/bench_*.json; do echo $f; head -c 300 $f; echo; done
tail -2 $O/*.err
