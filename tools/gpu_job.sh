# One GPU-box session: full GPU test suite, A/B of the compositor options, both bench arms, other workloads,
# sort micro-benchmark, then the profiler passes on the default configuration.  Outputs under gpurun_out/$TAG.
set -x
TAG=${TAG:-r02_v10}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > $O/gpu.txt
rm -f gpurun_out/grad_parity_stats.jsonl
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1
cp gpurun_out/grad_parity_stats.jsonl $O/ 2>/dev/null
B="timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline"
$B > $O/bench_base.json 2> $O/bench_base.err
GUAVA_OPT_TAKEN_MASKS=0 $B > $O/bench_nomasks.json 2> $O/bench_nomasks.err
$B > $O/bench_base2.json 2> $O/bench_base2.err
GUAVA_OPT_TAKEN_MASKS=0 $B > $O/bench_nomasks2.json 2> $O/bench_nomasks2.err
$B --eager > $O/bench_eager.json 2> $O/bench_eager.err
$B --P 202000 > $O/bench_202k.json 2> $O/bench_202k.err
GUAVA_OPT_TAKEN_MASKS=0 $B --P 202000 > $O/bench_202k_nomasks.json 2> $O/bench_202k_nomasks.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 --P 202000 --no-cpu-baseline > $O/bench_ref_202k.json 2> $O/bench_ref_202k.err
timeout 200 python bench.py --steps 20 --warmup 5 > $O/bench_own_driverlike.json 2> $O/bench_own_driverlike.err
timeout 300 python bench.py --workload C3pose --steps 100 --warmup 10 --no-cpu-baseline > $O/bench_pose.json 2> $O/bench_pose.err
timeout 300 python bench.py --workload C3pose --impl reference --steps 20 --warmup 5 > $O/bench_pose_ref.json 2> $O/bench_pose_ref.err
timeout 200 python bench.py --workload C2 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err
timeout 300 python bench.py --workload C4 --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err
timeout 300 python tools/sort_vs_cub.py > $O/sort_vs_cub.jsonl 2> $O/sort_vs_cub.err
timeout 300 python tools/bench_configs.py > $O/configs.jsonl 2> $O/configs.err
tail -15 $O/pytest_gpu.log
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$CMD > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:"render_|preprocess_|tile_|scatter_|pack_|image_dot" -s 30 -c 20 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $O
