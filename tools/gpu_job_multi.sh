# One multi-GPU box session (gpurun --gpus 8): BASELINE config 4 (240 frames split by frame) at N = 1/2/4/8 for the
# own arm and N = 1/8 for the reference build, the C3 reference arm at N = 2 (same-N sharding + all-reduce), and the
# concurrent pinned host->device bandwidth of the box.  Outputs under gpurun_out/$TAG.
set -x
TAG=${TAG:-r02_multi}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python bench.py --workload C4 --steps 3 --warmup 3 --no-cpu-baseline > $O/c4_own_n1.json 2> $O/c4_own_n1.err
for n in 2 4 8; do
  timeout 300 $TR --nproc-per-node $n --master-port $((29500+n)) bench.py --gpus $n --workload C4 --steps 3 --warmup 3 > $O/c4_own_n$n.json 2> $O/c4_own_n$n.err
done
timeout 300 python bench.py --workload C4 --impl reference --steps 3 --warmup 3 > $O/c4_ref_n1.json 2> $O/c4_ref_n1.err
timeout 300 $TR --nproc-per-node 8 --master-port 29600 bench.py --gpus 8 --workload C4 --impl reference --steps 3 --warmup 3 > $O/c4_ref_n8.json 2> $O/c4_ref_n8.err
timeout 300 $TR --nproc-per-node 2 --master-port 29601 bench.py --gpus 2 --impl reference --steps 20 --warmup 5 > $O/c3_ref_n2.json 2> $O/c3_ref_n2.err
timeout 300 $TR --nproc-per-node 2 --master-port 29602 bench.py --gpus 2 --steps 20 --warmup 5 > $O/c3_own_n2.json 2> $O/c3_own_n2.err
timeout 300 $TR --nproc-per-node 8 --master-port 29603 bench.py --gpus 8 --steps 20 --warmup 5 > $O/c3_own_n8.json 2> $O/c3_own_n8.err
timeout 120 $TR --nproc-per-node 8 --master-port 29604 tools/h2d_bandwidth.py > $O/h2d_n8.json 2> $O/h2d_n8.err
timeout 120 $TR --nproc-per-node 8 --master-port 29605 tools/h2d_bandwidth.py --bind > $O/h2d_n8_bind.json 2> $O/h2d_n8_bind.err
timeout 120 python tools/h2d_bandwidth.py > $O/h2d_n1.json 2> $O/h2d_n1.err
for f in $O/*.json; do echo $f; head -c 400 $f; echo; done
tail -2 $O/*.err
