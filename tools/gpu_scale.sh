#!/bin/bash
# Scaling visit (run with gpurun --gpus 8): the driver's launch line at N = 8, 4, 2 plus a run without NUMA binding.
set -u
O=gpurun_out/${1:-scale}
mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1; lscpu | grep -i "numa\|socket\|^CPU(s)" >> $O/topo.txt
run() { # N extra-args tag
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 50 --warmup 5 $2 2> $O/n$1$3.err | tail -1 > $O/n$1$3.json
  echo "N=$1 $3: $(cut -c1-200 $O/n$1$3.json)"; grep -o '"e2e": {[^}]*}' $O/n$1$3.json
}
run 8 "" ""
run 8 "--no-numa-bind" "_nobind"
run 4 "" ""
run 2 "" ""
python bench.py --steps 50 --warmup 5 --no-cpu-baseline 2> $O/n1.err | tail -1 > $O/n1.json; cut -c1-200 $O/n1.json; grep -o '"e2e": {[^}]*}' $O/n1.json
