# Two-GPU check of the torchrun paths (gpurun --gpus 2): C3 own / reference arms and C4 at N = 2.
set -x
TAG=${TAG:-r02_n2}
O=gpurun_out/$TAG; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 2 --master-port 29701 bench.py --gpus 2 --steps 20 --warmup 5 > $O/c3_own_n2.json 2> $O/c3_own_n2.err
timeout 300 $TR --nproc-per-node 2 --master-port 29702 bench.py --gpus 2 --impl reference --steps 20 --warmup 5 > $O/c3_ref_n2.json 2> $O/c3_ref_n2.err
timeout 300 $TR --nproc-per-node 2 --master-port 29703 bench.py --gpus 2 --workload C4 --steps 3 --warmup 3 > $O/c4_own_n2.json 2> $O/c4_own_n2.err
for f in $O/*.json; do echo $f; python -c "
import json; d=json.loads(open('$f').read().strip().splitlines()[-1]); print(d.get('n_gpus'), d.get('value'), d.get('e2e',{}).get('value'), d.get('clocks'))"; done
tail -n 3 $O/*.err
