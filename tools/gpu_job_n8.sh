# Eight-GPU check of the final code (gpurun --gpus 8): C3 own and reference arms at N = 8, as the driver's scaling run does.
set -x
TAG=${TAG:-r02_n8}
O=gpurun_out/$TAG; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
[ -n "$SKIP_REF" ] || timeout 300 $TR --nproc-per-node 8 --master-port 29801 bench.py --gpus 8 --impl reference --steps 20 --warmup 5 > $O/c3_ref_n8.json 2> $O/c3_ref_n8.err
timeout 300 $TR --nproc-per-node 8 --master-port 29802 bench.py --gpus 8 --steps 20 --warmup 5 > $O/c3_own_n8.json 2> $O/c3_own_n8.err
timeout 300 $TR --nproc-per-node 4 --master-port 29803 bench.py --gpus 4 --steps 20 --warmup 5 > $O/c3_own_n4.json 2> $O/c3_own_n4.err
for f in $O/*.json; do echo $f; python -c "
import json; d=json.loads(open('$f').read().strip().splitlines()[-1]); print(d.get('n_gpus'), d.get('value'), d.get('e2e',{}).get('value'), d.get('clocks'))"; done
tail -n 2 $O/*.err
