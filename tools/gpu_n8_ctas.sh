# N = 8: the step against the number of CTAs NCCL may hold (sharding.limit_collective_footprint; NCCL_MAX_CTAS in the environment wins)
O=gpurun_out/${TAG:-r02_n8ctas}; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
port=29810
for c in ${CTAS:-4 2 8}; do
  port=$((port+1))
  NCCL_MAX_CTAS=$c timeout 300 $TR --nproc-per-node 8 --master-port $port bench.py --gpus 8 --steps 50 --warmup 8 --no-cpu-baseline > $O/c3_n8_ctas$c.json 2> $O/c3_n8_ctas$c.err
  python -c "
import json; d=json.loads(open('$O/c3_n8_ctas$c.json').read().strip().splitlines()[-1]); print('NCCL_MAX_CTAS=$c', d.get('n_gpus'), d.get('value'), d.get('ms_per_step'), d.get('e2e',{}).get('value'))"
done
