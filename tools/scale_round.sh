#!/bin/bash
# Multi-GPU bench lines on ONE box: bash tools/scale_round.sh TAG N "C2 C4"   (run under gpurun --gpus N)
TAG=${1:-rXX}; N=${2:-2}; CFGS=${3:-"C2 C4"}
O=gpurun_out; mkdir -p $O
for C in $CFGS; do
  c=${C,,}
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus $N --config $C --steps 10 --warmup 3 > $O/${TAG}_scale_${c}_n${N}.json 2> $O/${TAG}_scale_${c}_n${N}.err
  tail -1 $O/${TAG}_scale_${c}_n${N}.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']
print('$C N=$N value', round(d['value'],2), d['unit'], 'ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],2), 'regions', d.get('timed_regions_ms_per_step'), 'per-rank compute', c.get('per_rank_compute_ms_without_exchange'), c.get('exchange'))"
done
