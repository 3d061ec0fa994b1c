#!/bin/bash
TAG=${1:-g8}; N=${2:-8}
OUT=gpurun_out; mkdir -p $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 3 --warmup 3 > $OUT/${TAG}_bench_C2_${N}gpu.json 2> $OUT/${TAG}_bench_C2_${N}gpu.err; echo "bench C2 rc=$?"
python -c "
import json; d=json.load(open('$OUT/${TAG}_bench_C2_${N}gpu.json')); print('C2 x$N', d['value'], 'e2e', d['e2e']['value']); print('dbshard', d.get('dbshard')); print('dbpool', d.get('dbpool'))"
tail -n 3 $OUT/${TAG}_bench_C2_${N}gpu.err
