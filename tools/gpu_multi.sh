#!/bin/bash
# N-GPU box: [tests,] the bench line with its dbshard leg, the one-process driver on 1 and N GPUs
TAG=${1:-gm}; N=${2:-2}; TESTS=${3:-1}; CLI_READS=${4:-2000000}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name --format=csv > $OUT/${TAG}_smi.csv 2>&1; nproc >> $OUT/${TAG}_smi.csv
if [ "$TESTS" = 1 ]; then
python -m pytest tests -m gpu -q -rs > $OUT/${TAG}_pytest_gpu_${N}gpus.log 2>&1; echo "pytest rc=$?"; tail -n 5 $OUT/${TAG}_pytest_gpu_${N}gpus.log
fi
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 3 --warmup 3 > $OUT/${TAG}_bench_C2_${N}gpu.json 2> $OUT/${TAG}_bench_C2_${N}gpu.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('$OUT/${TAG}_bench_C2_${N}gpu.json')); print('C2 x$N', d['value'], 'e2e', d['e2e']['value'], 'dbshard', (d.get('dbshard') or {}).get('value'), 'dbpool', (d.get('dbpool') or {}).get('value'))"
tail -n 3 $OUT/${TAG}_bench_C2_${N}gpu.err
python tools/cli_e2e.py --reads $CLI_READS --devices 1 --out $OUT/${TAG}_cli_e2e_C2_1gpu.json > /dev/null 2> $OUT/${TAG}_cli1.err; echo "cli 1 rc=$?"; cat $OUT/${TAG}_cli_e2e_C2_1gpu.json | cut -c1-500
python tools/cli_e2e.py --reads $CLI_READS --devices $N --out $OUT/${TAG}_cli_e2e_C2_${N}gpu.json > /dev/null 2> $OUT/${TAG}_cli$N.err; echo "cli $N rc=$?"; cat $OUT/${TAG}_cli_e2e_C2_${N}gpu.json | cut -c1-500
