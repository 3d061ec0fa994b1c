#!/bin/bash
TAG=${1:-pool2}; N=${2:-2}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi topo -m > $OUT/${TAG}_topo.txt 2>&1; head -n 6 $OUT/${TAG}_topo.txt
for V in 16 32; do
timeout 600 python bench.py --mode dbpool --gpus $N --workload C5 --reads 100000 --steps 2 --warmup 1 --variant $V > $OUT/${TAG}_dbpool_v$V.json 2> $OUT/${TAG}_dbpool_v$V.err; echo "variant $V rc=$?"; cut -c1-120 $OUT/${TAG}_dbpool_v$V.json
done
