#!/bin/bash
# 2-GPU box: the whole GPU suite (the two-GPU exchange tests run here), then database-sharded benches
TAG=${1:-g2}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name --format=csv > $OUT/${TAG}_smi.csv 2>&1
python -m pytest tests -m gpu -q -rs > $OUT/${TAG}_pytest_gpu_2gpus.log 2>&1; echo "pytest rc=$?"; tail -n 6 $OUT/${TAG}_pytest_gpu_2gpus.log
for WL in C2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --mode dbshard --workload $WL --reads 1000000 --steps 3 --warmup 1 > $OUT/${TAG}_bench_dbshard2_p2p_$WL.json 2> $OUT/${TAG}_bench_dbshard2_p2p_$WL.err; echo "dbshard $WL rc=$?"; cut -c1-300 $OUT/${TAG}_bench_dbshard2_p2p_$WL.json
done
