#!/bin/bash
# N-GPU box: database-sharded tests, then the C5 dbshard bench at N GPUs
TAG=${1:-ds}; N=${2:-2}; TESTS=${3:-1}
OUT=gpurun_out; mkdir -p $OUT
if [ "$TESTS" = 1 ]; then
python -m pytest tests/test_dbshard.py -m gpu -x -q -rs > $OUT/${TAG}_pytest_dbshard_${N}gpus.log 2>&1; echo "pytest rc=$?"; tail -n 5 $OUT/${TAG}_pytest_dbshard_${N}gpus.log
EPK_TEST_C5=1 timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k config5 > $OUT/${TAG}_pytest_C5.log 2>&1; echo "C5 test rc=$?"; tail -n 3 $OUT/${TAG}_pytest_C5.log
fi
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --mode dbshard --workload C5 --reads 262144 --steps 3 --warmup 1 > $OUT/${TAG}_bench_dbshard_C5_${N}gpu.json 2> $OUT/${TAG}_bench_dbshard_C5_${N}gpu.err; echo "dbshard C5 x$N rc=$?"; cut -c1-160 $OUT/${TAG}_bench_dbshard_C5_${N}gpu.json; tail -n 2 $OUT/${TAG}_bench_dbshard_C5_${N}gpu.err
