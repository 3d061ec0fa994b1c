#!/bin/bash
TAG=${1:-pool}; N=${2:-2}; READS=${3:-1000000}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_dbshard.py -m gpu -x -q -k "pooled" > $OUT/${TAG}_pytest_pooled_${N}gpus.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_pooled_${N}gpus.log
timeout 900 python bench.py --mode dbpool --gpus $N --workload C5 --reads $READS --steps 3 --warmup 1 > $OUT/${TAG}_bench_dbpool_C5_${N}gpu.json 2> $OUT/${TAG}_bench_dbpool_C5_${N}gpu.err; echo "dbpool C5 x$N rc=$?"; cut -c1-900 $OUT/${TAG}_bench_dbpool_C5_${N}gpu.json; tail -n 3 $OUT/${TAG}_bench_dbpool_C5_${N}gpu.err
