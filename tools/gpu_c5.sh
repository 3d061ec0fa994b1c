#!/bin/bash
# BASELINE config 5 proper on one GPU: generator parity (small shapes), the 31 GB database itself, its bench line
TAG=${1:-c5}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,memory.total --format=csv > $OUT/${TAG}_smi.csv 2>&1; free -g | head -2 >> $OUT/${TAG}_smi.csv
python -m pytest tests/test_parity_gpu.py tests/test_dbshard.py -m gpu -x -q -k "generated_on_the_device or load_filter or sharded_load or ambiguous or one_shard" > $OUT/${TAG}_pytest_generator.log 2>&1; echo "generator tests rc=$?"; tail -n 4 $OUT/${TAG}_pytest_generator.log
EPK_TEST_C5=1 timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k config5 -s > $OUT/${TAG}_pytest_C5.log 2>&1; echo "C5 test rc=$?"; tail -n 4 $OUT/${TAG}_pytest_C5.log
timeout 900 python bench.py --workload C5 --steps 3 --warmup 1 > $OUT/${TAG}_bench_C5.json 2> $OUT/${TAG}_bench_C5.err; echo "bench C5 rc=$?"; cut -c1-400 $OUT/${TAG}_bench_C5.json; tail -n 3 $OUT/${TAG}_bench_C5.err
