#!/bin/bash
TAG=${1:-r3l}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
EPK_TEST_C5=1 timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k config5 > $OUT/${TAG}_pytest_C5.log 2>&1; echo "C5 test rc=$?"; tail -n 4 $OUT/${TAG}_pytest_C5.log
python bench.py > $OUT/${TAG}_bench_C2.json 2> $OUT/${TAG}_bench_C2.err; echo "bench C2 rc=$?"; python -c "
import json; d=json.load(open('$OUT/${TAG}_bench_C2.json')); print('C2', d['value'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'cpu', d['cpu_baseline']['value'])"
timeout 900 python bench.py --workload C5 --steps 3 --warmup 1 > $OUT/${TAG}_bench_C5.json 2> $OUT/${TAG}_bench_C5.err; echo "bench C5 rc=$?"; python -c "
import json; d=json.load(open('$OUT/${TAG}_bench_C5.json')); print('C5', d['value'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'cpu', d['cpu_baseline']['value'])"
