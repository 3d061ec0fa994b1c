#!/bin/bash
# what the driver runs at round end, on one GPU: the GPU suite with -x, smoke(), the default bench line and the reference arm
TAG=${1:-head}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/ -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 $OUT/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 $OUT/${TAG}_smoke.log
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('$OUT/${TAG}_bench.json')); print('C2', d['value'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'cpu', d['cpu_baseline']['value'], 'launches', d['gpu_launches'])"
