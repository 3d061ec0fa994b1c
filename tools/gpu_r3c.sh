#!/bin/bash
TAG=${1:-r3c}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 $OUT/${TAG}_pytest_gpu.log
NCU=0 bash tools/gpu_phase.sh $TAG C2 1000000
[ -x tools/_bin/fp64_chains ] && tools/_bin/fp64_chains > $OUT/${TAG}_fp64_chains.jsonl 2>&1
