#!/bin/bash
# quick parity subset + phase clocks of every experimental build (no ncu)
TAG=${1:-q}; WL=${2:-C2}; N=${3:-1000000}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_parity_gpu.py -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 $OUT/${TAG}_pytest_gpu.log
NCU=${NCU:-0} bash tools/gpu_phase.sh $TAG $WL $N
