#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then after every other GPU test
TAG=${1:-na}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/${TAG}_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -n 4 $OUT/${TAG}_pytest_noalloc_alone.log
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "all rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
