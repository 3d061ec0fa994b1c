#!/bin/bash
TAG=${1:-host}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_host_cpp.py -m gpu -x -q > $OUT/${TAG}_pytest_host.log 2>&1; echo "pytest rc=$?"; tail -n 12 $OUT/${TAG}_pytest_host.log
