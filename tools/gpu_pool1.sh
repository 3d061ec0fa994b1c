#!/bin/bash
TAG=${1:-pool}
OUT=gpurun_out; mkdir -p $OUT
timeout 500 python -m pytest tests/test_dbshard.py -m gpu -x -q -k "pool" > $OUT/${TAG}_pytest_pooled.log 2>&1; echo "pytest rc=$?"; tail -n 25 $OUT/${TAG}_pytest_pooled.log | cut -c1-220
