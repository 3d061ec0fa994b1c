#!/bin/bash
TAG=${1:-pool}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_dbshard.py -m gpu -x -q -k pooled > $OUT/${TAG}_pytest_pooled.log 2>&1; echo "pytest rc=$?"; tail -n 12 $OUT/${TAG}_pytest_pooled.log
