#!/bin/bash
TAG=${1:-ds1}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_dbshard.py -m gpu -x -q -rs > $OUT/${TAG}_pytest_dbshard.log 2>&1; echo "pytest rc=$?"; tail -n 5 $OUT/${TAG}_pytest_dbshard.log
