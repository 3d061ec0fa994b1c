#!/bin/bash
TAG=${1:-ns}
OUT=gpurun_out; mkdir -p $OUT
EPIK_B200_LIB=tools/_bin/libepik_b200_nosync.so python -m pytest tests/test_parity_gpu.py tests/test_dbshard.py -m gpu -x -q > $OUT/${TAG}_pytest_nosync.log 2>&1; echo "pytest nosync rc=$?"; tail -n 3 $OUT/${TAG}_pytest_nosync.log
for V in ph0 ph0nosync nosync; do
  EPIK_B200_LIB=tools/_bin/libepik_b200_$V.so python tools/phase_clocks.py C2 1000000 > $OUT/${TAG}_$V.jsonl 2> $OUT/${TAG}_$V.err; echo "$V rc=$?"; tail -n 1 $OUT/${TAG}_$V.jsonl
done
python tools/phase_clocks.py C2 1000000 > $OUT/${TAG}_main.jsonl 2>/dev/null; tail -n 1 $OUT/${TAG}_main.jsonl
EPIK_B200_LIB=tools/_bin/libepik_b200_nosync.so python tools/phase_clocks.py C5r 200000 > $OUT/${TAG}_nosync_C5r.jsonl 2>/dev/null; tail -n 1 $OUT/${TAG}_nosync_C5r.jsonl
python tools/phase_clocks.py C5r 200000 > $OUT/${TAG}_main_C5r.jsonl 2>/dev/null; tail -n 1 $OUT/${TAG}_main_C5r.jsonl
