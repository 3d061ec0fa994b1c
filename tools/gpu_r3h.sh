#!/bin/bash
TAG=${1:-r3h}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
for V in 16 32 64; do
  EPK_VARIANT=$V EPIK_B200_LIB=tools/_bin/libepik_b200_ph0.so python tools/phase_clocks.py C2 1000000 > $OUT/${TAG}_ph0_v$V.jsonl 2> $OUT/${TAG}_ph0_v$V.err; echo "ph0 v$V rc=$?"; tail -n 1 $OUT/${TAG}_ph0_v$V.jsonl
  EPK_VARIANT=$V python tools/phase_clocks.py C2 1000000 > $OUT/${TAG}_main_v$V.jsonl 2> $OUT/${TAG}_main_v$V.err; echo "main v$V rc=$?"; tail -n 1 $OUT/${TAG}_main_v$V.jsonl
done
[ -x tools/_bin/op_rates ] && tools/_bin/op_rates > $OUT/${TAG}_op_rates.jsonl 2>&1
