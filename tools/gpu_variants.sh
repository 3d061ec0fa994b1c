#!/bin/bash
# timing of every experimental build under tools/_bin (and the shipped one) on one workload
TAG=${1:-var}; WL=${2:-C2}; N=${3:-1000000}
OUT=gpurun_out; mkdir -p $OUT
python tools/phase_clocks.py $WL $N > $OUT/${TAG}_main.jsonl 2>/dev/null; echo main; tail -n 1 $OUT/${TAG}_main.jsonl | cut -c1-200
for f in tools/_bin/libepik_b200_*.so; do
  v=$(basename $f .so); v=${v#libepik_b200_}
  EPIK_B200_LIB=$f python tools/phase_clocks.py $WL $N > $OUT/${TAG}_$v.jsonl 2> $OUT/${TAG}_$v.err; echo "$v rc=$?"; tail -n 1 $OUT/${TAG}_$v.jsonl | cut -c1-200
done
