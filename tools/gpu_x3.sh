#!/bin/bash
TAG=${1:-x3}
OUT=gpurun_out; mkdir -p $OUT
for i in 1 2; do
echo HEAD; python tools/phase_clocks.py C2 1000000 2>/dev/null | tail -n 1 | cut -c1-140
echo emitfast; EPIK_B200_LIB=tools/_bin/libepik_b200_emitfast.so python tools/phase_clocks.py C2 1000000 2>/dev/null | tail -n 1 | cut -c1-140
done
BENCH_ARGS="--workload C5" bash tools/gpu_round.sh ${TAG}_C5 200000 ncu
