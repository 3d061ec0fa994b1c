#!/bin/bash
# Round-end record on one GPU: tests, the bench line, the ncu launch list and full capture (tools/gpu_round.sh),
# then the other workloads at HEAD with cpu_baseline and e2e, and the reference arm.
TAG=${1:-fin}
OUT=gpurun_out; mkdir -p $OUT
bash tools/gpu_round.sh $TAG 1000000 all
for WL in C3 C4 C1p C5r C5; do
  timeout 900 python bench.py --workload $WL --steps 3 --warmup 3 > $OUT/${TAG}_bench_$WL.json 2> $OUT/${TAG}_bench_$WL.err; echo "bench $WL rc=$?"
  python -c "
import json; d=json.load(open('$OUT/${TAG}_bench_$WL.json')); print('$WL', round(d['value']), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],4), 'cpu', round(d['cpu_baseline']['value']))"
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference_arm.json 2> $OUT/${TAG}_bench_reference_arm.err; echo "reference arm rc=$?"; cut -c1-200 $OUT/${TAG}_bench_reference_arm.json
