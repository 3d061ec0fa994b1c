#!/bin/bash
TAG=${1:-x}
OUT=gpurun_out; mkdir -p $OUT
for WL in C1p C4 C5r C2; do
  N=200000; [ $WL = C1p ] && N=100000; [ $WL = C2 ] && N=1000000; [ $WL = C4 ] && N=1000000
  for V in 16 32; do
    echo "$WL variant $V"; EPK_VARIANT=$V python tools/phase_clocks.py $WL $N 2>/dev/null | tail -n 1 | cut -c1-140
  done
done
for WL in C1p C4; do
  python bench.py --workload $WL --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_$WL.json 2> $OUT/${TAG}_bench_$WL.err; echo "bench $WL rc=$?"
  python -c "
import json; d=json.load(open('$OUT/${TAG}_bench_$WL.json')); print('$WL', round(d['value']), 'e2e', round(d['e2e']['value']))"
done
