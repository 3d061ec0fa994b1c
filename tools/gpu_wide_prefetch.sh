#!/bin/bash
# the 16-warp launch with the L2 prefetch of a descriptor buffer's lists before its gather: none (default there) / lines / sectors
TAG=${1:-wpf}
OUT=gpurun_out; mkdir -p $OUT
for WL in C1p C4; do
for V in 0 32 48; do
  python bench.py --workload $WL --variant $V --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_bench_${WL}_variant$V.json 2> $OUT/${TAG}_bench_${WL}_variant$V.err
  python -c "
import json; d=json.loads(open('$OUT/${TAG}_bench_${WL}_variant$V.json').read().strip().split('\n')[-1]); print('$WL variant $V', round(d['value']/1e6, 3), 'M reads/s')"
done
done
