#!/bin/bash
# gpu_head.sh + the small-tree workloads: the 16-warp launch against the 8-warp one (--variant 1) on long reads
TAG=${1:-head}
OUT=gpurun_out; mkdir -p $OUT
bash tools/gpu_head.sh $TAG
for V in 0 1; do
  python bench.py --workload C1p --variant $V --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_bench_C1p_variant$V.json 2> $OUT/${TAG}_bench_C1p_variant$V.err
  python -c "
import json; d=json.loads(open('$OUT/${TAG}_bench_C1p_variant$V.json').read().strip().split('\n')[-1]); print('C1p variant $V', round(d['value']/1e6, 3), 'M reads/s')"
done
python bench.py --workload C4 --steps 5 --warmup 3 > $OUT/${TAG}_bench_C4.json 2> $OUT/${TAG}_bench_C4.err
python -c "
import json; d=json.loads(open('$OUT/${TAG}_bench_C4.json').read().strip().split('\n')[-1]); print('C4', round(d['value']/1e6, 3), 'M reads/s, e2e', round(d['e2e']['value']/1e6, 3))"
