#!/bin/bash
# end-to-end figures of the small-tree workloads and the default line after the bench's host-thread change
TAG=${1:-e2e}
OUT=gpurun_out; mkdir -p $OUT
for WL in C1p C4; do
  python bench.py --workload $WL --steps 5 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_$WL.json 2> $OUT/${TAG}_bench_$WL.err; echo "bench $WL rc=$?"
  python -c "
import json; d=json.loads(open('$OUT/${TAG}_bench_$WL.json').read().strip().split('\n')[-1]); print('$WL', round(d['value']/1e6, 3), 'M reads/s, e2e', round(d['e2e']['value']/1e6, 3))"
done
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('$OUT/${TAG}_bench.json')); print('C2', d['value'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'launches', d['gpu_launches'])"
