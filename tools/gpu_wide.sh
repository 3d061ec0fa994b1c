#!/bin/bash
# the 16-warp launch (small trees): shipped library and every build under tools/_bin on C1p (long reads) and C4
TAG=${1:-wide}
OUT=gpurun_out; mkdir -p $OUT
run() { # name lib
  for WL in C1p C4; do
    EPIK_B200_LIB=$2 python bench.py --workload $WL --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_$1_$WL.json 2> $OUT/${TAG}_$1_$WL.err
    python -c "
import json; d=json.loads(open('$OUT/${TAG}_$1_$WL.json').read().strip().split('\n')[-1]); print('$1', '$WL', round(d['value']/1e6, 3), 'M reads/s', d['ms_per_step'])" 2>&1 | tail -n 1
  done
}
run shipped ""
for f in tools/_bin/libepik_b200_*.so; do
  v=$(basename $f .so); v=${v#libepik_b200_}
  run $v $f
done
run shipped_again ""
# round 1's C1p record was taken on 1 M reads (profiles/r2z_bench_C1p_v12.json: 8.72 M reads/s)
python bench.py --workload C1p --reads 1000000 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_shipped_C1p_1M.json 2> $OUT/${TAG}_shipped_C1p_1M.err
python -c "
import json; d=json.loads(open('$OUT/${TAG}_shipped_C1p_1M.json').read().strip().split('\n')[-1]); print('shipped C1p 1M reads', round(d['value']/1e6, 3), 'M reads/s', d['ms_per_step'])" 2>&1 | tail -n 1
