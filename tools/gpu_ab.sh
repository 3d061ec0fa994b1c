#!/bin/bash
# A/B of kernel variants on one box: bash tools/gpu_ab.sh <tag> <reads> <workload> <variant> [<variant> ...]
TAG=$1; READS=$2; WL=$3; shift 3
mkdir -p gpurun_out
for V in "$@"; do
  python bench.py --workload $WL --reads $READS --variant $V --no-cpu-baseline --no-e2e > gpurun_out/${TAG}_${WL}_v$V.json 2> gpurun_out/${TAG}_${WL}_v$V.err
  echo "variant $V rc=$?"; python -c "import json; d=json.load(open('gpurun_out/${TAG}_${WL}_v$V.json')); print('$WL variant $V reads/s', d['value'], 'frac', d['roofline']['frac'])"
done
