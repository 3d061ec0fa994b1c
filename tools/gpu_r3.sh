#!/bin/bash
# round-2 experiments on one box: quick parity, A/B of the prefetch modes, the independent gather ceiling
TAG=${1:-r3}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log
for V in ${VARIANTS:-0 32 48}; do
  python bench.py --reads ${READS:-4000000} --variant $V --no-cpu-baseline --no-e2e > $OUT/${TAG}_C2_v$V.json 2> $OUT/${TAG}_C2_v$V.err
  echo "variant $V rc=$?"; python -c "import json; d=json.load(open('$OUT/${TAG}_C2_v$V.json')); print('C2 variant $V reads/s', d['value'], 'frac', d['roofline']['frac'], 'ceiling', d['roofline'].get('gather_ceiling'))"
done
timeout 300 python tools/microbench.py C2 > $OUT/${TAG}_microbench_C2.jsonl 2> $OUT/${TAG}_microbench_C2.err; echo "microbench rc=$?"; cat $OUT/${TAG}_microbench_C2.jsonl | cut -c1-400
