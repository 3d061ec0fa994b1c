#!/bin/bash
# One GPU-box visit: parity tests, the bench line, the ncu launch list and one full capture of the
# placement kernel. Usage (from the repo root, under gpurun):  bash tools/gpu_round.sh <tag> [reads_for_ncu]
# Everything lands in gpurun_out/<tag>_*; copy what should be judged into profiles/.
set -u
TAG=${1:-run}
NCU_READS=${2:-1000000}
MODE=${3:-all}          # all | ncu (skip tests and the full bench) | quick (tests + a 2 M-read bench, no ncu)
KREGEX=${4:-place_kernel}   # kernel base name; the capture skips the three warm-up launches (uint8/16/32 counts)
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/${TAG}_smi.csv 2>&1
nproc > $OUT/${TAG}_nproc.txt

if [ "$MODE" = all ]; then
echo "== pytest -m gpu"
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log

echo "== bench (full workload)"
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?"; tail -c 600 $OUT/${TAG}_bench.json
fi

if [ "$MODE" = quick ]; then
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log
python bench.py --reads 2000000 --no-cpu-baseline --no-e2e > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?"; python -c "import json; d=json.load(open('$OUT/${TAG}_bench.json')); print('reads/s', d['value'], 'frac', d['roofline']['frac'])"
exit 0
fi

echo "== ncu launch list"
CMD="python bench.py ${BENCH_ARGS:-} --reads $NCU_READS --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
$CMD > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"

echo "== ncu full capture of the placement kernel"
CMD2="python bench.py ${BENCH_ARGS:-} --reads $NCU_READS --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
$CMD2 > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$KREGEX" -s 3 -c 1 -f -o $OUT/${TAG}_prof $CMD2 > $OUT/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la $OUT | tail -20
