#!/bin/bash
# kernel experiment: phase clocks of the ablation builds (tools/_bin/libepik_b200_ph*.so), then one ncu full capture
TAG=${1:-ph}; WL=${2:-C2}; N=${3:-1000000}
OUT=gpurun_out; mkdir -p $OUT
python tools/phase_clocks.py $WL $N > $OUT/${TAG}_main.jsonl 2> $OUT/${TAG}_main.err; echo "main rc=$?"; tail -n 1 $OUT/${TAG}_main.jsonl
for f in tools/_bin/libepik_b200_ph*.so; do
  v=$(basename $f .so); v=${v#libepik_b200_}
  EPIK_B200_LIB=$f python tools/phase_clocks.py $WL $N > $OUT/${TAG}_$v.jsonl 2> $OUT/${TAG}_$v.err; echo "$v rc=$?"; tail -n 1 $OUT/${TAG}_$v.jsonl
done
if [ "${NCU:-1}" = 1 ]; then
CMD2="python bench.py --workload $WL --reads $N --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k "regex:place_kernel" -s 3 -c 1 -f -o $OUT/${TAG}_prof $CMD2 > $OUT/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
fi
