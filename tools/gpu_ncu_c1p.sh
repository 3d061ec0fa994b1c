#!/bin/bash
# one full ncu capture of the 16-warp launch on long reads (C1p, 16-bit counts)
TAG=${1:-ncu_c1p}
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --workload C1p --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
$CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:place_kernel.*unsigned short" -s 1 -c 1 -f -o $OUT/${TAG}_prof $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"; tail -n 3 $OUT/${TAG}_ncu_full.log; ls -la $OUT/${TAG}_prof* 2>/dev/null
