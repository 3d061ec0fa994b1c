#!/bin/bash
TAG=${1:-x2}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_parity_gpu.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 2 $OUT/${TAG}_pytest.log
for WL in C1p C4; do
  N=1000000; [ $WL = C1p ] && N=100000
  echo "$WL HEAD"; python tools/phase_clocks.py $WL $N 2>/dev/null | tail -n 1 | cut -c1-140
  echo "$WL wide launch looking ahead"; EPIK_B200_LIB=tools/_bin/libepik_b200_wideahead.so python tools/phase_clocks.py $WL $N 2>/dev/null | tail -n 1 | cut -c1-140
done
