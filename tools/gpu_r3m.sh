#!/bin/bash
TAG=${1:-r3m}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 $OUT/${TAG}_pytest_gpu.log
EPK_TEST_C5=1 timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k config5 > $OUT/${TAG}_pytest_C5.log 2>&1; echo "C5 test rc=$?"; tail -n 4 $OUT/${TAG}_pytest_C5.log
# sanitizer passes on the smallest shapes: named-barrier teams, aliased buffers, lane-local appends
for TOOL in memcheck racecheck; do
  timeout 1200 compute-sanitizer --tool $TOOL --log-file $OUT/${TAG}_sanitizer_$TOOL.log python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "kat1 or golden or ties" > $OUT/${TAG}_sanitizer_${TOOL}_pytest.log 2>&1
  echo "$TOOL rc=$?"; tail -n 3 $OUT/${TAG}_sanitizer_${TOOL}_pytest.log; grep -c "ERROR SUMMARY" $OUT/${TAG}_sanitizer_$TOOL.log; grep "ERROR SUMMARY" $OUT/${TAG}_sanitizer_$TOOL.log | sort | uniq -c | head
done
