#!/bin/bash
TAG=${1:-cli}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_host_cpp.py -m gpu -x -q > $OUT/${TAG}_pytest_host.log 2>&1; echo "pytest rc=$?"; tail -n 3 $OUT/${TAG}_pytest_host.log
for D in /dev/shm /tmp; do
python tools/cli_e2e.py --reads 4000000 --tmpdir $D --out $OUT/${TAG}_cli_e2e_C2$(echo $D | tr / _).json > /dev/null 2> $OUT/${TAG}_cli.err; echo "cli $D rc=$?"; python -c "
import json; d=json.load(open('$OUT/${TAG}_cli_e2e_C2$(echo $D | tr / _).json')); print(d['reads_per_s_file_to_file'], d['driver_stages'], d['jplace_sha256_without_head'][:16])"
done
