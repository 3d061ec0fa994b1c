#!/bin/bash
# the driver on two GPUs behind one process: host tests (the jplace of --devices all against --devices 1) and file to file
TAG=${1:-cli2g}
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_host_cpp.py -m gpu -x -q > $OUT/${TAG}_pytest_host_2gpus.log 2>&1; echo "pytest rc=$?"; tail -n 2 $OUT/${TAG}_pytest_host_2gpus.log
python tools/cli_e2e.py --reads 4000000 --tmpdir /dev/shm --devices 2 --out $OUT/${TAG}_cli_e2e_C2_2gpu.json > /dev/null 2> $OUT/${TAG}_cli.err; echo "cli rc=$?"
python -c "
import json; d=json.load(open('$OUT/${TAG}_cli_e2e_C2_2gpu.json')); print(round(d['reads_per_s_file_to_file']), d['driver_stages'], d['jplace_sha256_without_head'][:16])"
