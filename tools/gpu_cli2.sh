#!/bin/bash
# the driver file to file on tmpfs with different -j (the box has 16 host threads, shared by reader, packing and writer)
TAG=${1:-cli2}
OUT=gpurun_out; mkdir -p $OUT
for J in 16 8 12 16; do
python tools/cli_e2e.py --reads 4000000 --tmpdir /dev/shm --jobs $J --out $OUT/${TAG}_cli_e2e_C2_j$J.json > /dev/null 2> $OUT/${TAG}_cli.err; echo "cli -j $J rc=$?"
python -c "
import json; d=json.load(open('$OUT/${TAG}_cli_e2e_C2_j$J.json')); print(round(d['reads_per_s_file_to_file']), d['driver_stages'], d['jplace_sha256_without_head'][:16])"
done
