#!/bin/bash
# the driver file to file on tmpfs: passes of 32 ... 512 batches on the same files, with the CPU time of the process
TAG=${1:-cli3}
OUT=gpurun_out; mkdir -p $OUT
python tools/cli_e2e.py --reads 4000000 --tmpdir /dev/shm --device-batches 128,32,64,256,512,128 --out $OUT/${TAG}_cli_e2e_C2_device_batches.jsonl > /dev/null 2> $OUT/${TAG}_cli.err; echo "cli rc=$?"
python -c "
import json
for l in open('$OUT/${TAG}_cli_e2e_C2_device_batches.jsonl'):
    d = json.loads(l); print(d['device_batches'], round(d['reads_per_s_file_to_file']), d['process_cpu_s'], round(d['process_wall_s'], 2), d['driver_stages'], d['jplace_sha256_without_head'][:8])"
