#!/bin/bash
for i in 1 2; do
echo "HEAD (header prefetch)"; python tools/phase_clocks.py C5r 200000 2>/dev/null | tail -n 1 | cut -c1-140
echo "no header prefetch"; EPIK_B200_LIB=tools/_bin/libepik_b200_nohp.so python tools/phase_clocks.py C5r 200000 2>/dev/null | tail -n 1 | cut -c1-140
done
python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "teams or stream or generated" 2>&1 | tail -n 2
