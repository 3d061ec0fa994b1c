#!/bin/bash
for i in 1 2; do
for WL in C2 C4; do
echo "$WL HEAD (letter prefetch)"; python tools/phase_clocks.py $WL 1000000 2>/dev/null | tail -n 1 | cut -c1-140
echo "$WL without"; EPIK_B200_LIB=tools/_bin/libepik_b200_nolp.so python tools/phase_clocks.py $WL 1000000 2>/dev/null | tail -n 1 | cut -c1-140
done
done
