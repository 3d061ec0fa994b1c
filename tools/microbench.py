#!/usr/bin/env python
"""The independent random-gather ceiling (epk_gather_microbench) on a workload's database.
usage: python tools/microbench.py [workload=C2] [probes=40000000] [reps=3]   -> one JSON line per method"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from epik_b200 import capi, synth  # noqa: E402
from epik_b200.placer import Database  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "C2"
probes = int(sys.argv[2]) if len(sys.argv) > 2 else 40_000_000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
db = Database(synth.make_config_db(workload), device=0)
names = {0: "ld.global.nc.v4, 4 lists in flight per warp, 32 warps/SM",
         1: "ld.global.nc.v4, 8 lists in flight per warp, 32 warps/SM",
         2: "cp.async.bulk into shared memory (mbarrier), 4 x 2 KB stages per warp, 24 warps/SM"}
for method in (0, 1, 2):
    r = capi.Microbench()
    capi.check(capi.lib().epk_gather_microbench(db.handle, probes, 0xE91C0003, method, reps, C.byref(r)))
    print(json.dumps({"workload": workload, "method": method, "what": names[method], "ms_per_launch": r.ms_per_launch,
                      "probes": r.n_probes, "hits": r.n_hits, "entries": r.n_entries, "bytes": r.bytes,
                      "GBps": r.bytes / r.ms_per_launch / 1e6, "checksum": "%016x" % r.checksum}), flush=True)
