"""Kernel experiment: where a leader warp's cycles go, per phase of a query (needs the -DEPK_PHASE_CLOCKS build:
tools/build_variant.sh phase -DEPK_PHASE_CLOCKS; EPIK_B200_LIB=tools/_bin/libepik_b200_phase.so python tools/phase_clocks.py C2)."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from epik_b200 import capi, synth
from epik_b200.placer import Database, Placer

NAMES = ["claim+offsets", "letters+encode+probe", "descriptors", "gather+accumulate", "ambiguous", "sweep pass 1", "ratios+rows", "queries",
         "sweep bar", "sweep pass 2", "sweep final selection", "lookahead D"]


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "C2"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
    variant = int(os.environ.get("EPK_VARIANT", "0"))
    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[workload]
    flat = synth.make_config_db(workload)
    db = Database(flat, device=0)
    placer = Placer(db, 7, 0.01, variant=variant)
    data, off = synth.make_reads(alphabet, n, read_len, seed=5)
    dev = torch.device("cuda:0")
    d_seq = torch.from_numpy(data).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    K = 7
    d_out = {"n_place": torch.zeros(n, dtype=torch.int32, device=dev), "branch": torch.zeros(n * K, dtype=torch.int32, device=dev),
             "score": torch.zeros(n * K, dtype=torch.float32, device=dev), "lwr": torch.zeros(n * K, dtype=torch.float64, device=dev),
             "count": torch.zeros(n * K, dtype=torch.int32, device=dev)}
    stream = torch.cuda.Stream(device=dev)
    lib = capi.lib()
    out = (C.c_ulonglong * 12)()
    have = hasattr(lib, "epk_debug_phase_clocks")
    for rep in range(3):
        if have:
            lib.epk_debug_phase_clocks(out)
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record(stream)
        placer.place_device(d_seq, d_off, n, d_out, stream=stream)
        t1.record(stream)
        stream.synchronize()
        ms = t0.elapsed_time(t1)
        rec = {"workload": workload, "variant": variant, "reads": n, "ms": ms, "reads_per_s": n / ms * 1e3}
        if have:
            lib.epk_debug_phase_clocks(out)
            q = max(1, out[7])
            rec["cycles_per_query"] = {NAMES[i]: round(out[i] / q, 1) for i in range(12) if i != 7}
            rec["total_cycles_per_query"] = round(sum(out[i] for i in range(12) if i != 7) / q, 1)
        print(json.dumps(rec), flush=True)


main()
