#!/usr/bin/env python
"""Where the end-to-end time of a workload goes: epk_place from pinned host buffers without / with the distal and
pendant lengths, the grouping alone on the host threads, the device-resident launch, and epk_place with the pass cut
into 1, 2, 4, 8, 16 chunks. usage: python tools/e2e_probe2.py [workload] [reads]"""
import ctypes, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from epik_b200 import capi, synth
from epik_b200.placer import Database, PinnedArray, Placer

wl = sys.argv[1] if len(sys.argv) > 1 else "C1p"
alphabet, k, n_leaves, p_occ, m_len, L = synth.CONFIGS[wl]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
flat = synth.make_config_db(wl)
db = Database(flat)
h_seq = PinnedArray((n * L,), np.uint8)
h_off = PinnedArray((n + 1,), np.uint64)
h_seq.array[:] = synth.make_reads(alphabet, n, L, first=0)[0]
h_off.array[:] = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
dev = torch.device("cuda", 0)
threads = len(os.sched_getaffinity(0))


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


p = Placer(db)
for lengths in (False, True):
    out = p.alloc_result(n, pinned=True, lengths=lengths)
    ms = timed(lambda: p.place_packed(h_seq.array, h_off.array, out=out))
    print("epk_place, pinned, lengths=%s: %.2f ms (%.2f M reads/s)" % (lengths, ms, n / ms / 1e3), flush=True)
rep = np.zeros(n + 1, np.uint32)
for t in (1, 4, threads):
    ms = timed(lambda: capi.check(capi.lib().epk_group_batches(h_seq.array.ctypes.data_as(ctypes.c_void_p), h_off.array.ctypes.data_as(ctypes.c_void_p),
                                                               n, 2000, t, rep.ctypes.data_as(ctypes.c_void_p))))
    print("epk_group_batches on %d threads: %.2f ms (%.2f GB/s)" % (t, ms, n * L / ms / 1e6), flush=True)
d_seq = torch.from_numpy(h_seq.array).to(dev); d_off = torch.from_numpy(h_off.array.view(np.int64)).to(dev)
out = p.alloc_result(n, pinned=True, lengths=False)
d_out = {kk: torch.zeros(v.shape, dtype={np.dtype("uint32"): torch.int32, np.dtype("float32"): torch.float32, np.dtype("float64"): torch.float64}[v.dtype], device=dev) for kk, v in out.items() if kk != "_pinned"}
ms = timed(lambda: p.place_device(d_seq, d_off, n, d_out))
print("epk_place_device: %.2f ms (%.2f M reads/s)" % (ms, n / ms / 1e3), flush=True)
p.close()
for chunks in (1, 2, 4, 8, 16):
    p = Placer(db, chunk_reads=(n + chunks - 1) // chunks)
    ms = timed(lambda: p.place_packed(h_seq.array, h_off.array, out=out))
    print("epk_place in %d chunks: %.2f ms (%.2f M reads/s)" % (chunks, ms, n / ms / 1e3), flush=True)
    p.close()
