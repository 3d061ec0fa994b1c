#!/usr/bin/env python
"""The gap between epk_place (pinned host buffers, copies inside) and epk_place_device (reads resident in HBM):
the same C2 reads through both, and device-resident in launches of epk_place's chunk size (which separates the
cost of the launches from the cost of the copies). usage: python tools/e2e_probe.py [reads]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from epik_b200 import synth
from epik_b200.placer import Database, PinnedArray, Placer

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
flat = synth.make_config_db("C2")
db = Database(flat)
h_seq = PinnedArray((n * 150,), np.uint8)
h_off = PinnedArray((n + 1,), np.uint64)
for a in range(0, n, 1_000_000):
    m = min(1_000_000, n - a)
    h_seq.array[a * 150:(a + m) * 150] = synth.make_reads(0, m, 150, first=a)[0]
h_off.array[:] = np.arange(n + 1, dtype=np.uint64) * np.uint64(150)
dev = torch.device("cuda", 0)
d_seq = torch.from_numpy(h_seq.array).to(dev); d_off = torch.from_numpy(h_off.array.view(np.int64)).to(dev)
for chunk in (0,):
    p = Placer(db, chunk_reads=chunk)
    out = p.alloc_result(n, pinned=True, lengths=False)
    p.place_packed(h_seq.array, h_off.array, out=out)
    t0 = time.perf_counter()
    for _ in range(3):
        p.place_packed(h_seq.array, h_off.array, out=out)
    torch.cuda.synchronize()
    t = (time.perf_counter() - t0) / 3
    d_out = {k: torch.zeros(v.shape, dtype={np.dtype("uint32"): torch.int32, np.dtype("float32"): torch.float32, np.dtype("float64"): torch.float64}[v.dtype], device=dev) for k, v in out.items() if k != "_pinned"}
    p.place_device(d_seq, d_off, n, d_out); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        p.place_device(d_seq, d_off, n, d_out)
    torch.cuda.synchronize()
    td = (time.perf_counter() - t0) / 3
    print("chunk_reads %8d: epk_place %.1f ms (%.2f M reads/s), epk_place_device %.1f ms (%.2f M reads/s)" % (chunk, t * 1e3, n / t / 1e6, td * 1e3, n / td / 1e6), flush=True)
    p.close()

# the same reads through epk_place_device in launches of 262144 (no copies): per-launch cost of the kernel
p = Placer(db)
m = 262144
outs = {k: torch.zeros((m,) + tuple(v.shape[1:]), dtype=v.dtype, device=dev) for k, v in d_out.items()}
def chunked():
    for a in range(0, n, m):
        c = min(m, n - a)
        p.place_device(d_seq, d_off[a:a + c + 1], c, outs)
chunked(); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    chunked()
torch.cuda.synchronize()
print("device-resident in launches of %d: %.1f ms" % (m, (time.perf_counter() - t0) / 3 * 1e3), flush=True)
