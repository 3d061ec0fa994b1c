#!/usr/bin/env python
"""Process-level throughput of the drop-in driver: FASTA file in -> bin/epik/epik-dna -> jplace file out
(the contract of epik.py:85-96 / main.cpp:308-381), on a synthetic workload of bench.py.

    python tools/cli_e2e.py [--workload C2] [--reads 2000000] [--out gpurun_out/cli_e2e.json]

Reports the driver's own "Placement time" (main.cpp:370-373: parse + place + write, database load excluded), the
reference-style "Average speed" (main.cpp:348-353: time inside placer::place only), the wall time of the whole
process and the size of the jplace it wrote. Needs a CUDA device (the driver has no CPU path)."""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--reads", type=int, default=2_000_000)
    ap.add_argument("--out", default="")
    ap.add_argument("--device-batches", default="128", help="--device-batches of the driver; a comma list runs it once per value on the same files")
    ap.add_argument("--devices", default="1", help="--devices of the driver: GPUs behind the one process, or 'all'")
    ap.add_argument("--tmpdir", default=None, help="where the FASTA, the database file and the jplace go (e.g. /dev/shm: no file system in the way)")
    ap.add_argument("--keep-jplace", default="", help="copy the jplace here (to compare runs byte for byte)")
    ap.add_argument("--jobs", type=int, default=len(os.sched_getaffinity(0)), help="-j of the driver: host threads")
    args = ap.parse_args()
    from epik_b200 import flatdb, synth
    import __graft_entry__ as g
    g.build_cuda()
    g.build_host()
    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[args.workload]
    flat = synth.make_config_db(args.workload)
    tmp = tempfile.mkdtemp(prefix="epik_cli_", dir=args.tmpdir)
    db_path, q_path = os.path.join(tmp, "db.epkdb"), os.path.join(tmp, "q.fasta")
    flatdb.save_epkdb(flat, db_path, synth.tree_newick(flat.branch_length, flat.meta["parent"]))
    # FASTA: one header line + one sequence line per read, written in blocks
    with open(q_path, "wb") as f:
        step = 500_000
        for a in range(0, args.reads, step):
            n = min(step, args.reads - a)
            data, _ = synth.make_reads(alphabet, n, read_len, first=a)
            rows = data.reshape(n, read_len)
            heads = np.char.add(np.char.add(">r", np.arange(a, a + n).astype(str)), "\n").astype("S")
            block = bytearray()
            for i in range(n):
                block += heads[i]
                block += rows[i].tobytes()
                block += b"\n"
            f.write(block)
    exe = os.path.join(ROOT, "bin", "epik", "epik-aa" if alphabet else "epik-dna")
    import hashlib
    import resource
    lines = []
    for device_batches in [int(x) for x in str(args.device_batches).split(",")]:
        argv = [exe, "-d", db_path, "-q", q_path, "-j", str(args.jobs), "--omega", "1.5", "--mu", "1.0", "-o", tmp,
                "--device-batches", str(device_batches), "--devices", args.devices, q_path]
        before = resource.getrusage(resource.RUSAGE_CHILDREN)
        t0 = time.perf_counter()
        run = subprocess.run(argv, capture_output=True, text=True, env=dict(os.environ, EPK_DRIVER_TIMING="1"))
        wall = time.perf_counter() - t0
        after = resource.getrusage(resource.RUSAGE_CHILDREN)
        stages = [l for l in run.stderr.split("\n") if l.startswith("stages")]
        if run.returncode != 0:
            raise SystemExit("driver failed: " + run.stderr[-2000:])
        ms = int(re.search(r"Placement time: \d+s \((\d+) ms\)", run.stdout).group(1))
        speed = re.search(r"Average speed: ([^\n]+) seq/s", run.stdout).group(1)
        jplace = os.path.join(tmp, "placements_q.fasta.jplace")
        line = {
            "what": "bin/epik/%s: FASTA file -> placements -> jplace file" % os.path.basename(exe),
            "workload": args.workload, "reads": args.reads, "read_len": read_len, "jobs": args.jobs, "devices": args.devices,
            "device_batches": device_batches, "tmpdir": args.tmpdir or tempfile.gettempdir(),
            "placement_time_ms": ms, "reads_per_s_file_to_file": args.reads / (ms / 1e3),
            "driver_average_speed": speed, "process_wall_s": wall,
            # (of the whole process: loading the database and creating the CUDA context included)
            "process_cpu_s": {"user": after.ru_utime - before.ru_utime, "system": after.ru_stime - before.ru_stime},
            "fasta_bytes": os.path.getsize(q_path), "jplace_bytes": os.path.getsize(jplace),
            "db_file_bytes": os.path.getsize(db_path), "driver_stages": stages[0] if stages else None,
        }
        h = hashlib.sha256()
        with open(jplace, "rb") as f:
            for _ in range(4):
                f.readline()  # (the metadata block holds the invocation, which differs between runs)
            for block in iter(lambda: f.read(1 << 24), b""):
                h.update(block)
        line["jplace_sha256_without_head"] = h.hexdigest()
        lines.append(line)
        if len(lines) > 1:
            print(json.dumps(lines[-2]))
    line = lines[-1]
    print(json.dumps(line))
    if args.keep_jplace:
        import shutil
        shutil.copy(jplace, args.keep_jplace)
    if args.out:
        with open(args.out, "w") as f:
            for one in lines:
                f.write(json.dumps(one) + "\n")
    for p in (db_path, q_path, jplace):
        os.remove(p)
    os.rmdir(tmp)


if __name__ == "__main__":
    main()
