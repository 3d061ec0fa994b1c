#!/usr/bin/env python
"""Throughput of the placement hot path on the configuration BASELINE.json's metric is quoted on.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C2] [--reads R]

Workload (default C2 = BASELINE.json configs[1]): synthetic nucleotide database, k = 10, 4999 branches,
~33 M (branch, score) entries; 10 M synthetic 150-bp reads per GPU per step. A step = one pass of the hot
path (encode -> look-up -> accumulate -> top-k/LWR) over the whole read set.

  value  reads/s with reads, offsets and result arrays resident in HBM (epk_place_device, CUDA events)
  e2e    reads/s through epk_place with pinned HOST buffers, H2D and D2H inside the timed region
  roofline.achieved  algorithmic bytes per launch (L + 16*n_k + 8*sum(l) + 24*K per read)
                     / launch duration; peak = measured copy bandwidth of MEASURED_PEAKS.json
  cpu_baseline       the reference's place.cpp (oracle/_ref, OpenMP, all host cores) on a bounded sample

Multi-GPU (torchrun, one rank per GPU): reads are sharded across ranks, the database is replicated,
there is no data-path collective; NCCL only carries the barrier and the max-over-ranks of the time.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "placed queries/sec"
UNIT = "reads/s"


T0 = time.time()


def log(msg):
    """progress to stderr (stdout carries only the JSON line)"""
    print("[bench %6.1fs rank %s] %s" % (time.time() - T0, os.environ.get("RANK", "0"), msg), file=sys.stderr, flush=True)


REAL_STDOUT = None


def emit(line):
    out = REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C2", "C3", "C4", "C1p", "C5r", "C5"])
    ap.add_argument("--mode", default="replicated", choices=["replicated", "dbshard", "dbpool"],
                    help="replicated: queries sharded over the ranks, database in every HBM (default); "
                         "dbshard: database sharded by key over the ranks, every rank sees every query, partial rows "
                         "exchanged with an all-to-all (SURVEY.md 8e second row)")
    ap.add_argument("--dbshard-rows", type=int, default=32768, help="queries per exchange step in dbshard mode")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "a2a"],
                    help="dbshard mode: p2p = the partial launch stores its rows into the owner's HBM over NVLink (CUDA "
                         "IPC), a2a = NCCL all_to_all_single of staged rows")
    ap.add_argument("--reads", type=int, default=0, help="reads per GPU per step (default: the workload's)")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-dbshard", action="store_true", help="N > 1: skip the database-sharded leg (the `dbshard` key)")
    ap.add_argument("--no-dbpool", action="store_true", help="N > 1: skip the pooled-database leg (the `dbpool` key)")
    ap.add_argument("--dbpool-reads", type=int, default=1000000, help="reads per GPU per step of the pooled-database leg")
    ap.add_argument("--dbshard-reads", type=int, default=262144, help="reads per step of the database-sharded leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


WORKLOAD_READS = {"C2": 10_000_000, "C3": 10_000_000, "C4": 1_000_000, "C1p": 100_000, "C5r": 1_000_000, "C5": 2_000_000}
DEVICE_GENERATED = ("C5",)  # databases that exist only as the on-device generator makes them (epk_db_create_synthetic)


def workload_config(name, n_reads, world):
    from epik_b200 import synth
    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[name]
    flt = synth.LOAD_FILTERS.get(name)
    return {
        "workload": "%s: synthetic %s DB k=%d, %d branches, p_occ=%.2f, m=%.0f%s; %d x %d-letter reads per GPU per step"
                    % (name, "amino" if alphabet else "nucl", k, 2 * n_leaves - 1, p_occ, m,
                       " loaded with mu=%g omega=%g max_entries=%d" % flt if flt else "", n_reads, read_len),
        "reads_per_gpu": n_reads, "read_len": read_len, "k": k, "branches": 2 * n_leaves - 1,
        "keep_at_most": 7, "keep_factor": 0.01, "omega": 1.5,
        "parallelism": "query-sharded x%d, DB replicated" % world,
        "l2": "inputs larger than L2 (reads %.0f MB + DB per step vs 126 MB L2); no explicit flush" % (n_reads * read_len / 1e6),
    }


def make_reads_chunked(alphabet, n, length, first, out_bytes):
    from epik_b200 import synth
    step = 1_000_000
    for a in range(0, n, step):
        m = min(step, n - a)
        data, _ = synth.make_reads(alphabet, m, length, first=first + a)
        out_bytes[a * length:(a + m) * length] = data


class ClockSampler:
    """SM clock and throttle reasons during the timed region (pynvml, 100 ms period)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake_slowdown",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_reference(db, alphabet, read_len, seconds_target, steps=1, warmup=0, load_filter=None):
    """Times the reference's CPU placer (oracle/_ref if built, else the port) on a bounded sample.
    Returns (reads/s of the timed steps, kind, cores, sample description)."""
    from epik_b200 import synth
    from oracle import oracle as O
    kind = "reference" if O.available("reference", alphabet) else "port"
    cores = host_threads()
    gen = db if isinstance(db, synth.Gen2) else None
    if gen is not None:
        # a device-generated database: the reference is given the lists of the keys its sample of reads looks up and
        # nothing else (same look-ups, same hits, same postings; a smaller hash map than the real one - which flatters
        # the CPU a little: fewer cache misses per look-up)
        n = 4000  # (130 M postings of lists for these reads: what numpy restates in a few seconds)
        data, off = synth.make_reads(alphabet, n, read_len, first=0)
        sub = gen.lists(synth.keys_of_reads(alphabet, gen.k, data, off))
        o = O.Oracle(sub, 7, 0.01, cores, kind=kind)
        times = []
        for i in range(warmup + steps + 1):
            r = o.place_packed(data, off, threads=cores, batch_size=2000)
            if i >= warmup + 1:
                times.append(r["seconds"])
        value = n * len(times) / sum(times)
        sample = "%d reads of the workload (the first %d) against the %d lists they look up, batches of 2000 (main.cpp:214), " \
                 "timer around placer::place only (main.cpp:343-345), %d OpenMP threads" % (n, n, sub.n_keys, cores)
        return value, ("reference" if kind == "reference" else "port"), cores, sample, sum(times) / len(times)
    if load_filter is not None:
        # what i2l::load would hand the placer (main.cpp:277): the filter applied before the timed region
        from oracle import load_filter as LF
        mu, omega, max_entries = load_filter
        probe = synth.make_db(alphabet, db.kmer_size, 8, 0.001, 2.0, omega=omega, key_space_limit=4096)
        _, log_thr = O.Oracle(probe, kind=kind).thresholds()
        db = LF.apply(db, mu, omega, max_entries, float(log_thr))
    o = O.Oracle(db, 7, 0.01, cores, kind=kind)
    probe_n = 4000 * max(1, cores // 4)
    data, off = synth.make_reads(alphabet, probe_n, read_len, first=900_000_000)
    o.place_packed(data, off, threads=cores)  # warm the caches / thread pool
    r = o.place_packed(data, off, threads=cores)
    rate = probe_n / max(r["seconds"], 1e-9)
    n = int(max(probe_n, min(5_000_000, rate * seconds_target)))
    n -= n % 2000
    data, off = synth.make_reads(alphabet, n, read_len, first=0)
    times = []
    for i in range(warmup + steps):
        r = o.place_packed(data, off, threads=cores, batch_size=2000)
        if i >= warmup:
            times.append(r["seconds"])
    value = n * len(times) / sum(times)
    sample = "%d reads of the workload (the first %d), batches of 2000 (main.cpp:214), timer around placer::place only " \
             "(main.cpp:343-345), %d OpenMP threads" % (n, n, cores)
    return value, ("reference" if kind == "reference" else "port"), cores, sample, sum(times) / len(times)


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from epik_b200 import synth
    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[args.workload]
    n_reads = args.reads or WORKLOAD_READS[args.workload]
    db = synth.make_config_gen2(args.workload) if args.workload in DEVICE_GENERATED else synth.make_config_db(args.workload)
    total = max(args.steps + args.warmup, 1)
    # (EPK_BENCH_REF_SECONDS: the CPU tests of this arm run a one-second sample)
    seconds = float(os.environ.get("EPK_BENCH_REF_SECONDS", min(20.0, 150.0 / total)))
    value, kind, cores, sample, step_s = cpu_reference(db, alphabet, read_len, seconds_target=seconds,
                                                       steps=args.steps, warmup=args.warmup,
                                                       load_filter=synth.LOAD_FILTERS.get(args.workload))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, n_reads, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "kmer_lookups_per_sec": value * (read_len - k + 1),
        "note": "CPU arm: %s; each step is a bounded sample of the workload, not the full read set" % (
            "reference place.cpp compiled unmodified over the i2l shim (oracle/_ref)" if kind == "reference"
            else "oracle port (oracle/_ref not built)"),
    }
    emit(line)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from epik_b200 import synth
    from epik_b200.placer import Database, PinnedArray, Placer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (epik_b200 has no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[args.workload]
    n = args.reads or WORKLOAD_READS[args.workload]
    K = 7
    log("world %d, device %d: generating database %s" % (world, local, args.workload))
    flt = synth.LOAD_FILTERS.get(args.workload)
    if args.workload in DEVICE_GENERATED:
        flat = synth.make_config_gen2(args.workload)
        t_build = time.perf_counter()
        db = Database.synthetic(flat, device=local)
        log("generated and built on the device in %.1f s" % (time.perf_counter() - t_build))
    else:
        flat = synth.make_config_db(args.workload)
        db = Database(flat, device=local) if flt is None else Database(flat, mu=flt[0], omega=flt[1], max_entries=flt[2], device=local)
    placer = Placer(db, K, 0.01, variant=args.variant)

    log("database on the device: %d entries; generating %d reads" % (db.info["n_entries_loaded"], n))
    # ---- inputs: pinned host copies (for e2e) and device-resident copies (for value) ----
    h_seq = PinnedArray((n * read_len,), np.uint8)
    h_off = PinnedArray((n + 1,), np.uint64)
    make_reads_chunked(alphabet, n, read_len, rank * n, h_seq.array)
    h_off.array[:] = np.arange(n + 1, dtype=np.uint64) * np.uint64(read_len)
    d_seq = torch.from_numpy(h_seq.array).to(dev)
    d_off = torch.from_numpy(h_off.array.view(np.int64)).to(dev)
    d_out = {"n_place": torch.zeros(n, dtype=torch.int32, device=dev),
             "branch": torch.zeros(n * K, dtype=torch.int32, device=dev),
             "score": torch.zeros(n * K, dtype=torch.float32, device=dev),
             "lwr": torch.zeros(n * K, dtype=torch.float64, device=dev),
             "count": torch.zeros(n * K, dtype=torch.int32, device=dev)}
    stream = torch.cuda.Stream(device=dev)

    log("reads ready; timing the device-resident path")
    # ---- value: everything resident in HBM ----
    for _ in range(args.warmup):
        placer.place_device(d_seq, d_off, n, d_out, stream=stream)
    stream.synchronize()
    counters = placer.counters()
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    with torch.cuda.stream(stream):
        ev[0].record(stream)
        for i in range(args.steps):
            placer.place_device(d_seq, d_off, n, d_out, stream=stream)
            ev[i + 1].record(stream)
    stream.synchronize()
    barrier()
    clocks = sampler.stop()
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    total_s = max_over_ranks(sum(step_ms) / 1e3)
    value = world * n * args.steps / total_s
    launches = int(counters["n_launches"]) * args.steps

    # a checksum of the results so that the work cannot be optimised away, and a sanity check
    n_place_sum = int(d_out["n_place"].sum().item())
    assert n_place_sum > 0

    # ---- roofline of the placement kernel (the only kernel that moves data) ----
    bytes_per_launch = n * (read_len + 16 * (read_len - k + 1) + 24 * K) + 8 * int(counters["n_entries"])
    kernel_s = (sum(step_ms) / len(step_ms)) / 1e3
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    # DRAM bytes of the launch from the committed ncu capture (dram__bytes_read.sum + dram__bytes_write.sum), taken
    # on a shorter launch of the same workload and scaled by the number of reads (the traffic is per read: the
    # database is a few hundred MB against tens of GB moved)
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        t = json.load(open(tpath)).get(args.workload)
        if t:
            traffic = t["dram_bytes_per_read"] * n
            traffic_src = "%s; scaled from %d to %d reads" % (t["source"], t["captured_reads"], n)
    achieved = bytes_per_launch / kernel_s / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "kernel": "epk::place_kernel", "bytes_per_launch": bytes_per_launch,
                "launch_ms": kernel_s * 1e3, "peak_source": peak_src,
                "entries_per_read": counters["n_entries"] / n, "hits_per_read": counters["n_hits"] / n}

    # ---- the random-gather ceiling: the same launch with loads only (epk_gather_ceiling_device) ----
    d_sum = torch.zeros(n, dtype=torch.int32, device=dev)
    for _ in range(2):
        placer.gather_ceiling_device(d_seq, d_off, n, d_sum, stream=stream)
    stream.synchronize()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(2, min(args.steps, 5))
    with torch.cuda.stream(stream):
        g0.record(stream)
        for _ in range(reps):
            placer.gather_ceiling_device(d_seq, d_off, n, d_sum, stream=stream)
        g1.record(stream)
    stream.synchronize()
    gather_s = g0.elapsed_time(g1) / reps / 1e3
    gather_gbs = bytes_per_launch / gather_s / 1e9
    ceiling_entries = int(placer.counters()["n_entries"])
    if ceiling_entries == int(counters["n_entries"]):
        roofline["gather_ceiling"] = {
            "achieved": gather_gbs, "unit": "GB/s", "launch_ms": gather_s * 1e3, "frac_of_peak": gather_gbs / peak,
            "what": "same kernel, same reads: encode + probe + posting gathers, accumulator updates replaced by a register "
                    "XOR, no sweep (epk_gather_ceiling_device); same algorithmic bytes. A ceiling of this design (8 warps "
                    "per SM), not of the memory system: see gather_ceiling_independent"}
        roofline["frac_of_gather_ceiling"] = achieved / gather_gbs
    else:
        # the load-only launch exists for the queries with up to 255 windows only: it did not walk these reads
        roofline["gather_ceiling"] = None
        roofline["frac_of_gather_ceiling"] = None
    roofline["frac_of_8TBs_nominal"] = achieved / 8000.0
    del d_sum
    # ---- the independent ceiling: random k-mers against the same key table, lists gathered at full occupancy ----
    if rank == 0:
        from epik_b200 import capi
        import ctypes
        mb = capi.Microbench()
        probes = int(min(60_000_000, max(2_000_000, 4e9 / max(8.0 * counters["n_entries"] / max(counters["n_windows"], 1) + 16.0, 16.0))))
        capi.check(capi.lib().epk_gather_microbench(db.handle, probes, 0xE91C0003, 1, 3, ctypes.byref(mb)))
        ind = mb.bytes / mb.ms_per_launch / 1e6
        roofline["gather_ceiling_independent"] = {
            "achieved": ind, "unit": "GB/s", "launch_ms": mb.ms_per_launch, "frac_of_peak": ind / peak,
            "probes": int(mb.n_probes), "entries": int(mb.n_entries),
            "what": "epk_gather_microbench: uniformly random k-mers probed in the same key table, every hit's posting list "
                    "gathered with 128-bit loads, 8 lists in flight per warp, 32 warps per SM, no accumulators, no sweep; "
                    "bytes = 16 per probe + 8 per posting"}
        roofline["frac_of_independent_ceiling"] = achieved / ind

    log("device-resident: %.1f ms/step; timing the host path" % (total_s / args.steps * 1e3))
    # ---- e2e: pinned host buffers through epk_place, copies inside the timed region ----
    e2e = None
    if not args.no_e2e:
        # what placer::place does between taking a batch and handing back its placed_collection (place.cpp:201-275):
        # the grouping of identical sequences per --batch-size batch (place.cpp:207) - here on the host threads,
        # concurrently with the device pass, as the C++ facade does - the placement itself, and the distal / pendant
        # lengths of the kept placements (place.cpp:424-439, filled by epk_place after the copy out)
        import ctypes
        from epik_b200 import capi
        out = placer.alloc_result(n, pinned=True, lengths=True)
        group_rep = np.zeros(n, np.uint32)
        # (one host thread of the rank's share is left to the thread that issues the chunks of the pass: on short steps -
        # 100 k long reads take 13 ms - a launcher that waits for a core is what the step then measures)
        threads = max(1, host_threads() // max(1, world) - 1)

        def group():
            capi.check(capi.lib().epk_group_batches(h_seq.array.ctypes.data_as(ctypes.c_void_p), h_off.array.ctypes.data_as(ctypes.c_void_p),
                                                    n, 2000, threads, group_rep.ctypes.data_as(ctypes.c_void_p)))

        def one_pass():
            worker = threading.Thread(target=group)
            worker.start()
            placer.place_packed(h_seq.array, h_off.array, out=out)
            worker.join()

        e2e_warm = args.warmup
        for _ in range(e2e_warm):
            one_pass()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            one_pass()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        e2e_s = max_over_ranks(t1 - t0)
        barrier()
        assert int(out["n_place"].sum()) == n_place_sum, "host and device paths disagree"
        e2e = {"value": world * n * args.steps / e2e_s, "unit": UNIT,
               "h2d_bytes_per_step": int(n * read_len + (n + 1) * 8),
               "d2h_bytes_per_step": int(n * (4 + K * 20)),
               "ms_per_step": e2e_s / args.steps * 1e3,
               "api": "epk_place (C ABI), pinned host buffers, %d-read chunks on 3 streams, distal / pendant filled; "
                      "epk_group_batches (batches of 2000, place.cpp:207) on %d host threads alongside" % (1 << 18, threads)}

    db_info = {k2: int(v2) if isinstance(v2, (int, np.integer)) else float(v2) for k2, v2 in db.info.items()}
    # ---- CPU baseline: rank 0, N = 1 only ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, kind, cores, sample, _ = cpu_reference(flat, alphabet, read_len, seconds_target=15.0, load_filter=flt)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}

    # ---- N > 1: the one path with an exchange - BASELINE config 5 with the database SHARDED over the ranks ----
    dbshard_leg = None
    if world > 1 and not args.no_dbshard:
        del placer, db, d_seq, d_off, d_out
        h_seq.free() if hasattr(h_seq, "free") else None
        torch.cuda.empty_cache()
        dbshard_leg = time_dbshard(args, "C5", rank, world, local, dev, reads=args.dbshard_reads, steps=max(2, min(args.steps, 3)),
                                   warmup=1, transport=args.transport)

    # ---- N > 1, second form of "the database does not fit one GPU": pooled over the N HBMs, read over NVLink; every
    # rank holds an N-th of it and places its own reads
    dbpool_leg = None
    if world > 1 and not args.no_dbpool:
        torch.cuda.empty_cache()
        barrier()
        try:
            dbpool_leg = time_dbpool_ranks(args, "C5", rank, world, local, dev, args.dbpool_reads, steps=max(2, min(args.steps, 3)), warmup=1)
        except RuntimeError as exc:  # (raised on every rank together: RankPooledPlacer agrees on failures)
            dbpool_leg = {"error": str(exc)}
        barrier()

    total_reads = sum_over_ranks(float(n))
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_s / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args.workload, n, world),
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "kmer_lookups_per_sec": value * (read_len - k + 1), "reads_per_step_all_gpus": int(total_reads),
            "dtype_note": "scores accumulate in f32 (bit-identical to the reference), likelihood weight ratios in f64",
            "db": db_info,
        }
        if dbshard_leg is not None:
            line["dbshard"] = dbshard_leg
        if dbpool_leg is not None:
            line["dbpool"] = dbpool_leg
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def time_dbshard(args, workload, rank, world, local, dev, reads, steps, warmup, transport):
    """One measurement of database-sharded mode inside an initialised process group: rank r holds shard r of
    `workload`'s keys, every rank sees the same `reads` queries per step and finalises its slice of them. Returns the
    record of the `dbshard` key (rank 0) - reads of a step / max-over-ranks time."""
    import torch
    import torch.distributed as dist
    from epik_b200 import dbshard, synth
    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[workload]
    K = 7
    flat = synth.make_config_gen2(workload) if workload in DEVICE_GENERATED else synth.make_config_db(workload)
    rows = args.dbshard_rows
    sp = dbshard.ShardedPlacer(flat, rank, world, device=local, keep_at_most=K, transport=transport, max_rows=rows)
    h_seq = np.zeros(reads * read_len, np.uint8)
    make_reads_chunked(alphabet, reads, read_len, 0, h_seq)  # the SAME reads on every rank
    d_seq = torch.from_numpy(h_seq).to(dev)
    d_off = torch.from_numpy((np.arange(reads + 1, dtype=np.uint64) * np.uint64(read_len)).view(np.int64)).to(dev)

    def step():
        out = None
        for first in range(0, reads, rows):
            mrows = min(rows, reads - first)
            out, _, _ = sp.place_device(d_seq, d_off[first:first + mrows + 1], mrows)
        return out

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / 1e3], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    secs = float(t.item())
    placed = torch.tensor([float(out["n_place"].sum().item())], dtype=torch.float64, device=dev)
    dist.all_reduce(placed, op=dist.ReduceOp.SUM)
    row_bytes = sp.row_bytes
    info = sp.db.info
    sp.close()
    return {"metric": METRIC, "value": reads * steps / secs, "unit": UNIT, "n_gpus": world, "scaling": "strong",
            "workload": "%s: k=%d, %d branches, database sharded by key hash over %d GPUs (%.1f GB of postings per shard), "
                        "%d x %d-letter reads per step, every rank sees every read" % (
                            workload, k, 2 * n_leaves - 1, world, info["n_entries_loaded"] * 8 / 1e9, reads, read_len),
            "transport": transport, "ms_per_step": secs / steps * 1e3, "steps": steps, "queries_per_exchange": rows,
            "row_bytes": int(row_bytes), "exchange_bytes_per_query": row_bytes * (world - 1) / world,
            "ambiguous_reads": "replayed by the owner over the peers' tables" if sp.replays_ambiguous else "unsupported",
            "placed_last_pass": int(placed.item())}


def time_dbpool(args, workload, devices, reads_per_gpu, steps, warmup):
    """The database of `workload` POOLED over the HBM of `devices` (epk_db_pool: one process drives the GPUs, every GPU
    places its own reads with the fused kernel and gathers each posting list from whichever HBM holds it, over NVLink).
    Returns the record of the `dbpool` key: device-resident rate (one launch per device on its own stream, device
    events, the slowest device counts) and the end-to-end rate through epk_place_many from pinned host buffers."""
    import torch
    from epik_b200 import dbshard, synth
    from epik_b200.placer import PinnedArray, Placer
    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[workload]
    K, G, n = 7, len(devices), reads_per_gpu
    flat = synth.make_config_gen2(workload) if workload in DEVICE_GENERATED else synth.make_config_db(workload)
    t0 = time.perf_counter()
    pooled = dbshard.PooledPlacer(flat, devices, keep_at_most=K, variant=args.variant)
    build_s = time.perf_counter() - t0
    info = pooled.dbs[0].info
    h_seq = PinnedArray((G * n * read_len,), np.uint8)
    h_off = PinnedArray((G * n + 1,), np.uint64)
    make_reads_chunked(alphabet, G * n, read_len, 0, h_seq.array)
    h_off.array[:] = np.arange(G * n + 1, dtype=np.uint64) * np.uint64(read_len)
    local_off = torch.from_numpy((np.arange(n + 1, dtype=np.uint64) * np.uint64(read_len)).view(np.int64))
    state = []
    for g, dev_id in enumerate(devices):
        dev = torch.device("cuda", dev_id)
        d_seq = torch.from_numpy(h_seq.array[g * n * read_len:(g + 1) * n * read_len]).to(dev)
        d_off = local_off.to(dev)
        out = {"n_place": torch.zeros(n, dtype=torch.int32, device=dev), "branch": torch.zeros(n * K, dtype=torch.int32, device=dev),
               "score": torch.zeros(n * K, dtype=torch.float32, device=dev), "lwr": torch.zeros(n * K, dtype=torch.float64, device=dev),
               "count": torch.zeros(n * K, dtype=torch.int32, device=dev)}
        state.append((dev, d_seq, d_off, out, torch.cuda.Stream(device=dev)))

    def launch_all():
        for g, (dev, d_seq, d_off, out, stream) in enumerate(state):
            pooled.placers[g].place_device(d_seq, d_off, n, out, stream=stream)

    def sync_all():
        for dev, _, _, _, stream in state:
            stream.synchronize()

    for _ in range(warmup):
        launch_all()
    sync_all()
    events = []
    for dev, _, _, _, stream in state:
        with torch.cuda.device(dev):
            events.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
    for (e0, _), (dev, _, _, _, stream) in zip(events, state):
        e0.record(stream)
    for _ in range(steps):
        launch_all()
    for (_, e1), (dev, _, _, _, stream) in zip(events, state):
        e1.record(stream)
    sync_all()
    secs = max(e0.elapsed_time(e1) for e0, e1 in events) / 1e3
    placed = sum(int(out["n_place"].sum().item()) for _, _, _, out, _ in state)
    entries = sum(p.counters()["n_entries"] for p in pooled.placers)
    # end to end: one pass of G * n reads from pinned host memory through epk_place_many
    out = pooled.placers[0].alloc_result(G * n, pinned=True, lengths=True)
    pooled.place_packed(h_seq.array, h_off.array, out=out)
    t0 = time.perf_counter()
    for _ in range(steps):
        pooled.place_packed(h_seq.array, h_off.array, out=out)
    e2e_s = time.perf_counter() - t0
    assert int(out["n_place"].sum()) == placed
    rec = {"metric": METRIC, "value": G * n * steps / secs, "unit": UNIT, "n_gpus": G, "scaling": "weak",
           "e2e": {"value": G * n * steps / e2e_s, "unit": UNIT, "api": "epk_place_many (C ABI): one process, %d placers, pinned host buffers" % G},
           "workload": "%s: k=%d, %d branches, %.1f GB of postings pooled over %d GPUs (%.1f GB in each HBM), %d x %d-letter reads per GPU "
                       "per step, each GPU places its own reads" % (workload, k, 2 * n_leaves - 1, info["n_entries_loaded"] * 8 / 1e9, G,
                                                                    info["n_entries_loaded"] * 8 / 1e9 / G, n, read_len),
           "ms_per_step": secs / steps * 1e3, "steps": steps, "build_s": build_s,
           "nvlink_bytes_per_read": 8.0 * entries / (G * n) * (G - 1) / G,
           "how": "epk_db_pool: shards mapped into one address range (CUDA VMM), a key table per GPU pointing into it, the fused "
                  "kernel unchanged; results bit-identical to the unsharded database"}
    pooled.close()
    return rec


def time_dbpool_ranks(args, workload, rank, world, local, dev, reads_per_gpu, steps, warmup):
    """The pooled database with one PROCESS per GPU (inside an initialised process group): rank r holds shard r of
    `workload`'s keys in its HBM, maps everybody else's, and places ITS OWN reads with the fused kernel."""
    import torch
    import torch.distributed as dist
    from epik_b200 import dbshard, synth
    from epik_b200.placer import PinnedArray
    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[workload]
    K, n = 7, reads_per_gpu
    flat = synth.make_config_gen2(workload) if workload in DEVICE_GENERATED else synth.make_config_db(workload)
    t0 = time.perf_counter()
    rp = dbshard.RankPooledPlacer(flat, rank, world, device=local, keep_at_most=K)
    build_s = time.perf_counter() - t0
    info = rp.db.info
    h_seq = PinnedArray((n * read_len,), np.uint8)
    h_off = PinnedArray((n + 1,), np.uint64)
    make_reads_chunked(alphabet, n, read_len, rank * n, h_seq.array)
    h_off.array[:] = np.arange(n + 1, dtype=np.uint64) * np.uint64(read_len)
    d_seq = torch.from_numpy(h_seq.array).to(dev)
    d_off = torch.from_numpy(h_off.array.view(np.int64)).to(dev)
    out = {"n_place": torch.zeros(n, dtype=torch.int32, device=dev), "branch": torch.zeros(n * K, dtype=torch.int32, device=dev),
           "score": torch.zeros(n * K, dtype=torch.float32, device=dev), "lwr": torch.zeros(n * K, dtype=torch.float64, device=dev),
           "count": torch.zeros(n * K, dtype=torch.int32, device=dev)}
    stream = torch.cuda.Stream(device=dev)
    for _ in range(warmup):
        rp.placer.place_device(d_seq, d_off, n, out, stream=stream)
    stream.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        rp.placer.place_device(d_seq, d_off, n, out, stream=stream)
    e1.record(stream)
    stream.synchronize()
    entries = rp.placer.counters()["n_entries"]
    dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / 1e3], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    secs = float(t.item())
    # end to end: pinned host buffers through epk_place, every rank at once
    host_out = rp.placer.alloc_result(n, pinned=True, lengths=True)
    rp.placer.place_packed(h_seq.array, h_off.array, out=host_out)
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        rp.placer.place_packed(h_seq.array, h_off.array, out=host_out)
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    assert int(host_out["n_place"].sum()) == int(out["n_place"].sum().item()) > 0
    rec = {"metric": METRIC, "value": world * n * steps / secs, "unit": UNIT, "n_gpus": world, "scaling": "weak",
           "e2e": {"value": world * n * steps / e2e_s, "unit": UNIT, "api": "epk_place (C ABI) on every rank, pinned host buffers"},
           "workload": "%s: k=%d, %d branches, %.1f GB of postings pooled over %d GPUs, one process each (%.1f GB in each HBM), %d x "
                       "%d-letter reads per GPU per step, each rank places its own reads" % (
                           workload, k, 2 * n_leaves - 1, info["n_entries_loaded"] * 8 / 1e9, world,
                           info["n_entries_loaded"] * 8 / 1e9 / world, n, read_len),
           "ms_per_step": secs / steps * 1e3, "steps": steps, "build_s": build_s,
           "nvlink_bytes_per_read": 8.0 * entries / n * (world - 1) / world,
           "how": "epk_db_pool_share / epk_db_pool_join: every rank's shard exported as a file descriptor, mapped by all into one "
                  "address range (CUDA VMM), a key table per rank pointing into it, the fused kernel unchanged; posting lists are "
                  "gathered over NVLink from whichever HBM holds them; results bit-identical to the unsharded database"}
    rp.close()
    return rec


def run_dbpool(args):
    """--mode dbpool: ONE process (no torchrun), --gpus N devices."""
    import torch
    n_dev = min(args.gpus, torch.cuda.device_count())
    rec = time_dbpool(args, args.workload if args.workload != "C2" or args.reads else "C5", list(range(n_dev)),
                      args.reads or 1_000_000, args.steps, args.warmup)
    rec.update({"mode": "dbpool", "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "vs_baseline": None,
                "dtype": "f32", "data": "synthetic"})
    emit(rec)


def run_dbshard(args):
    """Database-sharded variant: rank r holds shard r of the keys, EVERY rank sees the whole read set of a step;
    value = reads of the step / time (strong scaling of the database, not of the queries)."""
    import torch
    import torch.distributed as dist
    from epik_b200 import dbshard, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    alphabet, k, n_leaves, p_occ, m, read_len = synth.CONFIGS[args.workload]
    n = args.reads or WORKLOAD_READS[args.workload]
    K = 7
    log("dbshard: world %d, generating database %s" % (world, args.workload))
    flat = synth.make_config_gen2(args.workload) if args.workload in DEVICE_GENERATED else synth.make_config_db(args.workload)
    if world > 1:
        sp = dbshard.ShardedPlacer(flat, rank, world, device=local, keep_at_most=K, transport=args.transport,
                                   max_rows=args.dbshard_rows)
        info = sp.db.info
    else:
        sp = dbshard.LocalShardedPlacer(flat, 1, device=local, keep_at_most=K)
        info = sp.dbs[0].info
    h_seq = np.zeros(n * read_len, np.uint8)
    make_reads_chunked(alphabet, n, read_len, 0, h_seq)  # the SAME reads on every rank
    d_seq = torch.from_numpy(h_seq).to(dev)
    d_off = torch.from_numpy((np.arange(n + 1, dtype=np.uint64) * np.uint64(read_len)).view(np.int64)).to(dev)
    rows = args.dbshard_rows
    row_bytes = (sp.placer if world > 1 else sp.placers[0]).partial_row_bytes()

    def step():
        got = 0
        for first in range(0, n, rows):
            mrows = min(rows, n - first)
            off = d_off[first:first + mrows + 1]
            if world > 1:
                out, f, l = sp.place_device(d_seq, off, mrows)
            else:
                pl = sp.placers[0]
                if getattr(sp, "_slab", None) is None:
                    sp._slab = torch.empty((rows, row_bytes), dtype=torch.uint8, device=dev)
                    sp._out = dbshard._alloc_out(torch, rows, K, dev)
                pl.place_partial_device(d_seq, off, mrows, [0, mrows], [sp._slab.data_ptr()])
                pl.place_finalize_device(sp._slab.data_ptr(), 1, rows * row_bytes, off.data_ptr(), mrows, sp._out)
                out = sp._out
            got += 1
        return out

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    secs = e0.elapsed_time(e1) / 1e3
    if world > 1:
        t = torch.tensor([secs], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = float(t.item())
    assert int(out["n_place"].sum().item()) > 0
    value = n * args.steps / secs
    if rank == 0:
        cfg = workload_config(args.workload, n, world)
        how = "stored by the partial launch straight into the owner's HBM (CUDA IPC peer memory over NVLink), one " \
              "1-element all_reduce per step as the barrier" if (world > 1 and args.transport == "p2p") else \
              "exchanged by all_to_all_single (NCCL)" if world > 1 else "left in local HBM (one shard)"
        cfg["parallelism"] = "database sharded by key hash x%d, every rank sees every query; partial rows (%d B per query) " \
                             "%s, %d queries per step" % (world, row_bytes, how, rows)
        cfg["reads_per_gpu"] = n
        emit({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
              "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
              "dtype": "f32", "data": "synthetic", "mode": "dbshard", "transport": args.transport if world > 1 else "local",
              "config": cfg, "clocks": clocks,
              "exchange_bytes_per_query": row_bytes * (world - 1) / max(world, 1),
              "db_shard": {k2: int(v2) if isinstance(v2, (int, np.integer)) else float(v2) for k2, v2 in info.items()}})
    if world > 1:
        sp.close()
        dist.destroy_process_group()


def main():
    # stdout carries ONE JSON line: anything libraries print there (NCCL announces its version on stdout) is
    # sent to stderr instead, and the line is written to the real stdout at the end
    global REAL_STDOUT
    sys.stdout.flush()
    REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.mode == "dbshard":
        run_dbshard(args)
    elif args.mode == "dbpool":
        run_dbpool(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
