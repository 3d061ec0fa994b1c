"""Host-side mirror of the reference's placer interface, on top of the C ABI.

    reference (C++)                                            here
    i2l::load(file, mu, omega, max_entries)  main.cpp:277      Database(flat_db, mu=..., omega=..., max_entries=...)
    epik::placer(db, tree, keep_at_most, keep_factor, threads) Placer(db, keep_at_most, keep_factor)
    placer.place(seq_records, threads) -> placed_collection    Placer.place(records) -> PlacedCollection
      place.h:94-103, place.cpp:201-275                        Placer.place_packed(bytes, offsets) -> arrays

All computing happens in libepik_b200.so (hand-written sm_100a kernels); this module only marshals
numpy / torch buffers across the ABI. Nothing here falls back to the CPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import capi
from .flatdb import ENTRY_DTYPE, FlatDB


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pack_sequences(seqs):
    """list of bytes/str -> (uint8 array, uint64 offsets[n+1])."""
    bs = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]
    off = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    data = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, np.uint8)
    return data, off


class PinnedArray:
    """numpy views over pinned host memory from epk_host_alloc (fast path of epk_place)."""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(np.atleast_1d(shape))
        n = int(np.prod(self.shape)) * self.dtype.itemsize
        self._p = capi.lib().epk_host_alloc(max(n, 1))
        if not self._p:
            raise MemoryError("epk_host_alloc(%d)" % n)
        buf = (C.c_char * max(n, 1)).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if getattr(self, "_p", None):
            self.array = None
            capi.lib().epk_host_free(self._p)
            self._p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Database:
    """HBM-resident phylo-k-mer database (epk_db). `mu`, `omega`, `max_entries` are the load-filter
    arguments of i2l::load (main.cpp:277); leave all three None to load everything. `shard` = (index, count)
    keeps only the keys of one shard (database-sharded mode, epk_db_create_shard)."""

    def __init__(self, db: FlatDB, mu: float | None = None, omega: float | None = None,
                 max_entries: int | None = None, device: int = 0, shard: tuple[int, int] | None = None):
        l = capi.lib()
        db.validate()
        self.flat = db
        keys = np.ascontiguousarray(db.keys, np.uint32)
        offsets = np.ascontiguousarray(db.offsets, np.uint64)
        entries = np.ascontiguousarray(db.entries, ENTRY_DTYPE)
        fv = None if db.filter_value is None else np.ascontiguousarray(db.filter_value, np.float32)
        bl = np.ascontiguousarray(db.branch_length, np.float64)
        sn = np.ascontiguousarray(db.subtree_num_nodes, np.uint64)
        st = np.ascontiguousarray(db.subtree_total_length, np.float64)
        desc = capi.DbDesc(int(db.alphabet), int(db.kmer_size), float(db.omega), len(keys), _ptr(keys), _ptr(offsets),
                           _ptr(entries), _ptr(fv), len(bl), _ptr(bl), _ptr(sn), _ptr(st))
        flt = None
        if mu is not None or omega is not None or max_entries is not None:
            flt = capi.LoadFilter(1.0 if mu is None else float(mu), float(db.omega if omega is None else omega),
                                  capi.UINT64_MAX if max_entries is None else int(max_entries))
        handle = C.c_void_p()
        fref = C.byref(flt) if flt is not None else None
        if shard is None:
            capi.check(l.epk_db_create(C.byref(desc), fref, int(device), C.byref(handle)))
        else:
            capi.check(l.epk_db_create_shard(C.byref(desc), fref, int(device), int(shard[0]), int(shard[1]), C.byref(handle)))
        self.shard = shard
        self.handle = handle
        self.device = int(device)
        info = capi.DbInfo()
        capi.check(l.epk_db_get_info(self.handle, C.byref(info)))
        self.info = {name: getattr(info, name) for name, _ in capi.DbInfo._fields_}

    def kmers(self, seq):
        """(pos, key, group, hit_len) of one sequence, computed on the device (epk_kmers)."""
        b = seq.encode() if isinstance(seq, str) else bytes(seq)
        cap = max(1, len(b) * 20)
        pos, key, group = (np.zeros(cap, np.uint32) for _ in range(3))
        hit = np.zeros(cap, np.int64)
        n = C.c_uint64(0)
        capi.check(capi.lib().epk_kmers(self.handle, b, len(b), cap, _ptr(pos), _ptr(key), _ptr(group), _ptr(hit), C.byref(n)))
        assert n.value <= cap
        return pos[:n.value], key[:n.value], group[:n.value], hit[:n.value]

    def close(self):
        if getattr(self, "handle", None):
            capi.lib().epk_db_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


@dataclass
class Placement:
    """epik::placement (place.h:45-56)."""
    branch_id: int
    score: float
    weight_ratio: float
    count: int
    distal_length: float
    pendant_length: float


@dataclass
class PlacedSequence:
    """epik::placed_sequence (place.h:59-68)."""
    sequence: bytes
    placements: list


@dataclass
class PlacedCollection:
    """epik::placed_collection (place.h:72-75): sequence -> headers of the records that carry it,
    and one PlacedSequence per unique sequence."""
    sequence_map: dict = field(default_factory=dict)
    placed_seqs: list = field(default_factory=list)


def _db_synthetic(cls, gen, mu: float | None = None, omega: float | None = None, max_entries: int | None = None,
                  device: int = 0, shard: tuple[int, int] | None = None):
    """epk_db_create_synthetic: the gen2 database of synth.Gen2 `gen`, generated on the device (never held raw)."""
    self = cls.__new__(cls)
    self.flat = None
    self.gen = gen
    bl = np.ascontiguousarray(gen.length, np.float64)
    sn = np.ascontiguousarray(gen.num, np.uint64)
    st = np.ascontiguousarray(gen.tot, np.float64)
    table = np.ascontiguousarray(gen.length_table, np.uint32)
    bits = int(len(table)).bit_length() - 1
    desc = capi.SynthDesc(int(gen.alphabet), int(gen.k), float(gen.omega), len(bl), _ptr(bl), _ptr(sn), _ptr(st),
                          int(gen.seed), int(gen.presence_threshold), _ptr(table), bits, float(gen.score_lo),
                          int(gen.key_space_limit))
    flt = None
    if mu is not None or omega is not None or max_entries is not None:
        flt = capi.LoadFilter(1.0 if mu is None else float(mu), float(gen.omega if omega is None else omega),
                              capi.UINT64_MAX if max_entries is None else int(max_entries))
    handle = C.c_void_p()
    idx, cnt = shard if shard is not None else (0, 1)
    capi.check(capi.lib().epk_db_create_synthetic(C.byref(desc), C.byref(flt) if flt is not None else None, int(device),
                                                  int(idx), int(cnt), C.byref(handle)))
    self.handle = handle
    info = capi.DbInfo()
    capi.check(capi.lib().epk_db_get_info(self.handle, C.byref(info)))
    self.info = {name: getattr(info, name) for name, _ in capi.DbInfo._fields_}
    self.device = int(device)
    self.shard = shard
    return self


Database.synthetic = classmethod(_db_synthetic)


def _db_pool(cls, shards):
    """epk_db_pool: shards[g] = shard g of len(shards) (Database(..., shard=(g, G)) or Database.synthetic, one per GPU of
    this process) -> one Database per shard's device over the pooled postings, each a WHOLE database to a placer.
    The shards are copied: close them afterwards."""
    handles = (C.c_void_p * len(shards))(*[s.handle for s in shards])
    out = (C.c_void_p * len(shards))()
    capi.check(capi.lib().epk_db_pool(C.cast(handles, C.c_void_p), len(shards), C.cast(out, C.c_void_p)))
    made = []
    for g, s in enumerate(shards):
        db = cls.__new__(cls)
        db.flat = getattr(s, "flat", None)
        db.gen = getattr(s, "gen", None)
        db.handle = C.c_void_p(out[g])
        db.device = s.device
        db.shard = None
        info = capi.DbInfo()
        capi.check(capi.lib().epk_db_get_info(db.handle, C.byref(info)))
        db.info = {name: getattr(info, name) for name, _ in capi.DbInfo._fields_}
        made.append(db)
    return made


Database.pool = classmethod(_db_pool)


def _db_pool_share(self):
    """epk_db_pool_share -> (fd, bytes, table_handle): what a peer PROCESS needs to map this shard."""
    fd, nbytes = C.c_int(-1), C.c_uint64(0)
    table = C.create_string_buffer(64)
    capi.check(capi.lib().epk_db_pool_share(self.handle, C.byref(fd), C.byref(nbytes), table))
    return fd.value, int(nbytes.value), table.raw


def _db_score_range(self):
    lo, hi, ok = C.c_float(), C.c_float(), C.c_uint32()
    entries, keys = C.c_uint64(), C.c_uint64()
    capi.check(capi.lib().epk_db_get_score_range(self.handle, C.byref(lo), C.byref(hi), C.byref(ok), C.byref(entries), C.byref(keys)))
    return lo.value, hi.value, int(ok.value), int(entries.value), int(keys.value)


def _db_pool_join(self, n_shards: int, fds, sizes, table_handles, score_range):
    """epk_db_pool_join + epk_db_pool_set_score_range: the pooled database of this process (a whole database to a placer)."""
    fd_arr = (C.c_int * n_shards)(*[int(f) for f in fds])
    size_arr = (C.c_uint64 * n_shards)(*[int(b) for b in sizes])
    out = C.c_void_p()
    capi.check(capi.lib().epk_db_pool_join(self.handle, n_shards, C.cast(fd_arr, C.c_void_p), C.cast(size_arr, C.c_void_p),
                                           b"".join(table_handles), C.byref(out)))
    db = type(self).__new__(type(self))
    db.flat, db.gen, db.handle, db.device, db.shard = getattr(self, "flat", None), getattr(self, "gen", None), out, self.device, None
    lo, hi, ok, entries, keys = score_range
    capi.check(capi.lib().epk_db_pool_set_score_range(db.handle, lo, hi, ok, entries, keys))
    info = capi.DbInfo()
    capi.check(capi.lib().epk_db_get_info(db.handle, C.byref(info)))
    db.info = {name: getattr(info, name) for name, _ in capi.DbInfo._fields_}
    return db


Database.pool_share = _db_pool_share
Database.score_range = _db_score_range
Database.pool_join = _db_pool_join


def _db_view(self) -> "capi.DbView":
    """epk_db_get_view: the device addresses of this database's key table and postings."""
    v = capi.DbView()
    capi.check(capi.lib().epk_db_get_view(self.handle, C.byref(v)))
    return v


Database.view = _db_view


class Placer:
    """epik::placer (place.h:81-140)."""

    def __init__(self, db: Database, keep_at_most: int = 7, keep_factor: float = 0.01, variant: int = 0,
                 chunk_reads: int = 0):
        self.db = db
        self.keep_at_most = int(keep_at_most)
        self.keep_factor = float(keep_factor)
        opts = capi.PlacerOpts(int(variant), int(chunk_reads))
        handle = C.c_void_p()
        capi.check(capi.lib().epk_placer_create(db.handle, self.keep_at_most, self.keep_factor, C.byref(opts), C.byref(handle)))
        self.handle = handle

    def close(self):
        if getattr(self, "handle", None):
            capi.lib().epk_placer_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ array interface -----
    def alloc_result(self, n: int, pinned: bool = False, lengths: bool = True):
        K = self.keep_at_most
        spec = [("n_place", (n,), np.uint32), ("branch", (n, K), np.uint32), ("score", (n, K), np.float32),
                ("lwr", (n, K), np.float64), ("count", (n, K), np.uint32)]
        if lengths:
            spec += [("distal", (n, K), np.float64), ("pendant", (n, K), np.float64)]
        if pinned:
            keep = {name: PinnedArray(shape, dt) for name, shape, dt in spec}
            out = {name: pa.array for name, pa in keep.items()}
            out["_pinned"] = keep
            return out
        return {name: np.zeros(shape, dt) for name, shape, dt in spec}

    def place_packed(self, data, offsets, out=None):
        """epk_place over packed sequences in HOST memory. Returns the arrays of epk_result."""
        data = np.ascontiguousarray(data, np.uint8) if not isinstance(data, np.ndarray) or data.dtype != np.uint8 else data
        offsets = np.ascontiguousarray(offsets, np.uint64) if not isinstance(offsets, np.ndarray) or offsets.dtype != np.uint64 else offsets
        n = len(offsets) - 1
        if out is None:
            out = self.alloc_result(n)
        res = capi.Result(_ptr(out["n_place"]), _ptr(out["branch"]), _ptr(out["score"]), _ptr(out["lwr"]),
                          _ptr(out["count"]), _ptr(out.get("distal")), _ptr(out.get("pendant")))
        dptr = _ptr(data) if data.size else None
        capi.check(capi.lib().epk_place(self.handle, dptr, _ptr(offsets), n, C.byref(res)))
        return out

    @staticmethod
    def place_packed_many(placers, data, offsets, out=None):
        """epk_place_many: one pass over several GPUs from this one process (one placer per device, database replicated);
        rows come back in input order, as from place_packed on one device."""
        data = np.ascontiguousarray(data, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        n = len(offsets) - 1
        if out is None:
            out = placers[0].alloc_result(n)
        res = capi.Result(_ptr(out["n_place"]), _ptr(out["branch"]), _ptr(out["score"]), _ptr(out["lwr"]),
                          _ptr(out["count"]), _ptr(out.get("distal")), _ptr(out.get("pendant")))
        handles = (C.c_void_p * len(placers))(*[p.handle for p in placers])
        capi.check(capi.lib().epk_place_many(C.cast(handles, C.c_void_p), len(placers), _ptr(data) if data.size else None,
                                             _ptr(offsets), n, C.byref(res)))
        return out

    def place_device(self, d_seq, d_off, n: int, d_out: dict, stream=None):
        """epk_place_device: everything already in HBM. Arguments are objects with .data_ptr()
        (torch tensors); asynchronous on `stream` (a torch.cuda.Stream or None)."""
        res = capi.Result(d_out["n_place"].data_ptr(), d_out["branch"].data_ptr(), d_out["score"].data_ptr(),
                          d_out["lwr"].data_ptr(), d_out["count"].data_ptr(), None, None)
        s = None if stream is None else C.c_void_p(stream.cuda_stream)
        capi.check(capi.lib().epk_place_device(self.handle, d_seq.data_ptr(), d_off.data_ptr(), int(n), C.byref(res), s))

    def trace_packed(self, data, offsets):
        """The placement kernel's own k-mer stream evidence (epk_placer_set_trace) for a packed batch: places the
        batch with everything resident in HBM and returns (trace [n, 4] uint32, results dict of numpy arrays)."""
        import torch
        n, K = len(offsets) - 1, self.keep_at_most
        dev = torch.device("cuda", self.db.device)
        d_seq = torch.from_numpy(np.ascontiguousarray(data, np.uint8)).to(dev) if len(data) else torch.zeros(1, dtype=torch.uint8, device=dev)
        d_off = torch.from_numpy(np.ascontiguousarray(offsets, np.uint64).view(np.int64)).to(dev)
        d_trace = torch.full((max(n, 1), 4), -1, dtype=torch.int32, device=dev)
        d_out = {"n_place": torch.zeros(max(n, 1), dtype=torch.int32, device=dev),
                 "branch": torch.zeros(max(n, 1) * K, dtype=torch.int32, device=dev),
                 "score": torch.zeros(max(n, 1) * K, dtype=torch.float32, device=dev),
                 "lwr": torch.zeros(max(n, 1) * K, dtype=torch.float64, device=dev),
                 "count": torch.zeros(max(n, 1) * K, dtype=torch.int32, device=dev)}
        capi.check(capi.lib().epk_placer_set_trace(self.handle, C.c_void_p(d_trace.data_ptr())))
        try:
            self.place_device(d_seq, d_off, n, d_out)
            torch.cuda.synchronize(dev)
        finally:
            capi.check(capi.lib().epk_placer_set_trace(self.handle, None))
        res = {k: v.cpu().numpy() for k, v in d_out.items()}
        return d_trace.cpu().numpy().view(np.uint32)[:n], res

    # ------------------------------------------------------------- database-sharded mode -----
    def partial_row_bytes(self) -> int:
        return int(capi.lib().epk_partial_row_bytes(self.handle))

    def place_partial_device(self, d_seq, d_off, n: int, owner_first, row_ptrs, stream=None):
        """epk_place_partial_device: the partial row of query i goes to row_ptrs[g] + (i - owner_first[g]) * row_bytes
        for the owner g with owner_first[g] <= i < owner_first[g + 1]. row_ptrs are device addresses (ints)."""
        first = np.ascontiguousarray(owner_first, np.uint32)
        ptrs = (C.c_void_p * len(row_ptrs))(*[C.c_void_p(int(p)) for p in row_ptrs])
        s = None if stream is None else C.c_void_p(stream.cuda_stream)
        capi.check(capi.lib().epk_place_partial_device(self.handle, d_seq.data_ptr(), d_off.data_ptr(), int(n),
                                                       len(row_ptrs), _ptr(first), ptrs, s))

    def place_finalize_device(self, rows_ptr: int, n_shards: int, shard_stride: int, d_off_ptr: int, n: int, d_out: dict,
                              stream=None, d_seq_ptr: int | None = None):
        """epk_place_finalize_seq_device over n queries whose rows of shard g start at rows_ptr + g * shard_stride.
        With d_seq_ptr (the base the offsets refer to) and shard views set, ambiguous windows are replayed here."""
        res = capi.Result(d_out["n_place"].data_ptr(), d_out["branch"].data_ptr(), d_out["score"].data_ptr(),
                          d_out["lwr"].data_ptr(), d_out["count"].data_ptr(), None, None)
        s = None if stream is None else C.c_void_p(stream.cuda_stream)
        seq = None if d_seq_ptr is None else C.c_void_p(int(d_seq_ptr))
        capi.check(capi.lib().epk_place_finalize_seq_device(self.handle, C.c_void_p(int(rows_ptr)), int(n_shards),
                                                            int(shard_stride), seq, C.c_void_p(int(d_off_ptr)), int(n),
                                                            C.byref(res), s))

    def set_shard_views(self, views) -> None:
        """epk_placer_set_shard_views: views[g] = capi.DbView of shard g with addresses valid on this placer's device."""
        arr = (capi.DbView * len(views))(*views)
        capi.check(capi.lib().epk_placer_set_shard_views(self.handle, len(views), C.cast(arr, C.c_void_p) if len(views) else None))

    def gather_ceiling_device(self, d_seq, d_off, n: int, d_checksum, stream=None):
        """epk_gather_ceiling_device: the load-only form of the placement launch (roofline measuring stick)."""
        s = None if stream is None else C.c_void_p(stream.cuda_stream)
        capi.check(capi.lib().epk_gather_ceiling_device(self.handle, d_seq.data_ptr(), d_off.data_ptr(), int(n),
                                                        d_checksum.data_ptr(), s))

    def counters(self) -> dict:
        c = capi.Counters()
        capi.check(capi.lib().epk_placer_counters(self.handle, C.byref(c)))
        return {name: getattr(c, name) for name, _ in capi.Counters._fields_}

    # ------------------------------------------------------- the reference's interface -----
    def place(self, records, batch_size: int = 0) -> PlacedCollection:
        """placer::place (place.cpp:201-275) for a list of (header, sequence) records: identical
        sequences are placed once and keep the list of their headers (place.cpp:73-81,207)."""
        headers = [h for h, _ in records]
        seqs = [s.encode() if isinstance(s, str) else bytes(s) for _, s in records]
        data, off = pack_sequences(seqs)
        n = len(seqs)
        out = self.place_packed(data, off)
        rep = np.zeros(n, np.uint32)
        if n:
            capi.check(capi.lib().epk_group_batch(_ptr(data) if data.size else None, _ptr(off), n, _ptr(rep)))
        result = PlacedCollection()
        for i in range(n):
            result.sequence_map.setdefault(seqs[i], []).append(headers[i])
            if rep[i] != i:
                continue
            m = int(out["n_place"][i])
            result.placed_seqs.append(PlacedSequence(seqs[i], [
                Placement(int(out["branch"][i, j]), float(out["score"][i, j]), float(out["lwr"][i, j]),
                          int(out["count"][i, j]), float(out["distal"][i, j]), float(out["pendant"][i, j]))
                for j in range(m)]))
        return result
