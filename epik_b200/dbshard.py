"""Database-sharded placement (second row): the k-mer keys are split by hash range over the
GPUs of a group, every GPU looks every query up in its shard, and the per-branch partial sums are brought
together on the GPU that owns the query before the threshold correction, the likelihood mass, the top-k and
the weight ratios (place.cpp:417-422, 164-199, 238-267) run there.

Needed only when a database exceeds one GPU's HBM; the default is query sharding (shard.py) with the
database replicated. The reference has nothing like it (one phylo_kmer_db in one address space,
main.cpp:277): what is preserved is the result, within float rounding of a re-associated sum.

Three transports for the partial rows (include/epik_b200.h, "database-sharded mode"):
  * "p2p"  — the fused form: every owner's slab is mapped into its peers (CUDA IPC over NVLink / NVSwitch) and
             the partial launch stores each row straight into the HBM of the GPU that finalises it - the
             launch IS the exchange, nothing is staged or copied. A one-element all_reduce on the stream is the
             only collective (it orders "all partial launches done" before the finalising launch); two slabs
             alternate so that a step's stores never land in a slab a peer is still reading;
  * "a2a"  — every rank writes the rows of all queries into a send buffer ordered by owner and one
             all_to_all_single (NCCL over NVLink; gloo in the CPU tests of the layout) delivers the slabs;
  * "local" — all shards live on ONE device (tests, and databases that merely do not fit one table build):
             the rows are written straight into the slab the finalising launch reads.
All computing is in libepik_b200.so; this module only sizes buffers and calls the collective.
"""
from __future__ import annotations

import numpy as np

from .placer import Database, Placer
from .shard import slice_for_rank

RESULT_FIELDS = ("n_place", "branch", "score", "lwr", "count")


def owner_bounds(n: int, world: int) -> np.ndarray:
    """owner_first[g] .. owner_first[g + 1]: the queries of a batch of n that rank g finalises."""
    return np.array([slice_for_rank(n, g, world)[0] for g in range(world)] + [n], dtype=np.uint32)


def exchange_rows(send, n: int, rank: int, world: int, group=None):
    """send: uint8 tensor [n, row_bytes], rows ordered by query (hence by owner). Returns the tensor
    [world, m, row_bytes] of the rows of this rank's m queries, one slab per shard, in shard order."""
    import torch
    import torch.distributed as dist
    bounds = owner_bounds(n, world)
    sizes = [int(bounds[g + 1] - bounds[g]) for g in range(world)]
    m = sizes[rank]
    recv = torch.empty((world * m, send.shape[1]), dtype=send.dtype, device=send.device)
    dist.all_to_all_single(recv, send, output_split_sizes=[m] * world, input_split_sizes=sizes, group=group)
    return recv.view(world, m, send.shape[1])


def _alloc_out(torch, n: int, K: int, device):
    return {"n_place": torch.zeros(n, dtype=torch.int32, device=device),
            "branch": torch.zeros((n, K), dtype=torch.int32, device=device),
            "score": torch.zeros((n, K), dtype=torch.float32, device=device),
            "lwr": torch.zeros((n, K), dtype=torch.float64, device=device),
            "count": torch.zeros((n, K), dtype=torch.int32, device=device)}


def _to_host(out: dict) -> dict:
    res = {k: v.cpu().numpy() for k, v in out.items()}
    for k in ("n_place", "branch", "count"):
        res[k] = res[k].view(np.uint32)
    return res


def _shard_database(flat, device: int, shard, load: dict):
    """One shard of `flat`: a FlatDB (host arrays) or a synth.Gen2 (generated on the device, shard by shard)."""
    from .synth import Gen2
    if isinstance(flat, Gen2):
        return Database.synthetic(flat, device=device, shard=shard, **load)
    return Database(flat, device=device, shard=shard, **load)


class LocalShardedPlacer:
    """All n_shards shards on one device: partial launches write into the slab the finalising launch reads."""

    def __init__(self, flat, n_shards: int, device: int = 0, keep_at_most: int = 7, keep_factor: float = 0.01, **load):
        self.dbs = [_shard_database(flat, device, (g, n_shards), load) for g in range(n_shards)]
        self.placers = [Placer(db, keep_at_most, keep_factor) for db in self.dbs]
        # the finalising placer sees every shard: windows with one ambiguous letter are replayed there (place.cpp:373-415)
        self.placers[0].set_shard_views([db.view() for db in self.dbs])
        self.n_shards = n_shards
        self.device = device
        self.keep_at_most = keep_at_most

    def place_packed(self, data, offsets, max_rows: int = 1 << 15):
        import torch
        dev = torch.device("cuda", self.device)
        n = len(offsets) - 1
        d_seq = torch.from_numpy(np.ascontiguousarray(data, np.uint8)).to(dev) if len(data) else torch.zeros(1, dtype=torch.uint8, device=dev)
        d_off = torch.from_numpy(np.ascontiguousarray(offsets, np.uint64).view(np.int64)).to(dev)
        out = _alloc_out(torch, n, self.keep_at_most, dev)
        row_bytes = self.placers[0].partial_row_bytes()
        step = max(1, min(n, max_rows))
        slab = torch.empty((self.n_shards, step, row_bytes), dtype=torch.uint8, device=dev)
        self.unsupported = 0
        self.short = 0
        for first in range(0, n, step):
            m = min(step, n - first)
            off_ptr = d_off.data_ptr() + 8 * first
            for g, pl in enumerate(self.placers):
                # absolute offsets: the sequence base stays, the offsets pointer moves
                pl.place_partial_device(d_seq, _Ptr(off_ptr), m, [0, m], [slab[g].data_ptr()])
            piece = {k: v[first:first + m] for k, v in out.items()}
            self.placers[0].place_finalize_device(slab.data_ptr(), self.n_shards, step * row_bytes, off_ptr, m, piece,
                                                  d_seq_ptr=d_seq.data_ptr())
            c = self.placers[0].counters()
            self.unsupported += c["n_unsupported"]
            self.short += c["n_short"]
        torch.cuda.synchronize(dev)
        return _to_host(out)


class _Ptr:
    """Minimal .data_ptr() carrier for an address inside a tensor."""

    def __init__(self, p: int):
        self._p = int(p)

    def data_ptr(self) -> int:
        return self._p


class ShardedPlacer:
    """One rank of a database-sharded group (torch.distributed, one process per GPU). Every rank passes the
    SAME batch; rank r gets back the placements of its slice [first, last) of the batch."""

    def __init__(self, flat, rank: int, world: int, group=None, device: int = 0, keep_at_most: int = 7,
                 keep_factor: float = 0.01, transport: str = "a2a", max_rows: int = 1 << 15, peer_databases: bool = True,
                 **load):
        if transport not in ("a2a", "p2p"):
            raise ValueError("transport must be 'a2a' or 'p2p'")
        self.db = _shard_database(flat, device, (rank, world), load)
        self.placer = Placer(self.db, keep_at_most, keep_factor)
        self.rank, self.world, self.group, self.device = rank, world, group, device
        self.keep_at_most = keep_at_most
        self.row_bytes = self.placer.partial_row_bytes()
        self.transport = transport
        self.max_rows = int(max_rows)
        self._send = None
        self.unsupported = 0
        self.short = 0
        self._slabs, self._peer = [], []
        self._peer_db = []
        self._turn = 0
        if transport == "p2p":
            self._open_peer_slabs()
        self.replays_ambiguous = self._open_peer_databases() if peer_databases else False

    # ------------------------------------------------------------------ fused exchange -----
    def _slab_rows(self) -> int:
        """rows of one shard's part of a slab: the largest slice a rank can own of max_rows queries"""
        return slice_for_rank(self.max_rows, 0, self.world)[1]

    def _open_peer_slabs(self):
        """Two slabs [world][slab_rows][row_bytes] in this rank's HBM, exported to every peer; the peers' slabs mapped here."""
        import ctypes as C
        import torch.distributed as dist
        from . import capi
        l = capi.lib()
        bytes_ = self.world * self._slab_rows() * self.row_bytes
        handles = []
        for _ in range(2):
            p = C.c_void_p()
            capi.check(l.epk_device_alloc(self.device, bytes_, C.byref(p)))
            self._slabs.append(p.value)
            h = C.create_string_buffer(64)
            capi.check(l.epk_ipc_export(C.c_void_p(p.value), h))
            handles.append(h.raw)
        everyone = [None] * self.world
        dist.all_gather_object(everyone, handles, group=self.group)
        for turn in range(2):
            mapped = []
            for r in range(self.world):
                if r == self.rank:
                    mapped.append(self._slabs[turn])
                    continue
                q = C.c_void_p()
                capi.check(l.epk_ipc_import(self.device, everyone[r][turn], C.byref(q)))
                mapped.append(q.value)
            self._peer.append(mapped)

    def _open_peer_databases(self) -> bool:
        """Maps every peer shard's key table and postings here (CUDA IPC over NVLink) and hands the views to the placer:
        the finalising launch then replays the ambiguous windows of its queries against the shard that holds each key.
        With either transport (the rows may still travel through NCCL). False if the peers cannot be mapped."""
        import ctypes as C
        import torch.distributed as dist
        from . import capi
        l = capi.lib()
        mine = self.db.view()
        handles = []
        for addr in (mine.table, mine.postings):
            h = C.create_string_buffer(64)
            capi.check(l.epk_ipc_export(C.c_void_p(addr), h))
            handles.append(h.raw)
        everyone = [None] * self.world
        dist.all_gather_object(everyone, (handles, int(mine.hash_mask), int(mine.hash_shift), int(mine.table_kind)), group=self.group)
        views, ok = [], True
        for r in range(self.world):
            if r == self.rank:
                views.append(mine)
                continue
            (h_table, h_post), mask, shift, kind = everyone[r]
            addrs = []
            for h in (h_table, h_post):
                q = C.c_void_p()
                if l.epk_ipc_import(self.device, h, C.byref(q)) != capi.EPK_OK:
                    ok = False
                    break
                addrs.append(q.value)
                self._peer_db.append(q.value)
            if not ok:
                break
            views.append(capi.DbView(addrs[0], addrs[1], mask, shift, kind))
        flags = [None] * self.world
        dist.all_gather_object(flags, ok, group=self.group)
        if not all(flags):
            return False
        self.placer.set_shard_views(views)
        return True

    def close(self):
        if not self._slabs and not self._peer and not self._peer_db:
            return
        from . import capi
        l = capi.lib()
        if self._peer_db:
            self.placer.set_shard_views([])
            for p in self._peer_db:
                l.epk_ipc_close(self.device, p)
            self._peer_db = []
        for turn, mapped in enumerate(self._peer):
            for r, p in enumerate(mapped):
                if r != self.rank:
                    l.epk_ipc_close(self.device, p)
        self._peer = []
        if self._slabs:
            import torch
            import torch.distributed as dist
            torch.cuda.synchronize(self.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)  # nobody unmaps-then-frees under a peer that still stores
        for p in self._slabs:
            l.epk_device_free(self.device, p)
        self._slabs = []

    def _place_p2p(self, d_seq, d_off, n: int, out):
        import torch
        import torch.distributed as dist
        if n > self.max_rows:
            raise ValueError("p2p transport: at most max_rows = %d queries per call" % self.max_rows)
        bounds = owner_bounds(n, self.world)
        first, last = int(bounds[self.rank]), int(bounds[self.rank + 1])
        turn = self._turn
        self._turn ^= 1
        stride = self._slab_rows() * self.row_bytes
        # owner g's slab, part `rank`: where this rank's rows of g's queries go
        dest = [self._peer[turn][g] + self.rank * stride for g in range(self.world)]
        self.placer.place_partial_device(d_seq, d_off, n, bounds, dest)
        if self._flag is None:
            self._flag = torch.zeros(1, dtype=torch.int32, device=d_off.device)
        dist.all_reduce(self._flag, group=self.group)  # stream-ordered: every rank's partial launch has completed
        m = last - first
        if out is None:
            out = _alloc_out(torch, m, self.keep_at_most, d_off.device)
        self.placer.place_finalize_device(self._slabs[turn], self.world, stride, d_off.data_ptr() + 8 * first, m, out,
                                          d_seq_ptr=d_seq.data_ptr())
        return out, first, last

    _flag = None

    def place_device(self, d_seq, d_off, n: int, out: dict | None = None):
        """d_seq (uint8) / d_off (int64 view of uint64 offsets, n + 1) on this rank's device; the whole batch.
        Returns (out, first, last): device tensors with the rows of queries [first, last)."""
        import torch
        if self.transport == "p2p":
            return self._place_p2p(d_seq, d_off, n, out)
        dev = d_off.device
        bounds = owner_bounds(n, self.world)
        first, last = int(bounds[self.rank]), int(bounds[self.rank + 1])
        if self._send is None or self._send.shape[0] < n:
            self._send = torch.empty((n, self.row_bytes), dtype=torch.uint8, device=dev)
        send = self._send[:n]
        base = send.data_ptr()
        self.placer.place_partial_device(d_seq, d_off, n, bounds, [base + int(bounds[g]) * self.row_bytes for g in range(self.world)])
        slabs = exchange_rows(send, n, self.rank, self.world, self.group)
        m = last - first
        if out is None:
            out = _alloc_out(torch, m, self.keep_at_most, dev)
        self.placer.place_finalize_device(slabs.data_ptr(), self.world, m * self.row_bytes, d_off.data_ptr() + 8 * first, m, out,
                                          d_seq_ptr=d_seq.data_ptr())
        return out, first, last

    def place_packed(self, data, offsets):
        """The whole batch in host memory (the same on every rank) -> (host arrays of this rank's queries, first, last).
        Batches larger than max_rows go through in pieces; the rank then owns a slice of every piece and
        (first, last) are lists."""
        import torch
        dev = torch.device("cuda", self.device)
        n = len(offsets) - 1
        d_seq = torch.from_numpy(np.ascontiguousarray(data, np.uint8)).to(dev) if len(data) else torch.zeros(1, dtype=torch.uint8, device=dev)
        d_off = torch.from_numpy(np.ascontiguousarray(offsets, np.uint64).view(np.int64)).to(dev)
        self.unsupported = self.short = 0
        pieces, firsts, lasts = [], [], []
        for at in range(0, max(n, 1), self.max_rows):
            m = min(self.max_rows, n - at)
            out, first, last = self.place_device(d_seq, d_off[at:at + m + 1], m)
            torch.cuda.synchronize(dev)
            c = self.placer.counters()
            self.unsupported += c["n_unsupported"]
            self.short += c["n_short"]
            pieces.append(_to_host(out))
            firsts.append(at + first)
            lasts.append(at + last)
        if len(pieces) == 1:
            return pieces[0], firsts[0], lasts[0]
        return {k: np.concatenate([p[k] for p in pieces]) for k in pieces[0]}, firsts, lasts


class PooledPlacer:
    """The database pooled over the HBM of `devices` (one process drives them; include/epik_b200.h "pooled database"):
    shard g of len(devices) is built on devices[g], the shards are mapped into one address range every GPU reads, and a
    pass is cut over one placer per device (epk_place_many) - each places its slice with the fused kernel, gathering
    every posting list from whichever HBM holds it. Results are the unsharded database's to the bit."""

    def __init__(self, flat, devices, keep_at_most: int = 7, keep_factor: float = 0.01, variant: int = 0, **load):
        self.devices = list(devices)
        G = len(self.devices)
        shards = [_shard_database(flat, dev, (g, G), load) for g, dev in enumerate(self.devices)]
        try:
            self.dbs = Database.pool(shards)
        finally:
            for s in shards:
                s.close()
        self.placers = [Placer(db, keep_at_most, keep_factor, variant=variant) for db in self.dbs]
        self.keep_at_most = keep_at_most

    def place_packed(self, data, offsets, out=None):
        return Placer.place_packed_many(self.placers, data, offsets, out=out)

    def close(self):
        for p in getattr(self, "placers", []):
            p.close() if hasattr(p, "close") else None
        self.placers = []
        for d in getattr(self, "dbs", []):
            d.close()
        self.dbs = []


class RankPooledPlacer:
    """One rank of a pool with one PROCESS per GPU (torchrun): shard `rank` of `world` is built on this rank's device,
    shared with the peers (its memory as a file descriptor over a unix socket, its key table as a CUDA IPC handle), and
    joined with theirs into one address range of this process. The rank then places ITS OWN queries against the whole
    database (`self.placer`, any entry point) - posting lists come from whichever HBM holds them, over NVLink.
    torch.distributed only carries the few bytes of bookkeeping (any backend: gloo will do)."""

    def __init__(self, flat, rank: int, world: int, group=None, device: int = 0, keep_at_most: int = 7,
                 keep_factor: float = 0.01, variant: int = 0, **load):
        import os
        import socket
        import tempfile
        import threading
        import torch.distributed as dist
        shard = _shard_database(flat, device, (rank, world), load)
        fd, nbytes, table = shard.pool_share()
        path = os.path.join(tempfile.gettempdir(), "epk_pool_%d_%d.sock" % (os.getpid(), rank))
        if os.path.exists(path):
            os.unlink(path)
        server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        server.bind(path)
        server.listen(world)
        everyone = [None] * world
        dist.all_gather_object(everyone, (nbytes, table, path) + tuple(shard.score_range()), group=group)

        def serve():
            for _ in range(world - 1):
                conn, _ = server.accept()
                socket.send_fds(conn, [b"fd"], [fd])
                conn.close()

        serving = threading.Thread(target=serve)
        serving.start()
        fds = []
        for r in range(world):
            if r == rank:
                fds.append(-1)
                continue
            client = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            client.connect(everyone[r][2])
            _, got, _, _ = socket.recv_fds(client, 16, 1)
            client.close()
            fds.append(got[0])
        serving.join()
        loaded = [e for e in everyone if e[6] > 0] or everyone
        score_range = (min(e[3] for e in loaded), max(e[4] for e in loaded), int(all(e[5] for e in everyone)),
                       sum(e[6] for e in everyone), sum(e[7] for e in everyone))
        # every process has read the tables it opened and holds its own references before anything is closed; and a
        # failure on one rank (GPUs that are not peers, descriptors that did not arrive) is everybody's: nobody is left
        # waiting in a collective
        failure = None
        try:
            self.db = shard.pool_join(world, fds, [e[0] for e in everyone], [e[1] for e in everyone], score_range)
        except Exception as exc:  # noqa: BLE001
            self.db = None
            failure = "rank %d: %s" % (rank, exc)
        failures = [None] * world
        dist.all_gather_object(failures, failure, group=group)
        for f in fds:
            if f >= 0:
                os.close(f)
        os.close(fd)
        server.close()
        os.unlink(path)
        shard.close()
        if any(failures):
            if self.db is not None:
                self.db.close()
                self.db = None
            raise RuntimeError("the pool could not be built: " + "; ".join(f for f in failures if f))
        self.placer = Placer(self.db, keep_at_most, keep_factor, variant=variant)
        self.rank, self.world, self.device, self.keep_at_most = rank, world, device, keep_at_most

    def close(self):
        if getattr(self, "placer", None) is not None:
            self.placer.close()
            self.placer = None
        if getattr(self, "db", None) is not None:
            self.db.close()
            self.db = None
