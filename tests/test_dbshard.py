"""Database-sharded mode (epik_b200/dbshard.py, include/epik_b200.h "database-sharded mode").

CPU part (gloo, world_size 2 and 3): the key -> shard rule exported by the library, the owner slices, and the
all-to-all layout of the partial rows; the rows themselves come from a numpy restatement of the exact
accumulation (place.cpp:349-371) over one shard, so that what is checked is: the shards partition the hits,
every row reaches its owner in shard order, and the summed rows equal the unsharded accumulators.
GPU part: the CUDA partial + finalize launches against the oracle (1e-5 relative: the per-branch sums are
re-associated by shard), and the one-shard case bit for bit against the fused kernel."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from epik_b200 import capi, dbshard, synth

from .helpers import assert_close_placements


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_rule_partitions_the_keys(built):
    l = capi.lib()
    keys = np.arange(0, 1 << 20, 7, dtype=np.uint32)
    for world in (1, 2, 3, 8):
        owner = np.array([l.epk_shard_of_key(int(k), world) for k in keys[:20000]])
        assert owner.min() >= 0 and owner.max() == world - 1
        # the python statement of the rule used below agrees with the library
        assert np.array_equal(owner, _shard_of(keys[:20000], world))
        share = np.bincount(owner, minlength=world) / len(owner)
        assert np.all(np.abs(share - 1.0 / world) < 0.02)


def test_owner_bounds():
    for n in (0, 1, 5, 1000):
        for world in (1, 2, 3, 8):
            b = dbshard.owner_bounds(n, world)
            assert b[0] == 0 and b[-1] == n and np.all(np.diff(b.astype(np.int64)) >= 0) and len(b) == world + 1


def _shard_of(keys, world):
    h = (keys.astype(np.uint64) * np.uint64(0x9E3779B1)) & np.uint64(0xFFFFFFFF)
    return (((h >> np.uint64(16)) * np.uint64(world)) >> np.uint64(16)).astype(np.int64)


def _rows_numpy(flat, data, off, shard, world, n_pad):
    """Partial rows {S f32[n_pad] | C u16[n_pad]} of every read over the keys of one shard (unambiguous reads)."""
    k = flat.kmer_size
    code = np.full(256, -1, np.int64)
    for i, ch in enumerate(b"ACGT"):
        code[ch] = i
    lookup = {int(key): i for i, key in enumerate(flat.keys)}
    n = len(off) - 1
    S = np.zeros((n, n_pad), np.float32)
    C = np.zeros((n, n_pad), np.uint16)
    for r in range(n):
        seq = code[data[int(off[r]):int(off[r + 1])]]
        for w in range(len(seq) - k + 1):
            key = 0
            for c in seq[w:w + k]:
                key = (key << 2) | int(c)
            i = lookup.get(key)
            if i is None or (world > 1 and _shard_of(np.array([key], np.uint32), world)[0] != shard):
                continue
            for e in flat.entries[int(flat.offsets[i]):int(flat.offsets[i + 1])]:
                S[r, e["branch"]] = np.float32(S[r, e["branch"]] + e["score"])
                C[r, e["branch"]] += 1
    return S, C


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    flat = synth.make_db(0, 5, 20, 0.7, 4.0, seed=3)
    data, off = synth.make_reads(0, 23, 40, seed=4)
    n = len(off) - 1
    n_pad = 128
    S, C = _rows_numpy(flat, data, off, rank, world, n_pad)
    rows = np.concatenate([S.view(np.uint8).reshape(n, -1), C.view(np.uint8).reshape(n, -1)], axis=1)
    slabs = dbshard.exchange_rows(torch.from_numpy(rows.copy()), n, rank, world).numpy()
    np.save(os.path.join(out_dir, "slabs%d.npy" % rank), slabs)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_rows_reach_their_owner_in_shard_order(built, tmp_path, world):
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    flat = synth.make_db(0, 5, 20, 0.7, 4.0, seed=3)
    data, off = synth.make_reads(0, 23, 40, seed=4)
    n, n_pad = len(off) - 1, 128
    S_full, C_full = _rows_numpy(flat, data, off, 0, 1, n_pad)
    per_shard = [_rows_numpy(flat, data, off, g, world, n_pad) for g in range(world)]
    assert np.array_equal(sum(c.astype(np.int64) for _, c in per_shard), C_full.astype(np.int64))  # the shards partition the hits
    bounds = dbshard.owner_bounds(n, world)
    for rank in range(world):
        slabs = np.load(str(tmp_path / ("slabs%d.npy" % rank)))
        first, last = int(bounds[rank]), int(bounds[rank + 1])
        assert slabs.shape == (world, last - first, 6 * n_pad)
        S = slabs[:, :, :4 * n_pad].copy().view(np.float32)
        C = slabs[:, :, 4 * n_pad:].copy().view(np.uint16)
        for g in range(world):
            assert np.array_equal(S[g], per_shard[g][0][first:last]) and np.array_equal(C[g], per_shard[g][1][first:last])
        total = np.zeros_like(S[0])
        for g in range(world):
            total = (total + S[g]).astype(np.float32)
        np.testing.assert_allclose(total, S_full[first:last], rtol=1e-5, atol=1e-6)
        assert np.array_equal(C.sum(axis=0), C_full[first:last])


# ------------------------------------------------------------------------------------------ GPU -----

def _oracle(flat):
    from oracle import oracle as O
    return O.Oracle(flat, kind="reference" if O.available("reference", flat.alphabet) else "port")


@pytest.mark.gpu
def test_one_shard_equals_the_fused_kernel_bit_for_bit(built):
    from epik_b200.placer import Database, Placer
    flat = synth.make_db(0, 8, 300, 0.6, 12.0, seed=21)
    data, off = synth.make_reads(0, 3000, 150, seed=22)
    want = Placer(Database(flat)).place_packed(data, off)
    sp = dbshard.LocalShardedPlacer(flat, 1)
    got = sp.place_packed(data, off, max_rows=1024)
    assert sp.unsupported == 0 and sp.short == 0
    for name in dbshard.RESULT_FIELDS:
        if name == "lwr":
            # the fused kernel's two-pass sweep and the finalising launch's one-pass sweep evaluate 10^s with different
            # polynomials (relative error 6e-12 / 2e-12) and add the FP64 terms of the likelihood mass in different
            # orders: equal far below the 1e-9 asked of the weight ratios, not to the bit
            np.testing.assert_allclose(got[name], want[name], rtol=1e-10, atol=0)
        else:
            assert np.array_equal(got[name], want[name]), name


@pytest.mark.gpu
@pytest.mark.parametrize("alphabet,k,leaves,p_occ,m,length,shards", [(0, 8, 300, 0.6, 12.0, 150, 3), (0, 10, 2500, 0.6, 32.0, 150, 2),
                                                                  (1, 6, 200, 0.05, 8.0, 100, 4), (0, 8, 100, 0.5, 6.0, 700, 2),
                                                                  (0, 8, 10000, 0.6, 40.0, 150, 2),   # N = 19 999: the partial launch runs in range teams
                                                                  (0, 8, 10000, 0.6, 40.0, 400, 3)])  # ... with 16-bit counts, a team of 8
def test_sharded_database_matches_the_oracle(built, alphabet, k, leaves, p_occ, m, length, shards):
    flat = synth.make_db(alphabet, k, leaves, p_occ, m, seed=31)
    data, off = synth.make_reads(alphabet, 2000, length, seed=32)
    sp = dbshard.LocalShardedPlacer(flat, shards)
    got = sp.place_packed(data, off, max_rows=777)
    assert sp.unsupported == 0 and sp.short == 0
    want = _oracle(flat).place_packed(data, off)
    assert_close_placements(got, want, what="dbshard x%d" % shards)
    loaded = sum(db.info["n_entries_loaded"] for db in sp.dbs)
    assert loaded == len(flat.entries)  # the shards partition the database


@pytest.mark.gpu
def test_sharded_load_filter_cuts_globally(built):
    from epik_b200.placer import Database
    flat = synth.make_db(0, 8, 300, 0.6, 12.0, seed=41)
    whole = Database(flat, mu=0.5, omega=1.5)
    parts = [Database(flat, mu=0.5, omega=1.5, shard=(g, 3)) for g in range(3)]
    assert sum(p.info["n_entries_loaded"] for p in parts) == whole.info["n_entries_loaded"]
    assert sum(p.info["n_keys_loaded"] for p in parts) == whole.info["n_keys_loaded"]


@pytest.mark.gpu
def test_short_reads_are_reported_and_ambiguous_ones_need_the_views(built):
    flat = synth.make_db(0, 8, 300, 0.6, 12.0, seed=21)
    seqs = [b"ACGTACGTACGTAACCGGTTACGT" * 4, b"ACGTACGTACGNAACCGGTTACGT" * 4, b"ACG", b"ACGTACGTACGTAACCGGTTACGT" * 4]
    from epik_b200.placer import pack_sequences
    data, off = pack_sequences(seqs)
    sp = dbshard.LocalShardedPlacer(flat, 2)
    got = sp.place_packed(data, off)
    assert sp.unsupported == 0 and sp.short == 1
    assert got["n_place"][1] > 0 and got["n_place"][2] == 0 and got["n_place"][0] > 0
    assert np.array_equal(got["branch"][0], got["branch"][3]) and np.array_equal(got["score"][0], got["score"][3])
    # without the views of the shards the owner cannot order the ambiguous keys: reported, not placed (as before)
    sp.placers[0].set_shard_views([])
    got = sp.place_packed(data, off)
    assert sp.unsupported == 1 and got["n_place"][1] == 0 and got["n_place"][0] > 0


@pytest.mark.gpu
@pytest.mark.parametrize("alphabet,k,leaves,p_occ,m,length,shards,amb", [
    (0, 8, 300, 0.6, 12.0, 150, 3, 0.02), (0, 10, 2500, 0.6, 32.0, 150, 2, 0.01), (1, 3, 200, 0.5, 8.0, 60, 4, 0.03),
    (0, 8, 5000, 0.9, 60.0, 300, 2, 0.01), (1, 6, 150, 0.02, 6.0, 60, 2, 0.05)],
    ids=["dna-3", "dna-C2-2", "aa-4", "dna-range-teams-2", "aa-hashed-2"])
def test_ambiguous_windows_are_replayed_by_the_owner(built, alphabet, k, leaves, p_occ, m, length, shards, amb):
    """place.cpp:373-415 in database-sharded mode: the finalising launch looks every resolved key up in the shard that
    holds it, in the reference's key order. Parity with the oracle at the mode's 1e-5 (re-associated float sums)."""
    flat = synth.make_db(alphabet, k, leaves, p_occ, m, seed=91, key_space_limit=(1 << 22) if (alphabet, k) == (1, 6) else None)
    data, off = synth.make_reads(alphabet, 1500, length, seed=92, amb_rate=amb)
    want = _oracle(flat).place_packed(data, off)
    sp = dbshard.LocalShardedPlacer(flat, shards)
    got = sp.place_packed(data, off, max_rows=512)
    assert sp.unsupported == 0 and sp.short == 0
    # (the sparse amino-acid case is full of EXACT ties: a branch named only by an ambiguous key gets a probability of
    # 1e-7 .. 1e-2 added to a log score of -360, which float32 mostly swallows; exact ties are ordered by branch id
    # here and by libstdc++'s partial_sort in the reference - DESIGN.md, "Ties")
    assert_close_placements(got, want, what="ambiguous, %d shards" % shards, max_reordered=0.1 if (alphabet, k) == (1, 6) else 0.002)
    # and with one shard the ambiguous path adds to the very sums of the fused kernel
    if shards == 2:
        from epik_b200.placer import Database, Placer
        fused = Placer(Database(flat)).place_packed(data, off)
        one = dbshard.LocalShardedPlacer(flat, 1).place_packed(data, off, max_rows=512)
        for name in ("n_place", "branch", "score", "count"):
            assert np.array_equal(one[name], fused[name]), name


@pytest.mark.gpu
@pytest.mark.parametrize("shards", [1, 2, 4])
def test_both_row_formats_give_the_same_placements(built, monkeypatch, shards):
    """Partial rows are dense (header | S | C) up to two shards and compact (bitmap | step counts | touched values) from
    three on; either format carries the same sums, so forcing one or the other changes no bit of the result - reads with
    ambiguous windows, a range-team tree and a one-warp tree included."""
    for leaves, length in ((300, 150), (5000, 200)):
        flat = synth.make_db(0, 8, leaves, 0.7, 24.0, seed=51)
        data, off = synth.make_reads(0, 700, length, seed=52, amb_rate=0.01)
        got = {}
        for fmt in ("dense", "compact"):
            monkeypatch.setenv("EPK_ROW_FORMAT", fmt)
            sp = dbshard.LocalShardedPlacer(flat, shards)
            got[fmt] = sp.place_packed(data, off, max_rows=256)
            assert sp.unsupported == 0
        monkeypatch.delenv("EPK_ROW_FORMAT")
        for name in dbshard.RESULT_FIELDS:
            assert np.array_equal(got["dense"][name], got["compact"][name]), (leaves, name)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", ["dna-one-warp", "dna-range-teams", "dna-filtered", "dna-generated"])
def test_pooled_database_gives_the_unsharded_results_bit_for_bit(built, shape):
    """epk_db_pool: the shards mapped into one address range, every GPU's key table pointing into it, the fused kernel
    unchanged on top - so nothing is re-associated and every bit of the unsharded database's placements is there,
    ambiguous reads, load filter and range teams included. One GPU: three shards on it; more GPUs: one shard each."""
    from epik_b200.placer import Database, Placer
    n_dev = min(torch.cuda.device_count(), 4)
    devices = list(range(n_dev)) if n_dev > 1 else [0, 0, 0]
    load = {}
    if shape == "dna-generated":
        flat = synth.Gen2(0, 9, 600, 0.6, 16.0, seed=61)
        whole = Database.synthetic(flat)
    else:
        leaves = 5000 if shape == "dna-range-teams" else 400
        flat = synth.make_db(0, 8, leaves, 0.7, 24.0, seed=61)
        if shape == "dna-filtered":
            load = dict(mu=0.6, omega=1.2)
        whole = Database(flat, **load)
    data, off = synth.make_reads(0, 4000, 150, seed=62, amb_rate=0.01)
    want = Placer(whole).place_packed(data, off)
    pooled = dbshard.PooledPlacer(flat, devices, **load)
    assert pooled.dbs[0].info["n_entries_loaded"] == whole.info["n_entries_loaded"]
    assert pooled.dbs[0].info["n_keys_loaded"] == whole.info["n_keys_loaded"]
    for turn in range(2):
        got = pooled.place_packed(data, off)
        for name in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant"):
            assert np.array_equal(got[name], want[name]), (shape, turn, name)
    # every pooled database is a whole database on its own, too
    alone = Placer(pooled.dbs[-1]).place_packed(data, off)
    assert np.array_equal(alone["score"], want["score"]) and np.array_equal(alone["branch"], want["branch"])
    pooled.close()


def _pool_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    device = rank % torch.cuda.device_count()
    torch.cuda.set_device(device)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # (bookkeeping only: the data path is loads over NVLink)
    flat = synth.make_db(0, 8, 5000, 0.7, 24.0, seed=81) if rank >= 0 else None
    data, off = synth.make_reads(0, 1200, 150, seed=82, amb_rate=0.01)
    rp = dbshard.RankPooledPlacer(flat, rank, world, device=device, mu=0.7, omega=1.3)
    first, last = len(off) - 1, 0
    for _ in range(2):
        first, last = int(dbshard.owner_bounds(len(off) - 1, world)[rank]), int(dbshard.owner_bounds(len(off) - 1, world)[rank + 1])
        got = rp.placer.place_packed(data[int(off[first]):int(off[last])], off[first:last + 1] - off[first])
    np.savez(os.path.join(out_dir, "pool%d.npz" % rank), first=first, last=last, entries=rp.db.info["n_entries_loaded"], **got)
    dist.barrier()
    rp.close()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_pool_with_one_process_per_shard(built, tmp_path):
    """The pool across PROCESSES (what torchrun gives: one per GPU; here three, sharing the GPUs that are there): memory
    as file descriptors over unix sockets, key tables as IPC handles, one address range per process. Every rank places
    its own slice against the whole database - bit for bit what one GPU with the unsharded, filtered database gives."""
    from epik_b200.placer import Database, Placer
    world = 3
    mp.spawn(_pool_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    flat = synth.make_db(0, 8, 5000, 0.7, 24.0, seed=81)
    data, off = synth.make_reads(0, 1200, 150, seed=82, amb_rate=0.01)
    whole = Database(flat, mu=0.7, omega=1.3)
    want = Placer(whole).place_packed(data, off)
    for rank in range(world):
        part = np.load(str(tmp_path / ("pool%d.npz" % rank)))
        first, last = int(part["first"]), int(part["last"])
        assert int(part["entries"]) == whole.info["n_entries_loaded"]
        for name in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant"):
            assert np.array_equal(part[name], want[name][first:last]), (rank, name)


def _nccl_worker(rank, world, port, out_dir, transport):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    flat = synth.make_db(0, 8, 300, 0.6, 12.0, seed=31)
    data, off = synth.make_reads(0, 2001, 150, seed=32, amb_rate=0.005)
    sp = dbshard.ShardedPlacer(flat, rank, world, device=rank, transport=transport, max_rows=4096)
    assert sp.replays_ambiguous  # the peers' tables are mapped: ambiguous windows are replayed by the owner
    for _ in range(3):  # several times: buffers are reused, the two peer slabs alternate
        got, first, last = sp.place_packed(data, off)
    assert sp.unsupported == 0
    sp.close()
    np.savez(os.path.join(out_dir, "part%d.npz" % rank), first=first, last=last, **got)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("transport", ["a2a", "p2p"])
def test_two_gpus_exchange(built, tmp_path, transport):
    """a2a: NCCL all_to_all_single of the row buffers; p2p: the partial launch stores into the owner's HBM (CUDA IPC)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = 2
    mp.spawn(_nccl_worker, args=(world, _free_port(), str(tmp_path), transport), nprocs=world, join=True)
    flat = synth.make_db(0, 8, 300, 0.6, 12.0, seed=31)
    data, off = synth.make_reads(0, 2001, 150, seed=32, amb_rate=0.005)
    want = _oracle(flat).place_packed(data, off)
    local = dbshard.LocalShardedPlacer(flat, world).place_packed(data, off)
    for rank in range(world):
        part = np.load(str(tmp_path / ("part%d.npz" % rank)))
        first, last = int(part["first"]), int(part["last"])
        piece = {k: want[k][first:last] for k in dbshard.RESULT_FIELDS}
        assert_close_placements({k: part[k] for k in dbshard.RESULT_FIELDS}, piece, what="rank %d" % rank)
        for k in dbshard.RESULT_FIELDS:  # the transport does not change a bit: same sums in the same shard order
            assert np.array_equal(part[k], local[k][first:last]), k


@pytest.mark.gpu
@pytest.mark.parametrize("n", [0, 1, 2, 5])
def test_tiny_and_empty_batches(built, n):
    """fewer queries than shards / owners, and no queries at all"""
    flat = synth.make_db(0, 8, 300, 0.6, 12.0, seed=21)
    data, off = synth.make_reads(0, max(n, 1), 150, seed=5)
    if n == 0:
        data, off = data[:0], off[:1]
    sp = dbshard.LocalShardedPlacer(flat, 3)
    got = sp.place_packed(data, off)
    assert got["n_place"].shape == (n,) and sp.unsupported == 0
    if n:
        want = _oracle(flat).place_packed(data, off)
        assert_close_placements(got, want, what="n=%d" % n)
