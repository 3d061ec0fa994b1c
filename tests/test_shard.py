"""Multi-rank host logic on CPU: world_size 2 and 3 over gloo. The per-rank placer is the CPU oracle port here
(there is no GPU in this container); what is under test is the slicing and the gather of epik_b200/shard.py:
N ranks must return, on rank 0, exactly the arrays one rank returns for the whole batch."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from epik_b200 import shard, synth


def test_slices_cover_the_batch_once():
    for n in (0, 1, 7, 8, 1000, 1001):
        for world in (1, 2, 3, 8):
            cuts = [shard.slice_for_rank(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            sizes = [l - f for f, l in cuts]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        shard.slice_for_rank(10, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as O
    flat = synth.make_db(0, 7, 60, 0.6, 6.0, seed=5)
    data, off = synth.make_reads(0, 101, 80, seed=6, amb_rate=0.01)
    off = off.copy()
    oracle = O.Oracle(flat, kind="port")
    res = shard.place_sharded(lambda d, o: oracle.place_packed(d, o), data, off, rank, world)
    if rank == 0:
        np.savez(out_path, **res)
    else:
        assert res is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_placement_equals_single_rank(built, tmp_path, world):
    from oracle import oracle as O
    out_path = str(tmp_path / "sharded.npz")
    mp.spawn(_worker, args=(world, _free_port(), out_path), nprocs=world, join=True)
    got = np.load(out_path)
    flat = synth.make_db(0, 7, 60, 0.6, 6.0, seed=5)
    data, off = synth.make_reads(0, 101, 80, seed=6, amb_rate=0.01)
    want = O.Oracle(flat, kind="port").place_packed(data, off)
    single = shard.place_sharded(lambda d, o: O.Oracle(flat, kind="port").place_packed(d, o), data, off)
    for name in shard.RESULT_FIELDS:
        assert np.array_equal(got[name], want[name]), name
        assert np.array_equal(single[name], want[name]), name
