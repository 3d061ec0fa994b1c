"""epk_place_device_noalloc: the placement entry point that allocates nothing (static device scratch,
no counterpart in the reference) against epk_place_device, through the C ABI.

Collected last on purpose (the name): a failure here must not hide another test."""
import ctypes as C

import numpy as np
import pytest

from epik_b200 import capi, synth
from epik_b200.placer import Database, Placer, pack_sequences

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _build(built):
    return built


def _device_batch(torch, dev, data, off, K):
    n = len(off) - 1
    d_seq = torch.from_numpy(np.ascontiguousarray(data, np.uint8)).to(dev)
    d_off = torch.from_numpy(np.ascontiguousarray(off, np.uint64).view(np.int64)).to(dev)
    out = {"n_place": torch.full((n,), 77, dtype=torch.int32, device=dev),
           "branch": torch.zeros(n * K, dtype=torch.int32, device=dev),
           "score": torch.zeros(n * K, dtype=torch.float32, device=dev),
           "lwr": torch.zeros(n * K, dtype=torch.float64, device=dev),
           "count": torch.zeros(n * K, dtype=torch.int32, device=dev)}
    return d_seq, d_off, out


def _noalloc(placer, d_seq, d_off, n, out, stream=None):
    res = capi.Result(out["n_place"].data_ptr(), out["branch"].data_ptr(), out["score"].data_ptr(), out["lwr"].data_ptr(),
                      out["count"].data_ptr(), None, None)
    s = None if stream is None else C.c_void_p(stream.cuda_stream)
    return capi.lib().epk_place_device_noalloc(placer.handle, d_seq.data_ptr(), d_off.data_ptr(), int(n), C.byref(res), s)


def _host(out, n, K):
    return {"n_place": out["n_place"].cpu().numpy().view(np.uint32),
            "branch": out["branch"].cpu().numpy().view(np.uint32).reshape(n, K),
            "score": out["score"].cpu().numpy().reshape(n, K),
            "lwr": out["lwr"].cpu().numpy().reshape(n, K),
            "count": out["count"].cpu().numpy().view(np.uint32).reshape(n, K)}


def _same(got, want, what):
    assert np.array_equal(got["n_place"], want["n_place"]), what + " n_place"
    valid = np.arange(want["branch"].shape[1])[None, :] < want["n_place"][:, None]
    assert np.array_equal(got["branch"][valid], want["branch"][valid]), what + " branch"
    assert np.array_equal(got["count"][valid], want["count"][valid]), what + " count"
    assert np.array_equal(got["score"][valid], want["score"][valid]), what + " score (bit for bit)"
    np.testing.assert_allclose(got["lwr"][valid], want["lwr"][valid], rtol=1e-9, atol=0, err_msg=what + " lwr")


def test_placement_without_allocation_matches_the_allocating_entry_point():
    """Two calls in a row on one placer (different batches: mixed lengths over all three count widths, ambiguous
    letters, a read of k - 1 letters), then a call on a placer of another tree size - the static scratch is shared and
    keeps the last call's contents, every call clears what it uses - each against epk_place_device."""
    import torch
    dev = torch.device("cuda", 0)
    K = 7
    flat = synth.make_db(0, 8, 300, 0.6, 12.0, seed=41)
    placer = Placer(Database(flat), K, 0.01)
    rng = np.random.default_rng(5)
    batches = []
    data, off = synth.make_reads(0, 3000, 150, seed=42, amb_rate=0.01)
    batches.append((data, off))
    seqs = ["".join(rng.choice(list("ACGTN"), p=[.24, .24, .24, .24, .04], size=int(L))) for L in (7, 8, 150, 262, 263, 900, 70000, 30, 150)]
    batches.append(pack_sequences(seqs * 3))
    for turn, (data, off) in enumerate(batches):
        n = len(off) - 1
        d_seq, d_off, want = _device_batch(torch, dev, data, off, K)
        placer.place_device(d_seq, d_off, n, want)
        torch.cuda.synchronize(dev)
        _, _, got = _device_batch(torch, dev, data, off, K)
        assert _noalloc(placer, d_seq, d_off, n, got) == capi.EPK_OK, capi.lib().epk_last_error()
        torch.cuda.synchronize(dev)
        _same(_host(got, n, K), _host(want, n, K), "call %d" % turn)
    # another tree size right after: other row length in the same scratch, and a side stream
    flat2 = synth.make_db(0, 7, 2000, 0.7, 20.0, seed=43)
    placer2 = Placer(Database(flat2), K, 0.01)
    data, off = synth.make_reads(0, 2500, 120, seed=44, amb_rate=0.005)
    n = len(off) - 1
    d_seq, d_off, want = _device_batch(torch, dev, data, off, K)
    placer2.place_device(d_seq, d_off, n, want)
    torch.cuda.synchronize(dev)
    _, _, got = _device_batch(torch, dev, data, off, K)
    stream = torch.cuda.Stream(device=dev)
    assert _noalloc(placer2, d_seq, d_off, n, got, stream) == capi.EPK_OK, capi.lib().epk_last_error()
    stream.synchronize()
    _same(_host(got, n, K), _host(want, n, K), "other tree")
    # nothing to do is not an error
    assert _noalloc(placer2, d_seq, d_off, 0, got) == capi.EPK_OK


def test_trees_beyond_the_static_scratch_are_refused():
    import torch
    dev = torch.device("cuda", 0)
    flat = synth.make_db(0, 6, 40000, 0.3, 4.0, seed=45)  # 79 999 nodes > EPK_NOALLOC_MAX_NODES
    placer = Placer(Database(flat))
    data, off = synth.make_reads(0, 10, 100, seed=46)
    d_seq, d_off, got = _device_batch(torch, dev, data, off, 7)
    assert _noalloc(placer, d_seq, d_off, 10, got) == capi.EPK_E_UNSUPPORTED
    assert "65536" in capi.lib().epk_last_error().decode()
    placer.place_device(d_seq, d_off, 10, got)  # the allocating entry point takes it
    torch.cuda.synchronize(dev)
    assert int(got["n_place"].sum().item()) > 0
