"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle and the
committed outputs of the reference (tests/golden/). Integer work (k-mer codes, hits, counts,
branch choices) must be bit-exact; scores within 1e-5 relative (they are in fact bit-identical for
the one-warp-per-query kernel, which is asserted separately); LWR within 1e-9 relative (FP64 sum in
a different order + CUDA's exp10 vs glibc's pow)."""
import numpy as np
import pytest

from epik_b200 import capi, flatdb, synth
from epik_b200.placer import Database, Placer, pack_sequences
from oracle import oracle as O
from tests import helpers

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _build(built):
    return built


def best_oracle(db, keep_at_most=7, keep_factor=0.01, threads=4):
    kind = "reference" if O.available("reference", int(db.alphabet)) else "port"
    return O.Oracle(db, keep_at_most, keep_factor, threads, kind=kind), kind


def test_extension_is_loaded_and_sees_the_gpu():
    assert capi.lib().epk_device_count() >= 1


def test_kat1_known_answers():
    """SURVEY.md Appendix D values, straight from the kernel."""
    db = Database(helpers.kat1_db())
    assert db.info["threshold"] == np.float32(0.052734375)
    p = Placer(db)
    r = p.place_packed(*pack_sequences(helpers.KAT1_SEQS))
    for i in (0, 3):
        assert r["n_place"][i] == 3 and list(r["branch"][i, :3]) == [2, 3, 5]
        np.testing.assert_allclose(r["score"][i, :3], [-0.25, -0.75930208, -1.09263539], rtol=1e-7)
        np.testing.assert_allclose(r["lwr"][i, :3], [0.31214008, 0.0966156451, 0.044845012], rtol=1e-7)
        assert list(r["count"][i, :3]) == [2, 1, 1]
        np.testing.assert_allclose(r["distal"][i, 0], 0.15)
        np.testing.assert_allclose(r["pendant"][i, 0], 0.483333, rtol=1e-5)
    assert r["n_place"][1] == 7 and list(r["branch"][1]) == [3, 4, 1, 0, 5, 2, 6]
    np.testing.assert_allclose(r["score"][1], -1.27790618, rtol=1e-7)
    np.testing.assert_allclose(r["lwr"][1], 0.1, rtol=1e-12)
    assert not r["count"][1].any() and not r["distal"][1].any() and not r["pendant"][1].any()
    assert r["n_place"][2] == 2 and list(r["branch"][2, :2]) == [7, 8]
    np.testing.assert_allclose(r["score"][2, :2], [0.0999774411, 0.0674062148], rtol=1e-6)
    np.testing.assert_allclose(r["lwr"][2, :2], [0.231972609, 0.215211503], rtol=1e-7)
    assert r["n_place"][4] == 7 and np.all(r["score"][4] == 0) and np.all(np.signbit(r["score"][4]))
    g = np.load(helpers.GOLDEN + "/reference_outputs.npz")
    want = {k: g["kat1/" + k] for k in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant")}
    helpers.assert_same_placements(r, want, what="kat1 golden")
    assert np.array_equal(r["score"], want["score"])


@pytest.mark.parametrize("case", helpers.load_json("cases.json"), ids=lambda c: c["name"])
def test_golden_cases(case):
    """Committed outputs of the reference's own place.cpp."""
    g = np.load(helpers.GOLDEN + "/reference_outputs.npz")
    flat = helpers.golden_case_db(case)
    data, off = synth.make_reads(case["alphabet"], case["n_reads"], case["read_len"], seed=case["seed"] + 1000,
                                 amb_rate=case["amb_rate"])
    db = Database(flat)
    assert [db.info["threshold"], db.info["log10_threshold"]] == [np.float32(x) for x in case["thresholds"]]
    assert db.info["n_entries_loaded"] == case["n_entries"] and db.info["n_keys_loaded"] == case["n_keys"]
    p = Placer(db, case["keep_at_most"], case["keep_factor"])
    r = p.place_packed(data, off)
    name = case["name"]
    want = {k: g[name + "/" + k] for k in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant")}
    helpers.assert_same_placements(r, want, what=name)
    # stronger than the contract: window-ordered accumulation makes the float scores bit-identical
    K = case["keep_at_most"]
    valid = np.arange(K)[None, :] < want["n_place"][:, None]
    assert np.array_equal(r["score"][valid], want["score"][valid]), "scores are not bit-identical"
    # k-mer codes and hits: bit-exact
    at = 0
    for i, n in enumerate(g[name + "/kmers_len"]):
        seq = bytes(data[int(off[i]):int(off[i + 1])])
        pos, key, group, hit = db.kmers(seq)
        want_k = g[name + "/kmers"][:, at:at + n]
        at += n
        assert np.array_equal(np.stack([pos, key, group, hit]).astype(np.int64), want_k), i


@pytest.mark.parametrize("alphabet,k,n_leaves,p_occ,m,read_len,amb", [
    (0, 10, 2500, 0.6, 32.0, 150, 0.0),     # C2-shaped (full-size tree, full database)
    (0, 10, 652, 0.6, 24.0, 1500, 0.001),   # C1': D652-shaped, long reads (several window chunks)
    (0, 12, 1200, 0.05, 20.0, 150, 0.005),  # k = 12: 24-bit direct table
    (0, 14, 300, 0.0002, 10.0, 200, 0.01),  # k = 14: hashed table for DNA
    (1, 6, 1000, 0.002, 16.0, 100, 0.01),   # C4-shaped: amino acids, hashed table
    (1, 4, 200, 0.3, 10.0, 60, 0.03),       # amino acids, direct table
    (0, 6, 3000, 0.9, 300.0, 150, 0.0),     # C5-shaped lists: more segment descriptors per chunk than the buffer holds
    (0, 5, 3000, 0.9, 4000.0, 300, 0.002),  # lists of thousands of postings: one lane's windows overflow the buffer
])
def test_oracle_parity_on_synthetic_shapes(alphabet, k, n_leaves, p_occ, m, read_len, amb):
    flat = synth.make_db(alphabet, k, n_leaves, p_occ, m, seed=1000 + k)
    n = 3000 if read_len <= 200 else 300
    data, off = synth.make_reads(alphabet, n, read_len, seed=77, amb_rate=amb)
    db = Database(flat)
    p = Placer(db)
    got = p.place_packed(data, off)
    oracle, kind = best_oracle(flat)
    want = oracle.place_packed(data, off, threads=4)
    helpers.assert_same_placements(got, want, what="%s k=%d" % (kind, k))
    valid = np.arange(7)[None, :] < want["n_place"][:, None]
    assert np.array_equal(got["score"][valid], want["score"][valid])
    c = p.counters()
    assert c["n_reads"] == n and c["n_windows"] == n * (read_len - k + 1) and c["n_short"] == 0
    # hits / entries of the exact windows of a few reads, from the oracle's k-mer stream
    for i in range(5):
        seq = bytes(data[int(off[i]):int(off[i + 1])])
        a, b = db.kmers(seq), oracle.kmers(seq)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)


def test_ragged_empty_and_edge_inputs():
    flat = synth.make_db(0, 6, 60, 0.5, 6.0, seed=5)
    db = Database(flat)
    p = Placer(db)
    oracle, kind = best_oracle(flat, threads=1)
    # no sequences at all
    r = p.place_packed(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert r["n_place"].shape == (0,)
    # ragged lengths from k-1 up, all-gap, all-N, lower case, duplicates
    rng = np.random.default_rng(3)
    seqs = ["ACGTA", "ACGTAC", "------", "NNNNNNNN", "acgtacgtacgt", "ACGTACGTACGT", "ACGTAC",
            "A" * 300, "ACGT" * 200, "ACGU" * 10, "ACG-TACGTAC..ACGTACGNACGTACRYACGT", "*" * 20]
    seqs += ["".join(rng.choice(list("ACGT"), size=int(L))) for L in rng.integers(5, 700, size=200)]
    got = p.place(list(enumerate(seqs)))
    want = oracle.place(seqs, threads=1)
    arr = p.place_packed(*pack_sequences(seqs))
    helpers.assert_same_placements(arr, want, what="ragged " + kind)
    # the reference-shaped result: one placed sequence per unique sequence, headers grouped
    assert len(got.placed_seqs) == len(set(seqs))
    assert got.sequence_map[b"ACGTAC"] == [1, 6]
    # a sequence shorter than k - 1 letters: the reference aborts; the ABI reports it
    with pytest.raises(capi.EpkError) as e:
        p.place_packed(*pack_sequences(["ACGTACGT", "ACG"]))
    assert e.value.status == capi.EPK_E_SHORT_SEQ
    # the placer is still usable afterwards
    arr = p.place_packed(*pack_sequences(seqs[:5]))
    helpers.assert_same_placements(arr, oracle.place(seqs[:5]), what="after error")


@pytest.mark.parametrize("keep_at_most,keep_factor", [(1, 0.01), (3, 0.5), (7, 0.0), (8, 1.0), (12, 0.01), (32, 0.001)])
def test_keep_at_most_and_keep_factor(keep_at_most, keep_factor):
    flat = synth.make_db(0, 7, 100, 0.6, 8.0, seed=21)
    data, off = synth.make_reads(0, 500, 90, seed=22, amb_rate=0.01)
    got = Placer(Database(flat), keep_at_most, keep_factor).place_packed(data, off)
    oracle, kind = best_oracle(flat, keep_at_most, keep_factor, threads=2)
    helpers.assert_same_placements(got, oracle.place_packed(data, off, threads=2), what=kind)


def test_long_reads_underflow_the_likelihood_mass():
    """10^score underflows FP64 for long queries: score_sum == 0 -> every LWR 0, nothing filtered
    (place.cpp:243-251)."""
    flat = synth.make_db(0, 10, 200, 0.3, 10.0, seed=31)
    data, off = synth.make_reads(0, 20, 6000, seed=32)
    got = Placer(Database(flat)).place_packed(data, off)
    oracle, kind = best_oracle(flat, threads=2)
    want = oracle.place_packed(data, off, threads=2)
    assert np.all(want["lwr"] == 0) and np.all(want["n_place"] == 7)
    helpers.assert_same_placements(got, want, what=kind)
    assert np.all(got["lwr"] == 0)


def test_pinned_and_pageable_host_buffers_and_chunking_agree():
    flat = synth.make_db(0, 8, 150, 0.6, 10.0, seed=41)
    data, off = synth.make_reads(0, 5000, 100, seed=42)
    db = Database(flat)
    base = Placer(db).place_packed(data, off)
    # small chunks: many launches, three slots in flight
    p = Placer(db, chunk_reads=257)
    r = p.place_packed(data, off)
    assert p.counters()["n_launches"] == 2 * ((5000 + 256) // 257)  # per chunk: classify, one class of reads
    # pinned in/out
    from epik_b200.placer import PinnedArray
    pd, po = PinnedArray(data.shape, np.uint8), PinnedArray(off.shape, np.uint64)
    pd.array[:] = data
    po.array[:] = off
    out = p.alloc_result(5000, pinned=True)
    r2 = p.place_packed(pd.array, po.array, out=out)
    for k in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant"):
        assert np.array_equal(base[k], r[k]), k
        assert np.array_equal(base[k], r2[k]), k


def test_device_resident_call_matches_host_call():
    import torch
    flat = synth.make_db(0, 8, 150, 0.6, 10.0, seed=41)
    data, off = synth.make_reads(0, 4000, 100, seed=43)
    p = Placer(Database(flat))
    base = p.place_packed(data, off)
    dev = torch.device("cuda:0")
    d_seq = torch.from_numpy(data).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    K = 7
    d_out = {"n_place": torch.zeros(4000, dtype=torch.int32, device=dev),
             "branch": torch.zeros(4000 * K, dtype=torch.int32, device=dev),
             "score": torch.zeros(4000 * K, dtype=torch.float32, device=dev),
             "lwr": torch.zeros(4000 * K, dtype=torch.float64, device=dev),
             "count": torch.zeros(4000 * K, dtype=torch.int32, device=dev)}
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        p.place_device(d_seq, d_off, 4000, d_out, stream=stream)
    stream.synchronize()
    assert np.array_equal(d_out["n_place"].cpu().numpy().view(np.uint32), base["n_place"])
    assert np.array_equal(d_out["branch"].cpu().numpy().view(np.uint32).reshape(-1, K), base["branch"])
    assert np.array_equal(d_out["score"].cpu().numpy().reshape(-1, K), base["score"])
    assert np.array_equal(d_out["lwr"].cpu().numpy().reshape(-1, K), base["lwr"])
    c = p.counters()
    assert c["n_reads"] == 4000 and c["n_launches"] == 4  # classify, three count widths


def test_database_validation_errors():
    flat = helpers.kat1_db()
    bad = helpers.kat1_db()
    bad.entries["branch"][1] = bad.entries["branch"][0]  # same branch twice in one list
    with pytest.raises(capi.EpkError) as e:
        Database(bad)
    assert e.value.status == capi.EPK_E_INVALID
    bad = helpers.kat1_db()
    bad.entries["branch"][0] = 10  # >= n_nodes
    with pytest.raises(capi.EpkError):
        Database(bad)
    bad = helpers.kat1_db()
    bad.keys[1] = bad.keys[0]
    with pytest.raises(capi.EpkError):
        Database(bad)
    with pytest.raises(capi.EpkError):
        Placer(Database(flat), keep_at_most=33)
    with pytest.raises(capi.EpkError):
        Database(flat, mu=1.5)  # check_mu, main.cpp:196-202


@pytest.mark.parametrize("n_leaves", [15001, 25001], ids=["long-read-on-30k-nodes", "50k-nodes"])
def test_accumulators_that_do_not_fit_shared_memory_live_in_hbm(n_leaves):
    """The shared-memory accumulators bound the tree (8 bytes per node with 32-bit counts: ~28 k nodes; 5 bytes with
    8-bit counts: ~44 k); the reference has no such bound (place.cpp:92-96). Beyond it the same kernel keeps a warp's
    accumulators in a row of global memory: a 70 000-letter read on a 30 001-node tree, and every read on a 50 001-node
    tree, are placed with the reference's results - scores bit for bit."""
    flat = synth.make_db(0, 8, n_leaves, 0.3, 4.0, seed=77)
    rng = np.random.default_rng(12)
    seqs = ["".join(rng.choice(list("ACGT"), size=150)) for _ in range(24)]
    seqs.insert(5, "".join(rng.choice(list("ACGT"), size=70000)))
    seqs.insert(9, "".join(rng.choice(list("ACGTN"), size=400)))   # ambiguous windows too
    seqs.append("ACGTACG")                                         # k - 1 letters: the fake placements
    data, off = pack_sequences(seqs)
    p = Placer(Database(flat))
    got = p.place_packed(data, off)
    c = p.counters()
    assert c["n_unsupported"] == 0 and c["n_short"] == 0 and c["n_reads"] == len(seqs)
    oracle, kind = best_oracle(flat, threads=2)
    want = oracle.place_packed(data, off, threads=2)
    helpers.assert_same_placements(got, want, what="accumulators in HBM " + kind)
    valid = np.arange(7)[None, :] < want["n_place"][:, None]
    assert np.array_equal(got["score"][valid], want["score"][valid])
    # several passes, chunks on the placer's three streams (each with its own rows): nothing is left behind in them
    many, moff = pack_sequences(seqs[:5] * 40)
    small = Placer(Database(flat), chunk_reads=16)
    again = small.place_packed(many, moff)
    for name in ("n_place", "branch", "score", "lwr", "count"):
        assert np.array_equal(again[name][:5], got[name][:5]) and np.array_equal(again[name][-5:], got[name][:5]), name


def test_hit_count_widths_and_mixed_lengths():
    """One kernel instantiation per width of the hit counts (uint8 up to 255 windows, uint16 up to
    65535, uint32 beyond): a batch mixing all three must come out as if placed one by one."""
    flat = synth.make_db(0, 8, 120, 0.6, 8.0, seed=51)
    rng = np.random.default_rng(9)
    lens = [7, 8, 100, 262, 263, 264, 300, 5000, 65542, 65543, 70000, 150, 65000]
    seqs = ["".join(rng.choice(list("ACGT"), size=L)) for L in lens]
    seqs += ["".join(rng.choice(list("ACGTN"), size=int(L), p=[0.245, 0.245, 0.245, 0.245, 0.02])) for L in rng.integers(200, 330, size=60)]
    p = Placer(Database(flat))
    got = p.place_packed(*pack_sequences(seqs))
    oracle, kind = best_oracle(flat, threads=4)
    want = oracle.place(seqs, threads=4)
    helpers.assert_same_placements(got, want, what="mixed widths " + kind)
    valid = np.arange(7)[None, :] < want["n_place"][:, None]
    assert np.array_equal(got["score"][valid], want["score"][valid])
    c = p.counters()
    assert c["n_unsupported"] == 0 and c["n_launches"] == 4 and c["n_windows"] == sum(L - 7 for L in lens) + sum(len(s) - 7 for s in seqs[len(lens):])


@pytest.mark.parametrize("tweak", ["positive", "tiny", "zero", "below_threshold"])
def test_databases_outside_the_short_arithmetic_forms(tweak):
    """Scores the fast sweep was not checked for (positive, denormal-range, far below the threshold)
    switch the kernel to IEEE division + library exp10; zeros stay on the fast path."""
    flat = synth.make_db(0, 7, 100, 0.6, 8.0, seed=61)
    e = flat.entries
    if tweak == "positive":
        e["score"][::97] = 0.25
    elif tweak == "tiny":
        e["score"][::89] = -1e-33
    elif tweak == "zero":
        e["score"][::5] = 0.0
    else:
        e["score"][::101] = -250.0
    data, off = synth.make_reads(0, 400, 120, seed=62, amb_rate=0.01)
    got = Placer(Database(flat)).place_packed(data, off)
    oracle, kind = best_oracle(flat, threads=2)
    want = oracle.place_packed(data, off, threads=2)
    helpers.assert_same_placements(got, want, what=tweak + " " + kind)
    valid = np.arange(7)[None, :] < want["n_place"][:, None]
    assert np.array_equal(got["score"][valid], want["score"][valid])


@pytest.mark.parametrize("n_nodes", [1500, 5001, 9001], ids=["16-warps", "8-warps", "range-teams"])
def test_masses_of_exact_ties_overflow_the_candidate_buffer(n_nodes):
    """Hundreds of branches with bit-equal scores at the top: more candidates reach the bar than the candidate buffer
    holds, so the second pass of the sweep has to squeeze; the kept set follows the documented tie rule (lower branch id
    first), the scores / counts / weight ratios are the reference's."""
    rng = np.random.default_rng(n_nodes)
    tied = sorted(int(b) for b in rng.choice(n_nodes, size=700, replace=False))
    better = [tied[5], tied[300], tied[650]]
    kmers = {"ACGT": [(b, -1.0) for b in tied], "CGTA": [(b, -0.5) for b in better], "TTTT": [(tied[10], -2.0)]}
    flat = flatdb.from_dict(flatdb.DNA, 4, 1.5, kmers, [0.01 * (i % 7 + 1) for i in range(n_nodes)], [1] * n_nodes,
                            [0.0] * n_nodes, encode=helpers.encode_dna)
    seqs = [b"ACGT", b"ACGTA", b"GGACGTGG", b"ACGTTTT", b"CGTA", b"GGGGGG"]
    data = np.frombuffer(b"".join(seqs), np.uint8).copy()
    off = np.zeros(len(seqs) + 1, np.uint64)
    off[1:] = np.cumsum([len(x) for x in seqs])
    for K in (7, 32):
        oracle, kind = best_oracle(flat, keep_at_most=K, keep_factor=0.0, threads=1)
        got = Placer(Database(flat), K, 0.0).place_packed(data, off)
        want = oracle.place_packed(data, off, threads=1)
        helpers.assert_same_placements(got, want, what="ties K=%d %s" % (K, kind))
        valid = np.arange(K)[None, :] < want["n_place"][:, None]
        assert np.array_equal(got["score"][valid], want["score"][valid])
        assert list(got["branch"][0]) == tied[:K]                                      # all tied: the K lowest ids
        assert list(got["branch"][1][:3]) == better and list(got["branch"][1][3:]) == [b for b in tied if b not in better][:K - 3]
        assert list(got["branch"][4][:3]) == better


@pytest.mark.parametrize("mu,omega,max_entries", [
    (0.5, 1.5, None),          # --mu 0.5
    (1.0, 1.5, 20000),         # --max-ram
    (0.5, 1.8, 30000),         # both, and a stricter omega (BASELINE config 3 combines them, main.cpp:277)
    (0.3, 1.2, None),          # a looser omega than the database's: nothing dropped by score
    (1.0, 2.5, None),          # omega only
    (0.0, 1.5, None),          # nothing loaded: every query gets the zero-hit placements
])
def test_load_filter_matches_its_restatement(mu, omega, max_entries):
    """--mu / --omega / --max-ram applied on the device at load time (device sort by filter value, prefix cut,
    per-entry threshold, compaction) against the numpy restatement in oracle/load_filter.py."""
    from oracle import load_filter
    flat = synth.make_db(0, 8, 200, 0.6, 10.0, seed=91)
    data, off = synth.make_reads(0, 1500, 120, seed=92, amb_rate=0.003)
    probe = synth.make_db(0, 8, 200, 0.001, 2.0, omega=omega, seed=91)   # only to ask the C++ oracle for the thresholds
    _, log_thr = best_oracle(probe)[0].thresholds()
    want_db = load_filter.apply(flat, mu, omega, max_entries, float(log_thr))
    db = Database(flat, mu=mu, omega=omega, max_entries=max_entries)
    assert db.info["n_entries_total"] == flat.n_entries
    assert db.info["n_entries_loaded"] == want_db.n_entries and db.info["n_keys_loaded"] == want_db.n_keys
    assert db.info["omega"] == np.float32(omega) and db.info["log10_threshold"] == log_thr
    if 0 < mu < 1 and max_entries is None:
        assert abs(want_db.n_entries - mu * flat.n_entries) < 0.2 * flat.n_entries
    got = Placer(db).place_packed(data, off)
    oracle, kind = best_oracle(want_db, threads=2)
    want = oracle.place_packed(data, off, threads=2)
    helpers.assert_same_placements(got, want, what="filter " + kind)
    valid = np.arange(7)[None, :] < want["n_place"][:, None]
    assert np.array_equal(got["score"][valid], want["score"][valid])
    # hits seen by the kernel = hits in the filtered database
    for i in range(3):
        seq = bytes(data[int(off[i]):int(off[i + 1])])
        for x, y in zip(db.kmers(seq), oracle.kmers(seq)):
            assert np.array_equal(x, y)


@pytest.mark.parametrize("n_leaves", [300, 10000], ids=["one-warp", "range-teams"])
def test_gather_ceiling_launch_moves_the_same_postings(n_leaves):
    """epk_gather_ceiling_device (the roofline's measuring stick) walks the same probes and posting lists as the
    placement launch: same windows / hits / entries counters, one checksum word per query, deterministic."""
    import torch
    flat = synth.make_db(0, 8, n_leaves, 0.6, 12.0 if n_leaves == 300 else 40.0, seed=11)
    data, off = synth.make_reads(0, 2000, 150, seed=12)
    p = Placer(Database(flat))
    p.place_packed(data, off)
    full = p.counters()
    dev = torch.device("cuda", 0)
    d_seq = torch.from_numpy(data).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    sums = [torch.zeros(len(off) - 1, dtype=torch.int32, device=dev) for _ in range(2)]
    for s in sums:
        p.gather_ceiling_device(d_seq, d_off, len(off) - 1, s)
    torch.cuda.synchronize()
    c = p.counters()
    assert (c["n_windows"], c["n_hits"], c["n_entries"]) == (full["n_windows"], full["n_hits"], full["n_entries"])
    assert torch.equal(sums[0], sums[1]) and int((sums[0] != 0).sum()) > 1900


@pytest.mark.parametrize("n_leaves,read_len,variant,k,m", [
    (6500, 150, 0, 8, 40.0),    # N = 12 999: three accumulator sets per SM, teams of 2 warps
    (10000, 150, 0, 8, 40.0),   # N = 19 999 (C5's tree): two sets, teams of 4
    (10000, 150, 1, 8, 40.0),   # the same with one warp per query
    (20000, 150, 0, 8, 40.0),   # N = 39 999: one set, a team of 8
    (10000, 400, 0, 8, 40.0),   # 16-bit hit counts: one set, a team of 8
    (10000, 150, 0, 5, 6000.0), # lists of thousands of postings: a member's sub-lists of one round of hits overflow its
                                # 120-descriptor buffer many times over
    (4300, 150, 0, 8, 300.0),   # N = 8599, just above the range-team threshold: four sets, teams of 2, long lists
])
def test_large_trees_sweep_in_teams(n_leaves, read_len, variant, k, m):
    """Several warps per query on large trees (place_kernel.cuh "teams" and "range teams"): same bits as the
    reference, including ambiguous windows and queries without hits."""
    flat = synth.make_db(0, k, n_leaves, 0.6 if k == 8 else 0.9, m, seed=2000 + n_leaves)
    n = 150 if m > 1000 else 1500 if read_len <= 200 else 400
    data, off = synth.make_reads(0, n, read_len, seed=78, amb_rate=0.002)
    p = Placer(Database(flat), variant=variant)
    got = p.place_packed(data, off)
    oracle, kind = best_oracle(flat)
    want = oracle.place_packed(data, off, threads=4)
    helpers.assert_same_placements(got, want, what="%s N=%d" % (kind, 2 * n_leaves - 1))
    valid = np.arange(7)[None, :] < want["n_place"][:, None]
    assert np.array_equal(got["score"][valid], want["score"][valid])
    # a second call reuses the accumulators the teams left behind
    again = p.place_packed(data, off)
    for name in ("n_place", "branch", "score", "lwr", "count"):
        assert np.array_equal(again[name], got[name]), name


def _ragged_reads(alphabet, lengths, seed, spoil=0.01):
    """Random reads of the given lengths with gaps, N / X and two-way ambiguity letters sprinkled in."""
    rng = np.random.default_rng(seed)
    plain = np.frombuffer(b"ACGT" if alphabet == 0 else b"RHKDESTNQCGPAILMFWYV", np.uint8)
    odd = np.frombuffer(b"-NRY" if alphabet == 0 else b"-XBZ", np.uint8)
    total = int(np.sum(lengths))
    data = plain[rng.integers(0, len(plain), total)].copy()
    hit = rng.random(total) < spoil
    data[hit] = odd[rng.integers(0, len(odd), int(hit.sum()))]
    off = np.concatenate([[0], np.cumsum(lengths)]).astype(np.uint64)
    return data, off


@pytest.mark.parametrize("shape", ["dna-direct", "aa-hashed", "dna-range-teams", "dna-k12"])
def test_hot_kernel_kmer_stream_matches_the_oracle(shape):
    """K-mer codes and database hits of the PLACEMENT kernel itself (its per-lane rolling encoder and its probes,
    epk_placer_set_trace), per read, against the oracle's k-mer stream (place.cpp:294-313): number of keys looked
    up, keys found, postings, and a checksum over (window position, code, keys of the window, list length).
    Lengths cover 1..8 windows per lane, the 256-window chunk boundary and multi-chunk reads; gaps and
    ambiguous letters fall on and around lane boundaries."""
    if shape == "aa-hashed":
        alphabet, k, flat = 1, 6, synth.make_db(1, 6, 150, 0.02, 6.0, seed=21, key_space_limit=1 << 22)
    elif shape == "dna-range-teams":
        alphabet, k, flat = 0, 8, synth.make_db(0, 8, 10000, 0.6, 24.0, seed=22)
    elif shape == "dna-k12":
        alphabet, k, flat = 0, 12, synth.make_db(0, 12, 300, 0.05, 8.0, seed=23)
    else:
        alphabet, k, flat = 0, 10, synth.make_db(0, 10, 600, 0.6, 12.0, seed=20)
    windows = [1, 2, 31, 32, 33, 63, 64, 65, 96, 97, 128, 141, 160, 161, 192, 224, 225, 255, 256, 257, 300, 511, 512, 513, 700]
    lengths = np.array([w + k - 1 for w in windows] * 8 + [150] * 200 + [k - 1], np.int64)
    data, off = _ragged_reads(alphabet, lengths, seed=31)
    db = Database(flat)
    p = Placer(db)
    trace, got = p.trace_packed(data, off)
    oracle, kind = best_oracle(flat)
    want = helpers.expected_trace(oracle, data, off)
    bad = np.nonzero((trace != want).any(axis=1))[0]
    assert len(bad) == 0, "reads %s: kernel %s, oracle %s" % (bad[:5], trace[bad[:5]], want[bad[:5]])
    assert want[:, 0].sum() > 0 and want[:, 1].sum() > 0 and (want[:, 0] != np.array(list(lengths - k + 1))).any()
    # and the placements of the same launch are the reference's
    ref = oracle.place_packed(data, off, threads=4)
    K = 7
    shaped = {"n_place": got["n_place"].view(np.uint32), "branch": got["branch"].view(np.uint32).reshape(-1, K),
              "score": got["score"].reshape(-1, K), "lwr": got["lwr"].reshape(-1, K),
              "count": got["count"].view(np.uint32).reshape(-1, K)}
    helpers.assert_same_placements(shaped, ref, what="trace launch " + kind, lengths=False)


@pytest.mark.parametrize("shape", ["dna", "dna-range-teams", "aa-hashed", "dna-filtered"])
def test_database_generated_on_the_device_is_its_host_restatement(shape, monkeypatch):
    """epk_db_create_synthetic (gen2, epik_b200/csrc/synth_gen.cuh) against synth.Gen2 in numpy: same key count, same
    entry count, and bit-identical placements whichever way the database got into HBM - streamed through many small key
    ranges or one, generated or uploaded; and the oracle agrees when it is fed only the lists the reads look up."""
    flt = {}
    if shape == "dna":
        gen, length = synth.Gen2(0, 8, 300, 0.6, 12.0, seed=5), 150
    elif shape == "dna-range-teams":
        gen, length = synth.Gen2(0, 8, 5000, 0.7, 40.0, seed=6), 200
    elif shape == "aa-hashed":
        gen, length = synth.Gen2(1, 6, 150, 0.02, 6.0, seed=7, key_space_limit=1 << 21), 60
    else:
        gen, length, flt = synth.Gen2(0, 8, 300, 0.6, 12.0, seed=8), 150, dict(mu=0.6, omega=1.2)
    data, off = synth.make_reads(gen.alphabet, 3000, length, seed=9)
    monkeypatch.setenv("EPK_BUILD_CHUNK_ENTRIES", "20000")
    device_db = Database.synthetic(gen, **flt)
    flat = gen.whole()
    if flt:
        flat.filter_value = None  # (the device generator has no filter values: file order)
    host_db = Database(flat, **flt)
    monkeypatch.delenv("EPK_BUILD_CHUNK_ENTRIES")
    one_range_db = Database(flat, **flt)
    for name in ("n_keys_loaded", "n_entries_loaded", "n_entries_total", "table_kind"):
        assert device_db.info[name] == host_db.info[name] == one_range_db.info[name], name
    assert device_db.info["n_entries_total"] == flat.n_entries
    got = Placer(device_db).place_packed(data, off)
    for other in (host_db, one_range_db):
        again = Placer(other).place_packed(data, off)
        for name in ("n_place", "branch", "score", "lwr", "count"):
            assert np.array_equal(got[name], again[name]), name
    if not flt:
        # the reference, fed the lists of the keys these reads contain and nothing else
        some = gen.lists(synth.keys_of_reads(gen.alphabet, gen.k, data, off))
        assert some.n_keys < flat.n_keys or shape == "dna"
        oracle, kind = best_oracle(some)
        want = oracle.place_packed(data, off, threads=4)
        helpers.assert_same_placements(got, want, what=shape + " " + kind, lengths=False)
        valid = np.arange(7)[None, :] < want["n_place"][:, None]
        assert np.array_equal(got["score"][valid], want["score"][valid])


def test_config5_proper_generated_on_the_device():
    """BASELINE config 5 itself: k = 12, 19 999 branches, ~3.9 G postings = 31 GB generated and built on the device
    (key table of 128 MB, posting offsets far beyond 2^32 bytes, lists sorted behind range headers), a sample of reads
    placed, and the reference run on the lists those reads look up. Takes a minute and 40 GB of HBM: only with
    EPK_TEST_C5=1 (tools/gpu_c5.sh runs it; log under profiles/)."""
    import os
    if os.environ.get("EPK_TEST_C5") != "1":
        pytest.skip("set EPK_TEST_C5=1 (needs ~40 GB of HBM and a minute)")
    gen = synth.make_config_gen2("C5")
    db = Database.synthetic(gen)
    assert db.info["n_entries_loaded"] == db.info["n_entries_total"] > 3_500_000_000 and db.info["table_slots"] == 4 ** 12
    assert db.info["device_bytes"] > 28 * (1 << 30)
    n = 2000
    data, off = synth.make_reads(0, n, 150, seed=77)
    p = Placer(db)
    got = p.place_packed(data, off)
    c = p.counters()
    some = gen.lists(synth.keys_of_reads(0, 12, data, off))
    # the kernel's own work counters against the restated lists: every window's key, each list as often as it is hit
    codes = np.searchsorted(np.frombuffer(b"ACGT", np.uint8), data).astype(np.uint64).reshape(n, 150)
    keys = np.zeros((n, 139), np.uint64)
    for j in range(12):
        keys = (keys << np.uint64(2)) | codes[:, j:j + 139]
    at = np.searchsorted(some.keys, keys.ravel())
    hit = (at < some.n_keys) & (some.keys[np.minimum(at, some.n_keys - 1)] == keys.ravel())
    lens = np.diff(some.offsets.astype(np.int64))
    assert c["n_hits"] == int(hit.sum()) > 0.85 * n * 139 and c["n_entries"] == int(lens[at[hit]].sum())
    oracle, kind = best_oracle(some, threads=8)
    want = oracle.place_packed(data, off, threads=8)
    helpers.assert_same_placements(got, want, what="C5 " + kind, lengths=False)
    valid = np.arange(7)[None, :] < want["n_place"][:, None]
    assert np.array_equal(got["score"][valid], want["score"][valid])
    # the same database in two shards (database-sharded mode): what the mode is for
    from epik_b200 import dbshard
    del p, db
    sp = dbshard.LocalShardedPlacer(gen, 2)
    assert sum(d.info["n_entries_loaded"] for d in sp.dbs) == sp.dbs[0].info["n_entries_total"]
    got2 = sp.place_packed(data, off, max_rows=1024)
    assert sp.unsupported == 0
    helpers.assert_close_placements(got2, want, what="C5 in two shards")


def test_one_process_several_placers_fill_one_result_in_input_order():
    """epk_place_many: the pass is cut into contiguous slices, one per placer (one per GPU when there are several; two
    placers on one device otherwise), and the caller's arrays come back exactly as epk_place on one device leaves them -
    including the rows a slice leaves unplaced (a too-short read) and the status that reports them."""
    import torch
    flat = synth.make_db(0, 8, 300, 0.6, 12.0, seed=11)
    data, off = synth.make_reads(0, 5001, 150, seed=12, amb_rate=0.01)
    n_dev = min(torch.cuda.device_count(), 4)
    devices = list(range(n_dev)) if n_dev > 1 else [0, 0, 0]
    dbs = {d: Database(flat, device=d) for d in set(devices)}
    placers = [Placer(dbs[d]) for d in devices]
    want = placers[0].place_packed(data, off)
    got = Placer.place_packed_many(placers, data, off)
    for name in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant"):
        assert np.array_equal(got[name], want[name]), name
    seqs = [bytes(data[int(off[i]):int(off[i + 1])]) for i in range(40)]
    seqs[33] = b"ACG"
    d2, o2 = pack_sequences(seqs)
    with pytest.raises(capi.EpkError) as e:
        Placer.place_packed_many(placers, d2, o2)
    assert e.value.status == capi.EPK_E_SHORT_SEQ and "device" in str(e.value)
