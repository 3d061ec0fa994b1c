"""CPU tests of the checker itself: the oracle (reference place.cpp over the i2l shim, and the plain
port) against the known-answer vectors of SURVEY.md Appendix D (KAT-1) and the committed golden
outputs of the reference (tests/golden/reference_outputs.npz, made by tests/golden/make_golden.py)."""
import numpy as np
import pytest

from epik_b200 import synth
from oracle import oracle as O
from tests import helpers


def kinds(alphabet=0):
    out = ["port"]
    if O.available("reference", alphabet):
        out.append("reference")
    return out


@pytest.fixture(scope="module", autouse=True)
def _build(built):
    return built


@pytest.mark.parametrize("kind", ["port", "reference"])
def test_kat1_known_answers(kind):
    """Values verified by hand against place.cpp's formulas in SURVEY.md Appendix D."""
    if not O.available(kind, 0):
        pytest.skip("oracle/_ref not built here")
    o = O.Oracle(helpers.kat1_db(), 7, 0.01, 1, kind=kind)
    thr, logthr = o.thresholds()
    assert thr == np.float32(0.052734375)
    assert abs(float(logthr) - (-1.27790618)) < 1e-7
    r = o.place(helpers.KAT1_SEQS[:4])
    # exact: ACGT (records 0 and 3 are the same sequence)
    for i in (0, 3):
        assert r["n_place"][i] == 3
        assert list(r["branch"][i, :3]) == [2, 3, 5]
        np.testing.assert_allclose(r["score"][i, :3], [-0.25, -0.75930208, -1.09263539], rtol=1e-7)
        np.testing.assert_allclose(r["lwr"][i, :3], [0.31214008, 0.0966156451, 0.044845012], rtol=1e-7)
        assert list(r["count"][i, :3]) == [2, 1, 1]
        np.testing.assert_allclose(r["distal"][i, 0], 0.15)
        np.testing.assert_allclose(r["pendant"][i, 0], 0.483333, rtol=1e-5)
    assert r["group_rep"][3] == r["group_rep"][0]
    # no hit: TTTTT -> 7 tied placements on branches 0..6, in std::partial_sort's order
    assert r["n_place"][1] == 7
    assert list(r["branch"][1]) == [3, 4, 1, 0, 5, 2, 6]
    np.testing.assert_allclose(r["score"][1], -1.27790618, rtol=1e-7)
    np.testing.assert_allclose(r["lwr"][1], 0.1, rtol=1e-12)
    assert not r["count"][1].any() and not r["distal"][1].any() and not r["pendant"][1].any()
    # ambiguous: RAA -> a probability added to a log score (Appendix C #4-#5)
    assert r["n_place"][2] == 2
    assert list(r["branch"][2, :2]) == [7, 8]
    np.testing.assert_allclose(r["score"][2, :2], [0.0999774411, 0.0674062148], rtol=1e-6)
    np.testing.assert_allclose(r["lwr"][2, :2], [0.231972609, 0.215211503], rtol=1e-7)
    np.testing.assert_allclose(r["distal"][2, 0], 0.4)
    np.testing.assert_allclose(r["pendant"][2, 0], 0.8375)
    # len == k - 1: zero windows, 7 tied placements with score -0
    r = o.place(["AC"])
    assert r["n_place"][0] == 7 and np.all(r["score"][0] == 0) and np.all(np.signbit(r["score"][0]))
    np.testing.assert_allclose(r["lwr"][0], 0.1, rtol=1e-12)


def test_golden_kat1_matches_port():
    g = np.load(helpers.GOLDEN + "/reference_outputs.npz")
    o = O.Oracle(helpers.kat1_db(), 7, 0.01, 1, kind="port")
    r = o.place(helpers.KAT1_SEQS)
    for k in ("n_place", "branch", "score", "lwr", "distal", "pendant"):
        assert np.array_equal(r[k], g["kat1/" + k]), k
    assert np.array_equal(r["count"], g["kat1/count"])


@pytest.mark.parametrize("case", helpers.load_json("cases.json"), ids=lambda c: c["name"])
def test_golden_cases(case):
    """Both oracles reproduce the committed outputs of the reference bit for bit (same compiler family)."""
    g = np.load(helpers.GOLDEN + "/reference_outputs.npz")
    db = helpers.golden_case_db(case)
    assert (db.n_keys, db.n_entries, db.n_nodes) == (case["n_keys"], case["n_entries"], case["n_nodes"])
    data, off = synth.make_reads(case["alphabet"], case["n_reads"], case["read_len"], seed=case["seed"] + 1000,
                                 amb_rate=case["amb_rate"])
    for kind in kinds(case["alphabet"]):
        o = O.Oracle(db, case["keep_at_most"], case["keep_factor"], 2, kind=kind)
        r = o.place_packed(data, off, threads=2)
        name = case["name"]
        assert np.array_equal(r["n_place"], g[name + "/n_place"]), kind
        want = {k: g[name + "/" + k] for k in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant")}
        if kind == "reference":
            for k in ("branch", "score", "lwr", "distal", "pendant"):
                assert np.array_equal(r[k], want[k]), (kind, k)
        else:
            # the port's likelihood mass is summed in first-touch order as well, so it is exact too;
            # fall back to the tie-aware rule only for branch order inside exact ties
            helpers.assert_same_placements(r, want, score_rtol=0, lwr_rtol=1e-15, what=kind)
        # k-mer streams of the first 8 reads
        at = 0
        for i, n in enumerate(g[name + "/kmers_len"]):
            seq = bytes(data[int(off[i]):int(off[i + 1])])
            pos, key, group, hit = o.kmers(seq)
            want_k = g[name + "/kmers"][:, at:at + n]
            at += n
            assert np.array_equal(np.stack([pos, key, group, hit]).astype(np.int64), want_k), (kind, i)


def test_port_matches_reference_on_random_inputs():
    """Wider randomised agreement of the two oracles, incl. gaps, ambiguity codes, duplicates."""
    if not O.available("reference", 0):
        pytest.skip("oracle/_ref not built here")
    for alphabet, k, seed in ((0, 5, 1), (0, 7, 2), (1, 2, 3), (1, 3, 4)):
        db = synth.make_db(alphabet, k, 25, 0.6, 5.0, seed=seed)
        data, off = synth.make_reads(alphabet, 400, 37, seed=seed + 50, amb_rate=0.06)
        # duplicates inside a batch
        data = np.concatenate([data, data[:37 * 10]])
        off = np.arange(411, dtype=np.uint64) * np.uint64(37)
        ref = O.Oracle(db, 7, 0.01, 1, kind="reference").place_packed(data, off, batch_size=100)
        port = O.Oracle(db, 7, 0.01, 1, kind="port").place_packed(data, off, batch_size=100)
        helpers.assert_same_placements(port, ref, score_rtol=0, lwr_rtol=1e-15, what="alphabet %d k %d" % (alphabet, k))
        assert np.array_equal(port["group_rep"], ref["group_rep"])


def test_short_sequences_kill_the_reference():
    """len <= k - 2 wraps size_t at place.cpp:291 (vector::reserve throws inside the OpenMP region):
    the port refuses them the same way the C ABI does (EPK_E_SHORT_SEQ) instead of aborting."""
    # documented behaviour only; running it against _ref would abort the test process
    o = O.Oracle(helpers.kat1_db(), 7, 0.01, 1, kind="port")
    r = o.place(["AC"])  # len == k - 1 is fine
    assert r["n_place"][0] == 7
