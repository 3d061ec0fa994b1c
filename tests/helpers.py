"""Shared pieces of the test-suite: the KAT-1 toy database, golden fixtures,
and the comparison rules (bit-exact integers, 1e-5 relative scores, tie-aware placement sets)."""
from __future__ import annotations

import json
import os

import numpy as np

from epik_b200 import flatdb, synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DNA_CODE = {"A": 0, "C": 1, "G": 2, "T": 3}


def encode_dna(kmer: str) -> int:
    key = 0
    for ch in kmer:
        key = (key << 2) | DNA_CODE[ch]
    return key


def kat1_db() -> flatdb.FlatDB:
    """The toy database of KAT-1: k = 3, omega = 1.5, 10 nodes."""
    kmers = {
        "ACG": [(2, -0.5), (3, -1.0)],
        "CGT": [(2, -0.25), (5, -2.0)],
        "AAA": [(7, -0.1)],
        "GAA": [(7, -0.2), (8, -0.3)],
    }
    n = 10
    return flatdb.from_dict(flatdb.DNA, 3, 1.5, kmers, [0.1 * (i + 1) for i in range(n)], [i + 1 for i in range(n)],
                            [0.5 * i for i in range(n)], encode=encode_dna)


KAT1_SEQS = ["ACGT", "TTTTT", "RAA", "ACGT", "AC"]


def load_json(name: str):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def golden_case_db(case: dict) -> flatdb.FlatDB:
    """Rebuilds the synthetic database a golden case was generated from (pure function of the seed)."""
    return synth.make_db(case["alphabet"], case["k"], case["n_leaves"], case["p_occ"], case["m"],
                         omega=case["omega"], seed=case["seed"])


def assert_same_placements(got: dict, want: dict, score_rtol: float = 1e-5, lwr_rtol: float = 1e-9,
                           lengths: bool = True, what: str = ""):
    """Parity rule of BASELINE.json: n_place and counts identical, branches identical except inside
    groups of exactly tied scores (documented tie rule: lower branch id first on the GPU, libstdc++
    partial_sort order in the reference), scores within 1e-5 relative, LWR within lwr_rtol."""
    n_place = np.asarray(want["n_place"])
    assert np.array_equal(np.asarray(got["n_place"]), n_place), what + " n_place"
    K = want["branch"].shape[1]
    valid = np.arange(K)[None, :] < n_place[:, None]
    gs, ws = np.asarray(got["score"]), np.asarray(want["score"])
    np.testing.assert_allclose(gs[valid], ws[valid], rtol=score_rtol, atol=0, err_msg=what + " score")
    np.testing.assert_allclose(np.asarray(got["lwr"])[valid], np.asarray(want["lwr"])[valid], rtol=lwr_rtol, atol=0,
                               err_msg=what + " lwr")
    gb, wb = np.asarray(got["branch"]), np.asarray(want["branch"])
    same = (gb == wb) | ~valid
    for i in np.nonzero(~same.all(axis=1))[0]:
        # rows that differ may only differ by a permutation inside runs of equal scores
        m = int(n_place[i])
        assert np.array_equal(ws[i, :m], gs[i, :m]) or np.allclose(ws[i, :m], gs[i, :m], rtol=score_rtol), what
        j = 0
        while j < m:
            e = j
            while e + 1 < m and ws[i, e + 1] == ws[i, j]:
                e += 1
            if e + 1 == m and m == K:
                # the run touches the cut: members outside the kept set may differ
                assert len(set(gb[i, j:e + 1])) == e + 1 - j, what
            else:
                assert sorted(gb[i, j:e + 1]) == sorted(wb[i, j:e + 1]), "%s row %d: branches differ beyond ties" % (what, i)
            j = e + 1
    cnt_ok = (np.asarray(got["count"]).astype(np.uint64) == np.asarray(want["count"]).astype(np.uint64)) | ~valid | ~(gb == wb)
    assert cnt_ok.all(), what + " count"
    if lengths and "distal" in got and "distal" in want:
        ok = (gb == wb) & valid
        np.testing.assert_allclose(np.asarray(got["distal"])[ok], np.asarray(want["distal"])[ok], rtol=1e-15, err_msg=what)
        np.testing.assert_allclose(np.asarray(got["pendant"])[ok], np.asarray(want["pendant"])[ok], rtol=1e-15, err_msg=what)


def assert_close_placements(got: dict, want: dict, score_rtol: float = 1e-5, lwr_rtol: float = 2e-4, what: str = "",
                            max_reordered: float = 0.002):
    """Parity rule for paths whose per-branch float sums are associated differently from the reference's
    (database-sharded mode): scores within 1e-5 relative position by position, weight ratios within
    lwr_rtol (10^s moves by 2.3 * |s| * 1e-5 relative when s moves by 1e-5 relative; the cases here have
    |s| < 10 per 1e-5), and the kept sets identical except where two candidates are within score_rtol of each
    other (a near tie may fall the other way); at most `max_reordered` of the rows may show such a swap."""
    gn, wn = np.asarray(got["n_place"]).astype(np.int64), np.asarray(want["n_place"]).astype(np.int64)
    K = want["branch"].shape[1]
    gs, ws = np.asarray(got["score"]), np.asarray(want["score"])
    gb, wb = np.asarray(got["branch"]), np.asarray(want["branch"])
    gl, wl = np.asarray(got["lwr"]), np.asarray(want["lwr"])
    swapped = 0
    for i in range(len(wn)):
        m = int(min(gn[i], wn[i]))
        np.testing.assert_allclose(gs[i, :m], ws[i, :m], rtol=score_rtol, atol=0, err_msg="%s row %d score" % (what, i))
        exact = gn[i] == wn[i] and np.array_equal(gb[i, :m], wb[i, :m])
        if exact:
            scale = np.maximum(1.0, np.abs(ws[i, :m]).astype(np.float64) * 2.4)
            assert np.all(np.abs(gl[i, :m] - wl[i, :m]) <= lwr_rtol * scale * np.abs(wl[i, :m]) + 1e-300), "%s row %d lwr" % (what, i)
            cnt_ok = np.asarray(got["count"])[i, :m].astype(np.int64) == np.asarray(want["count"])[i, :m].astype(np.int64)
            assert cnt_ok.all(), "%s row %d count" % (what, i)
            continue
        swapped += 1
        # a differing row: every position that differs must sit in a near tie, and a differing row count
        # must come from a weight ratio sitting on the keep_factor cut
        assert abs(int(gn[i]) - int(wn[i])) <= 1, "%s row %d n_place" % (what, i)
        for j in np.nonzero(gb[i, :m] != wb[i, :m])[0]:
            near = [jj for jj in (j - 1, j + 1) if 0 <= jj < wn[i] and abs(ws[i, jj] - ws[i, j]) <= score_rtol * abs(ws[i, j])]
            assert near or j == K - 1, "%s row %d: branches differ beyond near ties" % (what, i)
    assert swapped <= max(1, int(max_reordered * len(wn))), "%s: %d rows reordered" % (what, swapped)


def trace_terms(pos, key, group, length):
    """epk_trace_term (place_kernel.cuh::trace_term) over numpy arrays, modulo 2^32."""
    u = lambda a: np.asarray(a).astype(np.uint64)
    M = np.uint64(0xFFFFFFFF)
    h = ((u(pos) * np.uint64(0x9E3779B1)) ^ (u(key) * np.uint64(0x85EBCA6B)) ^ (u(group) * np.uint64(0xC2B2AE35))
         ^ (u(length) * np.uint64(0x27D4EB2F))) & M
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & M
    h ^= h >> np.uint64(12)
    h = (h * np.uint64(0x297A2D39)) & M
    return (h ^ (h >> np.uint64(15))).astype(np.uint32)


def expected_trace(oracle, data, off):
    """Per query {keys looked up, keys found, postings, checksum} from the oracle's k-mer stream (orc_kmers)."""
    n = len(off) - 1
    out = np.zeros((n, 4), np.uint32)
    for i in range(n):
        seq = bytes(data[int(off[i]):int(off[i + 1])])
        pos, key, group, hit = oracle.kmers(seq)
        found = np.maximum(np.asarray(hit, np.int64), 0)
        out[i] = (len(pos), int((np.asarray(hit) >= 0).sum()), int(found.sum() & 0xFFFFFFFF),
                  int(trace_terms(pos, key, group, found).sum(dtype=np.uint64) & np.uint64(0xFFFFFFFF)))
    return out
