"""Generates the golden vectors under tests/golden/ by running the REFERENCE's own placer
(oracle/_ref: /root/reference/epik/src/epik/place.cpp compiled unmodified over the i2l shim).
Needs /root/reference (run `make -C oracle ref` first); the outputs are committed so that the GPU
box, which has no /root/reference, can check against them.

    python tests/golden/make_golden.py

PARITY NOTE: these vectors pin accumulate / threshold default / top-k / LWR (the reference's code).
The k-mer codes and hits inside them come from the i2l *shim* (upstream i2l is absent): unpinned.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from epik_b200 import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402
from tests import helpers  # noqa: E402

CASES = [
    # name, alphabet, k, n_leaves, p_occ, m, omega, seed, n_reads, read_len, amb_rate, keep_at_most, keep_factor
    dict(name="dna_k6", alphabet=0, k=6, n_leaves=40, p_occ=0.5, m=6.0, omega=1.5, seed=101, n_reads=300, read_len=60,
         amb_rate=0.0, keep_at_most=7, keep_factor=0.01),
    dict(name="dna_k8_amb", alphabet=0, k=8, n_leaves=150, p_occ=0.6, m=10.0, omega=1.5, seed=102, n_reads=300,
         read_len=120, amb_rate=0.02, keep_at_most=7, keep_factor=0.01),
    dict(name="dna_k10_long", alphabet=0, k=10, n_leaves=300, p_occ=0.02, m=16.0, omega=1.25, seed=103, n_reads=40,
         read_len=1500, amb_rate=0.002, keep_at_most=5, keep_factor=0.05),
    dict(name="aa_k3_amb", alphabet=1, k=3, n_leaves=60, p_occ=0.4, m=8.0, omega=1.5, seed=104, n_reads=300,
         read_len=80, amb_rate=0.02, keep_at_most=7, keep_factor=0.01),
    dict(name="aa_k6", alphabet=1, k=6, n_leaves=100, p_occ=0.0005, m=8.0, omega=1.5, seed=105, n_reads=200,
         read_len=100, amb_rate=0.0, keep_at_most=3, keep_factor=0.01),
    dict(name="dna_k4_dense_k16", alphabet=0, k=4, n_leaves=30, p_occ=0.9, m=10.0, omega=1.0, seed=106, n_reads=200,
         read_len=50, amb_rate=0.05, keep_at_most=16, keep_factor=0.0),
]


def main():
    arrays, meta = {}, []
    # KAT-1
    o = O.Oracle(helpers.kat1_db(), 7, 0.01, 1, kind="reference")
    r = o.place(helpers.KAT1_SEQS[:4])  # "AC" (len k-1 = 2) goes separately: see below
    r2 = o.place(["AC"])
    for k in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant"):
        arrays["kat1/" + k] = np.concatenate([r[k], r2[k]])
    for case in CASES:
        db = helpers.golden_case_db(case)
        data, off = synth.make_reads(case["alphabet"], case["n_reads"], case["read_len"], seed=case["seed"] + 1000,
                                     amb_rate=case["amb_rate"])
        o = O.Oracle(db, case["keep_at_most"], case["keep_factor"], 1, kind="reference")
        r = o.place_packed(data, off, threads=1, batch_size=2000)
        for k in ("n_place", "branch", "score", "lwr", "count", "distal", "pendant", "group_rep"):
            arrays[case["name"] + "/" + k] = r[k]
        # the k-mer stream of the first 8 reads (codes + hits: bit-exact targets)
        streams = []
        for i in range(8):
            seq = bytes(data[int(off[i]):int(off[i + 1])])
            pos, key, group, hit = o.kmers(seq)
            streams.append(np.stack([pos.astype(np.int64), key.astype(np.int64), group.astype(np.int64), hit]))
        arrays[case["name"] + "/kmers"] = np.concatenate(streams, axis=1)
        arrays[case["name"] + "/kmers_len"] = np.asarray([s.shape[1] for s in streams], np.int64)
        c = dict(case)
        c.update(n_keys=db.n_keys, n_entries=db.n_entries, n_nodes=db.n_nodes,
                 thresholds=[float(x) for x in o.thresholds()])
        meta.append(c)
        print(case["name"], db.n_keys, db.n_entries, db.n_nodes, "amb reads:",
              int(sum(1 for i in range(case["n_reads"]) if set(bytes(data[int(off[i]):int(off[i + 1])])) - set(b"ACGT" if case["alphabet"] == 0 else b"RHKDESTNQCGPAILMFWYV"))))
    np.savez_compressed(os.path.join(HERE, "reference_outputs.npz"), **arrays)
    with open(os.path.join(HERE, "cases.json"), "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
