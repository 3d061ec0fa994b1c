"""TEST INFRASTRUCTURE ONLY — numpy restatement of the load-time filter of i2l::load(file, mu, omega, max_entries)
(call site epik/src/epik/main.cpp:252-277) on a flat database. PARITY UNPINNED: i2l is absent from the reference
tree; the rule restated here is the one recalled in SURVEY.md Appendix B and implemented on the device by
epik_b200/csrc/db.cu:

  * k-mers are taken in order of decreasing filter value (stable: equal values keep their given order; with no
    filter values the given order is the file order);
  * a k-mer is kept, whole, while the entries taken before it are fewer than min(mu * total, max_entries)
    (mu = 1 means no limit from mu);
  * of a kept k-mer only the entries with score >= log10(score_threshold(user omega, k)) survive; k-mers left
    without entries disappear;
  * afterwards db.omega() is the user's omega.
"""
from __future__ import annotations

import numpy as np


def apply(db, mu: float, omega: float, max_entries: int | None, log10_threshold: float):
    """-> a new flat database of the same class as `db`. log10_threshold: float32 log10 of the threshold for
    (omega, k), taken from the CPU oracle (Oracle(...).thresholds()) so that it is the C++ float arithmetic."""
    n_keys = len(db.keys)
    lengths = (db.offsets[1:] - db.offsets[:-1]).astype(np.int64)
    total = int(db.offsets[-1])
    budget = (1 << 64) - 1
    if mu < 1.0:
        budget = int(mu * float(total))
    if max_entries is not None:
        budget = min(budget, int(max_entries))
    order = np.arange(n_keys)
    if budget < total and db.filter_value is not None:
        order = np.argsort(-db.filter_value.astype(np.float32), kind="stable")
    before = np.zeros(n_keys, np.int64)
    if n_keys:
        before[1:] = np.cumsum(lengths[order])[:-1]
    keep_key = np.zeros(n_keys, bool)
    keep_key[order] = before < budget
    owner = np.repeat(np.arange(n_keys), lengths)
    keep_entry = keep_key[owner] & (db.entries["score"] >= np.float32(log10_threshold))
    new_len = np.bincount(owner[keep_entry], minlength=n_keys)
    alive = new_len > 0
    offsets = np.zeros(int(alive.sum()) + 1, np.uint64)
    np.cumsum(new_len[alive], out=offsets[1:])
    out = type(db)(db.alphabet, db.kmer_size, float(omega), db.keys[alive].copy(), offsets, db.entries[keep_entry].copy(),
                   db.branch_length, db.subtree_num_nodes, db.subtree_total_length,
                   None if db.filter_value is None else db.filter_value[alive].copy())
    out.meta = dict(getattr(db, "meta", {}), n_entries_total=total)
    return out
