"""Host-side description of a phylo-k-mer database as flat arrays.

This is the shape the C ABI takes (`epk_db_desc`, include/epik_b200.h): what the reference holds as
`i2l::phylo_kmer_db` — a hash map key -> std::vector<pkdb_value{branch, score}> plus the tree index
(call sites epik/src/epik/place.cpp:87,113-114,300; epik/src/epik/main.cpp:277-294) — flattened to
CSR form: `keys[i]` owns `entries[offsets[i]:offsets[i+1]]`.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

DNA, AA = 0, 1
ENTRY_DTYPE = np.dtype([("branch", np.uint32), ("score", np.float32)])


@dataclass
class FlatDB:
    alphabet: int                      # DNA (2 bits/letter) or AA (5 bits/letter)
    kmer_size: int
    omega: float
    keys: np.ndarray                   # uint32 [n_keys], unique k-mer codes
    offsets: np.ndarray                # uint64 [n_keys + 1]
    entries: np.ndarray                # ENTRY_DTYPE [offsets[-1]]
    branch_length: np.ndarray          # float64 [n_nodes], by post-order id
    subtree_num_nodes: np.ndarray      # uint64 [n_nodes]   (tree_index, place.cpp:113)
    subtree_total_length: np.ndarray   # float64 [n_nodes]  (tree_index, place.cpp:114)
    filter_value: np.ndarray | None = None   # float32 [n_keys]; larger = more informative
    meta: dict = field(default_factory=dict)

    @property
    def n_keys(self) -> int:
        return int(len(self.keys))

    @property
    def n_entries(self) -> int:
        return int(self.offsets[-1])

    @property
    def n_nodes(self) -> int:
        return int(len(self.branch_length))

    @property
    def alphabet_size(self) -> int:
        return 20 if self.alphabet == AA else 4

    @property
    def bits_per_letter(self) -> int:
        return 5 if self.alphabet == AA else 2

    def validate(self) -> None:
        assert self.keys.dtype == np.uint32 and self.offsets.dtype == np.uint64
        assert self.entries.dtype == ENTRY_DTYPE
        assert len(self.offsets) == len(self.keys) + 1 and self.offsets[0] == 0
        assert len(self.entries) == int(self.offsets[-1])
        assert len(self.subtree_num_nodes) == self.n_nodes == len(self.subtree_total_length)
        assert self.kmer_size * self.bits_per_letter <= 32


def from_dict(alphabet, kmer_size, omega, kmers: dict, branch_length, subtree_num_nodes, subtree_total_length,
              encode=None) -> FlatDB:
    """Small-case helper: {key or k-mer string: [(branch, score), ...]} -> FlatDB (insertion order kept)."""
    keys, offs, ents = [], [0], []
    for kmer, plist in kmers.items():
        keys.append(encode(kmer) if isinstance(kmer, str) else int(kmer))
        ents.extend(plist)
        offs.append(len(ents))
    e = np.zeros(len(ents), ENTRY_DTYPE)
    if ents:
        e["branch"] = [b for b, _ in ents]
        e["score"] = [s for _, s in ents]
    return FlatDB(alphabet, kmer_size, omega, np.asarray(keys, np.uint32), np.asarray(offs, np.uint64), e,
                  np.asarray(branch_length, np.float64), np.asarray(subtree_num_nodes, np.uint64),
                  np.asarray(subtree_total_length, np.float64))


# ------------------------------------------------------------------------- .epkdb container -----
# The on-disk form the C++ host layer reads (epik_b200/host/include/epik/flat_db.h, load_epkdb). The
# reference's `.ipk` format is defined only inside the absent i2l submodule.
EPKDB_MAGIC = b"EPKDB1\0\0"


def _pad8(n: int) -> bytes:
    return b"\0" * ((8 - n % 8) % 8)


def save_epkdb(db: FlatDB, path: str, newick: str = "") -> None:
    import struct
    db.validate()
    has_filter = db.filter_value is not None
    nw = newick.encode()
    with open(path, "wb") as f:
        f.write(EPKDB_MAGIC)
        f.write(struct.pack("<IIfIQQQQ", int(db.alphabet), int(db.kmer_size), float(db.omega), db.n_nodes, db.n_keys,
                            db.n_entries, int(has_filter), len(nw)))
        keys = np.ascontiguousarray(db.keys, np.uint32).tobytes()
        f.write(keys + _pad8(len(keys)))
        f.write(np.ascontiguousarray(db.offsets, np.uint64).tobytes())
        f.write(np.ascontiguousarray(db.entries, ENTRY_DTYPE).tobytes())
        if has_filter:
            fv = np.ascontiguousarray(db.filter_value, np.float32).tobytes()
            f.write(fv + _pad8(len(fv)))
        f.write(np.ascontiguousarray(db.branch_length, np.float64).tobytes())
        f.write(np.ascontiguousarray(db.subtree_num_nodes, np.uint64).tobytes())
        f.write(np.ascontiguousarray(db.subtree_total_length, np.float64).tobytes())
        f.write(nw)


def load_epkdb(path: str):
    """-> (FlatDB, newick string)."""
    import struct
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:8] != EPKDB_MAGIC:
        raise ValueError("not an .epkdb file: %s" % path)
    alphabet, k, omega, n_nodes, n_keys, n_entries, has_filter, nw_len = struct.unpack_from("<IIfIQQQQ", buf, 8)
    at = 8 + struct.calcsize("<IIfIQQQQ")

    def take(dtype, count, pad=False):
        nonlocal at
        nbytes = np.dtype(dtype).itemsize * count
        a = np.frombuffer(buf, dtype, count, at).copy()
        at += nbytes + (len(_pad8(nbytes)) if pad else 0)
        return a

    keys = take(np.uint32, n_keys, True)
    offsets = take(np.uint64, n_keys + 1)
    entries = take(ENTRY_DTYPE, n_entries)
    fv = take(np.float32, n_keys, True) if has_filter else None
    bl = take(np.float64, n_nodes)
    sn = take(np.uint64, n_nodes)
    st = take(np.float64, n_nodes)
    newick = buf[at:at + nw_len].decode()
    return FlatDB(alphabet, k, omega, keys, offsets, entries, bl, sn, st, fv), newick
