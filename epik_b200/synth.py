"""Seeded synthetic trees, phylo-k-mer databases and reads of the shapes BASELINE.json names.

Everything is a pure function of (seed, index) through a splitmix64-style integer mix, so the same
workload can be rebuilt anywhere (here, on the GPU box, by the CPU baseline) without shipping data.
Shapes follow SURVEY.md §8(d): a database is `p_occ` of the key space, posting-list lengths are
log-normal around `m`, the branches of a list are distinct and clustered on neighbouring post-order
ids, scores are uniform in [log10(threshold), -0.3).

The reference has no generator (it ships no fixtures at all); this module only feeds
the C ABI and the oracle with identical inputs.
"""
from __future__ import annotations

import numpy as np

from .flatdb import AA, DNA, ENTRY_DTYPE, FlatDB

SEED_TREE, SEED_DB, SEED_READS = 0xE91C0000, 0xE91C0001, 0xE91C0002
_U64 = np.uint64
_AA_ORDER = b"RHKDESTNQCGPAILMFWYV"
_DNA_ORDER = b"ACGT"


def mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on uint64 arrays (wrapping arithmetic)."""
    with np.errstate(over="ignore"):
        x = (x + _U64(0x9E3779B97F4A7C15)).astype(_U64)
        x = (x ^ (x >> _U64(30))) * _U64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> _U64(27))) * _U64(0x94D049BB133111EB)
        return x ^ (x >> _U64(31))


def _stream(seed: int, tag: int, idx: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        base = mix64(np.asarray([(seed << 8) ^ tag], dtype=_U64))[0]
        return mix64(idx.astype(_U64) * _U64(0xD6E8FEB86659FD93) + base)


def _unit(h: np.ndarray) -> np.ndarray:
    """uint64 hash -> float64 in (0, 1)."""
    return ((h >> _U64(11)).astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)


# ------------------------------------------------------------------------------------ tree -----
def make_tree(n_leaves: int, seed: int = SEED_TREE, mean_length: float = 0.05):
    """Random rooted binary tree numbered in post-order. Returns per-node arrays
    (branch_length, subtree_num_nodes, subtree_total_length, parent) with N = 2*n_leaves - 1.

    subtree_num_nodes counts the node itself; subtree_total_length sums the branch lengths strictly
    below the node — the two quantities place.cpp:113-120 divides to get the mean subtree branch
    length (what upstream puts in them is [UNVERIFIED], they are opaque inputs to the placer)."""
    n = 2 * n_leaves - 1
    u = _unit(_stream(seed, 1, np.arange(2 * n + 4)))
    length = -mean_length * np.log(_unit(_stream(seed, 2, np.arange(n))))
    num = np.ones(n, np.uint64)
    tot = np.zeros(n, np.float64)
    parent = np.full(n, -1, np.int64)
    stack, leaves_left, nid, step = [], n_leaves, 0, 0
    while nid < n:
        can_join = len(stack) >= 2
        join = can_join and (leaves_left == 0 or u[step] < 0.5)
        step += 1
        if join:
            r, l = stack.pop(), stack.pop()
            num[nid] = 1 + num[l] + num[r]
            tot[nid] = tot[l] + tot[r] + length[l] + length[r]
            parent[l] = parent[r] = nid
        else:
            leaves_left -= 1
        stack.append(nid)
        nid += 1
    assert len(stack) == 1
    return length, num, tot, parent


def tree_newick(length: np.ndarray, parent: np.ndarray) -> str:
    """The tree of make_tree as a jplace tree string: `(child,child)name:length{edge}`, edge numbers =
    post-order ids (what the reference obtains from i2l::io::to_newick(tree, true), main.cpp:297; the exact
    upstream spelling is [UNVERIFIED], the jplace standard only asks for `{edge_num}` after the length)."""
    n = len(length)
    kids = [[] for _ in range(n)]
    for i in range(n):
        if parent[i] >= 0:
            kids[int(parent[i])].append(i)
    text = [""] * n
    for i in range(n):  # post-order: children come before their parent
        inner = "(" + ",".join(text[c] for c in kids[i]) + ")" if kids[i] else "L%d" % i
        text[i] = "%s:%.6g{%d}" % (inner, float(length[i]), i)
        for c in kids[i]:
            text[c] = ""
    return text[n - 1] + ";"


# -------------------------------------------------------------------------------- database -----
def make_db(alphabet: int, k: int, n_leaves: int, p_occ: float, m: float, omega: float = 1.5,
            seed: int = SEED_DB, max_len: int | None = None, key_space_limit: int | None = None) -> FlatDB:
    length, num, tot, parent = make_tree(n_leaves, seed ^ 0x7EE)
    n_nodes = len(length)
    if alphabet == DNA:
        n_space = 4 ** k
        cand = np.arange(n_space, dtype=np.uint64)
    else:
        n_space = 20 ** k
        cand = np.arange(n_space, dtype=np.uint64)
    if key_space_limit is not None:
        cand = cand[:key_space_limit]
    present = _unit(_stream(seed, 3, cand)) < p_occ
    cand = cand[present]
    if alphabet == DNA:
        keys = cand.astype(np.uint32)
    else:
        idx, key = cand.copy(), np.zeros(len(cand), np.uint64)
        for i in range(k):
            key |= (idx % _U64(20)) << _U64(5 * i)
            idx //= _U64(20)
        keys = key.astype(np.uint32)
    kk = keys.astype(np.uint64)

    # list length: log-normal(ln m, 1) via Box-Muller, clamped to [1, max_len]
    u1, u2 = _unit(_stream(seed, 4, kk)), _unit(_stream(seed, 5, kk))
    z = np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    cap = max(1, min(n_nodes // 8, max_len if max_len else n_nodes))
    ell = np.clip(np.rint(np.exp(np.log(m) + z)), 1, cap).astype(np.uint64)
    offsets = np.zeros(len(keys) + 1, np.uint64)
    np.cumsum(ell, out=offsets[1:])
    n_entries = int(offsets[-1])

    # branches: start + running sum of strides in 1..7, modulo N  (distinct because span < N)
    owner = np.repeat(np.arange(len(keys), dtype=np.uint64), ell.astype(np.int64))
    j = np.arange(n_entries, dtype=np.uint64) - offsets[:-1][owner]
    eid = (kk[owner] << _U64(24)) ^ j
    stride = (_stream(seed, 6, eid) % _U64(7)) + _U64(1)
    run = np.cumsum(stride, dtype=np.uint64)
    first = offsets[:-1].astype(np.int64)
    run -= np.repeat(run[first] - stride[first], ell.astype(np.int64))   # restart the sum at every list
    start = _stream(seed, 7, kk) % _U64(n_nodes)
    branch = ((start[owner] + run) % _U64(n_nodes)).astype(np.uint32)

    thr = np.float32(omega / (20.0 if alphabet == AA else 4.0)) ** np.float32(k)
    lo = float(np.log10(thr))
    score = (lo + (-0.3 - lo) * _unit(_stream(seed, 8, eid))).astype(np.float32)
    score = np.minimum(score, np.float32(-0.3))

    entries = np.empty(n_entries, ENTRY_DTYPE)
    entries["branch"], entries["score"] = branch, score
    fv = _unit(_stream(seed, 9, kk)).astype(np.float32)
    db = FlatDB(alphabet, k, omega, keys, offsets, entries, length, num, tot, fv,
                meta=dict(p_occ=p_occ, m=m, n_leaves=n_leaves, seed=seed, n_space=n_space, parent=parent))
    return db


def make_reads(alphabet: int, n: int, length: int, seed: int = SEED_READS, first: int = 0,
               amb_rate: float = 0.0) -> tuple[np.ndarray, np.ndarray]:
    """n i.i.d. uniform reads of fixed length, record indices first..first+n-1.
    Returns (uint8 bytes [n*length], uint64 offsets [n+1]). With amb_rate > 0 that fraction of the
    letters is replaced by ambiguity codes / gaps (for the ambiguity-parity tests)."""
    order = np.frombuffer(_AA_ORDER if alphabet == AA else _DNA_ORDER, np.uint8)
    # 8 letters per 64-bit hash (one per byte); letter i of the stream comes from hash i // 8
    lo, hi = first * length, (first + n) * length
    h = _stream(seed, 10, np.arange(lo // 8, (hi + 7) // 8, dtype=np.uint64))
    byte = h.view(np.uint8).reshape(-1)[lo - 8 * (lo // 8): hi - 8 * (lo // 8)].astype(np.uint16)
    data = order[(byte * np.uint16(len(order))) >> np.uint16(8)]
    idx = np.arange(lo, hi, dtype=np.uint64) if amb_rate > 0 else None
    if amb_rate > 0:
        amb = np.frombuffer(b"BZJX-" if alphabet == AA else b"NRYSWKMBDHV-n", np.uint8)
        h2 = _stream(seed, 11, idx)
        hit = _unit(h2) < amb_rate
        data = np.where(hit, amb[((h2 >> _U64(3)) % _U64(len(amb))).astype(np.int64)], data)
    off = np.arange(n + 1, dtype=np.uint64) * _U64(length)
    return np.ascontiguousarray(data, np.uint8), off


# ------------------------------------------------------- gen2: the generator the device shares -----
# A database as a pure function of (seed, key, position in the list), every step integer arithmetic or one correctly
# rounded IEEE operation, so that epik_b200/csrc/synth_gen.cuh builds the same database on the device - BASELINE
# config 5 (32 GB) exists only there - and this module restates any subset of its lists for the reference.
GEN2_TABLE_BITS = 10


class Gen2:
    """Parameters of a gen2 database (what epk_synth_desc carries)."""

    def __init__(self, alphabet: int, k: int, n_leaves: int, p_occ: float, m: float, omega: float = 1.5, seed: int = SEED_DB,
                 key_space_limit: int = 0):
        self.alphabet, self.k, self.omega, self.seed, self.key_space_limit = alphabet, k, omega, seed, key_space_limit
        self.length, self.num, self.tot, self.parent = make_tree(n_leaves, seed ^ 0x7EE)
        self.n_nodes = len(self.length)
        self.presence_threshold = min((1 << 64) - 1, int(p_occ * float(1 << 64)))
        # list lengths: 2^GEN2_TABLE_BITS quantiles of log-normal(ln m, 1), clamped to [1, N / 8] (distinct branches)
        from scipy.special import ndtri
        q = (np.arange(1 << GEN2_TABLE_BITS, dtype=np.float64) + 0.5) / float(1 << GEN2_TABLE_BITS)
        cap = max(1, self.n_nodes // 8)
        self.length_table = np.clip(np.rint(np.exp(np.log(m) + ndtri(q))), 1, cap).astype(np.uint32)
        thr = np.float32(omega / (20.0 if alphabet == AA else 4.0)) ** np.float32(k)
        self.score_lo = float(np.log10(thr))
        self.n_space = (20 if alphabet == AA else 4) ** k
        if key_space_limit:
            self.n_space = min(self.n_space, key_space_limit)
        self.meta = dict(p_occ=p_occ, m=m, n_leaves=n_leaves, seed=seed, parent=self.parent, generator="gen2")

    def candidate_of(self, keys: np.ndarray) -> np.ndarray:
        """Index of a code in the device's enumeration of the key space (inverse of all_keys)."""
        kk = np.asarray(keys).astype(np.uint64)
        if self.alphabet == DNA:
            return kk
        cand, scale = np.zeros(len(kk), np.uint64), _U64(1)
        for i in range(self.k):
            cand += ((kk >> _U64(5 * i)) & _U64(31)) * scale
            scale *= _U64(20)
        return cand

    def present(self, keys: np.ndarray) -> np.ndarray:
        """Whether the database holds a list for these codes (inside the key space and drawn as present)."""
        inside = self.candidate_of(keys) < _U64(self.n_space)
        return inside & (_stream(self.seed, 3, np.asarray(keys).astype(_U64)) < _U64(self.presence_threshold))

    def all_keys(self) -> np.ndarray:
        """Codes of the whole key space, in the order the device enumerates them."""
        cand = np.arange(self.n_space, dtype=np.uint64)
        if self.alphabet == DNA:
            return cand.astype(np.uint32)
        idx, key = cand.copy(), np.zeros(len(cand), np.uint64)
        for i in range(self.k):
            key |= (idx % _U64(20)) << _U64(5 * i)
            idx //= _U64(20)
        return key.astype(np.uint32)

    def lists(self, keys: np.ndarray) -> FlatDB:
        """The database restricted to `keys` (absent ones are dropped; order kept): what the reference needs to
        place reads that look up only these keys."""
        keys = np.asarray(keys, np.uint32)
        keys = keys[self.present(keys)]
        kk = keys.astype(np.uint64)
        ell = self.length_table[(_stream(self.seed, 4, kk) >> _U64(64 - GEN2_TABLE_BITS)).astype(np.int64)].astype(np.uint64)
        offsets = np.zeros(len(keys) + 1, np.uint64)
        np.cumsum(ell, out=offsets[1:])
        n_entries = int(offsets[-1])
        owner = np.repeat(np.arange(len(keys), dtype=np.uint64), ell.astype(np.int64))
        j = np.arange(n_entries, dtype=np.uint64) - offsets[:-1][owner]
        eid = (kk[owner] << _U64(24)) ^ j
        stride = (_stream(self.seed, 6, eid) % _U64(7)) + _U64(1)
        run = np.cumsum(stride, dtype=np.uint64)
        if n_entries:
            first = offsets[:-1].astype(np.int64)
            run -= np.repeat(run[first] - stride[first], ell.astype(np.int64))
        start = _stream(self.seed, 7, kk) % _U64(self.n_nodes)
        branch = ((start[owner] + run) % _U64(self.n_nodes)).astype(np.uint32)
        u = ((_stream(self.seed, 8, eid) >> _U64(11)).astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)
        score = np.minimum((self.score_lo + (-0.3 - self.score_lo) * u).astype(np.float32), np.float32(-0.3))
        entries = np.empty(n_entries, ENTRY_DTYPE)
        entries["branch"], entries["score"] = branch, score
        return FlatDB(self.alphabet, self.k, self.omega, keys, offsets, entries, self.length, self.num, self.tot, None,
                      meta=dict(self.meta))

    def whole(self) -> FlatDB:
        """Every list (small shapes only: this is what the device generates without ever holding it raw)."""
        return self.lists(self.all_keys())


def keys_of_reads(alphabet: int, k: int, data: np.ndarray, off: np.ndarray) -> np.ndarray:
    """The distinct exact k-mer codes of unambiguous fixed-alphabet reads (what place_seq looks up for them)."""
    order = _AA_ORDER if alphabet == AA else _DNA_ORDER
    bits = 5 if alphabet == AA else 2
    lut = np.full(256, 255, np.uint8)
    for i, c in enumerate(order):
        lut[c] = i
    out = []
    for i in range(len(off) - 1):
        codes = lut[data[int(off[i]):int(off[i + 1])]].astype(np.uint64)
        assert (codes != 255).all(), "unambiguous letters only"
        if len(codes) < k:
            continue
        key = np.zeros(len(codes) - k + 1, np.uint64)
        for j in range(k):
            key = (key << _U64(bits)) | codes[j:len(codes) - k + 1 + j]
        out.append(key)
    return np.unique(np.concatenate(out)).astype(np.uint32) if out else np.zeros(0, np.uint32)


# ---------------------------------------------------------------- the BASELINE.json configs -----
CONFIGS = {
    # name: (alphabet, k, n_leaves, p_occ, m, read_len)
    "C1p": (DNA, 10, 652, 0.6, 24.0, 1500),    # D652-shaped stand-in for config 1 (N = 1303)
    "C2": (DNA, 10, 2500, 0.6, 32.0, 150),     # N = 4999, ~33 M entries
    "C3": (DNA, 10, 2500, 0.6, 32.0, 150),     # C2 loaded with --mu 0.5 --max-ram 4G (LOAD_FILTERS)
    "C4": (AA, 6, 1000, 0.05, 16.0, 100),      # N = 1999, 64 M-key space
    "C5": (DNA, 12, 10000, 0.9, 160.0, 150),   # N = 19999, ~4 G entries (device-generated only)
    "C5r": (DNA, 10, 10000, 0.9, 160.0, 150),  # C5's tree and per-read work (~33 k postings per read) over the
                                               # 4^10 key space: ~250 M entries, 2 GB (what numpy generates in a minute)
}


# (mu, omega, max_entries) of i2l::load for the configs that filter at load time; --max-ram 4G is
# 4 * 2^30 / sizeof(pkdb_value) entries (main.cpp:252-265)
LOAD_FILTERS = {"C3": (0.5, 1.5, (4 << 30) // 8)}


def make_config_gen2(name: str) -> Gen2:
    """The device-generated form of a config (C5: the only form it has)."""
    alphabet, k, n_leaves, p_occ, m, _ = CONFIGS[name]
    g = Gen2(alphabet, k, n_leaves, p_occ, m)
    g.meta["config"] = name
    return g


def make_config_db(name: str, **kw) -> FlatDB:
    alphabet, k, n_leaves, p_occ, m, _ = CONFIGS[name]
    db = make_db(alphabet, k, n_leaves, p_occ, m, **kw)
    db.meta["config"] = name
    return db


def tiny_db(alphabet: int = DNA, k: int = 4, n_leaves: int = 20, p_occ: float = 0.7, m: float = 4.0,
            seed: int = 7, **kw) -> FlatDB:
    """A database small enough for pure-Python checks."""
    return make_db(alphabet, k, n_leaves, p_occ, m, seed=seed, **kw)
