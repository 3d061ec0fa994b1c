"""TEST INFRASTRUCTURE ONLY — ctypes loader for the CPU oracles (see oracle/oracle_capi.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module. It is the checker, never the thing measured as the product or shipped.

  kind="reference": oracle/_ref/libepik_ref_{dna,aa}.so — the reference's own place.cpp
                    (epik/src/epik/place.cpp, compiled unmodified) over the i2l shim.
  kind="port":      oracle/libepik_port.so — our plain restatement of the same path.

PARITY: steps 3-4 of the path (accumulate + threshold default, top-k + LWR) are pinned by the
reference's own code; the k-mer codec / DB lookup / load filter belong to the absent phylo42/i2l
submodule and are UNPINNED.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

ENTRY_DTYPE = np.dtype([("branch", np.uint32), ("score", np.float32)])


class _Desc(C.Structure):
    _fields_ = [
        ("alphabet", C.c_uint32),
        ("kmer_size", C.c_uint32),
        ("omega", C.c_float),
        ("n_keys", C.c_uint64),
        ("keys", C.c_void_p),
        ("offsets", C.c_void_p),
        ("entries", C.c_void_p),
        ("n_nodes", C.c_uint32),
        ("branch_length", C.c_void_p),
        ("subtree_num_nodes", C.c_void_p),
        ("subtree_total_length", C.c_void_p),
    ]


def lib_path(kind: str, alphabet: int) -> str:
    if kind == "reference":
        return os.path.join(_HERE, "_ref", "libepik_ref_%s.so" % ("aa" if alphabet else "dna"))
    if kind == "port":
        return os.path.join(_HERE, "libepik_port.so")
    raise ValueError(kind)


def available(kind: str, alphabet: int = 0) -> bool:
    return os.path.exists(lib_path(kind, alphabet))


_LIBS: dict = {}


def _load(kind: str, alphabet: int):
    path = lib_path(kind, alphabet)
    if path in _LIBS:
        return _LIBS[path]
    lib = C.CDLL(path)
    lib.orc_alphabet.restype = C.c_int
    lib.orc_kind.restype = C.c_char_p
    lib.orc_last_error.restype = C.c_char_p
    lib.orc_create.restype = C.c_void_p
    lib.orc_create.argtypes = [C.POINTER(_Desc), C.c_uint64, C.c_double, C.c_uint64]
    lib.orc_destroy.argtypes = [C.c_void_p]
    lib.orc_place.restype = C.c_int
    lib.orc_place.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 9
    lib.orc_kmers.restype = C.c_uint64
    lib.orc_kmers.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_uint64] + [C.c_void_p] * 4
    lib.orc_thresholds.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    _LIBS[path] = lib
    return lib


def pack_sequences(seqs):
    """list of bytes/str -> (uint8 array, uint64 offsets[n+1])."""
    bs = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]
    off = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    data = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, np.uint8)
    return data, off


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    """A CPU placer over a flat database (any object with the attributes of epik_b200.flatdb.FlatDB)."""

    def __init__(self, db, keep_at_most=7, keep_factor=0.01, max_threads=1, kind="reference"):
        self.kind = kind
        self.lib = _load(kind, int(db.alphabet))
        self.keep_at_most = int(keep_at_most)
        self.max_threads = int(max_threads)
        # keep the arrays alive for the duration of create()
        keys = np.ascontiguousarray(db.keys, dtype=np.uint32)
        offsets = np.ascontiguousarray(db.offsets, dtype=np.uint64)
        entries = np.ascontiguousarray(db.entries, dtype=ENTRY_DTYPE)
        bl = np.ascontiguousarray(db.branch_length, dtype=np.float64)
        sn = np.ascontiguousarray(db.subtree_num_nodes, dtype=np.uint64)
        st = np.ascontiguousarray(db.subtree_total_length, dtype=np.float64)
        d = _Desc(int(db.alphabet), int(db.kmer_size), float(db.omega), len(keys), _ptr(keys), _ptr(offsets),
                  _ptr(entries), len(bl), _ptr(bl), _ptr(sn), _ptr(st))
        self.handle = self.lib.orc_create(C.byref(d), self.keep_at_most, float(keep_factor), self.max_threads)
        if not self.handle:
            raise RuntimeError("oracle: " + self.lib.orc_last_error().decode())

    def close(self):
        if getattr(self, "handle", None):
            self.lib.orc_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def thresholds(self):
        t, lt = C.c_float(), C.c_float()
        self.lib.orc_thresholds(self.handle, C.byref(t), C.byref(lt))
        return np.float32(t.value), np.float32(lt.value)

    def place_packed(self, data, offsets, threads=1, batch_size=2000):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n, K = len(offsets) - 1, self.keep_at_most
        out = {
            "n_place": np.zeros(n, np.uint32),
            "branch": np.zeros((n, K), np.uint32),
            "score": np.zeros((n, K), np.float32),
            "lwr": np.zeros((n, K), np.float64),
            "count": np.zeros((n, K), np.uint64),
            "distal": np.zeros((n, K), np.float64),
            "pendant": np.zeros((n, K), np.float64),
            "group_rep": np.zeros(n, np.uint32),
        }
        secs = C.c_double(0.0)
        if data.size == 0:
            data = np.zeros(1, np.uint8)
        rc = self.lib.orc_place(self.handle, _ptr(data), _ptr(offsets), n, int(threads), int(batch_size),
                                _ptr(out["n_place"]), _ptr(out["branch"]), _ptr(out["score"]), _ptr(out["lwr"]),
                                _ptr(out["count"]), _ptr(out["distal"]), _ptr(out["pendant"]),
                                _ptr(out["group_rep"]), C.addressof(secs))
        if rc != 0:
            raise RuntimeError("oracle: " + self.lib.orc_last_error().decode())
        out["seconds"] = secs.value
        return out

    def place(self, seqs, threads=1, batch_size=2000):
        data, off = pack_sequences(seqs)
        return self.place_packed(data, off, threads, batch_size)

    def kmers(self, seq):
        b = seq.encode() if isinstance(seq, str) else bytes(seq)
        cap = max(1, len(b) * 20)
        pos = np.zeros(cap, np.uint32)
        key = np.zeros(cap, np.uint32)
        group = np.zeros(cap, np.uint32)
        hit = np.zeros(cap, np.int64)
        n = self.lib.orc_kmers(self.handle, b, len(b), cap, _ptr(pos), _ptr(key), _ptr(group), _ptr(hit))
        assert n <= cap
        return pos[:n], key[:n], group[:n], hit[:n]
