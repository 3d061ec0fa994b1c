"""CPU tests of the drop-in boundary: libepik_b200.so loads, exports every symbol include/epik_b200.h
declares, and refuses to compute without a CUDA device (there is no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from epik_b200 import capi
from tests import helpers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", autouse=True)
def _build(built):
    return built


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "epik_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(epk_[a-z_0-9]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(capi.SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = capi.lib()
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert lib.epk_abi_version() == 1
    # extern "C", unmangled
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT (epk_[a-z_0-9]+)$", out, flags=re.M))
    assert set(declared_symbols()) <= exported


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", capi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "epik_b200.h"\nint main(void) { epk_db_desc d; (void)d; return (int)sizeof(epk_result) * 0; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                    "-o", str(tmp_path / "t.o")], check=True)


def test_struct_layouts_match_the_header(tmp_path):
    """ctypes mirrors vs the C compiler's sizeof/offsetof."""
    src = tmp_path / "s.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "epik_b200.h"\nint main(void) { printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", '
                   "sizeof(epk_db_desc), sizeof(epk_load_filter), sizeof(epk_db_info), sizeof(epk_placer_opts), "
                   "sizeof(epk_result), sizeof(epk_counters), offsetof(epk_db_desc, n_nodes), offsetof(epk_db_info, n_keys_loaded), "
                   "sizeof(epk_db_view), sizeof(epk_synth_desc), offsetof(epk_synth_desc, seed), offsetof(epk_synth_desc, score_lo), "
                   "sizeof(epk_microbench)); return 0; }\n")
    exe = tmp_path / "s"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(capi.DbDesc), C.sizeof(capi.LoadFilter), C.sizeof(capi.DbInfo), C.sizeof(capi.PlacerOpts),
                     C.sizeof(capi.Result), C.sizeof(capi.Counters), capi.DbDesc.n_nodes.offset,
                     capi.DbInfo.n_keys_loaded.offset, C.sizeof(capi.DbView), C.sizeof(capi.SynthDesc),
                     capi.SynthDesc.seed.offset, capi.SynthDesc.score_lo.offset, C.sizeof(capi.Microbench)]


def test_group_batch_is_first_occurrence():
    """group_by_sequence_content (place.cpp:73-81): host-side, needs no device."""
    seqs = [b"ACGT", b"TTTT", b"ACGT", b"", b"ACG", b"TTTT", b""]
    data = np.frombuffer(b"".join(seqs), np.uint8).copy()
    off = np.zeros(len(seqs) + 1, np.uint64)
    off[1:] = np.cumsum([len(s) for s in seqs])
    rep = np.zeros(len(seqs), np.uint32)
    capi.check(capi.lib().epk_group_batch(data.ctypes.data_as(C.c_void_p), off.ctypes.data_as(C.c_void_p), len(seqs),
                                          rep.ctypes.data_as(C.c_void_p)))
    assert list(rep) == [0, 1, 0, 3, 4, 1, 3]


def test_group_batches_groups_within_each_batch_on_threads():
    """epk_group_batches = epk_group_batch on every --batch-size slice (main.cpp:332-344), whatever the thread count."""
    rng = np.random.default_rng(3)
    pool = [bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), size=int(rng.integers(0, 12)))) for _ in range(40)]
    seqs = [pool[int(i)] for i in rng.integers(0, len(pool), size=1000)]
    data = np.frombuffer(b"".join(seqs), np.uint8).copy()
    off = np.zeros(len(seqs) + 1, np.uint64)
    off[1:] = np.cumsum([len(s) for s in seqs])
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for batch, threads in ((64, 1), (64, 7), (1000, 4), (0, 3), (1, 2)):
        rep = np.full(len(seqs), 0xFFFFFFFF, np.uint32)
        capi.check(capi.lib().epk_group_batches(p(data), p(off), len(seqs), batch, threads, p(rep)))
        size = batch or len(seqs)
        for first in range(0, len(seqs), size):
            want = np.zeros(min(size, len(seqs) - first), np.uint32)
            capi.check(capi.lib().epk_group_batch(p(data), p(off[first:]), len(want), p(want)))
            assert np.array_equal(rep[first:first + len(want)], want), (batch, threads, first)


def test_malformed_offsets_are_refused_before_anything_runs():
    """A CSR description whose offsets decrease (corrupt or hand-made database) is EPK_E_INVALID - checked on the host,
    before any kernel computes a list length from it (and before a device is even looked for)."""
    from epik_b200.placer import ENTRY_DTYPE
    keys = np.array([1, 2, 3], np.uint32)
    entries = np.zeros(4, ENTRY_DTYPE)
    one = np.ones(3, np.float64)
    num = np.ones(3, np.uint64)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for offsets, message in (([0, 3, 2, 4], "non-decreasing"), ([1, 2, 3, 4], "offsets[0]")):
        off = np.array(offsets, np.uint64)
        desc = capi.DbDesc(0, 4, 1.5, 3, p(keys), p(off), p(entries), None, 3, p(one), p(num), p(one))
        handle = C.c_void_p()
        status = capi.lib().epk_db_create(C.byref(desc), None, 0, C.byref(handle))
        assert status == capi.EPK_E_INVALID and message in capi.lib().epk_last_error().decode(), (offsets, status)


def test_no_cpu_fallback():
    """Without a CUDA device the product path must fail loudly, not compute on the host."""
    if capi.lib().epk_device_count() > 0:
        pytest.skip("a CUDA device is present")
    from epik_b200.placer import Database
    with pytest.raises(capi.EpkError) as e:
        Database(helpers.kat1_db())
    assert e.value.status == capi.EPK_E_CUDA
    assert "no CPU path" in str(e.value)


def test_product_does_not_touch_the_oracle():
    """Nothing under epik_b200/ may import, include, link or dlopen anything from oracle/."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "epik_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in text.lower().replace("oracle/", "oracle/") or "import oracle" not in text, f
                assert "from oracle" not in text and "import oracle" not in text and "libepik_port" not in text \
                    and "libepik_ref" not in text and "oracle_capi" not in text, os.path.join(dirpath, f)
    out = subprocess.run(["ldd", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "epik_port" not in out and "epik_ref" not in out
