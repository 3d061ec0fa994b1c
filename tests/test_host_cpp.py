"""The C++ host layer (epik_b200/host): jplace writer, FASTA reader, and — on a GPU — the epik-dna driver end
to end. The jplace layout is pinned by tests/golden/jplace_reference.jplace, written by the REFERENCE's own
jplace.cpp (tests/cpp/jplace_reference_main.cpp; regenerated and re-checked here whenever /root/reference
and a RapidJSON are around)."""
import glob
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from tests import helpers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
OURS = ["-I" + os.path.join(ROOT, "epik_b200", "host", "include"), "-I" + os.path.join(ROOT, "include")]
LINK = ["-pthread", "-L" + os.path.join(ROOT, "epik_b200"), "-lepik_b200", "-Wl,-rpath," + os.path.join(ROOT, "epik_b200")]


def compile_cpp(tmp_path, name, flags, sources, link=()):
    exe = str(tmp_path / name)
    # EPK_TEST_SANITIZE=address|thread|undefined: the host tests under a sanitizer (the threaded FASTA reader and jplace writer)
    sanitize = ["-fsanitize=" + os.environ["EPK_TEST_SANITIZE"], "-g"] if os.environ.get("EPK_TEST_SANITIZE") else []
    subprocess.run([CXX, "-std=c++17", "-O1", "-Wall", "-Wextra", "-I" + CPP] + sanitize + flags + ["-o", exe] + sources + list(link), check=True)
    return exe


def test_jplace_writer_is_byte_identical_to_the_reference_writer(built, tmp_path):
    exe = compile_cpp(tmp_path, "jp_ours", OURS, [os.path.join(CPP, "jplace_ours_main.cpp")], LINK)
    out = str(tmp_path / "ours.jplace")
    subprocess.run([exe, out], check=True)
    golden = open(os.path.join(helpers.GOLDEN, "jplace_reference.jplace"), "rb").read()
    assert open(out, "rb").read() == golden
    # the driver's way of writing - batches formatted on threads, a pass written with concurrent pwrites - gives the same bytes
    out2 = str(tmp_path / "ours_parallel.jplace")
    subprocess.run([exe, out2, "parallel"], check=True)
    assert open(out2, "rb").read() == golden
    # (the case data holds one float(1.8e308) = inf score: RapidJSON writes the separator and no value for it,
    # and so does this writer — the file is deliberately not valid JSON at that spot)
    # keep the golden file honest when the reference and a RapidJSON are available
    ref_src = "/root/reference/epik/src/epik/jplace.cpp"
    rj = glob.glob(os.path.join(sys.prefix, "lib", "python*", "site-packages", "tilelang", "3rdparty", "composable_kernel", "include"))
    if os.path.exists(ref_src) and rj and os.path.exists(os.path.join(rj[0], "rapidjson", "prettywriter.h")):
        ref_exe = str(tmp_path / "jp_ref")
        subprocess.run([CXX, "-std=c++17", "-O1", "-w", "-I" + CPP, "-I/root/reference/epik/include",
                        "-I" + os.path.join(ROOT, "oracle", "shim"), "-I" + rj[0], "-o", ref_exe,
                        os.path.join(CPP, "jplace_reference_main.cpp"), ref_src], check=True)
        ref_out = str(tmp_path / "ref.jplace")
        subprocess.run([ref_exe, ref_out], check=True)
        assert open(ref_out, "rb").read() == golden, "tests/golden/jplace_reference.jplace is stale"


def test_jplace_text_from_arrays_is_the_text_from_objects(built, tmp_path):
    """jplace_writer::format_arrays (the driver's path: straight from the arrays of a device pass) gives the bytes of
    jplace_writer::format on the placed_collection built from the same arrays, and append_batches (text and append in one
    stage, 1 and 5 threads, two passes) writes the file of format + append_many (tests/cpp/jplace_arrays_main.cpp)."""
    exe = compile_cpp(tmp_path, "jp_arrays", OURS, [os.path.join(CPP, "jplace_arrays_main.cpp")], LINK)
    out = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.startswith("ok "), out.stdout + out.stderr


def test_fasta_reader(tmp_path):
    exe = compile_cpp(tmp_path, "fasta", OURS, [os.path.join(CPP, "fasta_main.cpp")])
    text = (b"; a comment before the first record\n>r1 first read\nACGT\nACGT\n\n>r2\r\nAC GT\r\nNN--\r\n>empty\n>r4\tx  \n"
            b"acgt" + b"A" * 5000 + b"\n>last\nTTTT")
    path = tmp_path / "q.fasta"
    path.write_bytes(text)
    want = [("r1 first read", "ACGTACGT"), ("r2", "ACGTNN--"), ("empty", ""), ("r4\tx", "acgt" + "A" * 5000), ("last", "TTTT")]
    for batch_size, threads, block in ((1, 1, 1 << 25), (2, 1, 7), (3, 4, 1), (100, 3, 64), (2, 2, 5000), (4, 8, 1 << 25)):
        out = subprocess.run([exe, str(path), str(batch_size), str(threads), str(block)], check=True, capture_output=True).stdout.decode().split("\n")
        recs = [l.split("\t") for l in out if l and not l.startswith("bytes")]
        got = [("\t".join(r[1:-1]), r[-1]) for r in recs]
        assert got == want, (batch_size, threads, block)
        assert [int(r[0]) for r in recs] == [i // batch_size for i in range(len(want))]
        assert out[-2] == "bytes\t%d" % len(text)
    # many records over several blocks and threads: the same records as one thread reads
    big = tmp_path / "big.fasta"
    rng = np.random.default_rng(4)
    with open(big, "wb") as f:
        for i in range(20000):
            f.write(b">read%d some text\n" % i + bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), size=int(rng.integers(0, 300)))) + (b"\r\n" if i % 7 == 0 else b"\n"))
    one = subprocess.run([exe, str(big), "3000", "1"], check=True, capture_output=True).stdout
    for threads, block in ((4, 100000), (16, 1 << 25), (5, 333)):
        assert subprocess.run([exe, str(big), "3000", str(threads), str(block)], check=True, capture_output=True).stdout == one


def test_epkdb_round_trip(tmp_path):
    from epik_b200 import flatdb, synth
    db = synth.make_db(0, 6, 8, 0.5, 4.0, seed=3)
    nw = synth.tree_newick(db.branch_length, db.meta["parent"])
    assert nw.count("{") == db.n_nodes and nw.endswith("{%d};" % (db.n_nodes - 1))
    path = str(tmp_path / "t.epkdb")
    flatdb.save_epkdb(db, path, nw)
    back, nw2 = flatdb.load_epkdb(path)
    assert nw2 == nw
    for name in ("keys", "offsets", "entries", "filter_value", "branch_length", "subtree_num_nodes", "subtree_total_length"):
        assert np.array_equal(getattr(back, name), getattr(db, name)), name


def test_corrupt_epkdb_is_refused_by_the_loader(built, tmp_path):
    """The .epkdb header and offsets are not trusted: inconsistent offsets or counts larger than the file are an
    "Error: ..." exit like any other bad input (main.cpp:384-388), before anything is uploaded."""
    from epik_b200 import flatdb, synth
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "epik_b200", "host")], check=True)
    flat = synth.make_db(0, 6, 20, 0.5, 4.0, seed=5)
    good = str(tmp_path / "good.epkdb")
    flatdb.save_epkdb(flat, good, synth.tree_newick(flat.branch_length, flat.meta["parent"]))
    q_path = str(tmp_path / "q.fasta")
    open(q_path, "w").write(">a\nACGTACGTACGT\n")
    raw = bytearray(open(good, "rb").read())
    n_keys = int(np.frombuffer(raw[24:32], np.uint64)[0])
    keys_bytes = (4 * n_keys + 7) // 8 * 8
    exe = os.path.join(ROOT, "bin", "epik", "epik-dna")
    cases = {}
    bad = bytearray(raw)
    at = 56 + keys_bytes + 8 * 2                       # offsets[2] := a huge value -> decreasing at offsets[3]
    bad[at:at + 8] = np.uint64(1 << 40).tobytes()
    cases["offsets"] = (bad, "offsets")
    bad = bytearray(raw)
    bad[32:40] = np.uint64(1 << 50).tobytes()          # n_entries far beyond the file
    cases["counts"] = (bad, "exceed the file size")
    for name, (blob, message) in cases.items():
        path = str(tmp_path / (name + ".epkdb"))
        open(path, "wb").write(bytes(blob))
        r = subprocess.run([exe, "-d", path, "-q", q_path, "-o", str(tmp_path)], capture_output=True, text=True)
        assert r.returncode == 255 and "Error:" in r.stderr and message in r.stderr, (name, r.returncode, r.stderr)


@pytest.mark.gpu
@pytest.mark.parametrize("states", ["nucl", "amino"])
def test_driver_end_to_end(built, tmp_path, states):
    """bin/epik/epik-dna|aa with the argv the reference's epik.py builds (epik.py:85-96): the jplace must hold,
    per --batch-size batch, one record per unique sequence with the oracle's placements and all its names."""
    from epik_b200 import flatdb, synth
    from oracle import oracle as O
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "epik_b200", "host")], check=True)
    aa = states == "amino"
    flat = synth.make_db(1, 4, 150, 0.3, 8.0, seed=71) if aa else synth.make_db(0, 8, 150, 0.6, 10.0, seed=71)
    data, off = synth.make_reads(int(aa), 1000, 80, seed=72, amb_rate=0.005)
    seqs = [bytes(data[int(off[i]):int(off[i + 1])]).decode() for i in range(1000)]
    seqs[10] = seqs[3]; seqs[11] = seqs[3]; seqs[400] = seqs[3]  # duplicates inside and across batches of 300
    db_path, q_path = str(tmp_path / "db.epkdb"), str(tmp_path / "q.fasta")
    flatdb.save_epkdb(flat, db_path, synth.tree_newick(flat.branch_length, flat.meta["parent"]))
    with open(q_path, "w") as f:
        for i, s in enumerate(seqs):
            f.write(">read%d some description\n%s\n%s\n" % (i, s[:50], s[50:]))
    exe = os.path.join(ROOT, "bin", "epik", "epik-aa" if aa else "epik-dna")
    argv = [exe, "-d", db_path, "-q", q_path, "-j", "4", "--omega", "1.5", "--mu", "1.0", "-o", str(tmp_path), "--batch-size", "300",
            "--device-batches", "2", q_path]
    run = subprocess.run(argv, capture_output=True, text=True)
    assert run.returncode == 0, run.stderr
    assert "Placed 1000 sequences." in run.stdout and "Average speed:" in run.stdout
    jp = json.load(open(str(tmp_path / "placements_q.fasta.jplace")))
    assert jp["version"] == 3 and jp["fields"] == ["edge_num", "likelihood", "like_weight_ratio", "distal_length", "pendant_length"]
    assert jp["tree"].endswith("{%d};" % (flat.n_nodes - 1)) and jp["metadata"]["invocation"].startswith(exe)
    kind = "reference" if O.available("reference", int(aa)) else "port"
    want = O.Oracle(flat, 7, 0.01, 2, kind=kind).place(seqs, threads=2, batch_size=300)
    name_to_index = {"read%d some description" % i: i for i in range(1000)}
    seen = set()
    for rec in jp["placements"]:
        idx = sorted(name_to_index[n] for n, m in rec["nm"] if m == 1)
        assert len(idx) == len(rec["nm"])
        assert len({i // 300 for i in idx}) == 1 and len({seqs[i] for i in idx}) == 1  # one batch, one sequence
        seen.update(idx)
        i = idx[0]
        m = int(want["n_place"][i])
        assert len(rec["p"]) == m
        got_b = [p[0] for p in rec["p"]]
        if got_b != list(want["branch"][i, :m]):
            assert sorted(got_b) == sorted(want["branch"][i, :m]) and len(set(want["score"][i, :m])) < m  # ties only
        np.testing.assert_allclose([p[1] for p in rec["p"]], want["score"][i, :m].astype(np.float64), rtol=0, atol=0)
        np.testing.assert_allclose([p[2] for p in rec["p"]], want["lwr"][i, :m], rtol=1e-9)
        if got_b == list(want["branch"][i, :m]):
            np.testing.assert_allclose([p[3] for p in rec["p"]], want["distal"][i, :m], rtol=1e-15)
            np.testing.assert_allclose([p[4] for p in rec["p"]], want["pendant"][i, :m], rtol=1e-15)
    assert seen == set(range(1000))
    # the duplicates: reads 3, 10, 11 share a record (same batch), read 400 has its own
    groups = [sorted(name_to_index[n] for n, _ in rec["nm"]) for rec in jp["placements"]]
    assert [3, 10, 11] in groups and [400] in groups
    if not aa:
        # the same run with the database POOLED over three shards (--devices 3 --pool-database; on a box with fewer GPUs
        # the shards share them) and with it replicated on every GPU present: the jplace body must not change by a byte
        import torch
        body = open(str(tmp_path / "placements_q.fasta.jplace"), "rb").read().split(b'"tree"', 1)[1]
        variants = [["--devices", "3", "--pool-database"]]
        if torch.cuda.device_count() > 1:
            variants.append(["--devices", "all"])
        for extra in variants:
            run = subprocess.run(argv[:-1] + extra + [q_path], capture_output=True, text=True)
            assert run.returncode == 0, run.stderr
            assert open(str(tmp_path / "placements_q.fasta.jplace"), "rb").read().split(b'"tree"', 1)[1] == body, extra


@pytest.mark.gpu
def test_driver_errors_and_filters(built, tmp_path):
    from epik_b200 import flatdb, synth
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "epik_b200", "host")], check=True)
    flat = synth.make_db(0, 8, 150, 0.6, 10.0, seed=71)
    db_path, q_path = str(tmp_path / "db.epkdb"), str(tmp_path / "q.fasta")
    flatdb.save_epkdb(flat, db_path, synth.tree_newick(flat.branch_length, flat.meta["parent"]))
    open(q_path, "w").write(">a\nACGTACGTACGTACGTACGT\n")
    exe = os.path.join(ROOT, "bin", "epik", "epik-dna")
    base = [exe, "-d", db_path, "-q", q_path, "-o", str(tmp_path)]
    r = subprocess.run(base + ["--mu", "1.5"], capture_output=True, text=True)
    assert r.returncode == 255 and "Error: Mu has to a value in [0, 1]" in r.stderr          # main.cpp:196-202,384-388
    r = subprocess.run(base + ["--max-ram", "4"], capture_output=True, text=True)
    assert r.returncode == 255 and "Memory limit is too low" in r.stderr                      # main.cpp:259-262
    r = subprocess.run(base + ["--max-ram", "12x"], capture_output=True, text=True)
    assert r.returncode == 255 and "Unknown memory unit." in r.stderr
    r = subprocess.run([os.path.join(ROOT, "bin", "epik", "epik-aa")] + base[1:], capture_output=True, text=True)
    assert r.returncode == 255 and "nucleotide" in r.stderr
    r = subprocess.run(base + ["--mu", "0.5", "--max-ram", "0.5M"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Max-RAM provided: will be loaded not more than 65.5K phylo-k-mers." in r.stdout
    loaded = [l for l in r.stdout.split("\n") if l.startswith("Loaded ")][0]
    assert " of %s phylo-k-mers" % ("%.1fK" % (flat.n_entries / 1e3)) in loaded


@pytest.mark.gpu
def test_reference_driver_loop_against_the_facade(built, tmp_path):
    """main.cpp:294-366 written against epik_b200's headers with i2l-shaped db/tree types (tests/cpp/facade_main.cpp):
    epik::placer(db, tree, ...) flattens and uploads the database; the jplace equals the driver binary's."""
    from epik_b200 import flatdb, synth
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "epik_b200", "host")], check=True)
    exe = compile_cpp(tmp_path, "facade", OURS + ["-I" + os.path.join(ROOT, "oracle", "shim")],
                      [os.path.join(CPP, "facade_main.cpp")], LINK)
    flat = synth.make_db(0, 8, 150, 0.6, 10.0, seed=81)
    data, off = synth.make_reads(0, 700, 90, seed=82, amb_rate=0.005)
    db_path, q_path = str(tmp_path / "db.epkdb"), str(tmp_path / "q.fasta")
    flatdb.save_epkdb(flat, db_path, synth.tree_newick(flat.branch_length, flat.meta["parent"]))
    with open(q_path, "w") as f:
        for i in range(700):
            f.write(">r%d\n%s\n" % (i % 650, bytes(data[int(off[i % 650]):int(off[i % 650 + 1])]).decode()))
    a = str(tmp_path / "facade.jplace")
    r = subprocess.run([exe, db_path, q_path, a, "256"], capture_output=True, text=True)
    assert r.returncode == 0 and "placed 700" in r.stdout, r.stderr
    r = subprocess.run([os.path.join(ROOT, "bin", "epik", "epik-dna"), "-d", db_path, "-q", q_path, "-o", str(tmp_path),
                        "--batch-size", "256"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    ja, jb = json.load(open(a)), json.load(open(str(tmp_path / "placements_q.fasta.jplace")))
    assert ja["placements"] == jb["placements"] and ja["tree"] == jb["tree"] and len(ja["placements"]) > 600
