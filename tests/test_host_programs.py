"""The reference's entry points served by the C++ host layer (vlg_matching_b200/csrc/host) over the C ABI:
examples/vlg_matching.cpp and benchmark/gapped-matching's create_collection / gm_index / gm_search.
CPU part: the programs exist, link against libvlgb200.so and fail LOUDLY without a GPU (no CPU fallback).
GPU part: stdout identical to the reference's own example; num_results / checksum identical to the reference
matcher on the committed pattern set (tests/golden/gm_search_ref.json)."""
import json
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
BIN = os.path.join(ROOT, "vlg_matching_b200", "bin")
PROGS = ["vlg_matching.x", "gm_index-VLGB200.x", "gm_search-VLGB200.x", "create_collection.x", "generate-pattern.x"]


def run(args, **kw):
    return subprocess.run(args, capture_output=True, text=True, timeout=600, **kw)


def test_programs_are_built():
    for p in PROGS:
        assert os.access(os.path.join(BIN, p), os.X_OK), f"{p} missing: run __graft_entry__.build()"


def test_create_collection_is_host_only_and_bit_compresses(tmp_path):
    """text.TEXT = serialized int_vector<0> (u64 bit length, u8 width, packed words); \\0 \\n \\r \\f -> ' '."""
    raw = tmp_path / "raw.txt"
    raw.write_bytes(b"ab\ncd\0ef\rgh\fij")
    col = tmp_path / "col"
    r = run([os.path.join(BIN, "create_collection.x"), "-i", str(raw), "-c", str(col)])
    assert r.returncode == 0, r.stderr
    for sub in ("tmp", "index", "patterns", "results"):
        assert (col / sub).is_dir()
    blob = (col / "text.TEXT").read_bytes()
    bits = int.from_bytes(blob[:8], "little")
    width = blob[8]
    assert width == 7 and bits == 14 * 7          # max byte 'j' = 106 -> hi(106)+1 = 7 bits
    words = np.frombuffer(blob[9:], dtype="<u8")
    val = 0
    for i, w in enumerate(words):
        val |= int(w) << (64 * i)
    got = bytes((val >> (7 * i)) & 0x7F for i in range(14))
    assert got == b"ab cd ef gh ij"


def test_no_gpu_means_loud_failure(tmp_path):
    import vlg_matching_b200.capi as capi
    if capi.device_count() > 0:
        pytest.skip("GPU present")
    r = run([os.path.join(BIN, "vlg_matching.x")])
    assert r.returncode != 0 and "CUDA" in r.stderr
    assert "occ starting" not in r.stdout            # nothing was computed on the CPU


@pytest.mark.gpu
def test_example_stdout_matches_reference():
    want = open(os.path.join(HERE, "golden", "example_stdout.txt")).read()
    r = run([os.path.join(BIN, "vlg_matching.x")])
    assert r.returncode == 0, r.stderr
    assert r.stdout == want


@pytest.mark.gpu
def test_example_with_file_argument(tmp_path):
    """construct(idx, file, 1) + write_structure (examples/vlg_matching.cpp:34-37).  The int block then reads the
    same raw file as a serialized int_vector<> -- the reference aborts there; we report the error."""
    f = tmp_path / "t.txt"
    f.write_bytes(b"abracadabrasimsalabim")
    r = run([os.path.join(BIN, "vlg_matching.x"), str(f)])
    want = open(os.path.join(HERE, "golden", "example_stdout.txt")).read()
    byte_block = want.split("\ncount(97 99")[0]
    assert r.stdout.startswith(byte_block)
    assert (tmp_path / "t.txt.info.html").exists()
    assert r.returncode != 0 and "int_vector" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("batch", ["0", "1"])
def test_gm_pipeline_matches_reference_checksum(tmp_path, batch):
    from vlg_matching_b200 import synth
    g = json.load(open(os.path.join(HERE, "golden", "gm_search_ref.json")))
    text = synth.english_text(g["text"]["n"], g["text"]["seed"])
    assert synth.sha256(text) == g["sha256"]
    raw = tmp_path / "raw.txt"
    text.tofile(raw)
    col = tmp_path / "col"
    assert run([os.path.join(BIN, "create_collection.x"), "-i", str(raw), "-c", str(col)]).returncode == 0
    r = run([os.path.join(BIN, "gm_index-VLGB200.x"), "-c", str(col)])
    assert r.returncode == 0, r.stderr
    assert (col / "index" / "index-VLGB200-b200v2.sdsl").exists()
    pat = col / "patterns" / "p.txt"
    pat.write_text("\n".join(g["patterns"]) + "\nbroken.{5,1}pattern\n")   # the bad line is skipped (utils.hpp:94-99)
    r = run([os.path.join(BIN, "gm_search-VLGB200.x"), "-c", str(col), "-p", str(pat), "-b", batch])
    assert r.returncode == 0, r.stderr
    keys = dict(line[2:].split(" = ") for line in r.stdout.splitlines() if line.startswith("# ") and " = " in line)
    assert int(keys["num_results"]) == g["num_results"]
    assert int(keys["checksum"]) == g["checksum"]
    assert r.stdout.count("TIMING = ") == len(g["patterns"])
    for k in ("total_time_mus", "median_time_mus", "prep_total_time_mus"):
        assert k in keys
    assert "Skipped" in r.stderr


@pytest.mark.gpu
def test_generate_pattern_and_grid_pipeline(tmp_path):
    """The reference's pattern pipeline (benchmark/gapped-matching/Makefile:90-106): generate-pattern -> words file ->
    `xargs -n k` lines joined by the gap -> gm_search.  Words equal the numpy statement of generate-pattern.cpp;
    the k = 4 pattern file runs through gm_search -b 1 and its checksum equals the oracle's."""
    from oracle import pyoracle as O
    from vlg_matching_b200 import synth
    text = synth.make_text(2, 2 << 20)
    raw = tmp_path / "raw.txt"
    text.tofile(raw)
    r = run([os.path.join(BIN, "generate-pattern.x"), str(raw), "200", "3", "1"])
    assert r.returncode == 0, r.stderr
    words = r.stdout.split("\n")[:-1]
    assert [w.encode() for w in words] == synth.frequent_phrases(text, 200, 3)
    queries = synth.grid_queries([w.encode() for w in words], 4, (0, 100), 20)
    col = tmp_path / "col"
    assert run([os.path.join(BIN, "create_collection.x"), "-i", str(raw), "-c", str(col)]).returncode == 0
    assert run([os.path.join(BIN, "gm_index-VLGB200.x"), "-c", str(col)]).returncode == 0
    pat = col / "patterns" / "grid.txt"
    pat.write_bytes(b"\n".join(synth.to_benchmark_syntax(q) for q in queries) + b"\n")
    r = run([os.path.join(BIN, "gm_search-VLGB200.x"), "-c", str(col), "-p", str(pat), "-b", "1"])
    assert r.returncode == 0, r.stderr
    keys = dict(line[2:].split(" = ") for line in r.stdout.splitlines() if line.startswith("# ") and " = " in line)
    ref = O.search(text, queries, want_tuples=False, threads=8)
    assert int(keys["num_results"]) == sum(len(s) for _, s, _ in ref)
    assert int(keys["checksum"]) == sum(int(s.sum()) for _, s, _ in ref)


@pytest.mark.gpu
def test_gm_search_wcsearch3_gap_rule(tmp_path):
    """gm_search --wcsearch3-gaps = the reference's index_wcsearch3 rule (ONE bound |P_1| + gaps[0] for every level,
    index_wcsearch3.hpp:412-413) against tests/golden/wcsearch3_ref.json, generated from the reference's own
    index_wcsearch3: num_results / checksum over all expressible golden patterns, per text; without the flag the
    library's per-gap rule answers (a different checksum); the Python surface returns the same starts query by query;
    a pattern the rule cannot express in per-gap form is an error, not a silent empty result."""
    import numpy as np
    import vlg_matching_b200 as V
    from vlg_matching_b200 import capi, synth
    from test_oracle_golden import wcsearch3_cases
    from oracle import pyoracle as O
    capi.init(0)
    g, cases = wcsearch3_cases()
    for ci, (text, raw, rewritten, want) in enumerate(cases):
        idx = V.Index.build(text)
        r = idx.search(V.PackedBatch(rewritten), capi.WANT_STARTS)
        for q, w in enumerate(want):
            assert r.status(q) == 0 and r.starts(q).tolist() == w, (ci, q, raw[q])
        idx.free()
        rawf = tmp_path / f"raw{ci}.txt"
        text.tofile(rawf)
        col = tmp_path / f"col{ci}"
        assert run([os.path.join(BIN, "create_collection.x"), "-i", str(rawf), "-c", str(col)]).returncode == 0
        assert run([os.path.join(BIN, "gm_index-VLGB200.x"), "-c", str(col)]).returncode == 0
        pat = col / "patterns" / "w.txt"
        pat.write_bytes(b"\n".join(synth.to_benchmark_syntax(q) for q in raw) + b"\n")
        sums = {}
        for flag in ([], ["--wcsearch3-gaps"]):
            r = run([os.path.join(BIN, "gm_search-VLGB200.x"), "-c", str(col), "-p", str(pat), "-b", "1"] + flag)
            assert r.returncode == 0, r.stderr
            keys = dict(line[2:].split(" = ") for line in r.stdout.splitlines() if line.startswith("# ") and " = " in line)
            sums[bool(flag)] = (int(keys["num_results"]), int(keys["checksum"]))
        assert sums[True] == (sum(len(w) for w in want), sum(sum(w) for w in want))
        lib = O.search(text, raw, want_tuples=False, threads=8)
        assert sums[False] == (sum(len(s) for _, s, _ in lib), sum(int(s.sum()) for _, s, _ in lib))
        assert sums[False] != sums[True]
        if ci == 0:   # |P_2| = 6 > |P_1| + a_1 = 2: overlapping occurrences, not expressible
            pat2 = col / "patterns" / "bad.txt"
            pat2.write_bytes(b"th.{0,5}abcdef.{0,5}e\n")
            r = run([os.path.join(BIN, "gm_search-VLGB200.x"), "-c", str(col), "-p", str(pat2), "-b", "1", "--wcsearch3-gaps"])
            assert r.returncode != 0 and "wcsearch3 gap semantics" in r.stderr
