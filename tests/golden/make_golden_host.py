"""Golden outputs for the host-side drop-in programs, FROM THE UNMODIFIED REFERENCE (run in the build container):

    python tests/golden/make_golden_host.py

  example_stdout.txt   stdout of the reference's examples/vlg_matching.cpp (oracle/_ref/ref_example, no arguments)
  gm_search_ref.json   a 256 KiB synthetic text (seeded), 60 two-subpattern + 20 three-subpattern (uniform gap and
                       length) patterns in the benchmark syntax, and the num_results / checksum the reference matcher
                       reports for them (oracle/_ref/ref_vlg locate; checksum = sum of match starts, as
                       benchmark/gapped-matching/src/gm_search.cpp:110-114 computes it).
"""
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..", "..")
sys.path.insert(0, ROOT)
from oracle import pyoracle as O  # noqa: E402
from vlg_matching_b200 import synth  # noqa: E402


def main():
    assert O.have_reference()
    with tempfile.TemporaryDirectory() as d:
        out = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_example")], cwd=d, check=True, capture_output=True).stdout
    open(os.path.join(HERE, "example_stdout.txt"), "wb").write(out)

    n, seed = 256 << 10, 77
    text = synth.english_text(n, seed)
    qs = synth.sample_queries(text, 60, 2, (3, 7), [(10, 100), (100, 1000)], 78, word_aligned=True)
    qs += synth.sample_queries(text, 20, 3, (4, 4), [(5, 60)], 79, word_aligned=False)
    lib = [synth.to_library_syntax(q).decode("latin-1") for q in qs]
    res = O.reference_locate(text.tobytes(), lib)
    per = []
    for r in res:
        assert r[0] == "Q"
        per.append({"count": r[1], "checksum": sum(t[0] for t in r[2])})
    json.dump({"source": "oracle/_ref/ref_vlg locate (unmodified reference)", "text": {"generator": "synth.english_text", "n": n, "seed": seed},
               "sha256": synth.sha256(text), "patterns": [synth.to_benchmark_syntax(q).decode("latin-1") for q in qs],
               "per_pattern": per, "num_results": sum(p["count"] for p in per), "checksum": sum(p["checksum"] for p in per)},
              open(os.path.join(HERE, "gm_search_ref.json"), "w"), indent=1)
    print("num_results", sum(p["count"] for p in per))


if __name__ == "__main__":
    main()
