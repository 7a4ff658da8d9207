"""Golden vectors for the benchmark matcher's own gap rule, FROM THE REFERENCE'S index_wcsearch3 (run in the build
container):

    python tests/golden/make_golden_wcsearch3.py

  wcsearch3_ref.json   two seeded synthetic texts (English-like 96 KiB, DNA 48 KiB), queries with 2..5 subpatterns of
                       UNEQUAL lengths and UNEQUAL gaps in the benchmark syntax, and the match starts that
                       index_wcsearch3::search returns for them (oracle/_ref/ref_wcs3: the reference's
                       wild_card_match_iterator3 run over the reference's own wavelet tree, see oracle/ref_wcs3_driver.cpp).
                       index_wcsearch3 applies ONE bound, |P_1| + gaps[0], to every level
                       (benchmark/gapped-matching/include/index_wcsearch3.hpp:412-413), so these results DIFFER from the
                       library matcher's whenever lengths or gaps differ; `differs_from_library` records for how many
                       queries (the library's answer: oracle/_ref/ref_vlg locate on the same queries).
                       `representable` = every subpattern is at most |P_1| + a_1 long (else the rule lets
                       occurrences overlap, which the per-gap form cannot say).
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


def run_wcs3(text: bytes, lines):
    with tempfile.TemporaryDirectory() as d:
        tf, qf = os.path.join(d, "t"), os.path.join(d, "q")
        open(tf, "wb").write(text)
        open(qf, "wb").write(b"\n".join(lines) + b"\n")
        out = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_wcs3"), tf, qf], check=True, capture_output=True).stdout
    res = []
    for ln in out.decode().splitlines():
        f = ln.split()
        if f[0] == "X":          # the reference ran into its own undefined behaviour on this query (see ref_wcs3_driver.cpp)
            res.append(None)
            continue
        assert f[0] == "Q" and int(f[2]) == len(f) - 3, ln
        res.append([int(x) for x in f[3:]])
    assert len(res) == len(lines)
    return res


def main():
    assert O.have_reference()
    cases = []
    specs = [("english", synth.english_text(96 << 10, 301), 301, [(2, (1, 6), 401), (3, (1, 6), 402), (4, (2, 5), 403), (5, (1, 4), 404)],
              [(0, 12), (3, 40), (10, 100), (50, 400)]),
             ("dna", synth.dna_text(48 << 10, 302), 302, [(3, (1, 5), 411), (4, (2, 6), 412)], [(0, 8), (2, 30), (16, 64)])]
    differs = total = crashed = 0
    for name, text, seed, groups, gaps in specs:
        qs = []
        for k, lens, qseed in groups:
            qs += synth.sample_queries(text, 40, k, lens, gaps, qseed, word_aligned=False)
        lines = [synth.to_benchmark_syntax(q) for q in qs]
        got = run_wcs3(text.tobytes(), lines)
        lib = O.reference_locate(text.tobytes(), [synth.to_library_syntax(q).decode("latin-1") for q in qs])
        per = []
        for q, starts, lr in zip(qs, got, lib):
            if starts is None:
                crashed += 1
                continue
            subs, gp = q
            lo = len(subs[0]) + gp[0][0]
            rep = all(len(s) <= lo for s in subs[:-1])
            lib_starts = [t[0] for t in lr[2]] if lr[0] == "Q" else None
            d = lib_starts != starts
            differs += d
            total += 1
            per.append({"pattern": synth.to_benchmark_syntax(q).decode("latin-1"), "representable": rep, "starts": starts,
                        "differs_from_library": bool(d)})
        cases.append({"text": {"generator": "synth.%s_text" % name, "n": int(text.size), "seed": seed}, "sha256": synth.sha256(text),
                      "queries": per})
    json.dump({"source": "oracle/_ref/ref_wcs3 (the reference's index_wcsearch3::search, index_wcsearch3.hpp:399-422)",
               "differs_from_library": differs, "queries_total": total, "reference_crashed_on": crashed, "cases": cases},
              open(os.path.join(HERE, "wcsearch3_ref.json"), "w"), indent=None, separators=(",", ":"))
    print("crashed", crashed, "queries", total, "differ from the library matcher", differs,
          "representable", sum(q["representable"] for c in cases for q in c["queries"]),
          "matches", sum(len(q["starts"]) for c in cases for q in c["queries"]))


if __name__ == "__main__":
    main()
