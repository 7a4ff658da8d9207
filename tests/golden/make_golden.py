"""Generate golden vectors FROM THE UNMODIFIED REFERENCE (oracle/_ref/ref_vlg, built by oracle/Makefile
from /root/reference).  Run in the build container (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

Writes tests/golden/appendix_b.json (hand-picked vectors of SURVEY.md Appendix B + the six queries of
examples/vlg_matching.cpp:40-42) and tests/golden/random_ref.json (seeded random texts/queries).
Every "expect" field is the reference's own output; nothing is computed by the oracle or the product.
"""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import pyoracle as O  # noqa: E402

FIXED = [
    ("abracadabrasimsalabim", ["ac.{2,5}?a.{4,8}?b", "a.{0,10}?a.{0,10}?a", "foo.{0,10}?bar", "a", "abra",
                               "a.{1,2}b", "a.{3,2}?b", "a.{0,3}?", "br.{0,100}?im"]),
    ("aaaaaaaaaa", ["a.{0,0}?a", "a.{1,2}?a", "aa.{0,1}?aa", "a", "aa", "aaa", "a.{0,0}?a.{0,0}?a", "aaaaaaaaaa"]),
    ("aaaaaaaaaaa", ["a", "aa", "a.{0,0}?a"]),
    ("abababababab", ["ab.{0,2}?ab.{0,2}?ab", "b", "a", "ba.{1,1}?b"]),
    ("abcabcabcabc", ["a.{1,1}?c.{0,3}?b", "c.{0,0}?a", "abc.{0,0}?abc"]),
    ("xaxxbxxaxbxxb", ["a.{1,2}?b", "a.{0,1}?b", "x.{0,0}?x", "x"]),
    ("mississippi", ["ss.{0,3}?ss", "i.{0,2}?s.{0,2}?p", "i.{1,1}?i", "s.{0,9}?i.{0,9}?i", "i", "s", "p",
                     "i.{0,100}?i.{0,100}?i.{0,100}?i"]),
    ("the cat and the hat and the bat", ["the.{1,10}?and", "the.{1,10}?and.{1,10}?the", " ", "at.{0,20}?at",
                                         "the.{0,3}?at", "t.{0,0}?h.{0,0}?e"]),
]


def run(text, queries):
    res = O.reference_locate(text.encode("latin-1") if isinstance(text, str) else text, queries)
    out = []
    for q, r in zip(queries, res):
        if r[0] == "Q":
            out.append({"query": q, "count": r[1], "tuples": [list(t) for t in r[2]]})
        else:
            out.append({"query": q, "error": r[1]})
    return out


def main():
    assert O.have_reference(), "build oracle/_ref first: make -C oracle ref"
    fixed = [{"text": t, "cases": run(t, qs)} for t, qs in FIXED]
    with open(os.path.join(HERE, "appendix_b.json"), "w") as f:
        json.dump({"source": "oracle/_ref/ref_vlg locate (unmodified reference)", "vectors": fixed}, f, indent=1)

    rng = random.Random(20261104)
    rnd = []
    for t in range(48):
        sigma = rng.choice([1, 2, 2, 3, 4, 4, 8, 26])
        n = rng.choice([1, 2, 5, 20, 60, 200, 420, 1000, 2500])
        alpha = "abcdefghijklmnopqrstuvwxyz"[:sigma]
        text = "".join(rng.choice(alpha) for _ in range(n))
        qs = []
        for _ in range(40):
            k = rng.choice([1, 1, 2, 2, 2, 3, 3, 4, 6])
            parts = []
            for i in range(k):
                m = min(n, rng.choice([1, 1, 2, 2, 3, 4, 9, 12]))
                if rng.random() < 0.75:
                    o = rng.randrange(0, n - m + 1)
                    sub = text[o:o + m]
                else:
                    sub = "".join(rng.choice(alpha) for _ in range(m))
                parts.append(sub)
                if i + 1 < k:
                    a = rng.choice([0, 0, 1, 2, 3, 5, 7, 20, 100])
                    b = a + rng.choice([0, 0, 1, 2, 5, 11, 50, 300])
                    parts.append(".{%d,%d}?" % (a, b))
            qs.append("".join(parts))
        rnd.append({"text": text, "cases": run(text, qs)})
    with open(os.path.join(HERE, "random_ref.json"), "w") as f:
        json.dump({"source": "oracle/_ref/ref_vlg locate (unmodified reference)", "seed": 20261104, "vectors": rnd}, f)
    nq = sum(len(v["cases"]) for v in fixed + rnd)
    print("wrote", nq, "reference-generated cases")


if __name__ == "__main__":
    main()
