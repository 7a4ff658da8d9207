"""GPU, >= 2 devices: the NCCL-sharded search inside the library (vlgb_search_sharded) under torchrun, starts compared
for every query with one index over the whole text (tools/sharded_check.py).  Skipped on a one-GPU box; the host
logic of the exchange is covered on CPU by tests/test_sharded_gloo.py."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def n_gpus():
    from vlg_matching_b200 import capi
    return capi.device_count()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("per_rank,mib", [(1, 48), (2, 16)])
def test_torchrun_two_ranks_equal_single_index(per_rank, mib):
    if n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, SHARD_MIB=str(mib), SHARDS_PER_RANK=str(per_rank), QUERIES="300")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "tools", "sharded_check.py")]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "equal the single-index search on every rank" in out.stdout


def test_single_process_shard_chain_small():
    """One rank, several shards (no NCCL): the exact chain inside vlgb_search_sharded on a small text, all tuples."""
    import numpy as np
    import vlg_matching_b200 as V
    from vlg_matching_b200 import capi, sharded, synth
    capi.init(0)
    text = synth.make_text(2, 1 << 20)
    queries = synth.make_queries(2, text, 120) + synth.sample_queries(text, 60, 3, (1, 3), [(0, 40), (100, 1000)], 5, False)
    queries += [([b"e"], []), ([b" "], []), ([b"th"], [])]
    span = sharded.batch_span(queries)
    n = text.size
    whole = V.Index.build(text)
    b = V.PackedBatch(queries)
    rw = whole.search(b, capi.WANT_TUPLES)
    cuts = [0, n // 5, 2 * n // 5 + 1, 3 * n // 5, 4 * n // 5, n]
    shards = [V.Index.build(text[cuts[g]:min(n, cuts[g + 1] + span)], global_offset=cuts[g], owned_len=cuts[g + 1] - cuts[g], total_len=n)
              for g in range(5)]
    res, gc, st = capi.search_sharded(shards, b, capi.WANT_TUPLES)
    assert np.array_equal(gc, rw.counts())
    for q in range(len(queries)):
        assert np.array_equal(np.concatenate([r.tuples(q) for r in res], axis=0), rw.tuples(q)), (q, queries[q])


def test_gm_search_two_gpus_matches_reference_checksum(tmp_path):
    """gm_search --gpus 2 (one process, one host thread per GPU, NCCL exchange inside the library): num_results and
    checksum equal the reference matcher's on the committed pattern set (tests/golden/gm_search_ref.json)."""
    import json
    if n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    from vlg_matching_b200 import synth
    here = os.path.dirname(os.path.abspath(__file__))
    bin_dir = os.path.join(ROOT, "vlg_matching_b200", "bin")
    g = json.load(open(os.path.join(here, "golden", "gm_search_ref.json")))
    text = synth.english_text(g["text"]["n"], g["text"]["seed"])
    raw = tmp_path / "raw.txt"
    text.tofile(raw)
    col = tmp_path / "col"
    assert subprocess.run([os.path.join(bin_dir, "create_collection.x"), "-i", str(raw), "-c", str(col)]).returncode == 0
    pat = col / "patterns" / "p.txt"
    pat.write_text("\n".join(g["patterns"]) + "\n")
    r = subprocess.run([os.path.join(bin_dir, "gm_search-VLGB200.x"), "-c", str(col), "-p", str(pat), "--gpus", "2"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    keys = dict(line[2:].split(" = ") for line in r.stdout.splitlines() if line.startswith("# ") and " = " in line)
    assert int(keys["gpus"]) == 2
    assert int(keys["num_results"]) == g["num_results"] and int(keys["checksum"]) == g["checksum"]
