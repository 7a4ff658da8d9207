#!/usr/bin/env python
"""bench.py — the driver-facing benchmark of the VLG hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            our arm (one rank per GPU under torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...   the reference's own CPU matcher on the host cores

Headline workload (config.workload): BASELINE.json configs[1] — 200 MiB synthetic English-like byte text PER GPU,
1000 two-subpattern queries with gap [100,1000].  N > 1 is weak scaling over text: the whole text is N x 200 MiB,
cut into N contiguous shards that overlap by the maximum query span; every rank answers all 1000 queries on its
shard and the ranks exchange (count, cursor, first start) per query inside the library (vlgb_search_sharded: one
NCCL all-gather per round).  A step = one pass of the whole query batch over the resident text.

value  = effective text GB/s = Q * total_text_bytes / step time, text index resident in HBM, results left on the
         device (readable through vlgb_result_device_starts).
e2e    = same metric through the reference-facing call with HOST buffers: the pattern-file image in host memory is
         parsed, packed, searched and the match starts land in host memory inside the timed region.
config4 (extra block in the same JSON line) = BASELINE.json configs[3]: 8 GiB text in 8 shards of 1 GiB spread over
         the N ranks (8/N shards per GPU, cursors chained exactly inside a rank), 10 000 two-subpattern queries —
         the configuration the north-star's 8-GPU efficiency target is stated on (strong scaling: total work fixed).
rank_path (extra block, N = 1) = single rare-subpattern queries: latency of the scan path vs the rank path
         (suffix-array wavelet tree) on the same 200 MiB index.
The index (text upload / rank-structure build) is a one-time cost, reported as index_build_ms / rank_build_ms
(north_star: construction is reported separately).
"""
from __future__ import annotations

import argparse
import hashlib
import importlib.util
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

SHARD_BYTES = 200 << 20
NQ = 1000
QUERY_SOURCE_BYTES = 8 << 20  # queries are sampled from the first 8 MiB of shard 0 (every rank can make them)
WORKLOAD = "cfg2: 200MiB synthetic English-like text per GPU, 1000 two-subpattern queries, gap [100,1000]"
CFG4_SHARDS, CFG4_SHARD_BYTES, CFG4_NQ = 8, 1 << 30, 10000
REF_FALLBACK_SAMPLE = 32 << 20  # cpu_baseline leg when no cached full-size reference index exists on the box


def load_synth():
    """vlg_matching_b200/synth.py WITHOUT importing the package (its __init__ maps libvlgb200.so: the reference arm
    must not load the product library at all)."""
    spec = importlib.util.spec_from_file_location("vlgb_synth", os.path.join(ROOT, "vlg_matching_b200", "synth.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


_SYNTH = None


def synth():
    global _SYNTH
    if _SYNTH is None:
        _SYNTH = load_synth()
    return _SYNTH


def shard_text(g: int, n: int):
    return synth().english_text(n, 2 + 1000 * g)


def make_queries(nq=NQ):
    src = shard_text(0, QUERY_SOURCE_BYTES)
    return synth().sample_queries(src, nq, 2, (3, 7), [(100, 1000)], 12, word_aligned=True)


def cfg4_shard_text(g: int, n: int):
    return synth().english_text(n, 4 + 1000 * g)


def cfg4_queries(nq=CFG4_NQ):
    src = cfg4_shard_text(0, QUERY_SOURCE_BYTES)
    return synth().sample_queries(src, nq, 2, (3, 7), [(100, 1000)], 14, word_aligned=True)


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (B200_PROFILING.md, clocks line).  Sampled in-process
    through NVML (a few microseconds per sample): a `nvidia-smi` subprocess every 100 ms takes driver locks
    that stall the benchmark's own CUDA calls for milliseconds.  nvidia-smi is the fallback without pynvml."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        super().__init__(daemon=True)
        self.device, self.stop_flag = device, threading.Event()
        self.sm, self.sm_max, self.reasons, self.n = [], [], set(), 0
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices: map through CUDA_VISIBLE_DEVICES when it holds plain indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = device
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                phys = int(vis.split(",")[device])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
        except Exception:
            self.nv = None

    def sample(self):
        if self.nv is not None:
            nv = self.nv
            self.sm.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
            self.sm_max.append(int(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)))
            r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            for name, bit in (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown),
                              ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                              ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown),
                              ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap)):
                if r & bit:
                    self.reasons.add(name)
        else:
            out = subprocess.run(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-i",
                                  str(self.device)], capture_output=True, text=True, timeout=5).stdout.strip()
            r = [x.strip() for x in out.split(",")]
            self.sm.append(int(float(r[0])))
            self.sm_max.append(int(float(r[1])))
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)
        self.n += 1

    def run(self):
        while not self.stop_flag.is_set():
            try:
                self.sample()
            except Exception:
                pass
            self.stop_flag.wait(0.005 if self.nv is not None else 0.25)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(self.sm_max) if self.sm_max else None,
                "reasons": sorted(self.reasons), "samples": self.n, "source": "nvml" if self.nv is not None else "nvidia-smi"}


def kernel_source_hash() -> str:
    """sha256 over the kernel sources (csrc/*.cuh: every kernel of the search path lives in a header; engine.cu is the host
    side of the ABI): profiles/traffic.json records it, so that DRAM traffic captured for other kernels is not reported
    beside this run's times (tools/make_profiles.py writes the same hash)."""
    d = os.path.join(ROOT, "vlg_matching_b200", "csrc")
    h = hashlib.sha256()
    for f in sorted(os.listdir(d)):
        if f.endswith(".cuh"):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def bind_to_gpu_numa_node(device: int):
    """Pin this process to the CPUs next to its GPU (NVML's ideal affinity).  With one process per GPU and eight GPUs on
    two sockets, pinned result buffers otherwise land on whatever node the process started on and every device-to-host
    copy crosses the socket link (measured: 58 MB took 4 ms with 8 ranks copying at once, 1 ms alone)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        phys = int(vis.split(",")[device]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else device
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(phys))
    except Exception:
        pass


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------------------------- reference arm
def ref_index_path(text) -> str:
    """The reference index is cached per box: its build is the expensive part (minutes) and the driver runs the arm at
    every N of the scaling run."""
    h = hashlib.sha256(np.ascontiguousarray(text).tobytes()).hexdigest()[:24]
    return os.path.join(tempfile.gettempdir(), f"vlgb_ref_{h}.sdsl")


def cpu_reference_run(queries, steps, warmup, threads, sample_bytes, build_if_missing=True):
    """Time the reference's own CPU matcher (oracle/_ref/ref_vlg: the unmodified sdsl vlg_index built from
    /root/reference by oracle/Makefile) on the first `sample_bytes` of the bench text (the whole 200 MiB config when
    sample_bytes = SHARD_BYTES), all queries, all host threads sharing one index, W + K passes in one process.
    Falls back to the oracle port if the binary did not travel."""
    from oracle import pyoracle as O
    S = synth()
    text = shard_text(0, sample_bytes)
    n = text.size
    info = {}
    if O.have_reference():
        xf = ref_index_path(text)
        with tempfile.TemporaryDirectory() as d:
            qf = os.path.join(d, "q.txt")
            with open(qf, "wb") as f:
                for q in queries:
                    f.write(S.to_library_syntax(q) + b"\n")
            if not os.path.exists(xf):
                if not build_if_missing:
                    return None
                tf = os.path.join(d, "text.bin")
                text.tofile(tf)
                tmp = xf + f".{os.getpid()}.tmp"
                b = json.loads(subprocess.run([O.REF_BIN, "build", tf, tmp], check=True, capture_output=True, cwd=d).stdout)
                os.replace(tmp, xf)
                info["index_build_s"] = b["build_s"]
                with open(xf + ".json", "w") as f:
                    json.dump(b, f)
            elif os.path.exists(xf + ".json"):
                info["index_build_s"] = json.load(open(xf + ".json")).get("build_s")
                info["index_cached"] = True
            r = json.loads(subprocess.run([O.REF_BIN, "bench", xf, qf, str(threads), "--index", "--steps", str(steps), "--warmup",
                                           str(warmup)], check=True, capture_output=True, cwd=d).stdout)
            t = r["search_s"]
            info["num_results"], info["checksum"], info["index_load_s"] = r["num_results"], r["checksum"], r["build_s"]
        kind = "reference"
    else:
        times = []
        for i in range(warmup + steps):
            dt, m, cs = O.search_timed(text, queries, threads)
            if i >= warmup:
                times.append(dt)
            info["num_results"], info["checksum"] = m, cs
        t = float(np.mean(times))
        kind = "port"
    gbps = len(queries) * n / t / 1e9
    return {"kind": kind, "seconds_per_step": t, "gbps": gbps, "n": n, "info": info}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    queries = make_queries()
    sample = int(os.environ.get("VLGB_REF_SAMPLE_MIB", str(SHARD_BYTES >> 20))) << 20
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    r = cpu_reference_run(queries, steps, warmup, threads, sample)
    full = r["n"] == SHARD_BYTES
    what = (f"the whole {r['n'] >> 20} MiB text of ONE shard" if full else f"the first {r['n'] >> 20} MiB of the text")
    sample_txt = (f"{what}, all {len(queries)} queries, {threads} threads sharing one index, {warmup}+{steps} passes in one process; "
                  f"index build {r['info'].get('index_build_s')} s (one-time, cached in {tempfile.gettempdir()}, not timed)")
    workload = WORKLOAD if full else f"cfg2 SAMPLE: first {r['n'] >> 20} MiB of the 200MiB text, 1000 two-subpattern queries, gap [100,1000]"
    line = {"impl": "reference", "metric": "effective_text_GBps", "value": r["gbps"], "unit": "GB/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "queries_per_s": len(queries) / r["seconds_per_step"],
            "config": {"workload": workload, "text_bytes_total": r["n"], "queries": len(queries),
                       "note": "the CPU matcher runs ONE shard's text whatever --gpus is (it does not shard)",
                       "matches": r["info"].get("num_results"), "num_results": r["info"].get("num_results"),
                       "checksum": r["info"].get("checksum")},
            "cpu_baseline": {"value": r["gbps"], "unit": "GB/s", "cores": threads, "kind": r["kind"], "sample": sample_txt,
                             **r["info"]},
            "e2e": {"value": r["gbps"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the end-to-end leg")
    ap.add_argument("--no-config4", action="store_true", help="skip the 8 GiB / 10 000-query block")
    ap.add_argument("--keep-stage-timing", action="store_true", help="debug: leave the library's per-stage CUDA events on everywhere")
    ap.add_argument("--no-rank-path", action="store_true", help="skip the rare-pattern latency block")
    ap.add_argument("--shard-bytes", type=int, default=SHARD_BYTES, help="debug only: smaller shard")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    # stdout carries exactly ONE line (the JSON): anything libraries print there (NCCL's version banner) goes to stderr
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from vlg_matching_b200 import capi
    from vlg_matching_b200 import sharded

    S = synth()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if capi.device_count() < 1:
        raise SystemExit("no CUDA device: this benchmark has no CPU path")
    bind_to_gpu_numa_node(local)  # before the CUDA context and any pinned allocation exist
    torch.cuda.set_device(local)
    capi.init(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    sharded.comm_init(rank, world, dist if world > 1 else None)  # the library's own NCCL communicator
    warmup = max(args.warmup, 3)
    steps = args.steps
    shard_bytes = args.shard_bytes

    # ---- inputs
    queries = make_queries()
    span = sharded.batch_span(queries)
    own = shard_text(rank, shard_bytes)
    total = shard_bytes * world
    host_text = np.concatenate([own, shard_text(rank + 1, span)]) if rank + 1 < world else own
    t0 = time.perf_counter()
    index = capi.Index.build(host_text, global_offset=rank * shard_bytes, owned_len=shard_bytes, total_len=total)
    index_build_ms = (time.perf_counter() - t0) * 1e3
    query_strings = [S.to_benchmark_syntax(q) for q in queries]
    pattern_file = b"\n".join(query_strings) + b"\n"  # what gm_search reads: one pattern per line
    batch = capi.PackedBatch(queries)
    lib_stream = torch.cuda.ExternalStream(capi.lib.vlgb_stream_handle(), device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(lib_stream)
        for _ in range(k):
            step_fn()
        e1.record(lib_stream)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def allsum(x: int) -> int:
        if world == 1:
            return int(x)
        hi, lo = int(x) >> 32, int(x) & 0xFFFFFFFF  # exact in float64 up to 2^53 each
        t = torch.tensor([hi, lo], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return ((int(t[0].item()) << 32) + int(t[1].item())) & 0xFFFFFFFFFFFFFFFF

    state = {}

    def step_resident():
        rs, gcounts, st = capi.search_sharded([index], batch, capi.WANT_STARTS | capi.KEEP_ON_DEVICE)
        state["matches"], state["shard"] = int(gcounts.sum()), st
        state["stats"] = capi.last_stats()
        state["dev_ptr"] = rs[0].device_starts(0)
        rs[0].free()

    def parse_batch():
        return capi.ParsedBatch(pattern_file, capi.SYNTAX_BENCHMARK)  # pattern-file image in host memory -> CSR batch, one ABI call

    def e2e_search(flags, b=None):
        if b is None:
            b = parse_batch()
        rs, gcounts, _ = capi.search_sharded([index], b, capi.WANT_STARTS | flags)
        state["e2e_gcounts"] = gcounts
        return rs[0], len(b)

    from concurrent.futures import ThreadPoolExecutor
    parser = ThreadPoolExecutor(max_workers=1)

    def e2e_read(r, nq):
        # the step's result is READ on the host: first start of the first and of the last query (waits for the copy)
        return sum(int(r.starts(q)[:1].sum()) for q in (0, nq - 1))

    def keep_result(r):
        old = state.pop("e2e_result", None)
        if old is not None:
            old.free()
        state["e2e_result"] = r  # kept for the checksum below (read outside the timed region)

    def step_e2e_serial():
        # one call at a time: the call returns with the starts in host memory
        r, nq = e2e_search(0)
        e2e_read(r, nq)
        keep_result(r)

    def run_e2e_pipelined(k):
        # a serving loop: batch i + 1 is parsed and searched while the starts of batch i travel to the host
        # (VLGB_ASYNC_COPY); every result is read on the host before it is freed, the last one before the clock stops
        # the pattern file of step i + 1 is parsed by a helper thread while step i is searched (the parser is host code
        # behind the same C ABI; every step still parses its own file image, inside the timed region)
        prev = None
        nxt = parser.submit(parse_batch)
        for i in range(k):
            b = nxt.result()
            if i + 1 < k:
                nxt = parser.submit(parse_batch)
            r, nq = e2e_search(capi.ASYNC_COPY, b)
            if prev is not None:
                e2e_read(*prev)
                prev[0].free()
            prev = (r, nq)
        e2e_read(*prev)
        keep_result(prev[0])

    for _ in range(warmup):
        step_resident()
    sampler = ClockSampler(local)
    sampler.start()
    # per-stage device time over the timed region (events on the library's stream)
    agg = {}
    launches = 0
    exch = []

    def step_resident_acc():
        nonlocal launches
        step_resident()
        exch.append(state["shard"]["exchange_ms"])
        for name, s in state["stats"]["stages"].items():
            a = agg.setdefault(name, {"ms": 0.0, "bytes": 0, "launches": 0})
            a["ms"] += s["ms"]
            a["bytes"] += s["bytes"]
            a["launches"] += s["launches"]
            launches += s["launches"]

    # pass 1, the library's per-stage CUDA events ON: the stage table and the roofline come from this timed loop
    ms_total_staged = timed(step_resident_acc, steps)
    # pass 2, the same loop with the per-stage events OFF (vlgb_set_stage_timing: ~50 event records and one extra stream
    # synchronisation per search are measurement, not search): `value` / `ms_per_step` come from this one
    capi.check(capi.lib.vlgb_set_stage_timing(1 if args.keep_stage_timing else 0))
    for _ in range(2):
        step_resident()
    ms_total = timed(step_resident, steps)
    clocks = sampler.summary()
    checksum = num_results = None
    ms_e2e_serial = float("nan")
    if args.no_e2e:
        ms_e2e = float("nan")
        state["e2e_h2d"] = state["e2e_d2h"] = 0
    else:
        for _ in range(warmup):
            step_e2e_serial()
        ms_e2e_serial = timed(step_e2e_serial, steps)
        run_e2e_pipelined(warmup)
        ms_e2e = timed(lambda: run_e2e_pipelined(steps), 1)
        st = capi.last_stats()  # byte counters of the last search of the loop (the library counts what it copies)
        state["e2e_h2d"], state["e2e_d2h"] = st["h2d_bytes"], st["d2h_bytes"]
        state["e2e_matches"] = int(state["e2e_gcounts"].sum())
        # gm_search's checksum (gm_search.cpp:110-114): sum of all match starts, here over all ranks
        r = state["e2e_result"]
        cs = 0
        cnt = r.counts()
        for q in range(len(batch)):
            if cnt[q]:
                cs = (cs + int(r.starts(q).sum(dtype=np.uint64))) & 0xFFFFFFFFFFFFFFFF
        checksum = allsum(cs)
        num_results = allsum(int(cnt.sum()))
        r.free()

    ms_step = ms_total / steps
    value = NQ * total / (ms_step * 1e-3) / 1e9
    e2e_value = NQ * total / (ms_e2e / steps * 1e-3) / 1e9

    # ---- rank path: single rare-subpattern queries, scan vs suffix-array wavelet tree (N = 1 only)
    rank_block = None
    rare = []
    if world == 1 and not args.no_rank_path:
        info = index.attach_rank()
        rare = S.sample_queries(own[:QUERY_SOURCE_BYTES], 32, 2, (9, 12), [(100, 1000)], 21, word_aligned=True)
        lat = {}
        for name, path in (("scan", capi.PATH_SCAN), ("rank", capi.PATH_RANK), ("auto", capi.PATH_AUTO)):
            capi.set_path(path)
            bs = [capi.PackedBatch([q]) for q in rare]
            for b in bs[:4]:
                index.search(b, capi.WANT_STARTS).free()
            t0 = time.perf_counter()
            m = 0
            for b in bs:
                r = index.search(b, capi.WANT_STARTS)
                m += r.total()
                r.free()
            lat[name] = {"us_per_query": (time.perf_counter() - t0) / len(bs) * 1e6, "matches": int(m),
                         "path_used": capi.last_stats()["list_path"]}
        capi.set_path(capi.PATH_AUTO)
        rank_block = {"what": "32 single queries, two subpatterns of 9-12 bytes (rare), gap [100,1000], one query per call, "
                              "host wall clock per call incl. result copy",
                      "latency": lat, "rank_build_ms": info["build_ms"], "suffix_sort_ms": info["sa_ms"], "levels": info["levels"],
                      "doubling_rounds": info["doubling_rounds"], "rank_bytes": info["bytes"]}
    index.free()

    # ---- config 4: 8 GiB in 8 shards over the N ranks, 10 000 queries
    cfg4 = None
    if not args.no_config4 and CFG4_SHARDS % world == 0:
        q4 = cfg4_queries()
        span4 = sharded.batch_span(q4)
        per = CFG4_SHARDS // world
        shards = []
        t0 = time.perf_counter()
        for j in range(per):
            g = rank * per + j
            txt = cfg4_shard_text(g, CFG4_SHARD_BYTES)
            if g + 1 < CFG4_SHARDS:
                txt = np.concatenate([txt, cfg4_shard_text(g + 1, span4)])
            shards.append(capi.Index.build(txt, global_offset=g * CFG4_SHARD_BYTES, owned_len=CFG4_SHARD_BYTES,
                                           total_len=CFG4_SHARDS * CFG4_SHARD_BYTES))
            del txt
        build4_s = time.perf_counter() - t0
        b4 = capi.PackedBatch(q4)
        st4 = {}

        def step4():
            rs, gc, st = capi.search_sharded(shards, b4, capi.WANT_STARTS | capi.KEEP_ON_DEVICE)
            st4["matches"], st4["shard"] = int(gc.sum()), st
            for r in rs:
                r.free()

        step4()  # warm-up (arena growth)
        step4()  # and once more: the pools of result blocks settle in the second pass
        k4 = 2
        ms4 = timed(step4, k4) / k4
        # checksum of all match starts: one more pass with the starts in host memory
        rs, gc, st = capi.search_sharded(shards, b4, capi.WANT_STARTS)
        cs = 0
        for r in rs:
            cnt = r.counts()
            for q in range(len(b4)):
                if cnt[q]:
                    cs = (cs + int(r.starts(q).sum(dtype=np.uint64))) & 0xFFFFFFFFFFFFFFFF
            r.free()
        total4 = CFG4_SHARDS * CFG4_SHARD_BYTES
        cfg4 = {"workload": "cfg4: 8 GiB synthetic English-like text in 8 shards of 1 GiB (+ halo) over the ranks, 10000 two-subpattern "
                            "queries, gap [100,1000]", "scaling": "strong", "shards_per_gpu": per, "ms_per_step": ms4, "steps": k4,
                "warmup": 1, "value": CFG4_NQ * total4 / (ms4 * 1e-3) / 1e9, "unit": "GB/s", "queries_per_s": CFG4_NQ / (ms4 * 1e-3),
                "matches": st4["matches"], "num_results": int(gc.sum()), "checksum": allsum(cs), "exchange_ms": st4["shard"]["exchange_ms"],
                "rounds": st4["shard"]["rounds"], "shards_reentered": st4["shard"]["reentered_shards"],
                "text_gen_plus_upload_s": build4_s}
        for s in shards:
            s.free()

    if rank == 0:
        peak, peak_src = load_peaks()
        # dominant kernel = the stage with the largest device time
        top = max(agg.items(), key=lambda kv: kv[1]["ms"])
        name, a = top
        per_launch_ms = a["ms"] / max(a["launches"], 1)
        per_launch_bytes = a["bytes"] / max(a["launches"], 1)
        achieved = per_launch_bytes / (per_launch_ms * 1e-3) / 1e9 if per_launch_ms > 0 else 0.0
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                tj = json.load(open(tp))
                traffic = tj.get(name) if tj.get("source_hash") == kernel_source_hash() else None  # stale capture: say nothing
            except Exception:
                traffic = None
        stages = {k: {"ms_per_step": v["ms"] / steps, "launches_per_step": v["launches"] / steps,
                      "GBps": (v["bytes"] / (v["ms"] * 1e-3) / 1e9 if v["ms"] > 0 else 0.0)} for k, v in agg.items()}
        # the scan STAGE as a whole (count + offsets + fill): SURVEY §8(d) bytes of one pass = n + 4 B per occurrence
        scan_ms = sum(agg[k]["ms"] for k in ("scan_count", "scan_offsets", "scan_fill") if k in agg) / steps
        scan_bytes = host_text.size + 4 * state["stats"]["total_occurrences"]
        stage_sum = sum(v["ms"] for v in agg.values()) / steps
        ms_step_staged = ms_total_staged / steps
        line = {"metric": "effective_text_GBps", "value": value, "unit": "GB/s", "n_gpus": world, "steps": steps,
                "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u32", "data": "synthetic", "queries_per_s": NQ / (ms_step * 1e-3),
                "config": {"workload": WORKLOAD, "text_bytes_total": total, "queries": NQ, "sharding": f"text x{world}, overlap {span} B",
                           "l2": "text (200 MiB) and occurrence lists exceed the 126 MB L2; no flush needed",
                           "matches": state["matches"], "num_results": num_results, "checksum": checksum,
                           "shards_reentered": state["shard"]["reentered_shards"], "exchange_rounds": state["shard"]["rounds"],
                           "occurrences": state["stats"]["total_occurrences"], "scan_passes": state["stats"]["scan_passes"],
                           "list_path": state["stats"]["list_path"], "host_syncs_per_step": state["stats"]["host_syncs"]},
                "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": state["e2e_h2d"],
                        "d2h_bytes_per_step": state["e2e_d2h"], "ms_per_step": ms_e2e / steps,
                        "queries_per_s": NQ / (ms_e2e / steps * 1e-3),
                        "mode": "pipelined serving loop: pattern-file image in host memory -> vlgb_parse_lines -> "
                                "vlgb_search_sharded(WANT_STARTS | ASYNC_COPY) -> match starts in pinned host memory; the pattern file of step "
                                "i + 1 is parsed on a helper thread during step i, the "
                                "device-to-host copy of step i overlaps the search of step i + 1, every result is read on the "
                                "host before it is freed, the last copy is waited for inside the timed region",
                        "serial_ms_per_step": ms_e2e_serial / steps,
                        "serial_note": "the same without ASYNC_COPY: one blocking call per step (latency of a single call)"},
                "gpu_launches": launches, "index_build_ms": index_build_ms, "clocks": clocks,
                "exchange_ms": float(np.mean(exch)) if exch else 0.0, "stage_ms_sum": stage_sum, "host_gap_ms": ms_step_staged - stage_sum,
                "ms_per_step_stage_timers_on": ms_step_staged,
                "timing_note": "ms_per_step / value / e2e: per-stage CUDA events of the library off (vlgb_set_stage_timing(0)); "
                               "stages, stage_ms_sum, host_gap_ms, roofline: the same timed loop run once more with them on",
                "roofline": {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                             "bytes_per_launch": per_launch_bytes, "ms_per_launch": per_launch_ms,
                             "stage_frac": {"scan": scan_bytes / (scan_ms * 1e-3) / 1e9 / peak if scan_ms > 0 else None,
                                            "scan_ms": scan_ms, "scan_bytes": scan_bytes}},
                "stages": stages}
        if cfg4 is not None:
            line["config4"] = cfg4
        if rank_block is not None:
            line["rank_path"] = rank_block
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            r = cpu_reference_run(queries, 1, 0, threads, SHARD_BYTES, build_if_missing=False)  # cached by the reference arm?
            if r is None:
                r = cpu_reference_run(queries, 1, 0, threads, REF_FALLBACK_SAMPLE)
            what = "the whole 200 MiB text" if r["n"] == SHARD_BYTES else f"first {r['n'] >> 20} MiB of the text"
            line["cpu_baseline"] = {"value": r["gbps"], "unit": "GB/s", "cores": threads, "kind": r["kind"],
                                    "sample": f"{what}, all {NQ} queries, {threads} threads, 1 pass", **r["info"]}
            if r["n"] == SHARD_BYTES and r["kind"] == "reference" and rank_block is not None:
                # the rare single queries of the rank_path block through the reference, one thread, same 200 MiB index
                rr = cpu_reference_run(rare, 3, 1, 1, SHARD_BYTES, build_if_missing=False)
                if rr is not None:
                    line["cpu_baseline"]["rare_single_query_us"] = rr["seconds_per_step"] / len(rare) * 1e6
                    line["cpu_baseline"]["rare_matches"] = rr["info"].get("num_results")
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    capi.comm_destroy()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
