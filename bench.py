#!/usr/bin/env python3
"""bench.py — evidence pairs scored/s on BASELINE.json config 2 (synthetic 10k sessions x 40 Q/A turns
of Chinese text, ~2 KB/turn, 24 claims/session), one process per GPU.

  python bench.py --gpus N --steps K --warmup W            # this framework (CUDA, C-ABI)
  python bench.py --impl reference --gpus N ...            # reference's CPU path (oracle port, all host threads)

A step = one pass of the claim->fact binding path (dvb_bind_claims) over the rank's whole corpus.
`value`  : pairs/s with inputs resident in HBM (CUDA events on the launching stream, max over ranks).
`e2e`    : same metric through the host-buffer C-ABI call (pinned host -> device copies, kernel,
           results back to the host inside the timed region).
`roofline`: algorithmic bytes (all claim+fact UTF-8 bytes once + 16 B/claim) / kernel time.
Sessions shard across ranks with no data-path collective ("scaling": "weak": 10k sessions per GPU).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "evidence_pairs_scored_per_s"
UNIT = "pairs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sessions", type=int, default=10000, help="sessions per GPU (config 2: 10000)")
    ap.add_argument("--turns", type=int, default=40)
    ap.add_argument("--claims", type=int, default=24)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--index-chunks", type=int, default=200000,
                    help="chunks per GPU of the bounded config-3 sample behind the `index` key (0 = skip)")
    return ap.parse_args()


def workload_config(args, world):
    return {"workload": "config2: synthetic %d sessions/GPU x %d Q/A turns (~2 KB/turn) x %d claims, claim->fact weak binding"
                        % (args.sessions, args.turns, args.claims),
            "sessions_per_gpu": args.sessions, "turns": args.turns, "claims_per_session": args.claims,
            "parallelism": "session-sharded x%d (no collective)" % world,
            "l2_policy": "inputs (0.88 GB/GPU) larger than L2; no flush needed"}


def index_arm(ctx, args, rank, world=1):
    """BASELINE metric part 2, chunks indexed/s: a bounded config-3 sample (512-char chunks resident in HBM ->
    dvb_chunk_index_build -> a queryable positional index), plus the indexed top-8 of 1000 claim queries over it.
    Returns this rank's numbers; the caller reduces over ranks."""
    import torch
    from deep_vision_b200 import _lib, reference_material as rm, synth
    n = args.index_chunks
    data, off = synth.make_texts(n, 512, seed=synth.SEED + 1000 + rank)
    d_data, d_off = torch.from_numpy(data).cuda(), torch.from_numpy(off).cuda()
    builds, ix = [], None
    for _ in range(4):                                            # the first build grows the allocation pool
        if ix is not None:
            ix.close()
        torch.cuda.synchronize()
        ix = rm.ChunkIndex(device_ptrs=(d_data.data_ptr(), d_off.data_ptr(), n), index_base=rank * n, ctx=ctx)
        builds.append(ix.build_ms)
    st = ix.stats()
    qdata, qoff = synth.make_texts(1000, 45, seed=999)
    raw, qraw = data.tobytes(), qdata.tobytes()
    rng = np.random.default_rng(7)
    terms = []
    for q in range(1000):
        text = qraw[qoff[q]:qoff[q + 1]].decode("utf-8")
        if q % 2 == 0:                                            # half of every other query is copied from a chunk
            c = int(rng.integers(0, n)); src = raw[off[c]:off[c + 1]].decode("utf-8"); a = int(rng.integers(0, 488))
            text = text[:22] + src[a:a + 22]
        terms.append(rm._tokenize_reference_material_query(text))
    sharded = None
    if world > 1:
        # the queries are replicated (rank 0's), every rank answers them over its shard, ONE all-gather of nq x k
        # candidates over NCCL, merge on every rank (parallel.sharded_index_topk); host wall clock, max over ranks
        import hashlib
        import torch.distributed as dist
        from deep_vision_b200 import parallel
        box = [terms]
        dist.broadcast_object_list(box, src=0)
        terms = box[0]
        wall = []
        for _ in range(3):
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            gi, gs, gn = parallel.sharded_index_topk(ix, terms, 8)
            torch.cuda.synchronize()
            wall.append(1000.0 * (time.perf_counter() - t0))
        digest = int.from_bytes(hashlib.sha256(gi.tobytes() + gs.tobytes() + gn.tobytes()).digest()[:7], "little")
        lo_hi = torch.tensor([digest, -digest], dtype=torch.int64, device="cuda")
        dist.all_reduce(lo_hi, op=dist.ReduceOp.MAX)
        sharded = {"ms_1000_queries_allgather_merge": min(wall), "identical_on_all_ranks": bool(int(lo_hi[0]) == -int(lo_hi[1])),
                   "allgather_bytes": world * len(terms) * 8 * 12}
    q_ms = []
    for _ in range(3):
        ix.topk(terms, 8)
        q_ms.append(ctx.last_kernel_ms)
    ix.close()
    best = min(builds[1:])
    return {"chunks_per_gpu": n, "build_ms": best, "first_build_ms": builds[0], "postings": st["postings"],
            "algorithmic_bytes": st["text_bytes"] + 4 * st["postings"], "query_ms_1000_queries": min(q_ms), "sharded": sharded}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.lines = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram bytes per launch of k_bind_claims from the committed ncu --set full capture, if any."""
    p = ROOT / "profiles" / "bind_traffic.json"
    if p.exists():
        try:
            return json.loads(p.read_text()).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def cpu_baseline(corpus, threads, repeats=1, min_seconds=0.0):
    """Times the oracle port on `threads` host threads: `repeats` passes, then more passes until `min_seconds` of wall
    time were spent.  Returns (pairs/s over all passes, seconds, passes)."""
    from oracle import dv_oracle        # checker timed as the reported CPU baseline (allowed use)
    dv_oracle.build()
    packed = (corpus.claim_utf8, corpus.claim_off, corpus.fact_utf8, corpus.fact_off)
    total, passes = 0.0, 0
    while passes < repeats or total < min_seconds:
        t0 = time.perf_counter()
        cpu_baseline.last = dv_oracle.bind_claims(None, None, corpus.sess_claim_off, corpus.sess_fact_off, corpus.fact_quality, corpus.fact_dim,
                                                  corpus.fact_hard, corpus.claim_pref_dim, corpus.claim_threshold, packed=packed, n_threads=threads)
        total += time.perf_counter() - t0
        passes += 1
    return corpus.n_pairs * passes / total, total, passes


cpu_baseline.last = None


def run_reference(args, rank, world):
    """Reference arm: the reference's CPU algorithm (oracle port; the reference itself is Python that
    cannot travel to the GPU box) on all host threads, rank 0 only."""
    if rank != 0:
        return
    from deep_vision_b200 import synth
    cores = os.cpu_count() or 1
    n_sample = min(args.sessions, 1000 * cores)
    corpus = synth.make_sessions(n_sample, turns=args.turns, claims_per_session=args.claims, seed=synth.SEED)
    for _ in range(args.warmup):
        cpu_baseline(corpus, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_baseline(corpus, cores)
    dt = time.perf_counter() - t0
    value = corpus.n_pairs * args.steps / dt
    sample = "%d of %d sessions/step (%d pairs), oracle C port of web/server.py:24780-24864, %d threads" % (
        n_sample, args.sessions, corpus.n_pairs, cores)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/int32 + f64 score", "data": "synthetic",
            "config": workload_config(args, world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from deep_vision_b200 import _lib, synth
    from deep_vision_b200.packing import ptr

    corpus = synth.make_sessions(args.sessions, turns=args.turns, claims_per_session=args.claims, seed=synth.SEED + rank)
    ctx = _lib.Context.get(local_rank)
    lib, h = ctx.lib, ctx.handle
    hint = corpus.max_claim_group_bytes
    n_claims, n_facts = int(corpus.sess_claim_off[-1]), int(corpus.sess_fact_off[-1])
    host_in = [corpus.claim_utf8, corpus.claim_off, corpus.fact_utf8, corpus.fact_off, corpus.sess_claim_off,
               corpus.sess_fact_off, corpus.fact_quality, corpus.fact_dim, corpus.fact_hard, corpus.claim_pref_dim,
               corpus.claim_threshold]

    # ---------------- resident (device) arm
    dev_in = [torch.from_numpy(a).cuda() for a in host_in]
    d_best = torch.empty(n_claims, dtype=torch.int32, device="cuda")
    d_score = torch.empty(n_claims, dtype=torch.float64, device="cuda")
    d_bound = torch.empty(n_claims, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    ctx.set_timing(True)

    def step_resident():
        rc = lib.dvb_bind_claims(h, _lib.MEM_DEVICE, *[t.data_ptr() for t in dev_in[:6]], corpus.n_sessions,
                                 *[t.data_ptr() for t in dev_in[6:]], d_best.data_ptr(), d_score.data_ptr(),
                                 d_bound.data_ptr(), None, None, hint)
        ctx.check(rc, "dvb_bind_claims")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(args.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    l0 = ctx.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    ev0.record(stream)
    for _ in range(args.steps):
        step_resident()
        kernel_ms.append(ctx.last_kernel_ms)
    ev1.record(stream)
    barrier()
    elapsed_ms = reduce_max(ev0.elapsed_time(ev1))
    launches = ctx.launches - l0
    res_best = d_best.cpu().numpy()
    res_score, res_bound = d_score.cpu().numpy(), d_bound.cpu().numpy()

    # ---------------- end-to-end arm: pinned host buffers through the C-ABI (H2D + kernel + D2H timed)
    pinned = [torch.from_numpy(a).pin_memory() for a in host_in]
    pin_np = [t.numpy() for t in pinned]
    o_best = torch.empty(n_claims, dtype=torch.int32).pin_memory()
    o_score = torch.empty(n_claims, dtype=torch.float64).pin_memory()
    o_bound = torch.empty(n_claims, dtype=torch.uint8).pin_memory()

    def step_e2e():
        rc = lib.dvb_bind_claims(h, _lib.MEM_HOST, *[ptr(a) for a in pin_np[:6]], corpus.n_sessions,
                                 *[ptr(a) for a in pin_np[6:]], o_best.data_ptr(), o_score.data_ptr(), o_bound.data_ptr(),
                                 None, None, hint)
        ctx.check(rc, "dvb_bind_claims")

    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        step_e2e()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(e2e_steps):
        step_e2e()
    e1.record(stream)
    barrier()
    e2e_ms = reduce_max(e0.elapsed_time(e1))
    clocks = sampler.stop() if rank == 0 else None                # sampled over both timed regions (resident + end-to-end)
    assert np.array_equal(o_best.numpy(), res_best), "resident and end-to-end arms disagree"
    h2d = int(sum(a.nbytes for a in host_in)) - 64 * 2
    d2h = n_claims * (4 + 8 + 1)

    # ---------------- chunks indexed/s on a bounded config-3 sample (every rank indexes its own shard)
    index = None
    if args.index_chunks > 0:
        try:
            mine = index_arm(ctx, args, rank, world)
        except Exception as exc:                                  # the headline line must survive a failure here
            mine = {"error": repr(exc)[:200]}
        ok = reduce_max(1.0 if "error" in mine else 0.0) == 0.0
        if ok:
            b_ms, q_ms = reduce_max(mine["build_ms"]), reduce_max(mine["query_ms_1000_queries"])
            peak_i, _src = measured_peak()
            index = {"metric": "chunks_indexed_per_s", "value": world * mine["chunks_per_gpu"] / (b_ms / 1000.0), "unit": "chunks/s",
                     "workload": "config3 sample: %d chunks/GPU x 512 chars resident in HBM -> positional index (dvb_chunk_index_build)" % mine["chunks_per_gpu"],
                     "build_ms": b_ms, "first_build_ms": mine["first_build_ms"], "postings_per_gpu": mine["postings"],
                     "roofline": {"bound": "hbm", "achieved": mine["algorithmic_bytes"] / (b_ms / 1000.0) / 1e9, "peak": peak_i, "unit": "GB/s",
                                  "frac": mine["algorithmic_bytes"] / (b_ms / 1000.0) / 1e9 / peak_i,
                                  "algorithmic_bytes": mine["algorithmic_bytes"], "note": "UTF-8 in + 4 B per posting (SURVEY 8d); the build is 4 radix passes over 12 B postings"},
                     "query": {"effective_pairs_per_s": world * mine["chunks_per_gpu"] * 1000 / (q_ms / 1000.0), "ms_per_1000_queries": q_ms, "k": 8}}
            if mine.get("sharded"):
                sh = dict(mine["sharded"])
                sh["ms_1000_queries_allgather_merge"] = reduce_max(sh["ms_1000_queries_allgather_merge"])
                sh["effective_pairs_per_s"] = world * mine["chunks_per_gpu"] * 1000 / (sh["ms_1000_queries_allgather_merge"] / 1000.0)
                index["query"]["sharded_topk"] = sh
        else:
            index = {"error": mine.get("error", "failed on another rank")}

    pairs_all = corpus.n_pairs * world
    value = pairs_all * args.steps / (elapsed_ms / 1000.0)
    e2e_value = pairs_all * e2e_steps / (e2e_ms / 1000.0)
    k_ms = float(np.mean(kernel_ms))
    peak, peak_src = measured_peak()
    achieved = corpus.algorithmic_bytes / (k_ms / 1000.0) / 1e9

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32 + f64 score", "data": "synthetic",
                "config": workload_config(args, world),
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": ncu_traffic(), "kernel": "k_bind_claims", "kernel_ms": k_ms,
                             "algorithmic_bytes_per_launch": corpus.algorithmic_bytes,
                             "bytes_per_pair": corpus.algorithmic_bytes / corpus.n_pairs, "peak_source": peak_src,
                             "limiter": "instruction issue, not DRAM: 3.8 warp instructions per input byte, 69 % of the issue slots busy, "
                                        "DRAM traffic 1.03 x algorithmic (profiles/r01_bind_v6_config2_ncu_full.json; DESIGN.md section 5)"},
                "clocks": clocks,
                "pairs_per_step_per_gpu": corpus.n_pairs}
        if index is not None:
            line["index"] = index
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            v, secs, passes = cpu_baseline(corpus, cores, repeats=2, min_seconds=10.0)
            ref = cpu_baseline.last                           # the oracle's answer on the FULL workload checks the GPU's (checker use)
            same = bool(np.array_equal(ref["best_fact"], res_best) and np.array_equal(ref["bound"], res_bound)
                        and np.array_equal(ref["best_score"].view(np.uint64), res_score.view(np.uint64)))
            line["parity"] = {"vs": "oracle C port on all %d claims of the bench workload" % n_claims,
                              "best_fact_bound_and_fp64_score_bits_identical": same}
            if not same:
                print("bench.py: GPU results differ from the oracle on the bench workload", file=sys.stderr)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "all %d sessions (%d pairs) x %d passes, %.1f s wall, oracle C port of web/server.py:24780-24864, %d threads"
                                              % (corpus.n_sessions, corpus.n_pairs, passes, secs, cores)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
