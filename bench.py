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
    ap.add_argument("--index-chunks", type=int, default=1000000, help="config 3: chunks per GPU behind the `index` key (0 = skip)")
    ap.add_argument("--index-queries", type=int, default=100000, help="config 3: claim queries, top-8")
    ap.add_argument("--gate-candidates", type=int, default=1000000, help="config 4: candidate questions per GPU behind the `gate` key (0 = skip)")
    ap.add_argument("--gate-history", type=int, default=50000)
    ap.add_argument("--no-latency", action="store_true", help="skip the config-1 latency and Python-API arms")
    return ap.parse_args()


def workload_config(args, world):
    return {"workload": "config2: synthetic %d sessions/GPU x %d Q/A turns (~2 KB/turn) x %d claims, claim->fact weak binding"
                        % (args.sessions, args.turns, args.claims),
            "sessions_per_gpu": args.sessions, "turns": args.turns, "claims_per_session": args.claims,
            "parallelism": "session-sharded x%d (no collective)" % world,
            "l2_policy": "inputs (0.88 GB/GPU) larger than L2; no flush needed"}


def make_texts_blocked(n, chars, seed, block=125000):
    """synth.make_texts in blocks (bounded host memory): n texts of `chars` code points, packed CSR."""
    from deep_vision_b200 import synth
    parts, offs, base = [], [np.zeros(1, np.int64)], 0
    for i, b0 in enumerate(range(0, n, block)):
        d, o = synth.make_texts(min(block, n - b0), chars, seed=seed + 7919 * i)
        nb = int(o[-1])
        parts.append(d[:nb]); offs.append(o[1:] + base); base += nb
    parts.append(np.zeros(64, np.uint8))
    return np.concatenate(parts), np.concatenate(offs)


def index_arm(ctx, args, rank, world=1):
    """BASELINE metric part 2 on config 3: N chunks/GPU x 512 chars resident in HBM -> dvb_chunk_index_build (chunks
    indexed/s), then the indexed top-8 of Q claim queries (effective (query, chunk) pairs/s) with terms and candidates
    resident on the device; at N > 1 the sharded top-k (one all_gather_into_tensor + device merge).  Rank-local numbers;
    the caller reduces over ranks."""
    import torch
    from deep_vision_b200 import parallel, reference_material as rm, synth
    n, nq, k = args.index_chunks, args.index_queries, 8
    t0 = time.perf_counter()
    data, off = make_texts_blocked(n, 512, synth.SEED + 1000 + rank)
    gen_s = time.perf_counter() - t0
    d_data, d_off = torch.from_numpy(data).cuda(), torch.from_numpy(off).cuda()
    builds, ix = [], None
    for _ in range(4):                                            # the first build grows the allocation pool
        if ix is not None:
            ix.close()
        torch.cuda.synchronize()
        ix = rm.ChunkIndex(device_ptrs=(d_data.data_ptr(), d_off.data_ptr(), n), index_base=rank * n, ctx=ctx)
        builds.append(ix.build_ms)
    st = ix.stats()
    # queries: 30-60 chars, every other one half copied from a chunk of THIS rank's shard (replicated from rank 0 below)
    qdata, qoff = synth.make_texts(nq, 45, seed=999)
    raw, qraw = data.tobytes(), qdata.tobytes()
    rng = np.random.default_rng(7)
    src_chunk = rng.integers(0, n, size=nq); src_pos = rng.integers(0, 488, size=nq)
    queries = []
    for q in range(nq):
        text = qraw[qoff[q]:qoff[q + 1]].decode("utf-8")
        if q % 2 == 0:
            c = int(src_chunk[q]); a = int(src_pos[q])
            text = text[:22] + raw[off[c]:off[c + 1]].decode("utf-8")[a:a + 22]
        queries.append(text)
    del raw
    if world > 1:
        import torch.distributed as dist
        box = [queries]
        dist.broadcast_object_list(box, src=0)
        queries = box[0]
    # a6 on the device: query texts -> term symbol CSR (dvb_query_terms); the host extractor is timed on a sample for comparison
    t0 = time.perf_counter()
    sample_terms = [rm._tokenize_reference_material_query(t) for t in queries[:2000]]
    tok_s = (time.perf_counter() - t0) * nq / max(1, min(nq, 2000))
    rm.query_terms_batch(queries[:256], ctx=ctx, device=True)
    t0 = time.perf_counter()
    dev_terms = rm.query_terms_batch(queries, ctx=ctx, device=True)
    torch.cuda.synchronize()
    sym_s = time.perf_counter() - t0
    terms_kernel_ms = ctx.last_kernel_ms
    q_ms, li = [], None
    for _ in range(2):
        li, ls, ln = ix.topk_device(dev_terms, nq, k)
        q_ms.append(ctx.last_kernel_ms)
    li_h, ls_h, ln_h = li.cpu().numpy(), ls.cpu().numpy(), ln.cpu().numpy()
    sharded = None
    if world > 1:
        import hashlib
        import torch.distributed as dist
        sharded = {}
        for label, m in (("1000_queries", min(1000, nq)), ("all_queries", nq)):
            sub = rm.query_terms_batch(queries[:m], ctx=ctx, device=True) if m < nq else dev_terms
            local_ms, wall = [], []
            for _ in range(3):
                ix.topk_device(sub, m, k); local_ms.append(ctx.last_kernel_ms)
                dist.barrier(); torch.cuda.synchronize()
                t0 = time.perf_counter()
                gi, gs, gn = parallel.sharded_index_topk_device(ix, sub, m, k)
                torch.cuda.synchronize()
                wall.append(1000.0 * (time.perf_counter() - t0))
            digest = int.from_bytes(hashlib.sha256(gi.tobytes() + gs.tobytes() + gn.tobytes()).digest()[:7], "little")
            lo_hi = torch.tensor([digest, -digest], dtype=torch.int64, device="cuda")
            dist.all_reduce(lo_hi, op=dist.ReduceOp.MAX)
            sharded[label] = {"queries": m, "local_device_ms": min(local_ms), "ms_with_allgather_merge_and_d2h": min(wall),
                              "identical_on_all_ranks": bool(int(lo_hi[0]) == -int(lo_hi[1])), "allgather_bytes": world * m * k * 16}
    # parity: exact scores of a query slice over the WHOLE shard by the C oracle (all host threads), rank 0 only
    parity = None
    if rank == 0 and not args.no_cpu_baseline:
        from oracle import dv_oracle        # checker (allowed use)
        dv_oracle.build()
        cores = os.cpu_count() or 1
        ok, checked, t0 = True, 0, time.perf_counter()
        for q in range(0, nq, max(1, nq // 6))[:6]:
            tq = dv_oracle.query_terms(queries[q])
            if not tq:
                continue
            sc = dv_oracle.chunk_scores(None, tq, n_threads=cores, packed=(data, off))
            wi, ws = dv_oracle.topk(sc, k)
            nv = int(ln_h[q])
            ok = ok and (li_h[q][:nv] - rank * n).tolist() == wi.tolist() and ls_h[q][:nv].tolist() == ws.tolist()
            checked += 1
        parity = {"vs": "oracle C exact substring counts over all %d chunks for %d queries of the batch" % (n, checked),
                  "topk_indices_and_scores_identical": bool(ok), "seconds": time.perf_counter() - t0}
    ix.close()
    best = min(builds[1:])
    return {"chunks_per_gpu": n, "queries": nq, "build_ms": best, "first_build_ms": builds[0], "postings": st["postings"],
            "text_bytes": st["text_bytes"], "algorithmic_bytes": st["text_bytes"] + 4 * st["postings"],
            "query_ms": min(q_ms), "query_algorithmic_bytes_lower_bound": 8 * int(ls_h.sum()) + 16 * k * nq,
            "host_gen_s": gen_s, "host_tokenize_s_python_extractor_estimate": tok_s, "terms_call_s": sym_s, "terms_kernel_ms": terms_kernel_ms,
            "terms_match_host_extractor_on_sample": bool(rm.terms_from_symbols(*[a.cpu().numpy() for a in rm.query_terms_batch(queries[:2000], ctx=ctx, device=True)]) == sample_terms),
            "sharded": sharded, "parity": parity}


def _gather_segments(src, src_off, idx):
    """concatenation of src[src_off[i]:src_off[i+1]] for i in idx (vectorised) -> (values, offsets[len(idx)+1])"""
    lens = src_off[idx + 1] - src_off[idx]
    out_off = np.zeros(len(idx) + 1, np.int64); np.cumsum(lens, out=out_off[1:])
    pos = np.arange(int(out_off[-1]), dtype=np.int64) - np.repeat(out_off[:-1], lens) + np.repeat(src_off[idx], lens)
    return src[pos], out_off


def gate_arm(ctx, args, rank, world=1):
    """Config 4: question-similarity gate, C candidate questions/GPU (20-40 chars + '？') against a replicated history of
    H questions, char-bigram Jaccard >= 0.62, most recent match wins; 2 % of the candidates are a history question with
    one character edited.  Host buffers through dvb_question_gate (H2D, normalise + keys, inverted index, match, D2H)."""
    import torch
    from deep_vision_b200 import question_similarity as qs, synth
    nc, nh, thr = args.gate_candidates, args.gate_history, 0.62

    def questions(n, seed):
        rng = np.random.default_rng(seed)
        lens = rng.integers(20, 41, size=n)
        off = np.zeros(n + 1, np.int64); np.cumsum(lens + 1, out=off[1:])
        cp = synth.code_points(rng, int(off[-1]))
        cp[off[1:] - 1] = 0xFF1F
        return cp, off

    hcp, hoff = questions(nh, 31)
    ccp, coff = questions(nc, 1000 + rank)
    rng = np.random.default_rng(5 + rank)
    pick = np.nonzero(rng.random(nc) < 0.02)[0]
    src_h = rng.integers(0, nh, size=len(pick))
    both = np.concatenate([ccp, hcp]); both_off = np.concatenate([coff, hoff[1:] + coff[-1]])
    which = np.arange(nc, dtype=np.int64); which[pick] = nc + src_h
    cand_cp, cand_off = _gather_segments(both, both_off, which)
    cand_cp = cand_cp.copy()
    cand_cp[cand_off[pick]] = 0x4E00 + rng.integers(0, 3000, size=len(pick)).astype(np.uint32)     # one edit: the first character
    cu, cbo = synth.utf8_encode(cand_cp); co = cbo[cand_off]
    hu, hbo = synth.utf8_encode(hcp); ho = hbo[hoff]
    gate = qs.QuestionGate(ctx=ctx)
    wall, dev, out = [], [], None
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        out = gate.match(None, None, thr, packed=(cu, co, hu, ho))
        torch.cuda.synchronize(); wall.append(1000.0 * (time.perf_counter() - t0)); dev.append(ctx.last_kernel_ms)
    parity = None
    if rank == 0 and not args.no_cpu_baseline:
        from oracle import dv_oracle        # checker (allowed use)
        dv_oracle.build()
        og = dv_oracle.Gate(qs.DEFAULT_SYNONYMS, qs.DEFAULT_STOPWORDS)
        cb = cu.tobytes()
        sel = np.unique(np.concatenate([pick[:150], np.arange(0, nc, max(1, nc // 150))[:150]]))
        cu2, co2 = dv_oracle.pack([cb[co[i]:co[i + 1]].decode("utf-8") for i in sel])
        t0 = time.perf_counter()
        want = np.asarray(og.gate(None, None, thr, n_threads=os.cpu_count() or 1, packed=(cu2, co2, hu, ho)), np.int64)
        parity = {"vs": "oracle C gate: %d candidates (150 planted near-duplicates) x all %d history questions" % (len(sel), nh),
                  "match_index_identical": bool(np.array_equal(want, out[sel])), "planted_recalled": int((out[pick] >= 0).sum()),
                  "planted": int(len(pick)), "seconds": time.perf_counter() - t0}
    in_bytes = int(co[-1] + ho[-1])
    return {"candidates_per_gpu": nc, "history": nh, "threshold": thr, "call_ms": min(wall), "device_ms": min(dev),
            "matches": int((out >= 0).sum()), "input_bytes": in_bytes, "algorithmic_bytes_lower_bound": in_bytes + 8 * nc, "parity": parity}


def latency_arm(ctx):
    """Config 1 (the only shape a live request has): per-call latency of the three drop-in-bound functions on the
    fixtures session — 8 bindable claims x 31 facts (tests/golden/a2_bind.json, pack C1), the reference's own tail-recall
    document (tests/test_api_comprehensive.py:3754-3806, 8 chunks), 60 history questions (tests/golden/a8_a9_gate.json)."""
    import types
    from deep_vision_b200 import dropin
    gold = ROOT / "tests" / "golden"
    a2 = json.loads((gold / "a2_bind.json").read_text(encoding="utf-8"))
    g9 = json.loads((gold / "a8_a9_gate.json").read_text(encoding="utf-8"))
    mod = types.SimpleNamespace(load_reference_material_chunks=None)
    b = dropin.build_bindings(mod, ctx)
    pack = a2["packs"]["C1"]
    cases = [c for c in a2["cases"] if c["pack"] == "C1" and c["field"] in ("risks", "actions", "open_questions")]
    doc_text = "开头普通说明。" * 1800 + "\n\n尾部关键需求：必须支持离线审批和断网补录，并保留审批链路证据。"
    chunks = b["_build_reference_material_chunks"](doc_text)
    mod.load_reference_material_chunks = lambda doc: chunks
    logs = [{"question": h} for h in g9["history"]]

    def pct(us):
        us = np.sort(np.asarray(us))
        return {"p50_us": float(us[len(us) // 2]), "p99_us": float(us[min(len(us) - 1, int(len(us) * 0.99))]), "calls": int(len(us))}

    out, ok = {}, True
    t = []
    for rep in range(40):
        for c in cases:
            t0 = time.perf_counter(); r = b["infer_weak_evidence_refs_v3"](c["field"], c["item"], pack, c["min_score"]); t.append(1e6 * (time.perf_counter() - t0))
            ok = ok and r == c["result"]
    out["infer_weak_evidence_refs_v3"] = dict(pct(t[len(cases):]), shape="%d claims x %d facts, one call per claim" % (len(cases), len(pack["facts"])),
                                             equals_reference_golden=bool(ok))
    t = []
    for rep in range(200):
        t0 = time.perf_counter(); ctx_text, meta = b["select_reference_material_context"]({"doc_id": "d"}, query_text="离线审批能力验证", chunk_limit=4)
        t.append(1e6 * (time.perf_counter() - t0))
    out["select_reference_material_context"] = dict(pct(t[5:]), shape="%d chunks, 1 query" % len(chunks),
                                                   tail_chunk_selected=bool(meta["selected_chunks"] and meta["selected_chunks"][-1]["index"] == len(chunks) - 1))
    t, hits = [], 0
    for rep in range(3):
        for i, cand in enumerate(g9["candidates"]):
            t0 = time.perf_counter(); r = b["find_similar_interview_question_log"](logs, cand, g9["threshold"]); t.append(1e6 * (time.perf_counter() - t0))
            want = g9["match"][i]
            ok_i = (r is None and want < 0) or (r is not None and want >= 0 and r is logs[want])
            hits += 1 if ok_i else 0
    out["find_similar_interview_question_log"] = dict(pct(t[5:]), shape="%d history questions" % len(logs),
                                                     equals_reference_golden=bool(hits == len(t)))
    return out


def annot_arm(ctx, args):
    """Bulk re-index path (a11-a14): turns annotated/s through annotations.backfill_sessions on config-2-sized sessions (40
    turns each) — three launches per batch (text features, option containment, decision stage) + Python dict extraction and
    write-back — checked against the plain-Python restatement (oracle/dv_annot.py) on a sample, which is also the CPU row."""
    import copy
    import random
    sys.path.insert(0, str(ROOT / "tools"))
    from make_golden_annot import make_log
    from deep_vision_b200 import annotations as an
    rng = random.Random(5)
    ns = 1000
    sessions = [{"session_id": "s%d" % s_, "interview_log": [make_log(rng) for _ in range(args.turns)]} for s_ in range(ns)]
    an.backfill_sessions(copy.deepcopy(sessions[:32]), ctx=ctx)
    work = copy.deepcopy(sessions)
    l0 = ctx.launches
    t0 = time.perf_counter()
    got = an.backfill_sessions(work, ctx=ctx)
    dt = time.perf_counter() - t0
    out = {"metric": "turns_annotated_per_s", "value": ns * args.turns / dt, "unit": "turns/s", "sessions": ns, "turns": ns * args.turns,
           "seconds": dt, "kernel_launches": int(ctx.launches - l0),
           "workload": "%d synthetic sessions x %d interview-log turns (question, answer, options) through backfill_sessions" % (ns, args.turns)}
    if not args.no_cpu_baseline:
        from oracle import dv_annot          # checker (allowed use): plain-Python restatement of web/server.py:21221-21584
        sample = copy.deepcopy(sessions[:100])
        t0 = time.perf_counter()
        want = [dv_annot.backfill(s_) for s_ in sample]
        cpu = time.perf_counter() - t0
        out["parity"] = {"vs": "oracle/dv_annot.py on 100 sessions", "summaries_and_logs_identical": bool(got[:100] == want and work[:100] == sample)}
        out["cpu_baseline"] = {"value": 100 * args.turns / cpu, "unit": "turns/s", "cores": 1, "kind": "port", "sample": "100 sessions, plain Python restatement"}
    return out


def python_api_arm(ctx, args):
    """What a drop-in caller pays: evidence.bind_batch from Python dicts (packing, pageable numpy buffers, H2D, kernel,
    D2H, result dicts) on a bounded config-2 sample."""
    from deep_vision_b200 import evidence, synth
    ns = min(args.sessions, 500)
    c = synth.make_sessions(ns, turns=args.turns, claims_per_session=args.claims, seed=synth.SEED + 77)
    claims, facts = c.texts()
    fields = ["risks", "actions", "open_questions"]
    keys = {"risks": "risk", "actions": "action", "open_questions": "question"}
    requests = []
    for s_ in range(ns):
        f0, f1 = int(c.sess_fact_off[s_]), int(c.sess_fact_off[s_ + 1])
        pack = {"facts": [{"q_id": "Q%d" % (i - f0 + 1), "question": facts[i][:41], "answer": facts[i][41:], "dimension": synth.DIMENSIONS[int(c.fact_dim[i])],
                           "quality_score": float(c.fact_quality[i]), "hard_triggered": bool(c.fact_hard[i])} for i in range(f0, f1)]}
        for j in range(int(c.sess_claim_off[s_]), int(c.sess_claim_off[s_ + 1])):
            fld = fields[(j - int(c.sess_claim_off[s_])) * 3 // max(1, args.claims)]
            requests.append((fld, {keys[fld]: claims[j]}, pack, 0.46))
    evidence.bind_batch(requests, ctx=ctx)
    wall = []
    for _ in range(3):
        t0 = time.perf_counter(); res = evidence.bind_batch(requests, ctx=ctx); wall.append(time.perf_counter() - t0)
    pairs = ns * args.turns * args.claims
    return {"value": pairs / min(wall), "unit": UNIT, "sample": "%d sessions (%d pairs) as Python dicts through evidence.bind_batch, pageable buffers" % (ns, pairs),
            "ms_per_call": 1000.0 * min(wall), "bound_claims": int(sum(1 for r in res if r["refs"]))}


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
    """dram bytes per dvb_bind_claims call (k_bind_groups + k_bind_claim_tables + k_bind_facts) from the committed ncu --set full capture, if any."""
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
    emit_line(line)


_LINE_FD = None


def quiet_stdout():
    """stdout carries ONE line, the JSON line of rank 0: anything a library prints meanwhile (NCCL announces its version on
    stdout when the box sets NCCL_DEBUG) goes to stderr.  File descriptor 1 is pointed at stderr for the run; emit_line()
    writes to the saved descriptor."""
    global _LINE_FD
    if _LINE_FD is None:
        sys.stdout.flush()
        _LINE_FD = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _LINE_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
        return
    sys.stdout.flush()
    while data:
        data = data[os.write(_LINE_FD, data):]


def main():
    args = parse_args()
    quiet_stdout()
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

    # the corpus of the headline arm is no longer needed on the device
    del dev_in, pinned, pin_np
    torch.cuda.empty_cache()
    ctx.set_stream(None)                                          # the arms below run on the library's own streams

    def index_traffic():
        p_ = ROOT / "profiles" / "index_traffic.json"
        try:
            return json.loads(p_.read_text()).get("dram_bytes_per_build")
        except Exception:
            return None

    # ---------------- config 3: chunks indexed/s + indexed top-8 (every rank indexes its own shard)
    index = None
    if args.index_chunks > 0:
        try:
            mine = index_arm(ctx, args, rank, world)
        except Exception as exc:                                  # the headline line must survive a failure here
            mine = {"error": repr(exc)[:300]}
        ok = reduce_max(1.0 if "error" in mine else 0.0) == 0.0
        if ok:
            b_ms, q_ms = reduce_max(mine["build_ms"]), reduce_max(mine["query_ms"])
            peak_i, _src = measured_peak()
            nq = mine["queries"]
            index = {"metric": "chunks_indexed_per_s", "value": world * mine["chunks_per_gpu"] / (b_ms / 1000.0), "unit": "chunks/s",
                     "workload": "config3: %d chunks/GPU x 512 chars resident in HBM -> positional index (dvb_chunk_index_build); %d claim queries, top-8"
                                 % (mine["chunks_per_gpu"], nq),
                     "build_ms": b_ms, "first_build_ms": mine["first_build_ms"], "postings_per_gpu": mine["postings"],
                     "roofline": {"bound": "hbm", "achieved": mine["algorithmic_bytes"] / (b_ms / 1000.0) / 1e9, "peak": peak_i, "unit": "GB/s",
                                  "frac": mine["algorithmic_bytes"] / (b_ms / 1000.0) / 1e9 / peak_i,
                                  "algorithmic_bytes": mine["algorithmic_bytes"], "traffic": index_traffic(),
                                  "note": "UTF-8 in + 4 B per posting (SURVEY 8d) over the whole build (symbolise, emit, radix sort, heads)"},
                     "query": {"effective_pairs_per_s": world * mine["chunks_per_gpu"] * nq / (q_ms / 1000.0), "queries": nq, "device_ms": q_ms, "k": 8,
                               "ms_per_1000_queries": q_ms * 1000.0 / nq,
                               "roofline": {"bound": "hbm", "achieved": mine["query_algorithmic_bytes_lower_bound"] / (q_ms / 1000.0) / 1e9, "peak": peak_i,
                                            "unit": "GB/s", "frac": mine["query_algorithmic_bytes_lower_bound"] / (q_ms / 1000.0) / 1e9 / peak_i,
                                            "algorithmic_bytes": mine["query_algorithmic_bytes_lower_bound"], "traffic": None,
                                            "note": "lower bound: 8 B per term occurrence in the RETURNED top-k chunks + 16 B x k out (SURVEY 8d counts every chunk; the kernel is bound by dependent-load latency, DESIGN.md section 4)"},
                               "query_terms": {"call_s": mine["terms_call_s"], "kernel_ms": mine["terms_kernel_ms"],
                                               "python_extractor_s_estimate": mine["host_tokenize_s_python_extractor_estimate"],
                                               "equals_host_extractor_on_2000_queries": mine["terms_match_host_extractor_on_sample"],
                                               "note": "a6 (web/server.py:32037-32051) on the device: UTF-8 query texts -> term symbol CSR, dvb_query_terms"}},
                     "host_gen_s": mine["host_gen_s"], "parity": mine["parity"]}
            if mine.get("sharded"):
                sh = {}
                for label, d in mine["sharded"].items():
                    d = dict(d)
                    d["ms_with_allgather_merge_and_d2h"] = reduce_max(d["ms_with_allgather_merge_and_d2h"])
                    d["local_device_ms"] = reduce_max(d["local_device_ms"])
                    d["effective_pairs_per_s"] = world * mine["chunks_per_gpu"] * d["queries"] / (d["ms_with_allgather_merge_and_d2h"] / 1000.0)
                    d["efficiency"] = d["local_device_ms"] / d["ms_with_allgather_merge_and_d2h"]     # sharded rate / (N x the local rate)
                    sh[label] = d
                index["query"]["sharded_topk"] = sh
        else:
            index = {"error": mine.get("error", "failed on another rank")}

    # ---------------- config 4: question-similarity gate (history replicated, candidates partitioned, no exchange)
    gate = None
    if args.gate_candidates > 0:
        try:
            mine = gate_arm(ctx, args, rank, world)
        except Exception as exc:
            mine = {"error": repr(exc)[:300]}
        ok = reduce_max(1.0 if "error" in mine else 0.0) == 0.0
        if ok:
            c_ms, d_ms = reduce_max(mine["call_ms"]), reduce_max(mine["device_ms"])
            peak_g, _src = measured_peak()
            gate = {"metric": "gate_candidates_per_s", "value": world * mine["candidates_per_gpu"] / (c_ms / 1000.0), "unit": "candidates/s",
                    "workload": "config4: %d candidate questions/GPU vs %d history questions, char-bigram Jaccard >= %.2f, host buffers through dvb_question_gate"
                                % (mine["candidates_per_gpu"], mine["history"], mine["threshold"]),
                    "effective_pairs_per_s": world * mine["candidates_per_gpu"] * mine["history"] / (c_ms / 1000.0),
                    "call_ms": c_ms, "device_ms": d_ms, "matches_per_gpu": mine["matches"],
                    "roofline": {"bound": "hbm", "achieved": mine["algorithmic_bytes_lower_bound"] / (d_ms / 1000.0) / 1e9, "peak": peak_g, "unit": "GB/s",
                                 "frac": mine["algorithmic_bytes_lower_bound"] / (d_ms / 1000.0) / 1e9 / peak_g,
                                 "algorithmic_bytes": mine["algorithmic_bytes_lower_bound"], "traffic": None,
                                 "note": "inputs once + 8 B out per candidate; the postings term of SURVEY 8d is not counted (lower bound); limiter: issue / shared-memory, see profiles/"},
                    "parity": mine["parity"]}
        else:
            gate = {"error": mine.get("error", "failed on another rank")}

    latency = api = annot = None
    if rank == 0 and not args.no_latency:
        try:
            latency = latency_arm(ctx)
        except Exception as exc:
            latency = {"error": repr(exc)[:300]}
        try:
            api = python_api_arm(ctx, args)
        except Exception as exc:
            api = {"error": repr(exc)[:300]}
        try:
            annot = annot_arm(ctx, args)
        except Exception as exc:
            annot = {"error": repr(exc)[:300]}

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
                             "traffic": ncu_traffic(), "kernel": "k_bind_groups + k_bind_claim_tables + k_bind_facts (one dvb_bind_claims call)", "kernel_ms": k_ms,
                             "algorithmic_bytes_per_launch": corpus.algorithmic_bytes,
                             "bytes_per_pair": corpus.algorithmic_bytes / corpus.n_pairs, "peak_source": peak_src,
                             "limiter": "instruction issue / ALU pipe, not DRAM: 2.6 warp instructions per input byte; the fused kernel was bound by "
                                        "instruction FETCH (more than ~12 KB of hot code per SM halves the issue rate: experiments/icache), the split "
                                        "pair issues 0.6 (facts) and 0.5 (claim tables) per scheduler cycle; DRAM traffic 1.2 x algorithmic "
                                        "(profiles/r02_bind8_ncu_full.json, profiles/README.md; DESIGN.md section 5)"},
                "clocks": clocks,
                "pairs_per_step_per_gpu": corpus.n_pairs}
        if index is not None:
            line["index"] = index
        if gate is not None:
            line["gate"] = gate
        if latency is not None:
            line["latency"] = latency
        if api is not None:
            line["e2e_python_api"] = api
        if annot is not None:
            line["annot"] = annot
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
        emit_line(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
