#!/usr/bin/env python3
"""CPU rows (i) and (ii) of SURVEY.md 8(d): the REFERENCE'S OWN Python functions (AST-loaded from /root/reference/web/server.py
by oracle/ref_loader.py) on stated sub-samples of configs 2, 3 and 4 — one core, then `multiprocessing` over all cores.
Build container only (the reference does not exist on the GPU box).  Writes profiles/r02_reference_python_cpu.json.
  python tools/bench_reference_python.py"""
import json
import multiprocessing as mp
import os
import platform
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np  # noqa: E402
from deep_vision_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402

_NS = None


def ns():
    global _NS
    if _NS is None:
        _NS = ref_loader.load()
    return _NS


def _sessions(n, seed):
    c = synth.make_sessions(n, seed=seed)
    claims, facts = c.texts()
    out = []
    for s in range(n):
        f0, f1 = int(c.sess_fact_off[s]), int(c.sess_fact_off[s + 1])
        pack = {"facts": [{"q_id": "Q%d" % (i - f0 + 1), "question": facts[i][:41], "answer": facts[i][41:], "dimension": synth.DIMENSIONS[int(c.fact_dim[i])],
                           "quality_score": float(c.fact_quality[i]), "hard_triggered": bool(c.fact_hard[i])} for i in range(f0, f1)],
                "quality_snapshot": {"average_quality_score": 0.5}, "dimension_coverage": {}}
        items = [("risks", {"risk": claims[j]}) for j in range(int(c.sess_claim_off[s]), int(c.sess_claim_off[s + 1]))]
        out.append((pack, items))
    return out


def work_a2(arg):                                   # config 2: sessions of 24 claims x 40 facts
    n, seed = arg
    f = ns()["infer_weak_evidence_refs_v3"]
    pairs = 0
    for pack, items in _sessions(n, seed):
        for field, item in items:
            f(field, item, pack)
        pairs += len(items) * len(pack["facts"])
    return pairs


def work_a7(arg):                                   # config 3: count scoring loop of select_reference_material_context over chunks
    n_chunks, n_queries, seed = arg
    data, off = synth.make_texts(n_chunks, 512, seed=seed)
    raw = data.tobytes()
    chunks = [{"chunk_id": "c%d" % i, "index": i, "start": 0, "end": 512, "text": raw[off[i]:off[i + 1]].decode("utf-8")} for i in range(n_chunks)]
    qd, qo = synth.make_texts(n_queries, 45, seed=seed + 1)
    qraw = qd.tobytes()
    n = ns()
    for q in range(n_queries):
        text = qraw[qo[q]:qo[q + 1]].decode("utf-8")[:22] + chunks[q % n_chunks]["text"][100:122]
        ref_loader.select_context_with_chunks(n, chunks, query_text=text, chunk_limit=8)
    return n_chunks * n_queries


def work_a9(arg):                                   # config 4: candidates against a history
    n_cand, n_hist, seed = arg
    rng = np.random.default_rng(seed)
    def qs(k, sd):
        d, o = synth.make_texts(k, 30, seed=sd)
        r = d.tobytes()
        return [r[o[i]:o[i + 1]].decode("utf-8") + "？" for i in range(k)]
    hist = [{"question": t} for t in qs(n_hist, 31)]
    cands = qs(n_cand, seed)
    f = ns()["find_similar_interview_question_log"]
    for c in cands:
        f(hist, c, 0.62)
    return n_cand * n_hist


def work_a12(arg):                                  # bulk re-index body: annotation backfill per session
    n, seed = arg
    import copy
    import random
    sys.path.insert(0, str(ROOT / "tools"))
    from make_golden_annot import make_log
    rng = random.Random(seed)
    f = ns()["backfill_session_interview_log_evidence_annotations"]
    turns = 0
    for s in range(n):
        sess = {"session_id": "s", "interview_log": [make_log(rng) for _ in range(40)]}
        f(sess)
        turns += 40
    return turns


def timed(fn, args_list, procs):
    """the AST load of the reference (about a second) happens before the clock starts, in every process"""
    if procs == 1:
        ns()
        t0 = time.perf_counter()
        units = sum(fn(a) for a in args_list)
        return units, time.perf_counter() - t0
    with mp.Pool(procs, initializer=ns) as pool:
        pool.map(_noop, range(procs * 4))
        t0 = time.perf_counter()
        units = sum(pool.map(fn, args_list))
        return units, time.perf_counter() - t0


def _noop(_):
    return 0


def main():
    cores = os.cpu_count() or 1
    rows = []
    plans = [
        ("a2 infer_weak_evidence_refs_v3 (web/server.py:24780)", "config 2 sample: sessions of 24 claims x 40 facts (~2 KB)", "pairs", work_a2,
         [(6, 1)], [(6, 100 + i) for i in range(cores)]),
        ("a7 select_reference_material_context (web/server.py:32054)", "config 3 sample: 4 000 chunks x 512 chars, 40 queries, top-8", "pairs", work_a7,
         [(4000, 40, 5)], [(4000, 40, 50 + i) for i in range(cores)]),
        ("a9 find_similar_interview_question_log (web/server.py:20494)", "config 4 sample: 150 candidates x 2 000 history questions", "pairs", work_a9,
         [(150, 2000, 7)], [(150, 2000, 70 + i) for i in range(cores)]),
        ("a12 backfill_session_interview_log_evidence_annotations (web/server.py:21469)", "config 2 sample: sessions of 40 turns", "turns", work_a12,
         [(250, 3)], [(250, 30 + i) for i in range(cores)]),
    ]
    for name, sample, unit, fn, one, many in plans:
        u1, t1 = timed(fn, one, 1)
        un, tn = timed(fn, many, cores)
        rows.append({"function": name, "sample": sample, "unit": unit + "/s",
                     "one_core": {"cores": 1, "threads": 1, "seconds": round(t1, 3), "units": u1, "units_per_s": u1 / t1},
                     "all_cores": {"cores": cores, "processes": cores, "seconds": round(tn, 3), "units": un, "units_per_s": un / tn,
                                   "note": "seconds include synthetic input generation inside each worker"}})
        print(json.dumps(rows[-1], ensure_ascii=False), flush=True)
    out = {"what": "reference's own Python functions, AST-loaded unmodified from /root/reference/web/server.py, timed in the build container",
           "python_version": platform.python_version(), "machine": platform.processor() or platform.machine(), "cpu_count": cores, "rows": rows}
    (ROOT / "profiles" / "r02_reference_python_cpu.json").write_text(json.dumps(out, indent=1, ensure_ascii=False), encoding="utf-8")


if __name__ == "__main__":
    main()
