#!/usr/bin/env python3
"""Where a config-1 sized call spends its time: cProfile of the three drop-in-bound functions of bench.py's latency arm
(the ctypes call into libdvb200 shows up as one built-in).  python tools/prof_latency.py   (GPU box)"""
import cProfile, json, pstats, sys, types, io
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from deep_vision_b200 import _lib, dropin
ctx = _lib.Context.get(0)
gold = ROOT / "tests" / "golden"
a2 = json.loads((gold / "a2_bind.json").read_text(encoding="utf-8"))
g9 = json.loads((gold / "a8_a9_gate.json").read_text(encoding="utf-8"))
mod = types.SimpleNamespace(load_reference_material_chunks=None)
b = dropin.build_bindings(mod, ctx)
pack = a2["packs"]["C1"]
cases = [c for c in a2["cases"] if c["pack"] == "C1" and c["field"] in ("risks", "actions", "open_questions")]
chunks = b["_build_reference_material_chunks"]("开头普通说明。" * 1800 + "\n\n尾部关键需求：必须支持离线审批和断网补录，并保留审批链路证据。")
mod.load_reference_material_chunks = lambda doc: chunks
logs = [{"question": h} for h in g9["history"]]
def run(name, fn, n):
    for _ in range(20): fn()
    pr = cProfile.Profile(); pr.enable()
    for _ in range(n): fn()
    pr.disable()
    s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(9)
    print("=====", name, "x", n); print("\n".join(l for l in s.getvalue().split("\n") if l.strip())[:2400])
c0 = cases[0]
run("infer_weak_evidence_refs_v3", lambda: b["infer_weak_evidence_refs_v3"](c0["field"], c0["item"], pack, c0["min_score"]), 400)
run("select_reference_material_context", lambda: b["select_reference_material_context"]({"doc_id": "d"}, query_text="离线审批能力验证", chunk_limit=4), 400)
run("find_similar_interview_question_log", lambda: b["find_similar_interview_question_log"](logs, g9["candidates"][0], g9["threshold"]), 400)
