#!/usr/bin/env python3
"""Config-1 sized gate call (1 candidate x 60 history questions): wall time per call and the host-side checkpoints of
dvb_question_gate (DVB_GATE_TRACE=1) for the last call.  python tools/prof_small_gate.py   (GPU box)"""
import json, os, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from deep_vision_b200 import _lib, question_similarity as qs
ctx = _lib.Context.get(0)
g9 = json.loads((ROOT / "tests" / "golden" / "a8_a9_gate.json").read_text(encoding="utf-8"))
gate = qs.QuestionGate(ctx=ctx)
for _ in range(30):
    gate.match([g9["candidates"][0]], g9["history"], 0.62)
t0 = time.perf_counter()
for _ in range(300):
    gate.match([g9["candidates"][0]], g9["history"], 0.62)
print("us per match", 1e6 * (time.perf_counter() - t0) / 300)
os.environ["DVB_GATE_TRACE"] = "1"
gate.match([g9["candidates"][0]], g9["history"], 0.62)
