#!/usr/bin/env python3
"""Config-2 annotation rate (rows a11-a14): turns annotated/s through annotations.backfill_sessions — two kernel launches
per batch of sessions (text features + option containment) and the reference's decision tree on the host — next to
the oracle restatement (oracle/dv_annot.py, plain Python, one core) on a sample.
  python tools/bench_annot.py --sessions 2000 --turns 40"""
import argparse
import copy
import json
import random
import sys
import time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tools"))

ap = argparse.ArgumentParser()
ap.add_argument("--sessions", type=int, default=2000)
ap.add_argument("--turns", type=int, default=40)
ap.add_argument("--batch", type=int, default=2048)
ap.add_argument("--oracle-sessions", type=int, default=50)
args = ap.parse_args()
from make_golden_annot import make_log
from deep_vision_b200 import _lib, annotations as an

rng = random.Random(5)
sessions = [{"session_id": "s%d" % s, "interview_log": [make_log(rng) for _ in range(args.turns)]} for s in range(args.sessions)]
n_turns = args.sessions * args.turns
ctx = _lib.Context.get(0)
ctx.set_timing(True)
work = copy.deepcopy(sessions[:64]); an.backfill_sessions(work, ctx=ctx)          # warm-up
l0 = ctx.launches
work = copy.deepcopy(sessions)
t0 = time.time()
for i in range(0, len(work), args.batch):
    an.backfill_sessions(work[i:i + args.batch], ctx=ctx)
gpu_s = time.time() - t0
launches = ctx.launches - l0
# feature extraction alone (the GPU part + packing)
turns = []
for s in sessions:
    for log in s["interview_log"]:
        turns.append(an._log_texts(log))
t0 = time.time()
an.extract_turn_features(turns, ctx=ctx)
feat_s = time.time() - t0
from oracle import dv_annot
sample = copy.deepcopy(sessions[:args.oracle_sessions])
t0 = time.time()
for s in sample:
    dv_annot.backfill(s)
cpu_s = time.time() - t0
print(json.dumps({"sessions": args.sessions, "turns": n_turns, "backfill_s": gpu_s, "turns_per_s": n_turns / gpu_s, "kernel_launches": launches,
                  "feature_extraction_s": feat_s, "feature_turns_per_s": n_turns / feat_s,
                  "oracle_python_turns_per_s_one_core": args.oracle_sessions * args.turns / cpu_s}))
