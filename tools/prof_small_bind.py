#!/usr/bin/env python3
"""Latency of a one-report dvb_bind_claims call with host buffers (the reference's own call shape), with the one-copy staging
of small calls and with DVB_NO_SMALL_STAGE=1 (a copy per array).  python tools/prof_small_bind.py   (GPU box)"""
import os, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from deep_vision_b200 import _lib, evidence, synth

ctx = _lib.Context.get(0)
for claims, turns in ((1, 40), (24, 40), (100, 30)):
    c = synth.make_sessions(1, turns=turns, claims_per_session=claims, seed=3)
    a = (c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off,
         c.fact_quality, c.fact_dim, c.fact_hard, c.claim_pref_dim, c.claim_threshold)
    res = {}
    for mode in ("one copy", "per array"):
        if mode == "per array": os.environ["DVB_NO_SMALL_STAGE"] = "1"
        else: os.environ.pop("DVB_NO_SMALL_STAGE", None)
        for _ in range(50): got = evidence.bind_packed(ctx, *a, max_claim_group_bytes=c.max_claim_group_bytes)
        ts = []
        for _ in range(400):
            t0 = time.perf_counter(); got = evidence.bind_packed(ctx, *a, max_claim_group_bytes=c.max_claim_group_bytes); ts.append(time.perf_counter() - t0)
        res[mode] = (np.median(ts) * 1e6, got)
    assert np.array_equal(res["one copy"][1]["best_score"], res["per array"][1]["best_score"])
    print(f"claims {claims:4d} turns {turns}: one copy {res['one copy'][0]:.1f} us, per array {res['per array'][0]:.1f} us per call (p50 of 400)")
