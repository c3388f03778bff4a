#!/usr/bin/env python3
"""Where the microseconds of a one-report dvb_bind_claims call go: DVB_BIND_TRACE=1 host timestamps of three warm calls
(1 claim x 40 facts and 100 claims x 30 facts) and the wall time of the Python wrapper around them.  (GPU box)"""
import os, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from deep_vision_b200 import _lib, evidence, synth

ctx = _lib.Context.get(0)
for claims, turns in ((1, 40), (100, 30)):
    c = synth.make_sessions(1, turns=turns, claims_per_session=claims, seed=3)
    a = (c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off,
         c.fact_quality, c.fact_dim, c.fact_hard, c.claim_pref_dim, c.claim_threshold)
    for _ in range(50): evidence.bind_packed(ctx, *a, max_claim_group_bytes=c.max_claim_group_bytes)
    os.environ["DVB_BIND_TRACE"] = "1"
    for _ in range(3):
        t0 = time.perf_counter(); evidence.bind_packed(ctx, *a, max_claim_group_bytes=c.max_claim_group_bytes)
        print(f"claims {claims}: python wall {1e6 * (time.perf_counter() - t0):.0f} us", file=sys.stderr)
    os.environ.pop("DVB_BIND_TRACE")
