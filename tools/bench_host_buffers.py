#!/usr/bin/env python3
"""Host-buffer calls with PAGEABLE inputs (numpy arrays, what the ctypes shim passes for Python callers): config-2 bind and a
config-3 style index build, wall clock per call.  DVB_H2D_STAGED=0 switches the multi-threaded pinned staging off (A/B)."""
import json
import os
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from deep_vision_b200 import _lib, evidence, reference_material as rm, synth

n_sess = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
ctx = _lib.Context.get(0)
c = synth.make_sessions(n_sess, seed=synth.SEED)
args = (c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off, c.fact_quality, c.fact_dim, c.fact_hard,
        c.claim_pref_dim, c.claim_threshold)
bind = []
for _ in range(5):
    t0 = time.perf_counter()
    evidence.bind_packed(ctx, *args, max_claim_group_bytes=c.max_claim_group_bytes)
    bind.append(time.perf_counter() - t0)
data, off = synth.make_texts(300000, 512, seed=5)
build = []
for _ in range(4):
    t0 = time.perf_counter()
    ix = rm.ChunkIndex(packed=(data, off), ctx=ctx)
    build.append(time.perf_counter() - t0)
    ix.close()
print(json.dumps({"staged": os.environ.get("DVB_H2D_STAGED", "1") != "0", "bind_sessions": n_sess, "bind_input_MB": (c.fact_utf8.nbytes + c.claim_utf8.nbytes) / 1e6,
                  "bind_call_ms": [round(1e3 * x, 1) for x in bind], "bind_pairs_per_s": c.n_pairs / min(bind),
                  "index_text_MB": data.nbytes / 1e6, "index_build_call_ms": [round(1e3 * x, 1) for x in build],
                  "index_chunks_per_s": 300000 / min(build)}))
