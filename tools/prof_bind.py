#!/usr/bin/env python3
"""Small driver for ncu: a few dvb_bind_claims launches on a config-2 style corpus (resident inputs)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from deep_vision_b200 import _lib, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
kw = {}
if len(sys.argv) > 3: kw["claims_per_session"] = int(sys.argv[3])          # > 32: several claim groups per session
if len(sys.argv) > 4: kw["turns"] = int(sys.argv[4])
c = synth.make_sessions(n, seed=synth.SEED, **kw)
ctx = _lib.Context.get(0)
host = [c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off, c.fact_quality, c.fact_dim,
        c.fact_hard, c.claim_pref_dim, c.claim_threshold]
dev = [torch.from_numpy(a).cuda() for a in host]
nc = int(c.sess_claim_off[-1])
ob = torch.empty(nc, dtype=torch.int32, device="cuda"); os_ = torch.empty(nc, dtype=torch.float64, device="cuda")
obd = torch.empty(nc, dtype=torch.uint8, device="cuda")
ctx.set_timing(True)
for _ in range(reps):
    rc = ctx.lib.dvb_bind_claims(ctx.handle, _lib.MEM_DEVICE, *[t.data_ptr() for t in dev[:6]], c.n_sessions,
                                 *[t.data_ptr() for t in dev[6:]], ob.data_ptr(), os_.data_ptr(), obd.data_ptr(), None, None,
                                 c.max_claim_group_bytes)
    ctx.check(rc, "bind")
    print("kernel ms", ctx.last_kernel_ms, "GB/s", c.algorithmic_bytes / ctx.last_kernel_ms / 1e6)
print("launches", ctx.launches)
