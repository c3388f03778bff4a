#!/usr/bin/env python3
"""Config-4 measurement: question-similarity gate, 1 M candidate questions vs 50 k history (char-bigram Jaccard >= 0.62,
most recent match wins) through dvb_question_gate with host buffers (H2D, key build, inverted index, match, D2H).
  python tools/bench_gate.py --candidates 1000000 --history 50000
One process per GPU (torchrun): history replicated, candidates block-partitioned, no exchange."""
import argparse
import json
import os
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--candidates", type=int, default=1000000)
ap.add_argument("--history", type=int, default=50000)
ap.add_argument("--threshold", type=float, default=0.62)
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from deep_vision_b200 import _lib, parallel, question_similarity as qs, synth
from deep_vision_b200.packing import ptr


def questions(n, seed):
    """n questions of 20-40 synthetic characters + full-width '?': (code points, cp offsets [n+1])"""
    rng = np.random.default_rng(seed)
    lens = rng.integers(20, 41, size=n)
    off = np.zeros(n + 1, np.int64); np.cumsum(lens + 1, out=off[1:])
    cp = synth.code_points(rng, int(off[-1]))
    cp[off[1:] - 1] = 0xFF1F
    return cp, off


def pack(cp, off):
    data, boff = synth.utf8_encode(cp)
    return data, boff[off]


hcp, hoff = questions(args.history, 31)
lo, hi = parallel.shard_range(args.candidates, rank, world)
ccp, coff = questions(hi - lo, 1000 + rank)
# 2 % of the candidates are a history question with its first character replaced (one edit)
rng = np.random.default_rng(5 + rank)
parts, pos = [], 0
ncand = hi - lo
pick = np.nonzero(rng.random(ncand) < 0.02)[0]
src = rng.integers(0, args.history, size=len(pick))
pieces = []
sel = dict(zip(pick.tolist(), src.tolist()))
new_cp, new_off = [], [0]
for i in range(ncand):
    if i in sel:
        h = sel[i]; seg = hcp[hoff[h]:hoff[h + 1]].copy(); seg[0] = 0x4E00 + int(rng.integers(0, 3000))
    else:
        seg = ccp[coff[i]:coff[i + 1]]
    new_cp.append(seg); new_off.append(new_off[-1] + len(seg))
ccp = np.concatenate(new_cp); coff = np.asarray(new_off, np.int64)
cu, co = pack(ccp, coff)
hu, ho = pack(hcp, hoff)
ctx = _lib.Context.get(lr)
gate = qs.QuestionGate(ctx=ctx)
ctx.set_timing(True)
out = None
wall, dev = [], []
for _ in range(args.reps):
    torch.cuda.synchronize(); t0 = time.time()
    out = gate.match(None, None, args.threshold, packed=(cu, co, hu, ho))
    torch.cuda.synchronize(); wall.append(time.time() - t0); dev.append(ctx.last_kernel_ms)
hits = int((out >= 0).sum())
expected = len(pick)
w = min(wall)
if world > 1:
    t = torch.tensor([w], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); w = float(t.item())
if rank == 0:
    print(json.dumps({"world": world, "candidates": args.candidates, "history": args.history, "threshold": args.threshold,
                      "call_s_host_buffers": w, "device_ms": min(dev), "candidates_per_s": args.candidates / w,
                      "effective_pairs_per_s": args.candidates * args.history / w,
                      "matches_rank0": hits, "planted_rank0": expected, "input_bytes_rank0": int(cu.nbytes + hu.nbytes)}))
if world > 1:
    dist.destroy_process_group()
