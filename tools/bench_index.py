#!/usr/bin/env python3
"""Config-3 style measurement: chunks indexed/s (dvb_chunk_index_build, corpus resident in HBM) and effective
(query, chunk) pairs/s (dvb_chunk_index_topk), one process per GPU (torchrun for N > 1: contiguous shards + NCCL merge).
  python tools/bench_index.py --chunks 1000000 --queries 100000 --k 8"""
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
ap.add_argument("--chunks", type=int, default=1000000, help="global chunk count (sharded over ranks)")
ap.add_argument("--chars", type=int, default=512)
ap.add_argument("--queries", type=int, default=100000)
ap.add_argument("--k", type=int, default=8)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--qreps", type=int, default=0, help="query repetitions (default: --reps)")
ap.add_argument("--replicate", type=int, default=1,
                help="generate 1/R of the shard on the host and tile it R times in HBM (config-5 sized shards without minutes of numpy)")
ap.add_argument("--sidecar", default="", help="also time dvb_chunk_index_save / _load through this path (SURVEY 8(f)1)")
args = ap.parse_args()
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from deep_vision_b200 import _lib, parallel, reference_material as rm, synth
ctx = _lib.Context.get(lr)
lo, hi = parallel.shard_range(args.chunks, rank, world)
t0 = time.time()
n_gen = (hi - lo + args.replicate - 1) // args.replicate
data, off = synth.make_texts(n_gen, args.chars, seed=synth.SEED + 1000 + rank)
gen_s = time.time() - t0
# queries: 30-60 chars, half copied from a chunk of rank 0's shard (same on every rank: generated from seed)
qdata, qoff = synth.make_texts(args.queries, 45, seed=999)
qraw = qdata.tobytes()
rng = np.random.default_rng(7)
queries = [qraw[qoff[q]:qoff[q + 1]].decode("utf-8") for q in range(args.queries)]
if rank == 0:                                   # every other query: half of it copied from a chunk of rank 0's shard
    raw0 = data.tobytes()
    for q in range(0, args.queries, 2):
        c = int(rng.integers(0, n_gen)); s = raw0[off[c]:off[c + 1]].decode("utf-8"); a = int(rng.integers(0, max(1, args.chars - 24)))
        queries[q] = queries[q][:22] + s[a:a + 22]
    del raw0
if world > 1:                                   # queries are replicated: rank 0's list on every rank
    box = [queries]
    dist.broadcast_object_list(box, src=0)
    queries = box[0]
t0 = time.time()
terms = [rm._tokenize_reference_material_query(t) for t in queries]
tok_s = time.time() - t0
d_data = torch.from_numpy(data).cuda(); d_off = torch.from_numpy(off).cuda()
if args.replicate > 1:
    nb = int(off[-1])
    small = d_data
    d_data = torch.zeros(nb * args.replicate + 64, dtype=torch.uint8, device="cuda")
    d_data[: nb * args.replicate].view(args.replicate, nb).copy_(small[:nb][None, :].expand(args.replicate, nb))
    del small
    reps_off = (torch.arange(args.replicate, device="cuda", dtype=torch.int64) * nb)[:, None] + d_off[None, :-1]
    d_off = torch.cat([reps_off.reshape(-1)[: hi - lo], torch.tensor([0], device="cuda", dtype=torch.int64)])
    last = (hi - lo - 1) % n_gen
    d_off[-1] = d_off[-2] + int(off[last + 1] - off[last])
    del reps_off
    torch.cuda.empty_cache()
ctx.set_timing(True)
build_ms = []
ix = None
for _ in range(args.reps):
    if ix is not None:
        ix.close()
    torch.cuda.synchronize()
    ix = rm.ChunkIndex(device_ptrs=(d_data.data_ptr(), d_off.data_ptr(), hi - lo), index_base=lo, ctx=ctx)
    build_ms.append(ix.build_ms)
st = ix.stats()
tsym, toff, qto = rm.terms_to_symbols(terms)
oi = np.full((args.queries, args.k), -1, np.int64); os_ = np.zeros((args.queries, args.k), np.int32); on = np.zeros(args.queries, np.int32)
q_ms = []
for _ in range(args.qreps or args.reps):
    rc = ctx.lib.dvb_chunk_index_topk(ctx.handle, ix._h, _lib.MEM_HOST, rm.ptr(tsym), rm.ptr(toff), len(toff) - 1, rm.ptr(qto), args.queries,
                                      lo, args.k, rm.ptr(oi), rm.ptr(os_), rm.ptr(on))
    ctx.check(rc, "topk")
    q_ms.append(ctx.last_kernel_ms)
side = {}
if args.sidecar and rank == 0:
    p = Path(args.sidecar)
    t0 = time.time(); ix.save(p, bytes(32)); side["save_s"] = time.time() - t0
    side["file_bytes"] = p.stat().st_size
    t0 = time.time(); ix2 = rm.ChunkIndex.load(p, bytes(32), index_base=lo, ctx=ctx); side["load_s"] = time.time() - t0
    nqs = min(args.queries, 2000)
    a = ix2.topk(terms[:nqs], args.k)
    side["same_topk_as_built_index"] = bool(np.array_equal(a[0], oi[:nqs]) and np.array_equal(a[1], os_[:nqs]) and np.array_equal(a[2], on[:nqs]))
    side["save_GBps"] = side["file_bytes"] / side["save_s"] / 1e9; side["load_GBps"] = side["file_bytes"] / side["load_s"] / 1e9
    ix2.close(); p.unlink()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.time()
gi, gs, gn = parallel.merge_topk(oi, os_, args.k, ctx=ctx)
merge_s = time.time() - t0
same = True
if world > 1:                                   # max over ranks of the device times; merged result identical on every rank
    import hashlib
    t = torch.tensor([min(build_ms), min(q_ms), merge_s], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    max_build, max_query, max_merge = (float(x) for x in t.tolist())
    dg = int.from_bytes(hashlib.sha256(gi.tobytes() + gs.tobytes() + gn.tobytes()).digest()[:7], "little")
    t2 = torch.tensor([dg, -dg], dtype=torch.int64, device="cuda")
    dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    same = bool(int(t2[0]) == -int(t2[1]))
else:
    max_build, max_query, max_merge = min(build_ms), min(q_ms), merge_s
algo_build = st["text_bytes"] + 4 * st["postings"]            # UTF-8 in + 4 B per posting
algo_query = 8 * int(os_.sum()) + 16 * args.k * args.queries   # lower bound from the reported top-k only; see DESIGN.md
if rank == 0:
    b, qm = min(build_ms), min(q_ms)
    print(json.dumps({"world": world, "chunks_per_gpu": hi - lo, "queries": args.queries, "k": args.k, "stats": st,
                      "replicate": args.replicate, "segments_est": -(-st["text_bytes"] // (2000 << 20)), "build_ms": build_ms, "chunks_indexed_per_s_per_gpu": (hi - lo) / (b / 1e3), "build_GBps_algorithmic": algo_build / b / 1e6,
                      "query_ms": q_ms, "effective_pairs_per_s_per_gpu": (hi - lo) * args.queries / (qm / 1e3),
                      "merge_s_host_incl_allgather": merge_s,
                      "max_over_ranks": {"build_ms": max_build, "query_ms": max_query, "merge_s": max_merge,
                                         "chunks_indexed_per_s_all_gpus": args.chunks / (max_build / 1e3),
                                         "effective_pairs_per_s_all_gpus": args.chunks * args.queries / (max_query / 1e3 + max_merge)},
                      "merged_topk_identical_on_all_ranks": same, "allgather_bytes_per_rank": world * args.queries * args.k * 12, "n_terms": int(len(toff) - 1), "mean_hits": float(gn.mean()),
                      "host_gen_s": gen_s, "host_tokenize_s": tok_s, "sidecar": side}))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
