#!/usr/bin/env python3
"""torchrun check (N GPUs, NCCL): sharded chunk top-k through the real kernels equals the reference's selection.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/dist_topk_check.py"""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from deep_vision_b200 import _lib, parallel
from test_parallel import make_corpus, reference_picks
ctx = _lib.Context.get(lr)
chunks, queries = make_corpus(n_chunks=4099, nq=33, seed=21)
lo, hi = parallel.shard_range(len(chunks), rank, world)
got = parallel.sharded_chunk_select(chunks[lo:hi], lo, len(chunks), queries, 8, ctx=ctx)
want = reference_picks(chunks, queries, 8)
ok = got == want
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("dist_topk_check world=%d: %s (launches on rank0: %d)" % (world, "OK" if int(flag.item()) else "MISMATCH", ctx.launches))
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) else 1)
