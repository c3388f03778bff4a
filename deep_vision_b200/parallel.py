"""Multi-GPU host logic: one process per GPU, `torch.distributed` (NCCL over NVLink on GPUs, gloo in CPU tests).

SURVEY.md §8(e):
  * chunk corpora (a7 bulk, configs 3/5) shard by CONTIGUOUS global chunk index; every rank computes a local
    top-k under the reference's total order (score desc, global index asc) == sorted(key=(-score, index))
    (web/server.py:32083), ONE all-gather moves nq x k x (int64 index, int32 score) per rank, and the same
    `dvb_topk` kernel merges the gathered candidate lists on every rank.  The order is total and the ranges are
    disjoint, so the result does not depend on the world size.
  * sessions (a2, a11-a14) and gate candidates (a9) are independent units: block partition, no data-path
    collective; results are concatenated (object gather) only if the caller asks for them.

The C-ABI stays collective-free: gathered candidate lists go back into `dvb_topk` as plain arrays.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

from . import reference_material as rm


def _dist():
    import torch.distributed as dist
    return dist


def world(group=None) -> Tuple[int, int]:
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    try:
        dist = _dist()
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(group), dist.get_world_size(group)
    except Exception:
        pass
    return 0, 1


def shard_range(n: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous block partition of range(n): rank r owns [lo, hi); sizes differ by at most one."""
    n, rank, world_size = int(n), int(rank), max(1, int(world_size))
    base, extra = divmod(n, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def _all_gather_array(a: np.ndarray, group=None) -> List[np.ndarray]:
    """all_gather of equally-shaped numpy arrays (device tensors under NCCL, host tensors under gloo)."""
    import torch
    dist = _dist()
    _rank, ws = world(group)
    if ws == 1:
        return [a]
    backend = dist.get_backend(group)
    t = torch.from_numpy(np.ascontiguousarray(a))
    if backend == "nccl":
        t = t.cuda()
    outs = [torch.empty_like(t) for _ in range(ws)]
    dist.all_gather(outs, t, group=group)
    return [o.cpu().numpy() for o in outs]


def merge_topk(local_idx: np.ndarray, local_score: np.ndarray, k: int, group=None, ctx=None,
               topk_fn: Optional[Callable] = None) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """All-gather every rank's [nq, k] candidates (idx -1 / score 0 = empty) and merge them to the global top-k.
    Returns (idx[nq, k], score[nq, k], n_valid[nq]) — identical on every rank."""
    topk_fn = topk_fn or rm.topk
    local_idx = np.ascontiguousarray(local_idx, np.int64)
    local_score = np.ascontiguousarray(local_score, np.int32)
    idx_all = np.concatenate(_all_gather_array(local_idx, group), axis=1)          # [nq, world * k]
    score_all = np.concatenate(_all_gather_array(local_score, group), axis=1)
    # empty slots carry score 0 and are dropped by the positives-only rule of dvb_topk
    return topk_fn(score_all, k, index=idx_all, ctx=ctx)


def sharded_topk(local_score: np.ndarray, index_base: int, k: int, group=None, ctx=None,
                 topk_fn: Optional[Callable] = None) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Global top-k of a chunk corpus whose scores are sharded by contiguous index range.
    `local_score` is [nq, n_local] for global indices index_base .. index_base + n_local."""
    topk_fn = topk_fn or rm.topk
    k = int(k)
    local_score = np.ascontiguousarray(np.atleast_2d(local_score), np.int32)
    li, ls, _ln = topk_fn(local_score, k, index_base=int(index_base), ctx=ctx)
    return merge_topk(li, ls, k, group=group, ctx=ctx, topk_fn=topk_fn)


def sharded_chunk_select(chunk_texts_local: Sequence[str], index_base: int, n_total: int, queries_terms: Sequence[Sequence[str]],
                         k: int, group=None, ctx=None, scores_fn: Optional[Callable] = None,
                         topk_fn: Optional[Callable] = None) -> List[List[int]]:
    """Per query the global chunk indices the reference's selector would pick (web/server.py:32073-32099), already
    re-sorted by index (32087); the `[0, N//2, N-1]` fallback applies when no chunk on ANY rank scores > 0."""
    scores_fn = scores_fn or rm.chunk_scores
    nq = len(queries_terms)
    if len(chunk_texts_local):
        local = scores_fn(chunk_texts_local, queries_terms, ctx=ctx)
    else:
        local = np.zeros((nq, 0), np.int32)
    idx, _score, n_valid = sharded_topk(local, index_base, k, group=group, ctx=ctx, topk_fn=topk_fn)
    picked: List[List[int]] = []
    for q in range(nq):
        sel = sorted(int(i) for i in idx[q][: int(n_valid[q])])
        if not sel and n_total > 0:
            cand = [0]
            if n_total > 2:
                cand.append(n_total // 2)
            if n_total > 1:
                cand.append(n_total - 1)
            for i in cand:
                if i not in sel:
                    sel.append(i)
                if len(sel) >= max(1, k):
                    break
        picked.append(sel)
    return picked


def gather_objects(obj, group=None) -> list:
    """Concatenate per-rank Python results (e.g. bind results of a session shard) on every rank, in rank order."""
    rank, ws = world(group)
    if ws == 1:
        return [obj]
    out = [None] * ws
    _dist().all_gather_object(out, obj, group=group)
    return out


def sharded_bind(requests_by_session: Sequence[Sequence[tuple]], group=None, gather: bool = True, bind_fn: Optional[Callable] = None):
    """Sessions are independent: rank r binds sessions shard_range(S, r, W) with ONE kernel launch and no data-path
    collective.  `requests_by_session[s]` is the list of (field, item, evidence_pack, min_score) of session s.
    Returns the per-session result lists of all sessions (gather=True) or of the local shard."""
    from . import evidence
    bind_fn = bind_fn or evidence.bind_batch
    rank, ws = world(group)
    lo, hi = shard_range(len(requests_by_session), rank, ws)
    flat, sizes = [], []
    for s in range(lo, hi):
        flat.extend(requests_by_session[s])
        sizes.append(len(requests_by_session[s]))
    res = bind_fn(flat) if flat else []
    local, pos = [], 0
    for n in sizes:
        local.append(res[pos:pos + n])
        pos += n
    if not gather:
        return local
    merged: list = []
    for part in gather_objects(local, group):
        merged.extend(part)
    return merged


def sharded_gate(gate, candidates: Sequence[str], history: Sequence[str], threshold: float = 0.62, group=None, gather: bool = True):
    """a9 at config-4 scale: history replicated on every rank, candidates block-partitioned, no exchange."""
    rank, ws = world(group)
    lo, hi = shard_range(len(candidates), rank, ws)
    local = gate.match(candidates[lo:hi], history, threshold) if hi > lo else np.zeros(0, np.int64)
    if not gather:
        return local
    return np.concatenate([np.asarray(p, np.int64) for p in gather_objects(local, group)])


def sharded_index_topk(index, queries_terms: Sequence[Sequence[str]], k: int, group=None, topk_fn: Optional[Callable] = None):
    """Config 3/5: every rank holds a ChunkIndex over its contiguous shard (index.index_base = first global chunk index).
    Local indexed top-k -> one all-gather of nq x k candidates -> merge (dvb_topk).  Identical on every rank."""
    li, ls, _ln = index.topk(queries_terms, k)
    return merge_topk(li, ls, k, group=group, ctx=index.ctx, topk_fn=topk_fn)


def terms_to_device(queries_terms: Sequence[Sequence[str]], device=None):
    """terms_to_symbols() of a query set, moved to the GPU once (reused by every sharded_index_topk_device call on it)."""
    import torch
    tsym, toff, qoff = rm.terms_to_symbols(queries_terms)
    dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    return tuple(torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (tsym, toff, qoff))


def sharded_index_topk_device(index, dev_terms, nq: int, k: int, group=None):
    """Config 3/5 with the candidates kept on the device (web/server.py:32083-32099 over a sharded corpus): local indexed
    top-k (dvb_chunk_index_topk, DVB_MEM_DEVICE) -> ONE all_gather_into_tensor of nq x k x 16 B (global index i64, score
    widened to i64) over NCCL -> dvb_topk(DVB_MEM_DEVICE) on the gathered lists -> only the final [nq, k] crosses PCIe.
    Returns numpy (idx, score, n_valid), identical on every rank."""
    import torch
    dist = _dist()
    _rank, ws = world(group)
    li, ls, ln = index.topk_device(dev_terms, nq, k)
    if ws == 1 or nq == 0:
        return li.cpu().numpy(), ls.cpu().numpy(), ln.cpu().numpy()
    mine = torch.stack([li, ls.to(torch.int64)], dim=2).contiguous()              # [nq, k, 2] i64
    allc = torch.empty((ws, nq, k, 2), dtype=torch.int64, device=mine.device)
    dist.all_gather_into_tensor(allc, mine, group=group)
    idx_all = allc[..., 0].permute(1, 0, 2).reshape(nq, ws * k).contiguous()
    sc_all = allc[..., 1].permute(1, 0, 2).reshape(nq, ws * k).to(torch.int32).contiguous()
    oi = torch.empty((nq, k), dtype=torch.int64, device=mine.device)
    os_ = torch.empty((nq, k), dtype=torch.int32, device=mine.device)
    on = torch.empty((nq,), dtype=torch.int32, device=mine.device)
    ctx = index.ctx
    torch.cuda.current_stream(mine.device).synchronize()                           # the library runs on its own stream
    rc = ctx.lib.dvb_topk(ctx.handle, rm._lib.MEM_DEVICE, sc_all.data_ptr(), idx_all.data_ptr(), ws * k, nq, 0, k,
                          oi.data_ptr(), os_.data_ptr(), on.data_ptr())
    ctx.check(rc, "dvb_topk")
    return oi.cpu().numpy(), os_.cpu().numpy(), on.cpu().numpy()
