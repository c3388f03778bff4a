"""Multi-GPU host logic: shard-count invariance of the sharded top-k, session / candidate sharding.
CPU tests run world_size 2 over gloo with oracle stand-ins for the two kernels; the GPU test merges through dvb_topk."""
import os
import pickle
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


# ----------------------------------------------------------------------------- oracle stand-ins (tests only)
def oracle_scores(chunk_texts, queries_terms, ctx=None):
    from oracle import dv_oracle
    return np.stack([dv_oracle.chunk_scores(list(chunk_texts), list(t)) for t in queries_terms]) if len(queries_terms) else np.zeros((0, len(chunk_texts)), np.int32)


def oracle_topk(score, k, index=None, index_base=0, ctx=None):
    from oracle import dv_oracle
    score = np.atleast_2d(score)
    nq = score.shape[0]
    oi = np.full((nq, k), -1, np.int64); os_ = np.zeros((nq, k), np.int32); on = np.zeros(nq, np.int32)
    for q in range(nq):
        idx = np.atleast_2d(index)[q] if index is not None else np.arange(score.shape[1], dtype=np.int64) + index_base
        i, s = dv_oracle.topk(score[q], k, idx)
        oi[q, :len(i)] = i; os_[q, :len(i)] = s; on[q] = len(i)
    return oi, os_, on


def make_corpus(n_chunks=257, nq=9, seed=5):
    rng = np.random.default_rng(seed)
    words = ["离线", "审批", "能力", "验证", "数据", "平台", "对接", "流程", "abc", "x9", "系统", "方案", "指标", "风险", "，", "。", " "]
    chunks = ["".join(rng.choice(words, size=int(rng.integers(5, 40)))) for _ in range(n_chunks)]
    queries = [[str(w) for w in rng.choice(words[:13], size=int(rng.integers(1, 5)), replace=False)] for _ in range(nq - 1)]
    queries.append(["不存在的词"])                      # no positive chunk anywhere -> [0, N//2, N-1] fallback
    return chunks, queries


def reference_picks(chunks, queries, k):
    out = []
    for terms in queries:
        scores = [sum(c.lower().count(t) for t in terms) for c in chunks]                   # web/server.py:32075-32076
        ranked = sorted((i for i in range(len(chunks)) if scores[i] > 0), key=lambda i: (-scores[i], i))[:k]
        sel = sorted(ranked)
        if not sel:
            cand = [0] + ([len(chunks) // 2] if len(chunks) > 2 else []) + ([len(chunks) - 1] if len(chunks) > 1 else [])
            for i in cand:
                if i not in sel:
                    sel.append(i)
                if len(sel) >= k:
                    break
        out.append(sel)
    return out


def test_shard_range_partitions():
    from deep_vision_b200.parallel import shard_range
    for n in (0, 1, 7, 64, 1000003):
        for w in (1, 2, 3, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_single_process_matches_reference(oracle):
    from deep_vision_b200 import parallel
    chunks, queries = make_corpus()
    got = parallel.sharded_chunk_select(chunks, 0, len(chunks), queries, 8, scores_fn=oracle_scores, topk_fn=oracle_topk)
    assert got == reference_picks(chunks, queries, 8)


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, str(ROOT))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from deep_vision_b200 import parallel
    chunks, queries = make_corpus()
    lo, hi = parallel.shard_range(len(chunks), rank, world)
    picks = parallel.sharded_chunk_select(chunks[lo:hi], lo, len(chunks), queries, 8, scores_fn=oracle_scores, topk_fn=oracle_topk)
    # session sharding: each "session" is a list of requests; the stand-in binder tags them
    sessions = [[("risks", {"risk": "r%d_%d" % (s, j)}, {"facts": []}, 0.46) for j in range(s % 3 + 1)] for s in range(11)]
    bound = parallel.sharded_bind(sessions, bind_fn=lambda reqs: [{"refs": [r[1]["risk"]], "score": 1.0} for r in reqs])

    class FakeGate:
        def match(self, cands, hist, thr):
            return np.asarray([len(c) % 5 - 1 for c in cands], np.int64)
    gate = parallel.sharded_gate(FakeGate(), [("q" * i) for i in range(23)], ["h"], 0.62)
    with open(os.path.join(out_dir, "rank%d.pkl" % rank), "wb") as f:
        pickle.dump({"picks": picks, "bound": bound, "gate": gate.tolist()}, f)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_world2_gloo_shard_invariance(oracle, tmp_path):
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    chunks, queries = make_corpus()
    want = reference_picks(chunks, queries, 8)
    sessions = [["r%d_%d" % (s, j) for j in range(s % 3 + 1)] for s in range(11)]
    for rank in range(2):
        got = pickle.loads((tmp_path / ("rank%d.pkl" % rank)).read_bytes())
        assert got["picks"] == want                                     # identical on every rank, equal to 1-shard result
        assert [[r["refs"][0] for r in sess] for sess in got["bound"]] == sessions
        assert got["gate"] == [len("q" * i) % 5 - 1 for i in range(23)]


def test_dropin_install_uninstall_rebinds_only_existing_names():
    import types
    from deep_vision_b200 import dropin
    mod = types.ModuleType("fake_server")
    orig = lambda *a, **k: "orig"
    mod.infer_weak_evidence_refs_v3 = orig
    mod.select_reference_material_context = orig
    mod.QUESTION_SIMILARITY_STOPWORDS = {"的", "请"}
    done = dropin.install(mod)
    assert set(done) == {"infer_weak_evidence_refs_v3", "select_reference_material_context"}
    assert mod.infer_weak_evidence_refs_v3 is not orig and not hasattr(mod, "question_token_set")
    # invalid input keeps the reference's "return empty, never raise" contract without touching the GPU
    assert mod.infer_weak_evidence_refs_v3("needs", {}, {}) == {"refs": [], "score": 0.0}
    assert mod.infer_weak_evidence_refs_v3("risks", None, {"facts": [{}]}) == {"refs": [], "score": 0.0}
    dropin.install(mod)                                                  # idempotent: the saved original survives
    dropin.uninstall(mod)
    assert mod.infer_weak_evidence_refs_v3 is orig and mod.select_reference_material_context is orig


@pytest.mark.gpu
def test_gpu_sharded_topk_is_shard_count_invariant(gpu_ctx, oracle):
    from deep_vision_b200 import parallel, reference_material as rm
    chunks, queries = make_corpus(n_chunks=1031, nq=17, seed=11)
    want = reference_picks(chunks, queries, 8)
    full = rm.chunk_scores(chunks, queries, ctx=gpu_ctx)
    for world in (1, 2, 4, 8):
        li, ls = [], []
        for r in range(world):
            lo, hi = parallel.shard_range(len(chunks), r, world)
            i, s, _n = rm.topk(full[:, lo:hi], 8, index_base=lo, ctx=gpu_ctx)
            li.append(i); ls.append(s)
        idx, _sc, nv = rm.topk(np.concatenate(ls, axis=1), 8, index=np.concatenate(li, axis=1), ctx=gpu_ctx)   # the all-gather + merge
        for q in range(len(queries)):
            sel = sorted(int(x) for x in idx[q][: int(nv[q])])
            assert sel == (want[q] if want[q] and full[q].max() > 0 else [])
    got = parallel.sharded_chunk_select(chunks, 0, len(chunks), queries, 8, ctx=gpu_ctx)
    assert got == want
