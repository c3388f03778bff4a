"""GPU parity (B200): CUDA path through the C-ABI vs the CPU oracle and the golden fixtures."""
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent))

pytestmark = pytest.mark.gpu


def test_a1_keys_golden(gpu_ctx, oracle, golden):
    from deep_vision_b200 import evidence
    cases = golden("a1_tokenize.json")
    texts = [c["text"] for c in cases]
    got = evidence.tokenize_similarity_keys(texts, ctx=gpu_ctx)
    for c, keys in zip(cases, got):
        want = sorted(oracle.token_key(t) for t in c["tokens"])
        assert keys.tolist() == want, repr(c["text"][:40])
        # injective classes decode back to the reference's token strings
        for k in keys.tolist():
            s = evidence.decode_token_key(k)
            if s is not None:
                assert s in c["tokens"]


def test_a1_keys_synthetic_and_ragged(gpu_ctx, oracle):
    from deep_vision_b200 import evidence, synth
    data, off = synth.make_texts(300, 171, seed=7)
    raw = data.tobytes()
    texts = [raw[off[i]:off[i + 1]].decode("utf-8") for i in range(300)]
    texts += ["", "a", texts[0] * 9, "x" * 5000, "中" * 4000, texts[1][:5], "ab " * 3000]
    got = evidence.tokenize_similarity_keys(texts, ctx=gpu_ctx)
    for t, keys in zip(texts, got):
        assert keys.tolist() == oracle.tokenize_v3_keys(t).tolist(), repr(t[:30])


def test_a2_golden_cases(gpu_ctx, golden):
    from deep_vision_b200 import evidence
    g = golden("a2_bind.json")
    reqs = [(c["field"], c["item"], g["packs"][c["pack"]] if isinstance(c["pack"], str) else c["pack"], c["min_score"])
            for c in g["cases"]]
    got = evidence.bind_batch(reqs, ctx=gpu_ctx)
    for c, r in zip(g["cases"], got):
        assert r == c["result"], c["name"]
    # single-call form
    c = g["cases"][0]
    assert evidence.infer_weak_evidence_refs_v3(c["field"], c["item"], c["pack"], c["min_score"], ctx=gpu_ctx) == c["result"]


# (400, 40, 24) = 35 MB of fact text: the host-buffer call runs as >= 4 pipelined slices (H2D overlapped with scoring)
@pytest.mark.parametrize("n_sessions,turns,claims", [(64, 40, 24), (5, 70, 40), (3, 1, 1), (40, 7, 33), (400, 40, 24)])
def test_a2_synthetic_bit_exact(gpu_ctx, oracle, n_sessions, turns, claims):
    from deep_vision_b200 import evidence, synth
    c = synth.make_sessions(n_sessions, turns=turns, claims_per_session=claims, seed=11 + n_sessions)
    args = (c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off,
            c.fact_quality, c.fact_dim, c.fact_hard, c.claim_pref_dim, c.claim_threshold)
    got = evidence.bind_packed(gpu_ctx, *args, want_lens=True)
    ref = oracle.bind_claims(None, None, c.sess_claim_off, c.sess_fact_off, c.fact_quality, c.fact_dim, c.fact_hard,
                             c.claim_pref_dim, c.claim_threshold, packed=(c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off))
    assert np.array_equal(got["best_fact"], ref["best_fact"])
    assert np.array_equal(got["best_score"].view(np.uint64), ref["best_score"].view(np.uint64))   # bit-exact fp64
    assert np.array_equal(got["bound"], ref["bound"])
    assert np.array_equal(got["len_claim"], np.minimum(ref["len_claim"], 14))
    assert np.array_equal(got["len_fact"], np.minimum(ref["len_fact"], 18))
    # hint path gives the same answer
    got2 = evidence.bind_packed(gpu_ctx, *args, max_claim_group_bytes=c.max_claim_group_bytes)
    assert np.array_equal(got2["best_fact"], ref["best_fact"])


def _bind_vs_oracle(gpu_ctx, oracle, a, check_len_fact=True):
    from deep_vision_b200 import evidence
    cu, co, fu, fo, sco, sfo, fq, fd, fh, cp, ct = a
    got = evidence.bind_packed(gpu_ctx, *a, want_lens=True)
    ref = oracle.bind_claims(None, None, sco, sfo, fq, fd, fh, cp, ct, packed=(cu, co, fu, fo))
    if check_len_fact:
        has_claims = np.repeat(np.diff(sco) > 0, np.diff(sfo))
        assert np.array_equal(got["len_fact"][has_claims], np.minimum(ref["len_fact"], 18)[has_claims])
    assert np.array_equal(got["len_claim"], np.minimum(ref["len_claim"], 14))
    bad = np.nonzero(got["best_score"].view(np.uint64) != ref["best_score"].view(np.uint64))[0]
    assert bad.size == 0, [(int(i), got["best_score"][i], ref["best_score"][i]) for i in bad[:8]]
    assert np.array_equal(got["best_fact"], ref["best_fact"])
    assert np.array_equal(got["bound"], ref["bound"])
    return got


def test_a2_slab_waves_host_and_device_buffers(gpu_ctx, oracle, monkeypatch):
    """dvb_bind_claims hands the claim tables from k_bind_claim_tables to k_bind_facts through one slab per claim group and
    runs in waves of sessions when a call has more groups than slabs (1 GB of slabs in production; 64 slabs here): 150
    sessions x 2 groups = 5 waves, host buffers (offsets on the host) and device buffers (offsets fetched once) — same
    answers as the oracle and as the single-wave call."""
    import torch
    from deep_vision_b200 import _lib, evidence, synth
    c = synth.make_sessions(150, turns=6, claims_per_session=40, seed=5)
    a = (c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off,
         c.fact_quality, c.fact_dim, c.fact_hard, c.claim_pref_dim, c.claim_threshold)
    one = _bind_vs_oracle(gpu_ctx, oracle, a)
    monkeypatch.setenv("DVB_BIND_SLAB_BUDGET", "64")
    l0 = gpu_ctx.launches
    waves = _bind_vs_oracle(gpu_ctx, oracle, a)
    assert gpu_ctx.launches - l0 >= 2 * 5                           # five waves x two kernels
    assert np.array_equal(one["best_score"].view(np.uint64), waves["best_score"].view(np.uint64))
    dev = [torch.from_numpy(x).cuda() for x in a]
    nc = int(c.sess_claim_off[-1])
    ob = torch.empty(nc, dtype=torch.int32, device="cuda"); os_ = torch.empty(nc, dtype=torch.float64, device="cuda")
    obd = torch.empty(nc, dtype=torch.uint8, device="cuda")
    rc = gpu_ctx.lib.dvb_bind_claims(gpu_ctx.handle, _lib.MEM_DEVICE, *[t.data_ptr() for t in dev[:6]], c.n_sessions,
                                     *[t.data_ptr() for t in dev[6:]], ob.data_ptr(), os_.data_ptr(), obd.data_ptr(), None, None, 0)
    gpu_ctx.check(rc, "bind")
    torch.cuda.synchronize()
    assert np.array_equal(ob.cpu().numpy(), waves["best_fact"]) and np.array_equal(obd.cpu().numpy(), waves["bound"])
    assert np.array_equal(os_.cpu().numpy().view(np.uint64), waves["best_score"].view(np.uint64))


def test_a2_long_facts_chunk_parallel_tokenizer(gpu_ctx, oracle):
    """long facts (tests/bind_cases.py): cell/step boundaries, long runs, exotic case mappings (handed back to the general
    kernel), long ASCII tokens, 2-/4-byte characters — bit-exact against the oracle."""
    import bind_cases as bc
    _bind_vs_oracle(gpu_ctx, oracle, bc.long_facts(3))


def test_a2_every_pair_of_long_facts(gpu_ctx, oracle):
    """one fact per session, so best_score exposes the overlap of EVERY (claim, fact) pair (tests/bind_cases.py)."""
    import bind_cases as bc
    _bind_vs_oracle(gpu_ctx, oracle, bc.every_pair())


def test_a2_short_facts_and_mixed_scripts(gpu_ctx, oracle):
    import bind_cases as bc
    _bind_vs_oracle(gpu_ctx, oracle, bc.mixed_real_text())
    _bind_vs_oracle(gpu_ctx, oracle, bc.big_claim_groups())
    for n in (5, 30, 40):                                   # one table bucket: side list, side list overflow (hand-back)
        _bind_vs_oracle(gpu_ctx, oracle, bc.crowded_bucket(n))


def test_a2_general_kernel_same_results(gpu_ctx, oracle, monkeypatch):
    """DVB_BIND_KERNEL=general routes the call to k_bind_claims_general (the fallback of the fast kernel)."""
    from deep_vision_b200 import evidence, synth
    c = synth.make_sessions(48, turns=30, claims_per_session=40, seed=77)
    args = (c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off,
            c.fact_quality, c.fact_dim, c.fact_hard, c.claim_pref_dim, c.claim_threshold)
    fast = evidence.bind_packed(gpu_ctx, *args, want_lens=True)
    monkeypatch.setenv("DVB_BIND_KERNEL", "general")
    gen = evidence.bind_packed(gpu_ctx, *args, want_lens=True)
    for k in fast:
        assert np.array_equal(fast[k], gen[k]), k
    ref = oracle.bind_claims(None, None, c.sess_claim_off, c.sess_fact_off, c.fact_quality, c.fact_dim, c.fact_hard,
                             c.claim_pref_dim, c.claim_threshold, packed=(c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off))
    assert np.array_equal(gen["best_fact"], ref["best_fact"])
    assert np.array_equal(gen["best_score"].view(np.uint64), ref["best_score"].view(np.uint64))


def test_a2_one_group_of_a_session_handed_back(gpu_ctx, oracle):
    """The fast kernels take claim GROUPS as their unit of work (k_bind_groups numbers them): sessions of four groups where
    one group is handed back by the claim kernel, or the session by the fact kernel, come out as the oracle has them."""
    import bind_cases as bc
    l0 = gpu_ctx.launches
    _bind_vs_oracle(gpu_ctx, oracle, bc.handback_in_one_group())
    assert gpu_ctx.launches - l0 == 4                               # groups, claim tables, facts, general kernel for the two sessions


def test_a2_group_numbering_across_steps(gpu_ctx, oracle):
    """4300 sessions (> one 4096-session step of k_bind_groups), some without claims, some of two groups."""
    import bind_cases as bc
    _bind_vs_oracle(gpu_ctx, oracle, bc.many_tiny_sessions())


def test_a2_split_and_plain_fact_kernels_agree(gpu_ctx, oracle, monkeypatch):
    """A call of few claim groups runs k_bind_facts_split (facts of a group over several CTAs); DVB_BIND_NO_SPLIT=1 forces the
    plain kernel: both equal the oracle, bit for bit."""
    import bind_cases as bc
    for a in (bc.synthetic(3, 12, 40), bc.synthetic(40, 7, 33), bc.mixed_real_text()):
        split = _bind_vs_oracle(gpu_ctx, oracle, a)
        monkeypatch.setenv("DVB_BIND_NO_SPLIT", "1")
        plain = _bind_vs_oracle(gpu_ctx, oracle, a)
        monkeypatch.delenv("DVB_BIND_NO_SPLIT")
        assert np.array_equal(split["best_score"].view(np.uint64), plain["best_score"].view(np.uint64))
        assert np.array_equal(split["best_fact"], plain["best_fact"])


def test_a2_token_table_overflow_reruns_general(gpu_ctx, oracle):
    """> 704 distinct ASCII tokens in one claim group: the fast kernel raises `overflow`, the call re-runs exactly."""
    from deep_vision_b200 import packing
    import random
    rng = random.Random(9)
    abc = "abcdefghijklmnopqrstuvwxyz0123456789_"
    words = sorted({"".join(rng.choice(abc) for _ in range(3)) for _ in range(3000)})
    claims = [" ".join(rng.sample(words, 60)) for _ in range(25)]         # ~6 KB, ~1200 distinct tokens: fast path, then overflow
    facts = [" ".join(rng.sample(words, 300)) + " 中文测试数据" for _ in range(12)]
    cu, co = packing.pack_texts(claims)
    fu, fo = packing.pack_texts(facts)
    sco = np.array([0, len(claims)], np.int64); sfo = np.array([0, len(facts)], np.int64)
    fq = np.linspace(0.1, 0.9, len(facts)); fd = np.zeros(len(facts), np.int32); fh = np.zeros(len(facts), np.uint8)
    cp = np.full(len(claims), -1, np.int32); ct = np.full(len(claims), 0.46, np.float64)
    from deep_vision_b200 import evidence
    got = evidence.bind_packed(gpu_ctx, cu, co, fu, fo, sco, sfo, fq, fd, fh, cp, ct, want_lens=True)
    ref = oracle.bind_claims(None, None, sco, sfo, fq, fd, fh, cp, ct, packed=(cu, co, fu, fo))
    assert np.array_equal(got["best_fact"], ref["best_fact"])
    assert np.array_equal(got["best_score"].view(np.uint64), ref["best_score"].view(np.uint64))
    assert np.array_equal(got["len_claim"], np.minimum(ref["len_claim"], 14))


def test_a2_large_claim_table_global_fallback(gpu_ctx, oracle):
    """claims of one session too large for the shared-memory table -> global-memory table path"""
    from deep_vision_b200 import evidence, synth
    c = synth.make_sessions(2, turns=6, claims_per_session=30, claim_chars=4000, copy_frac=0.05, seed=5)
    args = (c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off,
            c.fact_quality, c.fact_dim, c.fact_hard, c.claim_pref_dim, c.claim_threshold)
    got = evidence.bind_packed(gpu_ctx, *args)
    ref = oracle.bind_claims(None, None, c.sess_claim_off, c.sess_fact_off, c.fact_quality, c.fact_dim, c.fact_hard,
                             c.claim_pref_dim, c.claim_threshold, packed=(c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off))
    assert np.array_equal(got["best_fact"], ref["best_fact"])
    assert np.array_equal(got["best_score"].view(np.uint64), ref["best_score"].view(np.uint64))


def test_concurrent_calls_on_one_ctx(gpu_ctx, oracle):
    """SURVEY §8(b) threading: gunicorn gthread workers share one ctx and ctypes drops the GIL during a call, so every
    entry point must be re-entrant (per-call stream + workspace lease).  8 threads x 6 rounds of different entry points on
    ONE ctx give exactly the results of the same calls made one after the other."""
    import threading
    from deep_vision_b200 import evidence, synth, question_similarity as qs, reference_material as rm
    corpora = [synth.make_sessions(12 + 5 * i, turns=20, claims_per_session=24, seed=300 + i) for i in range(4)]
    texts_d, texts_o = synth.make_texts(200, 120, seed=41)
    raw = texts_d.tobytes()
    texts = [raw[texts_o[i]:texts_o[i + 1]].decode("utf-8") for i in range(200)]
    gate = qs.QuestionGate(ctx=gpu_ctx)
    cand, hist = [t[:30] + "？" for t in texts[:120]], [t[:30] + "？" for t in texts[60:200]]
    queries = [rm._tokenize_reference_material_query(t[10:50]) for t in texts[:16]]

    def bind(i):
        c = corpora[i % 4]
        r = evidence.bind_packed(gpu_ctx, c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off,
                                 c.fact_quality, c.fact_dim, c.fact_hard, c.claim_pref_dim, c.claim_threshold)
        return (r["best_fact"].tobytes(), r["best_score"].tobytes(), r["bound"].tobytes())

    jobs = [lambda i=i: bind(i) for i in range(4)]
    jobs.append(lambda: tuple(k.tobytes() for k in evidence.tokenize_similarity_keys(texts, ctx=gpu_ctx)))
    jobs.append(lambda: gate.match(cand, hist, 0.62).tobytes())
    jobs.append(lambda: rm.chunk_scores(texts, queries, ctx=gpu_ctx).tobytes())
    jobs.append(lambda: tuple(a.tobytes() for a in rm.topk(rm.chunk_scores(texts, queries, ctx=gpu_ctx), 8, ctx=gpu_ctx)))
    serial = [j() for j in jobs]
    # the first corpus against the oracle, so "serial" itself is pinned
    c = corpora[0]
    ref = oracle.bind_claims(None, None, c.sess_claim_off, c.sess_fact_off, c.fact_quality, c.fact_dim, c.fact_hard,
                             c.claim_pref_dim, c.claim_threshold, packed=(c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off))
    assert serial[0][0] == ref["best_fact"].tobytes() and serial[0][1] == ref["best_score"].tobytes()
    errors, results = [], [[None] * 6 for _ in jobs]
    start = threading.Barrier(len(jobs))

    def worker(t):
        try:
            start.wait()
            for r in range(6):
                results[t][r] = jobs[(t + r) % len(jobs)]()
        except Exception as exc:          # noqa: BLE001
            errors.append(repr(exc))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(len(jobs))]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    for t in range(len(jobs)):
        for r in range(6):
            assert results[t][r] == serial[(t + r) % len(jobs)], (t, r)


def test_a2_degenerate_sessions(gpu_ctx, oracle):
    """zero sessions; sessions without claims / without facts; empty and white-space-only claim and fact texts."""
    from deep_vision_b200 import evidence, packing
    z64 = np.zeros(1, np.int64)
    e = evidence.bind_packed(gpu_ctx, np.zeros(16, np.uint8), z64, np.zeros(16, np.uint8), z64, z64, z64, np.zeros(0), np.zeros(0, np.int32),
                             np.zeros(0, np.uint8), np.zeros(0, np.int32), np.zeros(0))
    assert len(e["best_fact"]) == 0
    claims = ["", "   ", "预算有限 budget_2024", "，。；", "a", "审批流程太慢需要优化"]
    facts = ["", " \n ", "我们的预算有限，budget_2024 已经确定。审批流程太慢需要优化", "a b c", "流程"]
    # session 0: claims, no facts; 1: facts, no claims; 2: both; 3: nothing; 4: both again
    sco = np.array([0, 6, 6, 12, 12, 18], np.int64)
    sfo = np.array([0, 0, 5, 10, 10, 15], np.int64)
    cu, co = packing.pack_texts(claims * 3)
    fu, fo = packing.pack_texts(facts * 3)
    nf, nc = 15, 18
    fq = np.linspace(0.0, 1.0, nf); fd = (np.arange(nf) % 3).astype(np.int32); fh = (np.arange(nf) % 2).astype(np.uint8)
    cp = (np.arange(nc) % 4 - 1).astype(np.int32); ct = np.full(nc, 0.3, np.float64)
    got = evidence.bind_packed(gpu_ctx, cu, co, fu, fo, sco, sfo, fq, fd, fh, cp, ct, want_lens=True)
    ref = oracle.bind_claims(None, None, sco, sfo, fq, fd, fh, cp, ct, packed=(cu, co, fu, fo))
    assert np.array_equal(got["best_fact"], ref["best_fact"])
    assert np.array_equal(got["best_score"].view(np.uint64), ref["best_score"].view(np.uint64))
    assert np.array_equal(got["bound"], ref["bound"])
    assert np.array_equal(got["len_claim"], np.minimum(ref["len_claim"], 14))
    assert got["best_fact"][:6].tolist() == [-1] * 6                   # no facts to bind to


def _claim_chunk_case(n_chunks, seed):
    import random
    rng = random.Random(seed)
    han = [chr(0x4E00 + rng.randrange(1200)) for _ in range(260)]
    def cjk(n): return "".join(rng.choice(han) for _ in range(n))
    chunks = []
    for c in range(n_chunks):
        chunks.append("，".join(cjk(rng.randint(2, 30)) for _ in range(rng.randint(1, 12))) + rng.choice(["", " api_v2 sla 99", "\n\n预算有限 budget_2024"]))
    claims = []
    for i in range(23):
        src = rng.choice(chunks)
        a = rng.randrange(max(1, len(src) - 30))
        claims.append(cjk(rng.randint(0, 12)) + src[a:a + rng.randint(3, 40)] + rng.choice(["", " sla", " budget_2024", cjk(5)]))
    claims += ["", "   ", cjk(40), chunks[0], chunks[-1][:7]]
    return claims, chunks


@pytest.mark.gpu
def test_claim_chunk_binding_vs_oracle(gpu_ctx, oracle):
    """claims bound to document chunks with a2's overlap score — exhaustive for small documents, through the
    count-selector candidate stage (resident index, device a6 terms) for large ones — indices and fp64 scores bit-identical."""
    from deep_vision_b200 import evidence
    for n_chunks, mx, k, seed in ((12, 256, 4, 1), (70, 256, 8, 2), (90, 16, 4, 3), (300, 64, 32, 4)):
        claims, chunks = _claim_chunk_case(n_chunks, seed)
        got = evidence.bind_claims_to_chunks(claims, chunks, k, ctx=gpu_ctx, max_exhaustive=mx)
        want = oracle.claim_chunk_topk(claims, chunks, k, max_exhaustive=mx)
        assert [[r["index"] for r in row] for row in got] == [[r["index"] for r in row] for row in want]
        g = np.asarray([r["score"] for row in got for r in row]); w = np.asarray([r["score"] for row in want for r in row])
        assert np.array_equal(g.view(np.uint64), w.view(np.uint64))
        assert sum(len(r) for r in got) > len(claims)
    assert evidence.bind_claims_to_chunks([], ["x"], 4, ctx=gpu_ctx) == [] and evidence.bind_claims_to_chunks(["一二三"], [], 4, ctx=gpu_ctx) == [[]]
