"""Bulk a7 (configs 3/5): the positional inverted index gives the same scores / top-k as the scan scorer and the oracle."""
import numpy as np
import pytest

from test_parallel import make_corpus, reference_picks


def _ref_topk(chunks, terms, k, base=0):
    scores = [sum(c.lower().count(t) for t in terms) for c in chunks]
    ranked = sorted((i for i in range(len(chunks)) if scores[i] > 0), key=lambda i: (-scores[i], i))[:k]
    return [i + base for i in ranked], [scores[i] for i in ranked]


@pytest.mark.gpu
def test_gpu_index_topk_matches_python_count(gpu_ctx):
    from deep_vision_b200 import reference_material as rm
    chunks, queries = make_corpus(n_chunks=777, nq=25, seed=4)
    chunks += ["aaaa", "AAAAA_aaaa", "abababab x abab", "离线离线离线离线离", "x" * 3000 + "审批" * 700, "", "   ", "İstanbul Kelvin K"]
    chunks += ["xx" * 40 + "y" + "x" * 41 + " xxy xxz xx9 xx", ("审审批" * 900)[:2700], "批审" * 1200 + "审审审审审"]
    queries += [["aa"], ["abab", "ab"], ["离线离", "离线"], ["审批", "审批审批审批"], ["zzzz"], ["aaaa", "aa", "a_"], ["istanbul", "kelvin"], ["st"]]
    # 2-symbol terms walk all four next-symbol sub-lists of their pair; "xx"-type terms are counted per run
    queries += [["xx"], ["xx", "xxx", "xy"], ["审审"], ["批审", "审批"], ["审审", "审审审"], ["ab", "ba"]]
    ix = rm.ChunkIndex(chunks, index_base=1000, ctx=gpu_ctx)
    st = ix.stats()
    assert st["n_chunks"] == len(chunks) and st["postings"] > 0 and st["pair_keys"] > 0
    idx, sc, nv = ix.topk(queries, 8)
    for q, terms in enumerate(queries):
        wi, ws = _ref_topk(chunks, terms, 8, base=1000)
        assert idx[q][: nv[q]].tolist() == wi, (q, terms)
        assert sc[q][: nv[q]].tolist() == ws
    # same answer as the scan path
    full = rm.chunk_scores(chunks, queries, ctx=gpu_ctx)
    si, ss, sn = rm.topk(full, 8, index_base=1000, ctx=gpu_ctx)
    assert np.array_equal(si, idx) and np.array_equal(ss, sc) and np.array_equal(sn, nv)
    ix.close()
    empty = rm.ChunkIndex([], ctx=gpu_ctx)
    ei, es, en = empty.topk([["离线"]], 4)
    assert en.tolist() == [0] and ei.tolist() == [[-1] * 4]
    with pytest.raises(Exception):
        ix2 = rm.ChunkIndex(["abc"], ctx=gpu_ctx)
        import deep_vision_b200.reference_material as m
        tsym = np.asarray([0x8001], np.uint16)                     # a 1-symbol term is not something the reference emits
        ix2.ctx.check(ix2.ctx.lib.dvb_chunk_index_topk(ix2.ctx.handle, ix2._h, 0, m.ptr(tsym), m.ptr(np.asarray([0, 1], np.int32)), 1,
                                                      m.ptr(np.asarray([0, 1], np.int32)), 1, 0, 4, m.ptr(np.zeros(4, np.int64)),
                                                      m.ptr(np.zeros(4, np.int32)), m.ptr(np.zeros(1, np.int32))), "topk")


@pytest.mark.gpu
def test_gpu_index_shards_merge_to_single_shard_result(gpu_ctx):
    from deep_vision_b200 import parallel, reference_material as rm
    chunks, queries = make_corpus(n_chunks=2053, nq=21, seed=8)
    queries = [q for q in queries]
    one = rm.ChunkIndex(chunks, ctx=gpu_ctx)
    wi, ws, wn = one.topk(queries, 8)
    for world in (2, 4, 8):
        li, ls = [], []
        for r in range(world):
            lo, hi = parallel.shard_range(len(chunks), r, world)
            ix = rm.ChunkIndex(chunks[lo:hi], index_base=lo, ctx=gpu_ctx)
            i, s, _n = ix.topk(queries, 8)
            li.append(i); ls.append(s); ix.close()
        gi, gs, gn = rm.topk(np.concatenate(ls, axis=1), 8, index=np.concatenate(li, axis=1), ctx=gpu_ctx)
        assert np.array_equal(gi, wi) and np.array_equal(gs, ws) and np.array_equal(gn, wn)


@pytest.mark.gpu
def test_gpu_index_config3_sample_vs_oracle(gpu_ctx, oracle):
    """config-3 shaped sample: 20k synthetic 512-char chunks, 300 queries half-copied from chunks, top-8."""
    from deep_vision_b200 import reference_material as rm, synth
    n, nq = 20000, 300
    data, off = synth.make_texts(n, 512, seed=123)
    raw = data.tobytes()
    rng = np.random.default_rng(5)
    queries = []
    texts = {}
    for q in range(nq):
        c = int(rng.integers(0, n))
        t = texts.setdefault(c, raw[off[c]:off[c + 1]].decode("utf-8"))
        a = int(rng.integers(0, 480))
        queries.append(rm._tokenize_reference_material_query(t[a:a + int(rng.integers(15, 30))] + " " + t[5:12]))
    ix = rm.ChunkIndex(packed=(data, off), ctx=gpu_ctx)
    idx, sc, nv = ix.topk(queries, 8)
    for q in range(0, nq, 7):                                      # oracle: exact counts over the whole corpus (C, threaded)
        s = oracle.chunk_scores(None, queries[q], n_threads=8, packed=(data, off))
        wi, ws = oracle.topk(s, 8)
        assert idx[q][: nv[q]].tolist() == wi.tolist() and sc[q][: nv[q]].tolist() == ws.tolist()
    st = ix.stats()
    assert st["postings"] > n * 300


@pytest.mark.gpu
def test_gpu_index_work_items_do_not_split_chunks(gpu_ctx):
    """Posting lists longer than one work item (2048 postings) with chunks straddling item boundaries and a
    self-overlapping term: counts must still be per-chunk greedy (str.count)."""
    from deep_vision_b200 import reference_material as rm
    rng = np.random.default_rng(12)
    chunks = []
    for c in range(600):
        n = int(rng.integers(1, 400))
        chunks.append(("ab" * n) + ("x" if c % 3 else "") + ("ab" * int(rng.integers(0, 50))) + "离线" * int(rng.integers(0, 300)))
    queries = [["abab"], ["ab"], ["ababab", "bab"], ["离线离线离", "线离"], ["abx"]]
    ix = rm.ChunkIndex(chunks, ctx=gpu_ctx)
    idx, sc, nv = ix.topk(queries, 32)
    for q, terms in enumerate(queries):
        wi, ws = _ref_topk(chunks, terms, 32)
        assert idx[q][: nv[q]].tolist() == wi and sc[q][: nv[q]].tolist() == ws, terms


@pytest.mark.gpu
def test_gpu_index_segments_equal_single_segment(gpu_ctx, monkeypatch):
    """A shard larger than IX_SEG_BYTES is indexed in segments (bounded sort buffers); the running top-k carried from
    segment to segment must equal the one-segment answer, ties (score, index asc) included."""
    from deep_vision_b200 import reference_material as rm
    chunks, queries = make_corpus(n_chunks=1500, nq=40, seed=21)
    chunks[700] = "x" * 9000                                        # a chunk larger than a whole segment is its own segment
    chunks += ["离线审批"] * 40                                      # equal scores across segment borders: order by index
    queries += [["离线审批", "离线"], ["xx"], ["zzzzzz"]]
    one = rm.ChunkIndex(chunks, index_base=7, ctx=gpu_ctx)
    assert one.stats()["n_chunks"] == len(chunks)
    wi, ws, wn = one.topk(queries, 8)
    monkeypatch.setenv("DVB_IX_SEG_BYTES", "4096")
    seg = rm.ChunkIndex(chunks, index_base=7, ctx=gpu_ctx)
    monkeypatch.delenv("DVB_IX_SEG_BYTES")
    st1, st2 = one.stats(), seg.stats()
    assert st2["n_chunks"] == st1["n_chunks"] and st2["postings"] == st1["postings"] and st2["pair_keys"] > st1["pair_keys"]
    for k in (1, 8, 32):
        a = one.topk(queries, k); b = seg.topk(queries, k)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
    for q, terms in enumerate(queries[-3:]):
        ri, rs = _ref_topk(chunks, terms, 8, base=7)
        qq = len(queries) - 3 + q
        assert wi[qq][: wn[qq]].tolist() == ri and ws[qq][: wn[qq]].tolist() == rs
    one.close(); seg.close()


@pytest.mark.gpu
def test_gpu_index_large_corpus_query_slice_vs_oracle(gpu_ctx, oracle):
    """Parity at scale (SURVEY §8d protocol step 2): 400 k chunks x 512 chars (0.59 GB of text, 176 M postings — the
    config-3 corpus is 2.5 x this), 2 000 queries; a slice of the queries is scored exactly over the WHOLE corpus by the
    threaded C oracle and must give the same top-8 (indices, scores, order); shard-count invariance on the same corpus."""
    from deep_vision_b200 import reference_material as rm, synth
    n, nq = 400000, 2000
    data, off = synth.make_texts(n, 512, seed=synth.SEED + 1000)
    raw = data.tobytes()
    qdata, qoff = synth.make_texts(nq, 45, seed=999)
    qraw = qdata.tobytes()
    rng = np.random.default_rng(7)
    queries = []
    for q in range(nq):
        text = qraw[qoff[q]:qoff[q + 1]].decode("utf-8")
        if q % 2 == 0:
            c = int(rng.integers(0, n)); src = raw[off[c]:off[c + 1]].decode("utf-8"); a = int(rng.integers(0, 488))
            text = text[:22] + src[a:a + 22]
        queries.append(rm._tokenize_reference_material_query(text))
    ix = rm.ChunkIndex(packed=(data, off), ctx=gpu_ctx)
    idx, sc, nv = ix.topk(queries, 8)
    ix.close()
    for q in list(range(0, 24, 2)) + [501, 1000, 1999]:
        s = oracle.chunk_scores(None, queries[q], n_threads=16, packed=(data, off))
        wi, ws = oracle.topk(s, 8)
        assert idx[q][: nv[q]].tolist() == wi.tolist() and sc[q][: nv[q]].tolist() == ws.tolist(), q
    # two shards (contiguous halves, global indices) merged = one index
    half = n // 2
    li, ls = [], []
    for lo, hi in ((0, half), (half, n)):
        sub_off = off[lo:hi + 1] - off[lo]
        sub = np.concatenate([data[off[lo]:off[hi]], np.zeros(64, np.uint8)])
        part = rm.ChunkIndex(packed=(sub, sub_off), index_base=lo, ctx=gpu_ctx)
        i, s_, _n = part.topk(queries[:400], 8)
        li.append(i); ls.append(s_); part.close()
    gi, gs, gn = rm.topk(np.concatenate(ls, axis=1), 8, index=np.concatenate(li, axis=1), ctx=gpu_ctx)
    assert np.array_equal(gi, idx[:400]) and np.array_equal(gs, sc[:400]) and np.array_equal(gn, nv[:400])
