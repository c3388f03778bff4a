"""Rows a4-a7: host logic on CPU (oracle stand-ins for the CUDA calls) and CUDA parity on the GPU."""
import hashlib

import numpy as np
import pytest


def sha(s):
    return hashlib.sha256(s.encode("utf-8", "surrogatepass")).hexdigest()[:16]


def docs(golden):
    head = "开头普通说明。" * 1800
    tail = "\n\n尾部关键需求：必须支持离线审批和断网补录，并保留审批链路证据。"
    return {"tail_recall": head + tail, "blank": " \n\t ", "short": "短文档",
            "newline_heavy": ("段落一。\n\n" + "内容" * 333 + "\n") * 9, "no_newline": "字" * 5000,
            "ws_run": "甲" * 1700 + "\n" * 400 + "乙" * 2100, "random": golden("docs_random.json")["random"]}


def test_a6_host_terms_match_reference(golden):
    from deep_vision_b200 import reference_material as rm
    for case in golden("a6_query_terms.json"):
        assert rm._tokenize_reference_material_query(case["query"]) == case["terms"], repr(case["query"][:40])


def _oracle_standins(monkeypatch, oracle):
    from deep_vision_b200 import reference_material as rm

    def chunk_spans(contents, chunk_size, overlap, ctx=None):
        from deep_vision_b200.packing import pack_texts
        data, off = pack_texts(contents)
        out = []
        for text in contents:
            raw = text.encode("utf-8", "surrogatepass")
            cps = list(text)
            spans = []
            for (s, e, ts, te) in oracle.chunk_build(text, chunk_size, overlap):
                b0 = len("".join(cps[:ts]).encode("utf-8", "surrogatepass"))
                b1 = len("".join(cps[:te]).encode("utf-8", "surrogatepass"))
                spans.append((s, e, ts, te, b0, b1))
            out.append(spans)
            assert raw is not None
        return out, data, off

    def chunk_scores(texts, queries_terms, ctx=None, packed=None):
        return np.stack([oracle.chunk_scores(texts, terms) for terms in queries_terms])

    def topk(score, k, index=None, index_base=0, ctx=None):
        score = np.atleast_2d(score)
        oi = np.full((len(score), k), -1, np.int64)
        os_ = np.zeros((len(score), k), np.int32)
        on = np.zeros(len(score), np.int32)
        for q, row in enumerate(score):
            i, s = oracle.topk(row, k)
            oi[q, :len(i)] = i + index_base
            os_[q, :len(i)] = s
            on[q] = len(i)
        return oi, os_, on

    monkeypatch.setattr(rm, "chunk_spans", chunk_spans)
    monkeypatch.setattr(rm, "chunk_scores", chunk_scores)
    monkeypatch.setattr(rm, "topk", topk)
    return rm


def _check_chunks_and_select(rm, golden, ctx):
    d = docs(golden)
    built = {}
    for case in golden("a4_chunks.json"):
        if case["name"] not in d:
            continue
        content = d[case["name"]]
        chunks = rm._build_reference_material_chunks(content, ctx=ctx)
        built[case["name"]] = chunks
        assert len(chunks) == len(case["chunks"]), case["name"]
        for c, ref in zip(chunks, case["chunks"]):
            assert [c["start"], c["end"], sha(c["text"]), c["chunk_id"], c["index"], sha(c["preview"])] == ref
    n_sel = 0
    for case in golden("a7_select.json"):
        if case["doc"] not in d:
            continue
        content = d[case["doc"]]
        chunks = built[case["doc"]]
        text, meta = rm.select_reference_material_context({"content": content[:10000]}, query_text=case["query"],
                                                          max_chars=case["max_chars"], chunk_limit=case["chunk_limit"],
                                                          chunk_loader=lambda doc: chunks, ctx=ctx)
        assert meta == case["meta"], (case["doc"], case["query"])
        assert sha(text) == case["context_sha"] and len(text) == case["context_len"]
        n_sel += 1
    assert n_sel >= 30


def test_a4_a7_host_logic_with_oracle(golden, oracle, monkeypatch):
    rm = _oracle_standins(monkeypatch, oracle)
    _check_chunks_and_select(rm, golden, ctx=object())


def test_manifest_roundtrip(tmp_path, golden, oracle, monkeypatch):
    rm = _oracle_standins(monkeypatch, oracle)
    content = docs(golden)["newline_heavy"]
    man = rm.build_chunk_manifest(doc_id="d1", filename="a.md", content=content, ext=".md", ctx=object())
    assert man["version"] == 1 and man["content_sha256"] == hashlib.sha256(content.encode()).hexdigest()
    import json
    p = tmp_path / "chunks.json"
    p.write_text(json.dumps(man, ensure_ascii=False), encoding="utf-8")
    assert rm.load_chunks_from_manifest(p) == man["chunks"]
    assert rm.load_chunks_from_manifest(tmp_path / "missing.json") == []


# ------------------------------------------------------------------ GPU

@pytest.mark.gpu
def test_gpu_chunks_and_select_golden(gpu_ctx, golden):
    from deep_vision_b200 import reference_material as rm
    _check_chunks_and_select(rm, golden, ctx=gpu_ctx)


@pytest.mark.gpu
def test_gpu_symbolize_and_scores_vs_oracle(gpu_ctx, oracle, golden):
    from deep_vision_b200 import reference_material as rm
    texts = [c["text"] for c in golden("a1_tokenize.json")]
    syms = rm.symbolize(texts, ctx=gpu_ctx)
    for t, s in zip(texts, syms):
        low = oracle.lower(t)
        want = []
        for ch in low:
            o = ord(ch)
            if 0x4E00 <= o <= 0x9FFF:
                want.append(o - 0x4E00 + 1)
            elif "a" <= ch <= "z":
                want.append(0x8000 | (o - 96))
            elif "0" <= ch <= "9":
                want.append(0x8000 | (o - 48 + 27))
            elif ch == "_":
                want.append(0x8000 | 37)
            else:
                want.append(0)
        assert s.tolist() == want, repr(t[:30])
    queries = [["aa", "aaaa", "一二", "哈哈", "ab", "中中", "文文文"], rm._tokenize_reference_material_query("需求系统审批流程 abc_9 客户业务痛点方案数据"),
               rm._tokenize_reference_material_query(texts[45][:60]), []]
    got = rm.chunk_scores(texts, queries, ctx=gpu_ctx)
    for q, terms in enumerate(queries):
        assert got[q].tolist() == oracle.chunk_scores(texts, terms).tolist()


@pytest.mark.gpu
def test_gpu_topk_vs_oracle(gpu_ctx, oracle):
    from deep_vision_b200 import reference_material as rm
    rng = np.random.default_rng(5)
    for n, k in [(1, 1), (7, 4), (1000, 8), (70000, 8), (70000, 32), (33, 32)]:
        score = rng.integers(0, 6, size=(3, n)).astype(np.int32)
        score[1, :] = 0
        score[2, rng.integers(0, n, size=max(1, n // 3))] = 0
        oi, os_, on = rm.topk(score, k, index_base=1000, ctx=gpu_ctx)
        for q in range(3):
            wi, ws = oracle.topk(score[q], k)
            assert on[q] == len(wi)
            assert oi[q, :on[q]].tolist() == (wi + 1000).tolist()
            assert os_[q, :on[q]].tolist() == ws.tolist()
    # merge of candidate lists with explicit indices == top-k of the union
    score = rng.integers(1, 50, size=(2, 64)).astype(np.int32)
    idx = rng.permutation(10_000_000_000 + np.arange(128, dtype=np.int64)).reshape(2, 64)
    oi, os_, on = rm.topk(score, 8, index=idx, ctx=gpu_ctx)
    for q in range(2):
        order = sorted(range(64), key=lambda i: (-int(score[q, i]), int(idx[q, i])))[:8]
        assert oi[q].tolist() == [int(idx[q, i]) for i in order]


@pytest.mark.gpu
def test_gpu_chunk_build_batch_and_ragged(gpu_ctx, oracle):
    from deep_vision_b200 import reference_material as rm
    rng = np.random.default_rng(3)
    contents = ["", " ", "a", "中" * 799, "中" * 1801, "\n" * 5000, ("abc\n" * 2000), "x" * 100 + "\n\n" * 900 + "y" * 3000]
    for _ in range(12):
        n = int(rng.integers(0, 9000))
        chars = rng.choice(list("中文字符ab \n\t，。"), size=n, p=[.2, .2, .2, .2, .05, .05, .03, .03, .01, .02, .01])
        contents.append("".join(chars))
    for size, ov in [(1800, 160), (800, 160), (50, 40), (1000, 0)]:
        spans, data, off = rm.chunk_spans(contents, size, min(ov, size // 3), ctx=gpu_ctx)
        for text, sp in zip(contents, spans):
            want = oracle.chunk_build(text, size, ov)
            assert [(a, b, c, d) for (a, b, c, d, _e, _f) in sp] == want
            raw = text.encode("utf-8")
            for (_a, _b, ts, te, b0, b1) in sp:
                assert raw[b0:b1].decode("utf-8") == text[ts:te]


# ------------------------------------------------------------------ chunks.dvbidx sidecar

def test_sidecar_tag_identifies_content_and_chunking():
    from deep_vision_b200 import reference_material as rm
    chunks = [{"index": 0, "start": 0, "end": 1800, "text": "a"}, {"index": 1, "start": 1640, "end": 3000, "text": "b"}]
    t = rm.sidecar_tag("ab" * 32, chunks)
    assert len(t) == 32 and t == rm.sidecar_tag("ab" * 32, [dict(c) for c in chunks])
    assert t != rm.sidecar_tag("cd" * 32, chunks)                                     # other document
    assert t != rm.sidecar_tag("ab" * 32, chunks[:1])                                 # other chunk count
    moved = [dict(chunks[0]), dict(chunks[1], start=1600)]
    assert t != rm.sidecar_tag("ab" * 32, moved)                                      # re-chunked with another overlap
    with pytest.raises(ValueError):
        rm._tag32(b"short")


@pytest.mark.gpu
def test_gpu_sidecar_roundtrip_and_rejections(gpu_ctx, tmp_path, monkeypatch):
    from deep_vision_b200 import _lib, reference_material as rm
    from test_parallel import make_corpus
    chunks, queries = make_corpus(n_chunks=900, nq=30, seed=33)
    chunks += ["", "aaaa aa", "离线审批" * 30]
    queries += [["aa"], ["离线审批", "审批离线"]]
    tag = bytes(range(32))
    for seg_bytes in (None, "8192"):                                                   # one segment / many segments
        if seg_bytes:
            monkeypatch.setenv("DVB_IX_SEG_BYTES", seg_bytes)
        ix = rm.ChunkIndex(chunks, index_base=5, ctx=gpu_ctx)
        monkeypatch.delenv("DVB_IX_SEG_BYTES", raising=False)
        want = ix.topk(queries, 8)
        p = tmp_path / f"chunks_{seg_bytes}.dvbidx"
        ix.save(p, tag)
        assert p.exists() and not (tmp_path / (p.name + ".tmp")).exists()
        back = rm.ChunkIndex.load(p, tag, index_base=5, ctx=gpu_ctx)
        assert back.stats() == ix.stats()
        for a, b in zip(want, back.topk(queries, 8)):
            assert np.array_equal(a, b)
        back.close()
        rm.ChunkIndex.load(p, None, ctx=gpu_ctx).close()                              # NULL tag accepts any content
        with pytest.raises(_lib.DvbError, match="tag mismatch"):
            rm.ChunkIndex.load(p, bytes(32), ctx=gpu_ctx)
        raw = bytearray(p.read_bytes())
        bad = tmp_path / "bad.dvbidx"
        bad.write_bytes(raw[: len(raw) // 2])                                          # truncated
        with pytest.raises(_lib.DvbError):
            rm.ChunkIndex.load(bad, tag, ctx=gpu_ctx)
        flip = bytearray(raw); flip[len(flip) - 100] ^= 0x40                           # bit rot in the last array
        bad.write_bytes(flip)
        with pytest.raises(_lib.DvbError, match="checksum"):
            rm.ChunkIndex.load(bad, tag, ctx=gpu_ctx)
        bad.write_bytes(b"not an index" * 20)
        with pytest.raises(_lib.DvbError):
            rm.ChunkIndex.load(bad, tag, ctx=gpu_ctx)
        ix.close()
    with pytest.raises(_lib.DvbError):
        rm.ChunkIndex.load(tmp_path / "missing.dvbidx", tag, ctx=gpu_ctx)
    empty = rm.ChunkIndex([], ctx=gpu_ctx)
    empty.save(tmp_path / "e.dvbidx", tag)
    e2 = rm.ChunkIndex.load(tmp_path / "e.dvbidx", tag, ctx=gpu_ctx)
    assert e2.stats()["n_chunks"] == 0 and e2.topk([["离线"]], 4)[2].tolist() == [0]


@pytest.mark.gpu
def test_gpu_select_through_sidecar_cache_matches_golden(gpu_ctx, golden, tmp_path):
    """The resident / persisted index gives exactly the reference's selections (a7 goldens), first from a fresh build
    (sidecar written), then from a second cache that can only load the sidecar, then with a stale sidecar."""
    import json
    from deep_vision_b200 import reference_material as rm
    d = docs(golden)
    def hashable(text):                              # the reference's own manifest writer hashes content.encode("utf-8") (31995):
        try:                                         # documents with lone surrogates never get a manifest, hence no sidecar
            text.encode("utf-8")
            return True
        except UnicodeEncodeError:
            return False
    cases = [c for c in golden("a7_select.json") if c["doc"] in d and hashable(d[c["doc"]])]
    assert len(cases) >= 20
    dirs, recs = {}, {}
    for name in {c["doc"] for c in cases}:
        man = rm.build_chunk_manifest(doc_id=name, filename=name + ".md", content=d[name], ext=".md", ctx=gpu_ctx)
        (tmp_path / name).mkdir()
        (tmp_path / name / "chunks.json").write_text(json.dumps(man, ensure_ascii=False), encoding="utf-8")
        dirs[name] = tmp_path / name
        recs[name] = {"content": d[name][:10000], "full_content_sha256": man["content_sha256"], "name": name}

    called = set()

    def run(cache):
        def provider(doc, chunks):
            called.add(doc["name"])
            return cache.get(dirs[doc["name"]] / rm.SIDECAR_NAME, doc["full_content_sha256"], chunks)
        n = 0
        for case in cases:
            chunks = rm.load_chunks_from_manifest(dirs[case["doc"]] / "chunks.json")
            text, meta = rm.select_reference_material_context(recs[case["doc"]], query_text=case["query"], max_chars=case["max_chars"],
                                                              chunk_limit=case["chunk_limit"], chunk_loader=lambda doc: chunks,
                                                              ctx=gpu_ctx, index_provider=provider)
            assert meta == case["meta"], (case["doc"], case["query"])
            assert sha(text) == case["context_sha"] and len(text) == case["context_len"]
            n += 1
        return n

    c1 = rm.ChunkIndexCache(ctx=gpu_ctx)
    assert run(c1) >= 20
    assert len(called) >= 4 and c1.builds == len(called) and c1.loads == 0 and c1.hits > 0      # blank documents have no chunks
    assert all((dirs[n] / rm.SIDECAR_NAME).exists() for n in called)
    c1.clear()
    c2 = rm.ChunkIndexCache(ctx=gpu_ctx, write_back=False)
    run(c2)
    assert c2.builds == 0 and c2.loads == len(called)
    c2.clear()
    # a sidecar left over from other content is ignored and replaced
    names = sorted(called)
    (dirs[names[0]] / rm.SIDECAR_NAME).write_bytes((dirs[names[1]] / rm.SIDECAR_NAME).read_bytes())
    c3 = rm.ChunkIndexCache(ctx=gpu_ctx)
    run(c3)
    assert c3.builds == 1 and c3.loads == len(called) - 1
    c3.clear()
    c4 = rm.ChunkIndexCache(ctx=gpu_ctx, write_back=False)                                      # ... and was rewritten by c3
    run(c4)
    assert c4.builds == 0
    c4.clear()
    # explicit helper used by bulk jobs
    out = rm.save_chunk_index_sidecar(dirs[names[0]] / "chunks.json", ctx=gpu_ctx)
    assert out.name == rm.SIDECAR_NAME and out.stat().st_size > 256


@pytest.mark.gpu
def test_gpu_query_terms_golden_and_random_vs_oracle(gpu_ctx, oracle, golden):
    """a6 on the device (dvb_query_terms): the golden term lists of the reference, adversarial and random texts against the
    oracle, host-buffer and device-resident calls, and the result fed straight into the indexed top-k."""
    import random
    from deep_vision_b200 import reference_material as rm
    cases = golden("a6_query_terms.json")
    texts = [c["query"] for c in cases]
    want = [c["terms"] for c in cases]
    rng = random.Random(11)
    cjk = [chr(0x4E00 + rng.randrange(3000)) for _ in range(200)]
    special = ["İ", "K", "Ａ", "ß", " ", "\n", "，", "。", "_", "-", "A", "Z", "q", "0", "9", "😀", "é", "〇", "鿿", "一"]
    def rnd(n, p):
        return "".join(rng.choice(cjk) if rng.random() < p else rng.choice("abcXY019_") if rng.random() < 0.5 else rng.choice(special) for _ in range(n))
    extra = ["", " ", "一", "一二", "ab", "a", "一二三四五六", "一二三四五" * 7, "ab" * 50, "一" * 100, "abc一二三四五六七def", ("一二三四五六七八九十" * 9)[:85] + " xyz xyz",
             "".join(cjk[i % 200] for i in range(300)), "一二三四五六 一二三四五六 三四五六"]
    extra += [rnd(rng.choice([3, 9, 25, 60, 300, 1200]), rng.choice([0.6, 0.95, 0.2])) for _ in range(400)]
    texts += extra
    want += [oracle.query_terms(t) for t in extra]
    assert all(rm._tokenize_reference_material_query(t) == w for t, w in zip(texts, want))
    tsym, toff, qoff = rm.query_terms_batch(texts, ctx=gpu_ctx)
    assert rm.terms_from_symbols(tsym, toff, qoff) == want
    assert max(len(w) for w in want) == 80                                      # the cap is exercised
    d_tsym, d_toff, d_qoff = rm.query_terms_batch(texts, ctx=gpu_ctx, device=True)
    assert np.array_equal(d_tsym.cpu().numpy().view(np.uint16), tsym) and np.array_equal(d_toff.cpu().numpy(), toff) and np.array_equal(d_qoff.cpu().numpy(), qoff)
    # same CSR as the host path builds from the term strings, and usable as is by the index
    hs, ho, hq = rm.terms_to_symbols(want)
    assert np.array_equal(hs[: len(tsym)] if len(hs) >= len(tsym) else hs, tsym[: len(hs)]) and np.array_equal(ho, toff) and np.array_equal(hq, qoff)
    chunks = [t for t in texts if t.strip()][:200]
    ix = rm.ChunkIndex(chunks, ctx=gpu_ctx)
    keep = [i for i, w in enumerate(want) if w][:64]
    a = ix.topk([want[i] for i in keep], 8)
    sub = rm.query_terms_batch([texts[i] for i in keep], ctx=gpu_ctx, device=True)
    b = ix.topk_device(sub, len(keep), 8)
    assert np.array_equal(a[0], b[0].cpu().numpy()) and np.array_equal(a[1], b[1].cpu().numpy()) and np.array_equal(a[2], b[2].cpu().numpy())
    ix.close()
    e = rm.query_terms_batch([], ctx=gpu_ctx)
    assert len(e[2]) == 1 and int(e[2][0]) == 0
