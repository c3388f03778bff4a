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
