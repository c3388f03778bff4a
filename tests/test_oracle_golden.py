"""CPU suite: the C oracle and the host logic against fixtures generated from the REAL reference
(tools/make_golden.py).  No GPU."""
import hashlib

import numpy as np
import pytest

from conftest import oracle_bind_packed


def sha(s):
    return hashlib.sha256(s.encode("utf-8", "surrogatepass")).hexdigest()[:16]


def test_a1_tokens_match_reference(oracle, golden):
    for case in golden("a1_tokenize.json"):
        assert sorted(oracle.tokenize_v3(case["text"])) == case["tokens"], repr(case["text"][:40])


def test_a1_keys_are_injective_on_golden(oracle, golden):
    """distinct tokens <-> distinct keys, and the oracle key function agrees with the batch key API."""
    for case in golden("a1_tokenize.json"):
        keys = oracle.tokenize_v3_keys(case["text"])
        assert len(keys) == len(case["tokens"])
        assert sorted(oracle.token_key(t) for t in case["tokens"]) == keys.tolist()


def test_a6_query_terms(oracle, golden):
    for case in golden("a6_query_terms.json"):
        assert oracle.query_terms(case["query"]) == case["terms"], repr(case["query"][:40])


def _docs(golden):
    head = "开头普通说明。" * 1800
    tail = "\n\n尾部关键需求：必须支持离线审批和断网补录，并保留审批链路证据。"
    return {"tail_recall": head + tail, "blank": " \n\t ", "short": "短文档",
            "newline_heavy": ("段落一。\n\n" + "内容" * 333 + "\n") * 9, "no_newline": "字" * 5000,
            "ws_run": "甲" * 1700 + "\n" * 400 + "乙" * 2100, "random": golden("docs_random.json")["random"]}


def test_a4_chunks(oracle, golden):
    docs = _docs(golden)
    seen = 0
    for case in golden("a4_chunks.json"):
        if case["name"] not in docs:
            continue                       # fixture_md lives in the reference tree (checked in test_reference_diff)
        content = docs[case["name"]]
        assert sha(content) == case["content_sha"]
        got = oracle.chunk_build(content)
        assert len(got) == len(case["chunks"])
        for (s, e, ts, te), ref in zip(got, case["chunks"]):
            assert [s, e, sha(content[ts:te])] == ref[:3]
        seen += 1
    assert seen >= 6


def test_a7_scores_and_topk(oracle, golden):
    docs = _docs(golden)
    for case in golden("a7_select.json"):
        if case["doc"] not in docs or case["meta"]["mode"] != "chunk_selection":
            continue
        content = docs[case["doc"]]
        spans = oracle.chunk_build(content)
        chunks = [content[ts:te] for (_s, _e, ts, te) in spans]
        terms = oracle.query_terms(case["query"])
        assert terms[:12] == case["meta"]["query_terms"]
        scores = oracle.chunk_scores(chunks, terms)
        k = max(1, case["chunk_limit"] or 4)
        idx, _sc = oracle.topk(scores, k)
        if len(idx):
            want = [c["index"] for c in case["meta"]["selected_chunks"]]
            # the reference may clip the tail of the selection by max_chars; compare the prefix
            assert sorted(idx.tolist())[:len(want)] == want


def test_a8_a9_gate(oracle, golden):
    g = golden("a8_a9_gate.json")
    gate = oracle.Gate([tuple(x) for x in g["synonyms"]], g["stopwords"])
    for q, norm in g["normalized"]:
        assert gate.normalize(q) == norm, repr(q)
    for q, toks in g["token_sets"]:
        assert len(gate.keys(q)) == len(toks)
    got = gate.gate(g["candidates"], g["history"], g["threshold"])
    assert got.tolist() == g["match"]
    assert (np.array(g["match"]) >= 0).sum() > 5


def test_a10_solution_jaccard(oracle, golden):
    for case in golden("a10_solution_jaccard.json"):
        j, _i, _u = oracle.solution_jaccard(case["left_clean"], case["right_clean"])
        assert round(j, 3) == case["similarity"]


def test_a2_host_logic_with_oracle_scorer(oracle, golden, monkeypatch):
    """infer_weak_evidence_refs_v3 host mirror (packing, dimension preference, thresholds, rounding)
    with the CUDA call replaced by the oracle: every golden case must equal the reference's dict."""
    from deep_vision_b200 import evidence
    monkeypatch.setattr(evidence, "bind_packed", oracle_bind_packed)
    g = golden("a2_bind.json")
    n_bound = 0
    for case in g["cases"]:
        pack = g["packs"][case["pack"]] if isinstance(case["pack"], str) else case["pack"]
        res = evidence.infer_weak_evidence_refs_v3(case["field"], case["item"], pack, case["min_score"], ctx=object())
        assert res == case["result"], case["name"]
        n_bound += bool(res["refs"])
    assert n_bound >= 10
    # batch form == single calls
    reqs = [(c["field"], c["item"], g["packs"][c["pack"]] if isinstance(c["pack"], str) else c["pack"], c["min_score"])
            for c in g["cases"]]
    assert evidence.bind_batch(reqs, ctx=object()) == [c["result"] for c in g["cases"]]


def test_f3_claim_chunk_host_logic_with_oracle_scorer(oracle, monkeypatch):
    """SURVEY 8(f)3 host logic (session layout, key quantisation, top-k order, drop-in shape) with the oracle standing in
    for the two kernels; the exhaustive path only (the candidate stage needs the resident index)."""
    import types
    import conftest
    from deep_vision_b200 import dropin, evidence, reference_material as rm
    from test_reference_material import _oracle_standins
    monkeypatch.setattr(evidence, "bind_packed", conftest.oracle_bind_packed)
    _oracle_standins(monkeypatch, oracle)
    chunks = ["我们必须支持离线审批和断网补录，并保留审批链路证据。", "预算有限，budget_2024 已经确定。审批流程太慢需要优化", "开头普通说明。" * 5, "离线审批"]
    claims = ["离线审批能力验证需要支持断网补录", "budget_2024 预算有限", "", "完全无关的内容", "离线审批"]
    got = evidence.bind_claims_to_chunks(claims, chunks, 3, ctx=object())
    assert got == oracle.claim_chunk_topk(claims, chunks, 3)
    assert [r["index"] for r in got[0]][:1] == [0] and got[2] == [] and got[3] == []
    mod = types.SimpleNamespace(load_reference_material_chunks=lambda doc: [{"chunk_id": "chunk-%04d" % i, "index": i, "start": 0, "end": len(t), "text": t}
                                                                            for i, t in enumerate(chunks)])
    done = dropin.install(mod, ctx=object(), claim_chunk_binding=True)
    r = mod.infer_reference_chunk_refs_v3("risks", {"risk": claims[0]}, {"doc_id": "d"}, 2)
    assert [c["index"] for c in r["chunks"]] == [x["index"] for x in got[0][:2]] and r["chunks"][0]["chunk_id"] == "chunk-0000"
    assert r["score"] == round(got[0][0]["score"], 3)
    assert mod.infer_reference_chunk_refs_v3("needs", {"title": "x"}, {}, 2) == {"chunks": [], "score": 0.0}
    assert "infer_reference_chunk_refs_v3" in done
    dropin.uninstall(mod)
    assert not hasattr(mod, "infer_reference_chunk_refs_v3")


def test_dropin_gate_tables_follow_in_place_changes():
    """dropin._gate caches the device tables on the CONTENT of the module's synonym dict / stop-word set: an in-place change
    (what tests and hot-config do; the reference iterates the live objects per call, web/server.py:20464-20468) rebuilds them."""
    import types
    from deep_vision_b200 import dropin
    mod = types.SimpleNamespace(QUESTION_SIMILARITY_SYNONYMS={"痛点": "问题"}, QUESTION_SIMILARITY_STOPWORDS={"请", "你"})
    g1 = dropin._gate(mod, ctx=object())
    assert dropin._gate(mod, ctx=object()) is g1                      # unchanged content: cached
    mod.QUESTION_SIMILARITY_STOPWORDS.add("们")                       # same object, new content
    g2 = dropin._gate(mod, ctx=object())
    assert g2 is not g1 and set(g2.stopwords) == {"请", "你", "们"}
    mod.QUESTION_SIMILARITY_SYNONYMS["诉求"] = "需求"
    g3 = dropin._gate(mod, ctx=object())
    assert g3 is not g2 and ("诉求", "需求") in g3.synonyms
