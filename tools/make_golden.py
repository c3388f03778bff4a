#!/usr/bin/env python3
"""Generate tests/golden/*.json by running the REAL reference functions (AST-executed from
/root/reference/web/server.py via oracle/ref_loader.py) on fixed inputs.

Run in the build container only (the reference mount does not exist on the GPU box):
    PYTHONHASHSEED=0 python tools/make_golden.py
The fixtures are committed; tests compare the C oracle and the CUDA path against them.
"""
from __future__ import annotations

import hashlib
import json
import os
import random
import re
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from oracle import ref_loader  # noqa: E402

GOLD = ROOT / "tests" / "golden"
FIX = ref_loader.REFERENCE_ROOT / "tests" / "fixtures" / "report_solution"

CJK = [chr(0x4E00 + i) for i in range(80)] + list("需求系统审批流程客户业务痛点方案数据接口订单风险上线成本预算")
SPECIAL = ["İ", "K", "Ａ", "Σ", "ς", "ǅ", "ß", "　", " ", "\n", "\n\n", "\t", "，", "。", "？", "_", "-", "A", "Z", "q",
           "0", "9", "😀", "̇", "'", "é", "É", "Ⅷ", "①", "²", "℃", "\ud800", "〇", "々", "鿿", "一", "䷿", "ꀀ"]


def rnd_text(rng: random.Random, n: int, p_cjk=0.6, p_ascii=0.25) -> str:
    out = []
    for _ in range(n):
        r = rng.random()
        if r < p_cjk:
            out.append(rng.choice(CJK))
        elif r < p_cjk + p_ascii:
            out.append(rng.choice("abcdeXYZ0189_"))
        else:
            out.append(rng.choice(SPECIAL))
    return "".join(out)


def adversarial_texts(rng: random.Random):
    texts = ["", " ", "\n", "a", "ab", "A_b9", "一", "一二", "aİb", "İİ", "aKb", "K9", "ΑΣ ΑΣΑ Σ", "ａｂｃ ＡＢ",
             "hello世界你好ab", "x" * 11, "y" * 10, "z" * 300, "ab" * 600, "中" * 12, "中" * 13, "文" * 25,
             "".join(CJK[i % 50] for i in range(12)), "".join(CJK[i % 50] for i in range(13)),
             "".join(CJK[i % 50] for i in range(14)), "".join(CJK[i % 50] for i in range(24)),
             "".join(CJK[i % 50] for i in range(25)), "".join(CJK[i % 50] for i in range(26)),
             "".join(CJK[i % 70] for i in range(700)), "哈哈哈哈哈", "aaaa aa", "a一b二c三", "一a二b三c",
             "\ud800一二", "😀一二😀三四", "é一二É", "℃ 25℃ 温度", "Ⅷ①②", "TEST_case-42 v2.0", "  leading and trailing  "]
    for n in (3, 7, 15, 16, 17, 31, 32, 33, 63, 64, 65, 170, 171, 172, 511, 512, 513, 530, 1100):
        texts.append(rnd_text(rng, n))
    for n in (40, 200, 600):
        texts.append(rnd_text(rng, n, p_cjk=0.95, p_ascii=0.03))
        texts.append(rnd_text(rng, n, p_cjk=0.05, p_ascii=0.85))
    return texts


def sha(s: str) -> str:
    return hashlib.sha256(s.encode("utf-8", "surrogatepass")).hexdigest()[:16]


def dump(name: str, obj):
    GOLD.mkdir(parents=True, exist_ok=True)
    p = GOLD / name
    p.write_text(json.dumps(obj, ensure_ascii=True, indent=0, sort_keys=True), encoding="utf-8")
    print("wrote", p, p.stat().st_size, "bytes")


def parse_transcript(md: str):
    """31 Q/A turns of the fixture report's appendix -> fact dicts."""
    tail = md[md.index("## 附录：完整访谈记录"):]
    blocks = re.split(r"\n---\n", tail)
    facts = []
    dim_keys = {"客户需求": "customer_needs", "业务流程": "business_process", "技术约束": "tech_constraints", "项目约束": "project_constraints"}
    for blk in blocks:
        m = re.search(r"\*\*【(.+?)】问题 (\d+)：(.+?)\*\*", blk, re.S)
        if not m:
            continue
        picked = re.findall(r"<div>☑ (.+?)</div>", blk)
        other = re.findall(r"<div>(?!<strong>)(?!☐)(?!☑)(.+?)</div>", blk)
        answer = "；".join(picked + other)
        idx = int(m.group(2))
        facts.append({
            "q_id": f"Q{idx}", "question": m.group(3).strip(), "answer": answer,
            "dimension": dim_keys.get(m.group(1), m.group(1)), "dimension_name": m.group(1),
            "quality_score": round(0.35 + 0.02 * ((idx * 7) % 29), 2), "hard_triggered": idx % 9 == 0,
        })
    return facts


def main():
    if os.environ.get("PYTHONHASHSEED") != "0":
        print("re-exec with PYTHONHASHSEED=0 for a stable stop-word order")
        os.environ["PYTHONHASHSEED"] = "0"
        os.execv(sys.executable, [sys.executable] + sys.argv)
    ns = ref_loader.load()
    rng = random.Random(20260101)

    # ---- a1
    texts = adversarial_texts(rng)
    dump("a1_tokenize.json", [{"text": t, "tokens": sorted(ns["_tokenize_similarity_text_v3"](t))} for t in texts])

    # ---- a6
    queries = texts[:40] + ["离线审批能力验证", "需要对接哪些外部系统 ERP CRM erp", "一二三四五六七八九十" * 3, "ab一二三四五六cd"]
    dump("a6_query_terms.json", [{"query": q, "terms": ns["_tokenize_reference_material_query"](q)} for q in queries])

    # ---- a4 (+a7 on the same docs)
    docs = []
    head = "开头普通说明。" * 1800
    tail = "\n\n尾部关键需求：必须支持离线审批和断网补录，并保留审批链路证据。"
    md = (FIX / "historical-interview-report.md").read_text(encoding="utf-8")
    doc_inputs = {"tail_recall": head + tail, "fixture_md": md, "blank": " \n\t ", "short": "短文档",
                  "newline_heavy": ("段落一。\n\n" + "内容" * 333 + "\n") * 9, "no_newline": "字" * 5000,
                  "ws_run": "甲" * 1700 + "\n" * 400 + "乙" * 2100, "random": rnd_text(rng, 7000)}
    a4, a7 = [], []
    for name, content in doc_inputs.items():
        chunks = ns["_build_reference_material_chunks"](content)
        a4.append({"name": name, "content_sha": sha(content), "n_chars": len(content),
                   "chunks": [[c["start"], c["end"], sha(c["text"]), c["chunk_id"], c["index"], sha(c["preview"])] for c in chunks]})
        for q, mc, cl in (("离线审批能力验证", 0, 0), ("访谈 需求 AI 访谈系统的核心功能", 1200, 3), ("zzzz不存在的词", 0, 2), ("", 0, 0),
                          ("内容内容 段落", 50, 8)):
            ctx_text, meta = ref_loader.select_context_with_chunks(ns, chunks, query_text=q, max_chars=mc, chunk_limit=cl, content=content[:10000])
            a7.append({"doc": name, "query": q, "max_chars": mc, "chunk_limit": cl, "context_sha": sha(ctx_text),
                       "context_len": len(ctx_text), "meta": meta})
    dump("a4_chunks.json", a4)
    dump("a7_select.json", a7)
    # the document inputs that are not reproducible from a formula
    dump("docs_random.json", {"random": doc_inputs["random"]})

    # ---- a8 / a9
    stop = list(ns["QUESTION_SIMILARITY_STOPWORDS"])
    syn = [[k, v] for k, v in ns["QUESTION_SIMILARITY_SYNONYMS"].items()]
    words = ["当前", "目前", "主要", "最", "核心", "请", "你", "们", "痛点", "诉求", "最关键", "最大的", "头疼", "影响推进", "业务", "系统",
             "方案", "A", "b", "？", " ", "，", "的", "一", "个", "下", "是什么", "哪些", "方面", "对接", "外部", "流程", "İ", "Σ", "　"]
    qs = ["当前最核心的业务痛点是什么？", "目前最大的业务问题主要是什么？", "当前最让你们头疼的痛点是什么？", "当前最影响推进的痛点是什么？",
          "需要对接哪些外部系统？", "", "  ", "的", "一的个方案", "A", "？？"]
    qs += ["".join(rng.choice(words) for _ in range(rng.randint(0, 14))) for _ in range(120)]
    hist = qs[:60]
    cands = qs[40:] + [q + "？" for q in qs[:20]]
    logs = [{"question": h} for h in hist]
    matches = []
    for c in cands:
        r = ns["find_similar_interview_question_log"](logs, c, 0.62)
        matches.append(-1 if r is None else max(i for i, l in enumerate(logs) if l is r))
    dump("a8_a9_gate.json", {
        "stopwords": stop, "synonyms": syn, "threshold": 0.62,
        "normalized": [[q, ns["normalize_interview_question_text"](q)] for q in qs],
        "token_sets": [[q, sorted(ns["question_token_set"](q))] for q in qs[:40]],
        "history": hist, "candidates": cands, "match": matches,
        "pairs": [[a, b, bool(ns["is_similar_interview_question"](a, b))] for a, b in zip(qs[:30], qs[1:31])],
    })

    # ---- a10
    sols = [rnd_text(rng, rng.randint(0, 60)) for _ in range(40)] + ["<b>方案</b> **重点** [链接](http://x) # 标题\n- 列表", "&amp;&lt;方案一&gt;", "AB", "a"]
    pairs = []
    for i in range(len(sols)):
        l = sols[i]
        r = sols[(i * 7 + 3) % len(sols)] if i % 3 else l[: len(l) // 2] + rnd_text(rng, 8)
        pairs.append({"left": l, "right": r, "left_clean": ns["clean_solution_text"](l), "right_clean": ns["clean_solution_text"](r),
                      "similarity": ns["_solution_jaccard_similarity"](l, r)})
    dump("a10_solution_jaccard.json", pairs)

    # ---- a2: reference's own known-answer case + fixture session (config 1) + synthetic packs
    item = {"action": "alpha beta gamma delta epsilon zeta eta theta", "owner": "", "timeline": "", "metric": ""}
    fact = {"q_id": "Q1", "question": "alpha beta gamma delta epsilon iota kappa lambda mu nu xi omicron", "answer": "", "quality_score": 0.55}
    dense = {"facts": [fact], "unknowns": [], "blindspots": [],
             "quality_snapshot": {"average_quality_score": 0.55, "total_formal_questions": 3,
                                  "answer_mode_distribution": {"pick_only": 1, "pick_with_reason": 2, "text_only": 0, "mixed": 0},
                                  "evidence_intent_distribution": {"low": 1, "medium": 1, "high": 1}}}
    sparse = {"facts": [fact], "unknowns": [{"q_id": "Q1", "reason": "待补采"}],
              "blindspots": [{"dimension": "业务流程", "aspect": f"盲区{i}"} for i in range(10)],
              "quality_snapshot": {"average_quality_score": 0.55, "total_formal_questions": 3,
                                   "answer_mode_distribution": {"pick_only": 3, "pick_with_reason": 0, "text_only": 0, "mixed": 0},
                                   "evidence_intent_distribution": {"low": 3, "medium": 0, "high": 0}}}
    cases = [{"name": "ref_test_dense", "field": "actions", "item": item, "pack": dense, "min_score": 0.46},
             {"name": "ref_test_sparse", "field": "actions", "item": item, "pack": sparse, "min_score": 0.46}]
    sol = json.loads((FIX / "deepvision-tech-solution-report.solution.json").read_text(encoding="utf-8"))
    facts = parse_transcript(md)
    assert len(facts) == 31, len(facts)
    coverage = {k: {"name": n, "missing_aspects": []} for n, k in
                (("客户需求", "customer_needs"), ("业务流程", "business_process"), ("技术约束", "tech_constraints"), ("项目约束", "project_constraints"))}
    c1_pack = {"facts": facts, "unknowns": [], "blindspots": [], "dimension_coverage": coverage,
               "quality_snapshot": sol.get("quality_snapshot", {})}
    draft = sol["draft"]
    for field in ("risks", "actions", "open_questions", "needs"):
        for i, it in enumerate(draft.get(field, [])):
            cases.append({"name": f"c1_{field}_{i}", "field": field, "item": it, "pack": "C1", "min_score": 0.46})
    # synthetic packs: claims partially copied from facts, odd inputs, ties
    for s in range(6):
        n_f = [1, 3, 40, 70, 5, 12][s]
        sfacts = []
        for f in range(n_f):
            sfacts.append({"q_id": ["Q%d" % (f + 1), " q%d " % (f + 1), "X%d" % f, "Q"][0 if f % 11 else (f // 11) % 4],
                           "question": rnd_text(rng, 30), "answer": rnd_text(rng, rng.choice([0, 5, 60, 400])),
                           "dimension": rng.choice(list(coverage)), "dimension_name": rng.choice(["客户需求", "业务流程"]),
                           "quality_score": rng.choice([0.0, 0.3, 0.55, 0.9, 1.7, -1, "0.4", None, "bad"]),
                           "hard_triggered": rng.random() < 0.2})
        if s == 4:
            sfacts += [dict(sfacts[0], q_id="Q90"), dict(sfacts[0], q_id="Q91")]      # exact ties: first wins
        pack = {"facts": sfacts, "unknowns": [{}] * (s % 3), "blindspots": [{}] * (s * 2), "dimension_coverage": coverage,
                "quality_snapshot": {"average_quality_score": [0.55, 0.25, 0.31, 0.8, 0.5, 0.29][s], "total_formal_questions": n_f,
                                     "answer_mode_distribution": {"pick_only": n_f // 2}, "evidence_intent_distribution": {"low": n_f // 3}}}
        for c in range(8 if n_f < 60 else 40):
            field = ["risks", "actions", "open_questions"][c % 3]
            src = rng.choice(sfacts)
            body = (src["question"] + src["answer"])
            a = rng.randint(0, max(0, len(body) - 10))
            copied = body[a:a + rng.choice([0, 4, 12, 30])]
            keys = {"risks": ["risk", "impact", "mitigation"], "actions": ["action", "owner", "timeline", "metric"],
                    "open_questions": ["question", "reason", "impact", "suggested_follow_up"]}[field]
            it = {k: "" for k in keys}
            it[keys[0]] = rnd_text(rng, 10) + copied
            it[keys[1]] = rnd_text(rng, rng.choice([0, 6]))
            if c % 4 == 0:
                it["dimension"] = rng.choice(list(coverage))
            if c % 5 == 0:
                it["title"] = "业务流程 梳理"
            cases.append({"name": f"syn{s}_{c}", "field": field, "item": it, "pack": f"S{s}", "min_score": [0.46, 0.3, None, 0.6][c % 4]})
        cases.append({"name": f"syn{s}_packdef", "pack_def": pack, "pack": f"S{s}"})
    packs = {"C1": c1_pack}
    out_cases = []
    for c in cases:
        if "pack_def" in c:
            packs[c["pack"]] = c["pack_def"]
    for c in cases:
        if "pack_def" in c:
            continue
        pack = packs[c["pack"]] if isinstance(c["pack"], str) else c["pack"]
        kwargs = {} if c["min_score"] is None else {"min_score": c["min_score"]}
        res = ns["infer_weak_evidence_refs_v3"](c["field"], c["item"], pack, **kwargs)
        out_cases.append(dict(c, result=res))
    dump("a2_bind.json", {"packs": packs, "cases": out_cases})


if __name__ == "__main__":
    main()
