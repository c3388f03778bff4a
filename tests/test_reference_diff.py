"""Live differential tests against the reference's OWN functions (AST-executed from /root/reference/web/server.py by
oracle/ref_loader.py): the oracle (C restatement) and the host mirrors of the package on >= 10 k generated inputs, the
reference's repair pass (a3) driven through the drop-in rebinding, and the chunk manifest (a5) against the file the
reference writes.  Build container only: skipped where the reference is not mounted (the GPU box).  CUDA calls of the
host mirrors are replaced by the oracle stand-ins the other CPU tests use, so this file checks HOST logic + oracle."""
import copy
import json
import random
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tools"))
from oracle import ref_loader  # noqa: E402

pytestmark = [pytest.mark.reference,
              pytest.mark.skipif(not ref_loader.available(), reason="reference not mounted (GPU box)")]

CJK = [chr(0x4E00 + i) for i in range(120)] + list("需求系统审批流程客户业务痛点方案数据接口订单风险上线成本预算离线补录")
SPECIAL = ["İ", "K", "Ａ", "Σ", "ς", "ǅ", "ß", "　", " ", "\n", "\n\n", "\t", "，", "。", "？", "！", "、", "_", "-", "A", "Z", "q",
           "0", "9", "😀", "̇", "'", "\"", "é", "É", "Ⅷ", "①", "²", "℃", "〇", "々", "鿿", "一", "䷿", "ꀀ", "（", "）", "【", "】", "“", "”"]
WORDS = ["当前", "目前", "主要", "最", "核心", "请", "你", "们", "痛点", "诉求", "最关键", "最大的", "哪些", "是什么", "方面", "一下"]


def rnd(rng, n, p_cjk=0.6, p_ascii=0.25):
    out = []
    for _ in range(n):
        r = rng.random()
        out.append(rng.choice(CJK) if r < p_cjk else rng.choice("abcdeXYZ0189_") if r < p_cjk + p_ascii
                   else rng.choice(WORDS) if r < p_cjk + p_ascii + 0.05 else rng.choice(SPECIAL))
    return "".join(out)


@pytest.fixture(scope="module")
def ref():
    return ref_loader.load()


def test_fuzz_a1_a6_a8_a10_tokenizers(ref, oracle):
    from deep_vision_b200 import question_similarity as qs, reference_material as rm
    rng = random.Random(4242)
    n = 0
    for i in range(4000):                                                     # a1: token STRINGS
        t = rnd(rng, rng.choice([0, 1, 2, 5, 13, 14, 30, 80, 200, 700]), rng.choice([0.6, 0.95, 0.1]))
        assert set(oracle.tokenize_v3(t)) == ref["_tokenize_similarity_text_v3"](t), repr(t[:60])
        n += 1
    for i in range(3000):                                                     # a6: ordered term list, oracle and host mirror
        t = rnd(rng, rng.choice([0, 3, 9, 25, 60, 300]), rng.choice([0.6, 0.95]))
        want = ref["_tokenize_reference_material_query"](t)
        assert oracle.query_terms(t) == want and rm._tokenize_reference_material_query(t) == want, repr(t[:60])
        n += 1
    stop = list(ref["QUESTION_SIMILARITY_STOPWORDS"])                           # this process's set order (web/server.py:20467)
    og = oracle.Gate(list(ref["QUESTION_SIMILARITY_SYNONYMS"].items()), stop)
    assert dict(qs.DEFAULT_SYNONYMS) == dict(ref["QUESTION_SIMILARITY_SYNONYMS"]) and set(qs.DEFAULT_STOPWORDS) == set(stop)
    for i in range(3000):                                                     # a8: normalised text
        t = rnd(rng, rng.choice([0, 1, 4, 12, 40, 120]), 0.5, 0.15)
        assert og.normalize(t) == ref["normalize_interview_question_text"](t), repr(t)
        n += 1
    for i in range(1500):                                                     # a10: rounded Jaccard of cleaned texts
        a = rnd(rng, rng.choice([0, 1, 2, 3, 20, 200]), 0.6, 0.3)
        b = a[: len(a) // 2] + rnd(rng, rng.choice([0, 2, 30]), 0.6, 0.3) if rng.random() < 0.6 else rnd(rng, 25)
        ca, cb = ref["clean_solution_text"](a), ref["clean_solution_text"](b)
        assert qs.clean_solution_text(a) == ca
        assert round(oracle.solution_jaccard(ca, cb)[0], 3) == ref["_solution_jaccard_similarity"](ca, cb), (ca, cb)
        n += 1
    assert n >= 10000


def test_fuzz_a4_chunks_and_a9_gate(ref, oracle):
    rng = random.Random(99)
    for i in range(300):                                                      # a4: spans and texts
        parts = []
        for _ in range(rng.randint(0, 12)):
            parts.append(rnd(rng, rng.choice([1, 50, 400, 1200, 2500]), 0.8, 0.1) + rng.choice(["\n", "\n\n", " \n ", "", "\n\n\n", "\t"]))
        doc = "".join(parts)
        want = ref["_build_reference_material_chunks"](doc)
        got = oracle.chunk_build(doc)
        assert [(c["start"], c["end"]) for c in want] == [(s, e) for (s, e, _ts, _te) in got]
        assert [c["text"] for c in want] == [doc[ts:te] for (_s, _e, ts, te) in got]
    stop = list(ref["QUESTION_SIMILARITY_STOPWORDS"])
    og = oracle.Gate(list(ref["QUESTION_SIMILARITY_SYNONYMS"].items()), stop)
    for thr in (0.0, 0.3, 0.62, 0.9, 1.0):                                      # a9: most recent similar log, 600 candidates per threshold
        hist = [rnd(rng, rng.choice([0, 1, 2, 8, 20, 35]), 0.7, 0.1) + "？" for _ in range(40)]
        cands = []
        for _ in range(600):
            h = rng.choice(hist)
            cands.append(h if rng.random() < 0.2 else (h[: rng.randint(0, len(h))] + rnd(rng, rng.randint(0, 6), 0.8, 0.1) + h[rng.randint(0, len(h)):])
                         if rng.random() < 0.6 else rnd(rng, rng.randint(0, 30)))
        got = og.gate(cands, hist, thr)
        logs = [{"question": h} for h in hist]
        for c, g in zip(cands, got):
            w = ref["find_similar_interview_question_log"](logs, c, thr)
            assert (w is None and g < 0) or (w is not None and g >= 0 and w is logs[int(g)]), (c, thr)


def _random_pack(rng, n_f):
    coverage = {k: {"name": nm, "missing_aspects": [rnd(rng, 4, 1.0, 0.0)] if rng.random() < 0.5 else []} for nm, k in
                (("客户需求", "customer_needs"), ("业务流程", "business_process"), ("技术约束", "tech_constraints"), ("项目约束", "project_constraints"))}
    facts = []
    for f in range(n_f):
        facts.append({"q_id": rng.choice(["Q%d" % (f + 1)] * 8 + [" q%d " % (f + 1), "X%d" % f, "Q", 17]),
                      "question": rnd(rng, 30), "answer": rnd(rng, rng.choice([0, 5, 60, 400])),
                      "dimension": rng.choice(list(coverage) + ["other"]), "dimension_name": rng.choice(["客户需求", "业务流程", ""]),
                      "quality_score": rng.choice([0.0, 0.3, 0.55, 0.9, 1.7, -1, "0.4", None, "bad", round(rng.random(), 2)]),
                      "hard_triggered": rng.random() < 0.2})
    if n_f > 2 and rng.random() < 0.3:
        facts += [dict(facts[0], q_id="Q90"), dict(facts[0], q_id="Q91")]            # exact ties: the first wins
    return {"facts": facts, "unknowns": [{}] * rng.randint(0, n_f), "blindspots": [{}] * rng.randint(0, 12), "dimension_coverage": coverage,
            "quality_snapshot": {"average_quality_score": rng.choice([0.55, 0.25, 0.30, 0.31, 0.8, 0.5, 0.29, "x"]), "total_formal_questions": rng.randint(0, n_f + 2),
                                 "answer_mode_distribution": {"pick_only": rng.randint(0, n_f)}, "evidence_intent_distribution": {"low": rng.randint(0, n_f)}}}


def _random_item(rng, field, pack):
    keys = {"risks": ["risk", "impact", "mitigation"], "actions": ["action", "owner", "timeline", "metric"],
            "open_questions": ["question", "reason", "impact", "suggested_follow_up"], "needs": ["title", "description"]}[field]
    it = {k: "" for k in keys}
    src = rng.choice(pack["facts"]) if pack["facts"] else {"question": "", "answer": ""}
    body = str(src.get("question", "")) + str(src.get("answer", ""))
    a = rng.randint(0, max(0, len(body) - 10))
    it[keys[0]] = rnd(rng, rng.choice([0, 10])) + body[a:a + rng.choice([0, 4, 12, 30])]
    it[keys[1]] = rnd(rng, rng.choice([0, 6]))
    if rng.random() < 0.25:
        it["dimension"] = rng.choice(list(pack["dimension_coverage"]) + ["nope"])
    if rng.random() < 0.2:
        it["title"] = "业务流程 梳理"
    return it


def test_fuzz_a2_host_logic_vs_reference(ref, oracle, monkeypatch):
    import conftest
    from deep_vision_b200 import evidence
    monkeypatch.setattr(evidence, "bind_packed", conftest.oracle_bind_packed)
    rng = random.Random(7)
    n = 0
    for s in range(120):
        pack = _random_pack(rng, rng.choice([0, 1, 3, 7, 12, 40]))
        for c in range(6):
            field = rng.choice(["risks", "actions", "open_questions", "needs"])
            item = _random_item(rng, field, pack)
            ms = rng.choice([0.46, 0.3, None, 0.6, 0.95, 0.1])
            kw = {} if ms is None else {"min_score": ms}
            want = ref["infer_weak_evidence_refs_v3"](field, item, pack, **kw)
            got = evidence.infer_weak_evidence_refs_v3(field, item, pack, **kw, ctx=object())
            assert got == want, (field, item, ms)
            n += 1
    assert n >= 600


class _NamespaceModule:
    """The AST-loaded reference namespace seen as a module: getattr / setattr go to the dict the functions resolve names in."""

    def __init__(self, ns):
        object.__setattr__(self, "_ns", ns)

    def __getattr__(self, name):
        try:
            return object.__getattribute__(self, "_ns")[name]
        except KeyError:
            raise AttributeError(name)

    def __setattr__(self, name, value):
        object.__getattribute__(self, "_ns")[name] = value


A3_NAMES = ['CUSTOM_REPORT_ALLOWED_SOURCES', 'CUSTOM_REPORT_COMPONENTS', 'CUSTOM_SECTION_SOURCE_HINTS', 'REPORT_TEMPLATE_ALIAS',
            'REPORT_TEMPLATE_ASSESSMENT_V1', 'REPORT_TEMPLATE_CUSTOM_V1', 'REPORT_TEMPLATE_STANDARD_V1',
            'REPORT_V3_BLINDSPOT_ACTION_REQUIRED_BALANCED', 'REPORT_V3_BLINDSPOT_ACTION_REQUIRED_QUALITY', 'REPORT_V3_PROFILE',
            'REPORT_V3_UNKNOWNS_TO_OPEN_QUESTIONS_ENABLED', 'REPORT_V3_UNKNOWNS_TO_OPEN_QUESTIONS_MAX_ITEMS',
            '_BRACKET_Q_REF_LIST_PATTERN', '_INLINE_EVIDENCE_MARKER_PATTERN', '_PAREN_Q_REF_LIST_PATTERN',
            '_build_analysis_blindspot_status_line_v3', '_build_blindspot_open_question_v3', '_build_blindspot_pending_action_v3',
            '_build_default_custom_report_schema', '_build_long_horizon_action_v3', '_build_overview_blindspot_status_line_v3',
            '_build_sparse_session_explicit_action_v3', '_collect_actionability_candidate_indexes_v3', '_collect_text_corpus_for_items_v3',
            '_deduplicate_structured_list_v3', '_demote_item_to_open_question_v3', '_extract_blindspot_aspect_from_text_v3',
            '_extract_issue_field_index_v3', '_infer_action_owner_timeline_metric_v3', '_infer_analysis_key_for_blindspot_v3',
            '_infer_custom_section_config', '_infer_default_action_metric_v3', '_infer_default_action_owner_v3',
            '_infer_default_action_timeline_v3', '_is_action_timeline_specific_v3', '_is_evidence_index_claim_orphan_v3',
            '_normalize_action_metrics_v3', '_normalize_actionable_timelines_v3', '_normalize_dimension_hint_v3',
            '_pick_evidence_refs_for_dimension_v3', '_pick_strong_action_seed_refs_v3', '_reinforce_long_horizon_actions_v3',
            '_stabilize_sparse_actions_v3', 'apply_deterministic_report_repairs_v3', 'is_custom_report_source_empty_v3',
            'normalize_custom_report_schema', 'normalize_report_profile_choice', 'normalize_report_template_name',
            'resolve_custom_report_source_value_v3', 'strip_inline_evidence_markers', 'summarize_issue_types_v3', 'validate_report_draft_v3']


def _load_a3_namespace():
    """apply_deterministic_report_repairs_v3 (web/server.py:25241-25612), _reinforce_long_horizon_actions_v3 (25039-25093),
    validate_report_draft_v3 and every helper they reach, from the read-only mount; the config getters are stubbed to
    their defaults, the two helpers server.py imports come from the reference's own report_generation_runtime.py."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("dv_ref_rgr", str(ref_loader.REFERENCE_ROOT / "web" / "server_modules" / "report_generation_runtime.py"))
    rgr = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rgr)
    import os
    overrides = {"_cfg_get": lambda name, default: default, "_cfg_bool": lambda name, default: default,
                 "_cfg_text": lambda name, default="": default, "_cfg_int": lambda name, default: default,
                 "_cfg_float": lambda name, default: default, "os": os,
                 "module_classify_action_timeline_bucket_v3": rgr._classify_action_timeline_bucket_v3,
                 "module_is_action_metric_measurable_v3": rgr._is_action_metric_measurable_v3}
    ns = ref_loader.load(extra_names=A3_NAMES, overrides=overrides)
    assert not ns["__missing__"], ns["__missing__"]
    return ns


def _a3_cases():
    """Inputs of the reference's own tests for the repair pass (tests/test_security_regression.py:3232-3338, 3394-3445,
    3642-3760 restated as data) plus generated drafts."""
    pack = {"facts": [{"q_id": "Q1", "dimension": "business_process", "dimension_name": "业务流程", "question": "跨部门协同现状", "answer": "目前职责边界不清晰", "quality_score": 0.72},
                      {"q_id": "Q2", "dimension": "business_process", "dimension_name": "业务流程", "question": "角色分工是否明确", "answer": "还不明确", "quality_score": 0.68}],
            "quality_snapshot": {"average_quality_score": 0.45},
            "dimension_coverage": {"business_process": {"name": "业务流程", "missing_aspects": ["角色分工"]}}}
    draft = {"overview": "概述", "needs": [], "analysis": {}, "visualizations": {}, "solutions": [],
             "risks": [{"risk": "跨部门协同阻塞", "impact": "交付延期", "mitigation": "建立周会", "evidence_refs": []}],
             "actions": [{"action": "明确角色分工", "owner": "项目经理", "timeline": "2周内", "metric": "职责清单发布", "evidence_refs": []}],
             "open_questions": [], "evidence_index": [{"claim": "结论A", "confidence": "high", "evidence_refs": []}]}
    issues = [{"type": "no_evidence", "target": "risks[0]", "message": "缺少证据"}, {"type": "no_evidence", "target": "actions[0]", "message": "缺少证据"},
              {"type": "no_evidence", "target": "evidence_index[0]", "message": "缺少证据"}]
    cases = [(draft, pack, issues, ""), (draft, pack, issues, "quality"),
             (draft, pack, [{"type": "not_actionable", "target": "actions", "message": "缺少长期运营复盘机制"}], "balanced")]
    rng = random.Random(31)
    for s in range(25):
        p = _random_pack(rng, rng.choice([2, 6, 14, 30]))
        for f in p["facts"]:
            f["q_id"] = "Q%d" % (p["facts"].index(f) + 1)
        d = {"overview": "概述", "needs": [], "analysis": {}, "visualizations": {}, "solutions": [], "evidence_index": []}
        iss = []
        for field in ("risks", "actions", "open_questions"):
            d[field] = []
            for i in range(rng.randint(0, 5)):
                it = _random_item(rng, field, p)
                it["evidence_refs"] = rng.choice([[], [], ["Q1"], ["Q99"], "Q2"])
                d[field].append(it)
                if rng.random() < 0.7:
                    iss.append({"type": rng.choice(["no_evidence", "no_evidence", "not_actionable"]), "target": "%s[%d]" % (field, i),
                                "message": rng.choice(["缺少证据", "缺少长期运营复盘机制", "时间不具体"])})
        cases.append((d, p, iss, rng.choice(["", "balanced", "quality"])))
    return cases


def test_a3_repair_pass_through_the_dropin(oracle, monkeypatch):
    """The reference's OWN repair loop with the drop-in bound equals the reference-only run, and the a2 calls of one
    pass are scored in ONE launch (plus single calls only for items the pass rewrote before binding them)."""
    import conftest
    from deep_vision_b200 import dropin, evidence
    launches = []

    def counting_bind_packed(ctx, *a, **kw):
        launches.append(int(a[4][-1]))                                    # claims of this launch
        return conftest.oracle_bind_packed(ctx, *a, **kw)
    monkeypatch.setattr(evidence, "bind_packed", counting_bind_packed)
    plain = _load_a3_namespace()
    bound_ns = _load_a3_namespace()
    assert bound_ns is not plain
    mod = _NamespaceModule(bound_ns)
    done = dropin.install(mod, ctx=object())
    assert "apply_deterministic_report_repairs_v3" in done and "infer_weak_evidence_refs_v3" in done
    per_item_calls = []
    ref_infer = plain["infer_weak_evidence_refs_v3"]
    plain["infer_weak_evidence_refs_v3"] = lambda *a, **kw: (per_item_calls.append(1), ref_infer(*a, **kw))[1]
    total_ref_calls = total_launches = 0
    for draft, pack, issues, profile in _a3_cases():
        per_item_calls.clear(); launches.clear()
        want = plain["apply_deterministic_report_repairs_v3"](copy.deepcopy(draft), copy.deepcopy(pack), copy.deepcopy(issues), profile)
        got = bound_ns["apply_deterministic_report_repairs_v3"](copy.deepcopy(draft), copy.deepcopy(pack), copy.deepcopy(issues), profile)
        assert json.dumps(got, sort_keys=True, ensure_ascii=False, default=str) == json.dumps(want, sort_keys=True, ensure_ascii=False, default=str)
        total_ref_calls += len(per_item_calls); total_launches += len(launches)
        if per_item_calls:
            assert launches and launches[0] >= 1                          # the pass pre-scored its candidates in one launch
    assert total_ref_calls > 20 and total_launches < total_ref_calls       # fewer launches than the reference makes a2 calls
    # validate_report_draft_v3 runs on the repaired draft in the same namespace (the reference re-validates after repairs)
    draft, pack, issues, profile = _a3_cases()[0]
    repaired = bound_ns["apply_deterministic_report_repairs_v3"](copy.deepcopy(draft), pack, issues, profile)
    w = plain["apply_deterministic_report_repairs_v3"](copy.deepcopy(draft), pack, issues, profile)
    assert bound_ns["validate_report_draft_v3"](repaired["draft"] if "draft" in repaired else repaired, pack) == \
        plain["validate_report_draft_v3"](w["draft"] if "draft" in w else w, pack)
    dropin.uninstall(mod)
    assert bound_ns["infer_weak_evidence_refs_v3"].__module__ != "deep_vision_b200.dropin"


def test_a5_manifest_equals_the_file_the_reference_writes(ref, oracle, monkeypatch, tmp_path):
    """save_reference_material_fulltext_artifacts (web/server.py:31973-32017) writes chunks.json; build_chunk_manifest must
    produce the same JSON, and load_chunks_from_manifest must read the reference's file like the reference's loader."""
    from test_reference_material import _oracle_standins
    from deep_vision_b200 import reference_material as rm
    _oracle_standins(monkeypatch, oracle)
    ns = ref_loader.load(extra_names=["save_reference_material_fulltext_artifacts", "load_reference_material_chunks"],
                         overrides={"_reference_material_storage_dir": lambda sid, did: tmp_path / sid / did,
                                    "_relative_data_ref": lambda p: str(p), "_resolve_data_ref": lambda r: Path(r)})
    rng = random.Random(3)
    docs = ["开头普通说明。" * 1800 + "\n\n尾部关键需求：必须支持离线审批和断网补录，并保留审批链路证据。", "短文档",
            "".join(rnd(rng, rng.choice([30, 900, 2600]), 0.8, 0.1) + rng.choice(["\n", "\n\n", "", " "]) for _ in range(9))]
    for i, content in enumerate(docs):
        meta = ns["save_reference_material_fulltext_artifacts"](session_id="s1", doc_id="d%d" % i, filename="文档%d.md" % i, content=content, ext=".md")
        path = Path(meta["chunk_manifest_ref"])
        theirs = json.loads(path.read_text(encoding="utf-8"))
        ours = rm.build_chunk_manifest(doc_id="d%d" % i, filename="文档%d.md" % i, content=content, ext=".md", ctx=object())
        assert json.dumps(ours, ensure_ascii=False) == json.dumps(theirs, ensure_ascii=False)        # key order included
        assert rm.load_chunks_from_manifest(path) == ns["load_reference_material_chunks"]({"chunk_manifest_ref": str(path)})
        assert meta["chunk_count"] == len(ours["chunks"])
    assert ns["save_reference_material_fulltext_artifacts"](session_id="s", doc_id="e", filename="f", content=" \n", ext=".md") == {}


def test_unicode_table_version_is_checked_at_install(monkeypatch):
    import unicodedata
    import warnings
    from deep_vision_b200 import _lib, dropin
    assert _lib.load().dvb_unicode_version().decode() == unicodedata.unidata_version
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        assert dropin.check_unicode_tables() is True
    monkeypatch.setattr(unicodedata, "unidata_version", "99.0.0")
    with pytest.warns(RuntimeWarning):
        assert dropin.check_unicode_tables() is False
    with pytest.raises(RuntimeError):
        dropin.check_unicode_tables(strict=True)


def test_fuzz_f4_slim_facts_live(ref):
    """the reference's select_slimmed_facts_for_prompt (loaded once per settings row) against the host
    extraction of the package + the oracle restatement on fresh random evidence packs."""
    import make_golden_slim as mg
    from deep_vision_b200 import prompt_facts as pf
    from oracle import dv_slim
    rng = random.Random(99)
    packs = [(mg.rand_pack(rng, rng.choice([1, 4, 9, 17, 40, 90])), rng.choice([1, 3, 8, 24, 60, None])) for _ in range(400)]
    n = 0
    for st in mg.SETTINGS[:3]:
        fn = ref_loader.load(extra_names=mg.NAMES, overrides=st)["select_slimmed_facts_for_prompt"]
        for pack, limit in packs:
            facts = [f for f in pack.get("facts", []) if isinstance(f, dict)]
            want = [next(i for i, f in enumerate(facts) if f is o) for o in fn(pack, limit)]
            q, fl, sg, dm, ky = pf.extract_fact_arrays(pack, facts)
            got = dv_slim.slim_select(q, fl, sg, dm, ky, max(1, int(limit or 1)), dim_quota=max(1, st["REPORT_V3_EVIDENCE_DIM_QUOTA"]),
                                      min_quality=st["REPORT_V3_EVIDENCE_MIN_QUALITY"], keep_hard=st["REPORT_V3_EVIDENCE_KEEP_HARD_TRIGGERED"],
                                      dedup=st["REPORT_V3_EVIDENCE_DEDUP_ENABLED"])
            assert got == want
            n += 1
    assert n == 1200


def test_a2_session_cache_follows_in_place_changes(ref, oracle, monkeypatch):
    """evidence._session_for keeps the packed facts of an evidence pack between calls (the reference calls
    infer_weak_evidence_refs_v3 once per draft item on the same pack).  Every way of changing the pack in place — a fact's
    text, quality, q_id, dimension, a fact replaced, appended, removed, a value swapped for an equal one of another type —
    must show in the next call exactly as it does in the reference."""
    import conftest
    from deep_vision_b200 import evidence
    monkeypatch.setattr(evidence, "bind_packed", conftest.oracle_bind_packed)
    evidence._SESSION_CACHE.clear()
    built = []
    real_session = evidence._Session
    class Counting(real_session):
        __slots__ = ()
        def __init__(self, pack):
            built.append(1)
            real_session.__init__(self, pack)
    monkeypatch.setattr(evidence, "_Session", Counting)
    rng = random.Random(11)
    pack = _random_pack(rng, 12)
    items = [(f, _random_item(rng, f, pack)) for f in ("risks", "actions", "open_questions") for _ in range(4)]

    def check():
        for field, item in items:
            want = ref["infer_weak_evidence_refs_v3"](field, item, pack)
            got = evidence.infer_weak_evidence_refs_v3(field, item, pack, ctx=object())
            assert got == want, (field, item)

    check()
    assert len(built) == 1                                           # twelve calls, one packing
    check()
    assert len(built) == 1
    facts = pack["facts"]
    edits = [
        lambda: facts[0].__setitem__("answer", facts[0]["answer"] + " 离线审批 断网补录"),
        lambda: facts[1].__setitem__("quality_score", 0.99),
        lambda: facts[2].__setitem__("q_id", "nope"),
        lambda: facts[3].__setitem__("dimension", "customer_needs" if facts[3].get("dimension") != "customer_needs" else "tech_constraints"),
        lambda: facts[4].__setitem__("hard_triggered", not bool(facts[4].get("hard_triggered"))),
        lambda: facts.__setitem__(5, dict(facts[5], question="预算 审批 流程 " + str(facts[5].get("question", "")))),
        lambda: facts.append(dict(facts[0], q_id="Q99")),
        lambda: facts.pop(1),
        lambda: facts[0].__setitem__("quality_score", 1),              # int 1, then True: equal values, different objects
        lambda: facts[0].__setitem__("quality_score", True),
        lambda: facts[0].pop("dimension_name", None),
    ]
    for n, edit in enumerate(edits, start=2):
        edit()
        check()
        assert len(built) == n, n                                      # repacked once per edit, not per call
    # a value that can change without changing identity is never cached
    facts[0]["answer"] = ["列表", "答案"]
    before = len(built)
    check()
    assert len(built) == before + len(items)
    facts[0]["answer"].append("预算有限 需要 审批")
    check()
    assert id(pack) not in evidence._SESSION_CACHE
    # a new pack object at a recycled address cannot hit the old entry: the entry is validated by the identity of its values
    evidence._SESSION_CACHE.clear()
    other = _random_pack(rng, 7)
    evidence.infer_weak_evidence_refs_v3("risks", items[0][1], other, ctx=object())
    ent = evidence._SESSION_CACHE[id(other)]
    evidence._SESSION_CACHE[id(pack)] = ent                            # as if `pack` had been allocated where `other` was
    facts[0]["answer"] = "普通答案"
    check()
