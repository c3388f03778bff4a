#!/usr/bin/env python3
"""Generate tests/golden/a11_a13_annotations.json by running the REAL reference functions (AST-loaded from
/root/reference/web/server.py by oracle/ref_loader.py) on seeded synthetic interview logs + hand-made edge cases.
Run in the build container only (the mount does not exist on the GPU box):  python tools/make_golden_annot.py"""
import copy
import json
import random
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import ref_loader  # noqa: E402

WORDS = ["看情况", "可能", "差不多", "都可以", "不知道", "需要", "不需要", "可以", "不可以", "支持", "不支持", "已经", "还没", "有", "没有", "必须", "可选",
         "比如", "例如", "当", "如果", "场景", "案例", "因为", "所以", "导致", "影响", "瓶颈", "高峰期", "夜间", "审批", "接口", "流程", "订单",
         "客服", "销售", "运营", "风控", "研发", "流转", "复核", "上线", "对接", "集成", "分钟", "小时", "天", "周", "月", "成本", "预算",
         "；", "，", "。", "（", ")", "：", ":", "/", "-", " ", "　", "\n", "3", "12", "５", "②", "一", "系统", "平台", "数据", "我们", "目前",
         "主要", "问题", "处理", "人工", "效率", "abc", "API", "好的", "是的", "对", "嗯", "行", "同意", "没问题", "应该要", "如"]
QWORDS = ["原因", "为什么", "场景", "影响", "频次", "角色", "流程", "主要是因为", "哪种情况", "瓶颈", "您", "目前", "最大", "的", "是什么", "？", "哪些", "如何"]
DIMS = ["customer_needs", "business_process", "tech_constraints", "project_constraints", "", "Tech_Constraints", " custom "]


def text(rng, words, lo, hi):
    return "".join(rng.choice(words) for _ in range(rng.randint(lo, hi)))


def make_log(rng):
    options = []
    if rng.random() < 0.7:
        options = [text(rng, WORDS, 1, 8) for _ in range(rng.randint(1, 5))]
        if rng.random() < 0.1:
            options.append("")
        if rng.random() < 0.1:
            options.append("  " + options[0] + " ")
    r = rng.random()
    if options and r < 0.3:
        answer = rng.choice(options)
    elif options and r < 0.5:
        answer = "；".join(rng.sample(options, min(len(options), rng.randint(1, 3)))) + text(rng, WORDS, 0, 6)
    elif r < 0.6:
        answer = rng.choice(["好的", "是的", "可以", "没问题", "需要", "对", "嗯", " 行 ", "同意", "没有", "不需要", "", "  "])
    else:
        answer = text(rng, WORDS, 1, 60)
    if rng.random() < 0.15:
        answer = rng.choice([" ", "\n", "　", "\t"]) + answer + rng.choice([" ", "\n", " "])
    log = {"dimension": rng.choice(DIMS), "question": text(rng, QWORDS, 2, 10), "answer": answer, "options": options}
    if rng.random() < 0.4:
        log["is_follow_up"] = rng.random() < 0.5
    if rng.random() < 0.3:
        log["follow_up_round"] = rng.choice([0, 1, 2])
    if rng.random() < 0.5:
        log[rng.choice(["question_multi_select", "multi_select"])] = rng.random() < 0.6
    if rng.random() < 0.3:
        log["answer_mode"] = rng.choice(["pick_only", "pick_with_reason", "PICK_ONLY ", "bogus", ""])
        log["evidence_intent"] = rng.choice(["low", "medium", "high", "HIGH", "bogus", ""])
    if rng.random() < 0.2:
        log["requires_rationale"] = rng.random() < 0.5
    if rng.random() < 0.2:
        log["hard_triggered"] = rng.random() < 0.5
    if rng.random() < 0.15:
        log["quality_score"] = round(rng.random(), 2)
        log["follow_up_signals"] = ["too_short"]
    if rng.random() < 0.05:
        log["options"] = "not-a-list"
    return log


# ---- stand-ins for the app-level helpers build_report_evidence_pack calls (the reference's own test stubs the same
# functions: tests/test_security_regression.py:4215-4219).  Shared by the golden generator and tests/test_annotations.py.
class PackStubs:
    REPORT_TEMPLATE_CUSTOM_V1 = "custom_v1"
    DIMS = {
        "customer_needs": {"name": "客户需求", "key_aspects": ["核心痛点", "期望价值", "使用场景", "用户角色"]},
        "business_process": {"name": "业务流程", "key_aspects": ["关键流程", "角色分工"]},
        "tech_constraints": {"name": "技术约束", "key_aspects": []},
        "project_constraints": {"name": "项目约束"},
    }

    @staticmethod
    def get_dimension_info_for_session(session):
        return PackStubs.DIMS

    @staticmethod
    def get_interview_mode_config(session):
        return {"formal_questions_per_dim": 3, "max_formal_questions_per_dim": 4} if session.get("mode") != "quick" else {"formal_questions_per_dim": 2}

    @staticmethod
    def get_dimension_missing_aspects(session, dim_key):
        return list((session.get("missing", {}) or {}).get(dim_key, []))

    @staticmethod
    def resolve_report_template_for_session(session):
        return session.get("template", "standard_v1")

    @staticmethod
    def normalize_custom_report_schema(schema, fallback_sections=None):
        return ({"sections": list(fallback_sections or []), "schema": schema}, [])


def pack_session(rng, s):
    logs = [make_log(rng) for _ in range(rng.randint(0, 16))]
    logs = [log for log in logs if isinstance(log.get("options"), list)]
    for log in logs:
        if rng.random() < 0.1:
            log["answer"] = None if rng.random() < 0.5 else 42
    sess = {"session_id": "p%03d" % s, "topic": "主题%d" % s, "interview_log": logs,
            "dimensions": {k: {"coverage": rng.choice([0, 35, 70, 100, 140])} for k in list(PackStubs.DIMS)[: rng.randint(0, 4)]},
            "missing": {"customer_needs": rng.sample(["核心痛点", "期望价值", "使用场景", "用户角色", "其他"], rng.randint(0, 5)),
                        "business_process": rng.sample(["关键流程", "角色分工"], rng.randint(0, 2))},
            "scenario_config": {"report": {"type": rng.choice(["standard", "assessment"]), "sections": ["a", "b"], "schema": {"x": 1}}}}
    if s % 5 == 0:
        sess["template"] = "custom_v1"
    if s % 7 == 0:
        sess["mode"] = "quick"
    return sess


def main():
    ns = ref_loader.load()
    rng = random.Random(20260111)
    sessions = []
    for s in range(60):
        logs = [make_log(rng) for _ in range(rng.randint(0, 14))]
        if s % 9 == 0:
            logs.insert(rng.randint(0, len(logs)), "not-a-dict")
        sessions.append({"session_id": "s%03d" % s, "topic": "t%d" % s, "interview_log": logs})
    sessions.append({"session_id": "no-log"})
    sessions.append({"session_id": "bad-log", "interview_log": "x"})
    cases = []
    for sess in sessions:
        for refresh, overwrite in ((True, False), (False, True)):
            work = copy.deepcopy(sess)
            summary = ns["backfill_session_interview_log_evidence_annotations"](work, refresh_quality=refresh, overwrite_contract=overwrite)
            cases.append({"input": sess, "refresh_quality": refresh, "overwrite_contract": overwrite, "summary": summary,
                          "logs_after": work.get("interview_log") if isinstance(work, dict) else None})
    depth = []
    for _ in range(250):
        log = make_log(rng)
        opts = log["options"] if isinstance(log["options"], list) else []
        kw = dict(question=log["question"], answer=log["answer"], dimension=log["dimension"], options=opts,
                  is_follow_up=bool(log.get("is_follow_up", False)), multi_select=rng.random() < 0.5,
                  answer_mode=rng.choice(["", "pick_only", "pick_with_reason"]), requires_rationale=rng.random() < 0.4,
                  evidence_intent=rng.choice(["", "low", "medium", "high"]))
        ev = ns["evaluate_answer_depth"](**kw)
        fr = rng.choice([0, 1, 2])
        ql = ns["evaluate_answer_quality"](ev, log["answer"], kw["is_follow_up"], fr)
        cls = ns["classify_answer_evidence_class_v3"](ev["signals"], ql["quality_score"], kw["evidence_intent"])
        depth.append({"kwargs": kw, "follow_up_round": fr, "depth": ev, "quality": ql, "class": cls})
    from datetime import datetime, timezone
    ov = {k: getattr(PackStubs, k) for k in ("REPORT_TEMPLATE_CUSTOM_V1", "get_dimension_info_for_session", "get_interview_mode_config",
                                             "get_dimension_missing_aspects", "resolve_report_template_for_session",
                                             "normalize_custom_report_schema")}
    ov.update(datetime=datetime, timezone=timezone)
    ns2 = ref_loader.load(extra_names=["build_report_evidence_pack"], overrides=ov)
    packs = []
    for s in range(40):
        sess = pack_session(rng, s)
        before = copy.deepcopy(sess)
        pack = ns2["build_report_evidence_pack"](sess)
        assert sess == before
        pack.pop("generated_at")
        packs.append({"input": sess, "pack": pack})
    out = {"generator": "tools/make_golden_annot.py", "backfill": cases, "depth": depth, "packs": packs}
    p = ROOT / "tests" / "golden" / "a11_a13_annotations.json"
    p.write_text(json.dumps(out, ensure_ascii=False, separators=(",", ":")), encoding="utf-8")
    print("wrote", p, p.stat().st_size, "bytes;", len(cases), "backfill cases,", len(depth), "depth cases,", len(packs), "packs")


if __name__ == "__main__":
    main()
