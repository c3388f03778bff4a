#!/usr/bin/env python3
"""Golden vectors for SURVEY §8(f)2: calculate_dimension_saturation / get_dimension_missing_aspects
(web/server.py:19399-19515, 19583-19589) and build_report_evidence_pack with the REAL get_dimension_missing_aspects
(web/server.py:22711), produced by the reference's own functions (oracle/ref_loader.py) in the build container.
  python tools/make_golden_saturation.py   ->  tests/golden/f2_saturation.json"""
import copy
import json
import random
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))
from oracle import ref_loader                      # noqa: E402
from make_golden_annot import PackStubs, make_log  # noqa: E402


class SatStubs(PackStubs):
    """App configuration the saturation path reads through the server module."""
    DIMS = {
        "customer_needs": {"name": "客户需求", "key_aspects": ["核心痛点", "期望价值", "使用场景", "用户角色"]},
        "business_process": {"name": "业务流程", "key_aspects": ["关键流程", "角色分工", "触发事件", "异常处理"]},
        "tech_constraints": {"name": "技术约束", "key_aspects": ["部署方式", "系统集成", "性能要求", "安全合规", "安全合规", "", " 部署", "离线审批 "]},
        "project_constraints": {"name": "项目约束", "key_aspects": ["预算范围", "时间节点", "资源限制", "优先级", "自定义方面"]},
        "empty_aspects": {"name": "无方面", "key_aspects": []},
        "no_aspects": {"name": "缺省"},
    }

    @staticmethod
    def get_dimension_info_for_session(session):
        return SatStubs.DIMS

    @staticmethod
    def get_interview_mode_config(session):
        mode = session.get("mode")
        if mode == "quick":
            return {"formal_questions_per_dim": 2, "quality_thresholds": {"high": 0.55, "medium": 0.3}}
        if mode == "deep":
            return {"formal_questions_per_dim": 4, "max_formal_questions_per_dim": 6, "quality_thresholds": {"high": 0.95, "medium": 0.75, "low": 0.5}}
        return {"formal_questions_per_dim": 3, "max_formal_questions_per_dim": 4}


FRAGMENTS = ["我们的痛点是审批太慢", "希望有收益", "比如月底结账的情况下", "当系统失败时需要例外处理", "因为预算有限", "所以优先上云",
             "而不是本地部署", "相比之下更重要的是安全", "首先要对接ERP接口", "并发大概500", "响应要在２秒内", "约³个团队", "费用；周期、人力",
             "谁负责", "各部门职责", "没有想法", "不清楚", "核心痛点", "使用场景很多", "离线审批 能力", "新 部署方案", "自定义方面已确认",
             "what ever", "", " ", "启动流程的步骤", "加密和权限", "重要且紧急", "何时开始"]


def sat_session(rng, s):
    dims = list(SatStubs.DIMS) + ["unknown_dim"]
    logs = []
    for _ in range(rng.randint(0, 18)):
        log = make_log(rng)
        if not isinstance(log.get("options"), list):
            log["options"] = []
        log["dimension"] = rng.choice(dims)
        log["answer"] = "，".join(rng.sample(FRAGMENTS, rng.randint(0, 5))) + ("x" * rng.choice([0, 0, 0, 120, 400]))
        log["question"] = rng.choice(["", "请描述" + rng.choice(FRAGMENTS), log.get("question") or ""])
        if rng.random() < 0.1:
            del log["answer"]
        logs.append(log)
    sess = {"session_id": "f%03d" % s, "topic": "主题%d" % s, "interview_log": logs,
            "dimensions": {k: {"coverage": rng.choice([0, 35, 70, 100])} for k in list(SatStubs.DIMS)[: rng.randint(0, 6)]},
            "scenario_config": {"report": {"type": "standard"}}}
    if s % 4 == 1:
        sess["mode"] = "quick"
    if s % 4 == 2:
        sess["mode"] = "deep"
    return sess


def main():
    from datetime import datetime, timezone
    ov = {k: getattr(SatStubs, k) for k in ("REPORT_TEMPLATE_CUSTOM_V1", "get_dimension_info_for_session", "get_interview_mode_config",
                                            "resolve_report_template_for_session", "normalize_custom_report_schema")}
    ov.update(datetime=datetime, timezone=timezone)
    ns = ref_loader.load(extra_names=["SATURATION_THRESHOLDS", "get_mode_saturation_thresholds", "calculate_dimension_saturation",
                                      "get_dimension_missing_aspects", "build_report_evidence_pack"], overrides=ov)
    assert not ns["__missing__"], ns["__missing__"]
    rng = random.Random(20260212)
    cases, packs = [], []
    for s in range(70):
        sess = sat_session(rng, s)
        before = copy.deepcopy(sess)
        per_dim = {}
        for dim in list(SatStubs.DIMS) + ["unknown_dim"]:
            per_dim[dim] = {"saturation": ns["calculate_dimension_saturation"](sess, dim),
                            "missing": ns["get_dimension_missing_aspects"](sess, dim)}
        assert sess == before
        cases.append({"input": sess, "per_dimension": per_dim})
        if s < 30:
            pack = ns["build_report_evidence_pack"](sess)
            assert sess == before
            pack.pop("generated_at")
            packs.append({"input": sess, "pack": pack})
    out = {"generator": "tools/make_golden_saturation.py", "cases": cases, "packs": packs}
    p = ROOT / "tests" / "golden" / "f2_saturation.json"
    p.write_text(json.dumps(out, ensure_ascii=False, separators=(",", ":")), encoding="utf-8")
    levels = {}
    for c in cases:
        for d in c["per_dimension"].values():
            levels[d["saturation"]["level"]] = levels.get(d["saturation"]["level"], 0) + 1
    print("wrote", p, p.stat().st_size, "bytes;", len(cases), "sessions x 7 dimensions,", len(packs), "packs; levels", levels)


if __name__ == "__main__":
    main()
