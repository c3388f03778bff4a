"""TEST INFRASTRUCTURE ONLY — CPU restatement of the per-turn evidence annotation (rows a11-a13 of SURVEY.md §8).

Plain-Python restatement (small cases only: it is string logic, ~13 k turns/s) of
  evaluate_answer_depth        web/server.py:21221-21401
  _score_rich_option_answer    web/server.py:20318-20367
  evaluate_answer_quality      web/server.py:20370-20426
  classify_answer_evidence_class_v3   web/server.py:22552-22565
  _infer_legacy_log_capture_contract  web/server.py:21419-21466
  backfill_session_interview_log_evidence_annotations   web/server.py:21469-21584
It works directly on Python strings (`in`, `str.strip`, `str.isdigit`), i.e. not through the feature masks the product
uses, so it checks both the CUDA feature kernels and the host decision tree.  Pinned against the reference's own outputs:
tests/golden/a11_a13_annotations.json (tools/make_golden_annot.py runs the AST-loaded reference functions).
Never imported by the product package; only tests/ and __graft_entry__.smoke() may use it.
"""
from __future__ import annotations

import copy

CAUSE = ("导致", "影响", "因为", "所以", "瓶颈", "阻塞", "误拦截", "投诉", "风险", "失败", "延迟")
CONTEXT = ("高峰期", "夜间", "白天", "审批", "接口", "流程", "场景", "订单", "工单", "客户", "用户")
ROLE = ("客服", "销售", "运营", "风控", "研发", "产品", "实施", "财务", "法务", "商家")
PROCESS = ("流转", "复核", "上线", "对接", "集成", "回访", "分配", "排查", "协作", "审核")
SENSITIVITY = {"customer_needs": 0.8, "business_process": 0.6, "tech_constraints": 0.5, "project_constraints": 0.4}
HARD = {"vague_expression", "generic_answer", "contradiction_detected"}
MODES = {"pick_only", "pick_with_reason"}
INTENTS = {"low", "medium", "high"}
VAGUE = ["看情况", "不一定", "可能", "或许", "大概", "差不多", "到时候", "再说", "还没想好", "不确定", "看具体", "根据情况", "待定", "以后再说",
         "暂时不清楚", "目前还不好说", "都可以", "都行", "随便", "无所谓", "差不多", "一般", "不太了解", "没想过", "不知道", "说不好", "很难说"]
GENERIC = ["好的", "是的", "可以", "没问题", "需要", "应该要", "对", "嗯", "行", "同意", "没有", "不需要"]


def has_any(text, words):
    return any(w in text for w in words)


def norm_mode(v, fallback="pick_only"):
    v = str(v or "").strip().lower()
    if v in MODES:
        return v
    f = str(fallback or "").strip().lower()
    return f if f in MODES else "pick_only"


def norm_intent(v, fallback="low"):
    v = str(v or "").strip().lower()
    if v in INTENTS:
        return v
    f = str(fallback or "").strip().lower()
    return f if f in INTENTS else "low"


def rich_option(question, answer, options, dimension):
    """-> (is_rich_option, exact_option)"""
    text = str(answer or "").strip()
    opts = [str(o or "").strip() for o in (options or []) if str(o or "").strip()]
    if not text or not opts:
        return False, False
    exact = text in opts
    matched = [o for o in opts if o in text]
    if not exact and not matched:
        return False, False
    primary = text if exact else max(matched, key=len)
    q = str(question or "").strip()
    dim = str(dimension or "").strip().lower()
    score = int(exact) + int(len(primary) >= 10) + int(len(primary) >= 16)
    score += int(has_any(primary, ("（", ")", "：", ":", "/", "-", "，")))
    score += int(has_any(primary, CAUSE)) + int(has_any(primary, CONTEXT)) + int(has_any(primary, ROLE)) + int(has_any(primary, PROCESS))
    score += int(any(ch.isdigit() for ch in primary))
    score += int(has_any(q, ("原因", "场景", "影响", "频次", "角色", "流程", "瓶颈")))
    if dim in {"tech_constraints", "project_constraints"} and has_any(primary, ("分钟", "小时", "天", "周", "月", "成本", "预算", "上线")):
        score += 1
    return bool(score >= 4 or (exact and score >= 3 and len(primary) >= 12)), exact


def answer_depth(question, answer, dimension, options=None, is_follow_up=False, multi_select=False, answer_mode="",
                 requires_rationale=False, evidence_intent=""):
    signals = []
    ans = answer.strip()
    n = len(ans)
    sens = SENSITIVITY.get(dimension, 0.5)
    mode = norm_mode(answer_mode, "pick_with_reason" if requires_rationale else "pick_only")
    need_why = bool(requires_rationale or mode == "pick_with_reason")
    intent = norm_intent(evidence_intent, "high" if need_why else "low")
    pick_only = mode == "pick_only" and not need_why
    rich, exact = rich_option(question, ans, options, dimension)
    thr = int(20 + sens * 20)
    if options:
        thr = min(thr, (10 if multi_select else 6) if pick_only else (18 if multi_select else 12))
    if n < thr and not rich:
        signals.append("too_short")
    if has_any(ans, VAGUE):
        signals.append("vague_expression")
    if ans in GENERIC:
        signals.append("generic_answer")
    if options:
        informative = n >= 10 or rich or has_any(ans, ("（", ")", "：", ":", "如", "例如", "比如", "因为"))
        if not pick_only and exact and n < 16 and not informative:
            signals.append("option_only")
        if rich:
            signals.append("rich_option_answer")
    numbers = any(c.isdigit() for c in ans)
    if dimension in ["tech_constraints", "project_constraints"] and not numbers and n < 60:
        signals.append("no_quantification")
    if multi_select and options and "；" not in ans and len(options) >= 3 and not rich:
        if sum(1 for o in options if o in ans) <= 1 and n < 30:
            signals.append("single_selection")
    pairs = [("需要", "不需要"), ("可以", "不可以"), ("支持", "不支持"), ("已经", "还没"), ("有", "没有"), ("必须", "可选")]
    if any(a in ans and b in ans for a, b in pairs):
        signals.append("contradiction_detected")
    enough = []
    if n > 80:
        enough.append("detailed_answer")
    if "；" in ans and n > 40:
        enough.append("multi_point_answer")
    if numbers and n > 30:
        enough.append("quantified_answer")
    if rich:
        enough.append("rich_option_answer")
    w = {"too_short": 0.4, "vague_expression": 0.5, "generic_answer": 0.8, "option_only": 0.3, "no_quantification": 0.2,
         "single_selection": 0.2, "contradiction_detected": 0.6}
    score = sum(w.get(s, 0.1) for s in signals)
    score *= sens
    ew = {"detailed_answer": 0.5, "multi_point_answer": 0.3, "quantified_answer": 0.2, "rich_option_answer": 0.25 if intent == "high" else 0.18}
    score -= sum(ew.get(s, 0) for s in enough)
    hard = any(s in HARD for s in signals)
    reason_map = {"too_short": "回答过于简短，需要补充具体细节", "vague_expression": "回答包含模糊表述，需要明确具体要求",
                  "generic_answer": "回答过于笼统，需要深入了解具体需求", "option_only": "当前答案方向明确，但仍需补一个关键依据以支撑后续结论",
                  "no_quantification": "缺少量化指标，需要明确具体数据要求", "single_selection": "只选择了单一选项，需要了解是否还有其他需求",
                  "contradiction_detected": "回答中存在前后冲突，需要澄清真实约束"}
    rs = [reason_map[s] for s in signals if s in reason_map]
    reason = rs[0] if rs else "需要进一步了解详细需求"
    if score >= 0.4:
        needs, suggest = True, False
    elif score >= 0.15 and not enough:
        needs, suggest = False, True
    else:
        needs, suggest, reason = False, False, None
    return {"needs_follow_up": needs, "suggest_ai_eval": suggest, "reason": reason, "signals": signals, "hard_triggered": hard,
            "has_numbers": numbers, "follow_up_score": round(score, 2)}


def answer_quality(ev, answer, is_follow_up, follow_up_round):
    signals = ev.get("signals", [])
    hard = ev.get("hard_triggered", False)
    q = 0.5
    n = len((answer or "").strip())
    if n >= 120:
        q += 0.2
    elif n >= 80:
        q += 0.1
    if ev.get("has_numbers"):
        q += 0.1
    if has_any(answer or "", ["比如", "例如", "当", "如果", "场景", "案例"]):
        q += 0.1
    if "rich_option_answer" in signals:
        q += 0.12
    pen = {"too_short": 0.2, "vague_expression": 0.25, "generic_answer": 0.3, "option_only": 0.2, "no_quantification": 0.1,
           "single_selection": 0.1, "contradiction_detected": 0.25}
    q -= sum(pen.get(s, 0.05) for s in signals)
    if is_follow_up and follow_up_round >= 1 and not hard:
        q += 0.05
    q = max(0.0, min(1.0, q))
    qs = []
    if q >= 0.75:
        qs.append("high_quality")
    if ev.get("has_numbers"):
        qs.append("quantified")
    if n >= 80:
        qs.append("detailed")
    if "rich_option_answer" in signals:
        qs.append("rich_option")
    for s in signals:
        if s not in qs:
            qs.append(s)
    return {"quality_score": round(q, 2), "quality_signals": qs, "hard_triggered": bool(hard)}


def evidence_class(signals, quality_score, evidence_intent):
    sig = {str(s or "").strip().lower() for s in (signals or []) if str(s or "").strip()}
    intent = norm_intent(evidence_intent, "low")
    try:
        q = float(quality_score)
    except Exception:
        q = 0.0
    q = max(0.0, min(1.0, q))
    if "rich_option_answer" in sig:
        return "rich_option"
    if sig & {"vague_expression", "generic_answer", "contradiction_detected"}:
        return "pending_follow_up" if intent in {"medium", "high"} else "weak_inferred"
    if sig & {"too_short", "option_only", "no_quantification", "single_selection"}:
        return "pending_follow_up" if intent == "high" else "weak_inferred"
    if q < 0.35 and intent in {"medium", "high"}:
        return "weak_inferred"
    return "explicit"


def legacy_contract(log, prior):
    dim = str(log.get("dimension", "") or "").strip()
    question = str(log.get("question", "") or "").strip()
    options = log.get("options", []) if isinstance(log.get("options", []), list) else []
    follow = bool(log.get("is_follow_up", False) or int(log.get("follow_up_round", 0) or 0) > 0)
    hard = bool(log.get("hard_triggered", False))
    formal = len([p for p in (prior or []) if not p.get("is_follow_up", False)])
    if follow or hard:
        c = {"answer_mode": "pick_with_reason", "requires_rationale": True, "evidence_intent": "high"}
    elif formal >= 2 or dim.lower() in {"tech_constraints", "project_constraints"}:
        c = {"answer_mode": "pick_with_reason", "requires_rationale": True, "evidence_intent": "medium"}
    else:
        c = {"answer_mode": "pick_only", "requires_rationale": False, "evidence_intent": "low"}
    asks = has_any(question, ("原因", "为什么", "场景", "影响", "频次", "角色", "流程", "主要是因为", "哪种情况"))
    rich = any(len(str(o or "").strip()) >= 10 and has_any(str(o or ""), CAUSE + CONTEXT + ROLE + PROCESS) for o in options)
    if asks or rich:
        c["answer_mode"], c["requires_rationale"] = "pick_with_reason", True
        if follow or asks:
            c["evidence_intent"] = "high"
        elif c["evidence_intent"] == "low":
            c["evidence_intent"] = "medium"
    return c


KEYS = ("answer_mode", "requires_rationale", "evidence_intent", "follow_up_signals", "quality_score", "quality_signals",
        "hard_triggered", "needs_follow_up", "answer_evidence_class")


def backfill(session, refresh_quality=True, overwrite_contract=False):
    zero = {"changed": False, "logs_total": 0, "logs_updated": 0, "field_updates": {}}
    if not isinstance(session, dict) or not isinstance(session.get("interview_log", []), list):
        return zero
    logs = session.get("interview_log", [])
    changed, updated, fields, hist = False, 0, {}, {}
    for log in logs:
        if not isinstance(log, dict):
            continue
        dim = str(log.get("dimension", "") or "").strip()
        prior = hist.setdefault(dim, [])
        inf = legacy_contract(log, prior)
        rm_, ri = str(log.get("answer_mode", "") or "").strip().lower(), str(log.get("evidence_intent", "") or "").strip().lower()
        explicit = rm_ in MODES and ri in INTENTS
        mode = norm_mode(rm_, inf["answer_mode"]) if (explicit and not overwrite_contract) else inf["answer_mode"]
        why = bool(log.get("requires_rationale", False) or mode == "pick_with_reason")
        if overwrite_contract or not explicit:
            why = bool(inf["requires_rationale"] or mode == "pick_with_reason")
        intent = norm_intent(log.get("evidence_intent", "") if (explicit and not overwrite_contract) else inf["evidence_intent"],
                             "high" if why else ("medium" if mode == "pick_with_reason" else "low"))
        before = {k: copy.deepcopy(log.get(k)) for k in KEYS}
        log["answer_mode"], log["requires_rationale"], log["evidence_intent"] = mode, why, intent
        options = log.get("options", []) if isinstance(log.get("options", []), list) else []
        ev = answer_depth(str(log.get("question", "") or ""), str(log.get("answer", "") or ""), dim, options,
                          bool(log.get("is_follow_up", False)), bool(log.get("question_multi_select", log.get("multi_select", False))),
                          mode, why, intent)
        ql = answer_quality(ev, str(log.get("answer", "") or ""), bool(log.get("is_follow_up", False)), int(log.get("follow_up_round", 0) or 0))
        cls = evidence_class(ev["signals"], ql["quality_score"], intent)
        if refresh_quality or "follow_up_signals" not in log:
            log["follow_up_signals"], log["needs_follow_up"] = list(ev["signals"]), bool(ev["needs_follow_up"])
        if refresh_quality or "quality_score" not in log:
            log["quality_score"], log["quality_signals"], log["hard_triggered"] = ql["quality_score"], list(ql["quality_signals"]), bool(ql["hard_triggered"])
        log["answer_evidence_class"] = cls
        after = {k: log.get(k) for k in KEYS}
        if before != after:
            changed, updated = True, updated + 1
            for k in KEYS:
                if before.get(k) != after[k]:
                    fields[k] = fields.get(k, 0) + 1
        prior.append(log)
    return {"changed": changed, "logs_total": len([x for x in logs if isinstance(x, dict)]), "logs_updated": updated, "field_updates": fields}
