"""Per-turn evidence annotation and the bulk re-index path (rows a11-a14 of SURVEY.md §8).

Host-side mirror of the reference interface (web/server.py unless noted):
  * `evaluate_answer_depth(question, answer, dimension, options=None, ...)`          21221
  * `evaluate_answer_quality(eval_result, answer, is_follow_up, follow_up_round)`    20370
  * `classify_answer_evidence_class_v3(signals, quality_score, evidence_intent)`     22552
  * `backfill_session_interview_log_evidence_annotations(session, *, refresh_quality=True, overwrite_contract=False)`  21469
  * `backfill_session_files(session_files, *, apply_changes=False, backup_dir=None, server_module=None)`
                                                             scripts/migrate_session_evidence_annotations.py:64

What runs where: every string scan of those functions — ~100 `keyword in text` tests, the exact-match lists,
`any(c.isdigit())`, `len(text.strip())`, `option in answer` — is done for ALL turns of ALL sessions of a batch by two
CUDA kernels (`dvb_text_features`, `dvb_contains`; csrc/annotate.cu).  The host then runs the reference's small
decision tree per turn over the resulting bit masks, in the reference's fp64 operation order (sum() left to right),
because the contract inference is sequential per session (it counts the PRIOR formal questions of the same dimension,
web/server.py:21432).  There is no CPU text scanner here.
"""
from __future__ import annotations

import copy
import json
import shutil
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from .packing import pack_texts, ptr

# ----------------------------------------------------------------------------- the reference's literals, as pattern groups
RICH_OPTION_CAUSE_KEYWORDS = ("导致", "影响", "因为", "所以", "瓶颈", "阻塞", "误拦截", "投诉", "风险", "失败", "延迟")       # 20304
RICH_OPTION_CONTEXT_KEYWORDS = ("高峰期", "夜间", "白天", "审批", "接口", "流程", "场景", "订单", "工单", "客户", "用户")    # 20307
RICH_OPTION_ROLE_KEYWORDS = ("客服", "销售", "运营", "风控", "研发", "产品", "实施", "财务", "法务", "商家")               # 20310
RICH_OPTION_PROCESS_KEYWORDS = ("流转", "复核", "上线", "对接", "集成", "回访", "分配", "排查", "协作", "审核")            # 20313
VAGUE_INDICATORS = ("看情况", "不一定", "可能", "或许", "大概", "差不多", "到时候", "再说", "还没想好", "不确定", "看具体", "根据情况",
                    "待定", "以后再说", "暂时不清楚", "目前还不好说", "都可以", "都行", "随便", "无所谓", "一般", "不太了解", "没想过",
                    "不知道", "说不好", "很难说")                                                                          # 21271-21280
GENERIC_ANSWERS = ("好的", "是的", "可以", "没问题", "需要", "应该要", "对", "嗯", "行", "同意", "没有", "不需要")          # 21286-21289
INFORMATIVE_TOKENS = ("（", ")", "：", ":", "如", "例如", "比如", "因为")                                                  # 21299
CONTRADICTION_PAIRS = (("需要", "不需要"), ("可以", "不可以"), ("支持", "不支持"), ("已经", "还没"), ("有", "没有"), ("必须", "可选"))  # 21322-21329
SCENARIO_KEYWORDS = ("比如", "例如", "当", "如果", "场景", "案例")                                                          # 20387
OPTION_PUNCT_TOKENS = ("（", ")", "：", ":", "/", "-", "，")                                                              # 20343
DIMENSION_UNIT_TOKENS = ("分钟", "小时", "天", "周", "月", "成本", "预算", "上线")                                           # 20357
QUESTION_RICH_MARKERS = ("原因", "场景", "影响", "频次", "角色", "流程", "瓶颈")                                             # 20355
PACK_VAGUE_TERMS = ("看情况", "不确定", "都可以", "不知道", "可能", "暂时不清楚", "以后再说", "差不多")                                  # 22576
QUESTION_EVIDENCE_MARKERS = ("原因", "为什么", "场景", "影响", "频次", "角色", "流程", "主要是因为", "哪种情况")               # 21441-21443

G_VAGUE, G_INFORMATIVE, G_SEMICOLON = 0, 1, 2
G_CONTRA_L, G_CONTRA_R = 3, 9                     # six bits each
G_SCENARIO, G_GENERIC, G_OPT_PUNCT = 15, 16, 17
G_CAUSE, G_CONTEXT, G_ROLE, G_PROCESS, G_UNITS = 18, 19, 20, 21, 22
G_Q_RICH, G_Q_EVIDENCE, G_PACK_VAGUE = 24, 25, 26

DIMENSION_FOLLOW_UP_SENSITIVITY = {"customer_needs": 0.8, "business_process": 0.6, "tech_constraints": 0.5,
                                   "project_constraints": 0.4}                                                           # 19013
HARD_FOLLOW_UP_SIGNALS = {"vague_expression", "generic_answer", "contradiction_detected"}                                # 19249
QUESTION_ANSWER_MODES = {"pick_only", "pick_with_reason"}                                                                # 19624
QUESTION_EVIDENCE_INTENTS = {"low", "medium", "high"}


def _pattern_table():
    groups: List[Tuple[int, Sequence[str], int]] = [
        (G_VAGUE, VAGUE_INDICATORS, 0), (G_INFORMATIVE, INFORMATIVE_TOKENS, 0), (G_SEMICOLON, ("；",), 0),
        (G_SCENARIO, SCENARIO_KEYWORDS, 0), (G_GENERIC, GENERIC_ANSWERS, 1), (G_OPT_PUNCT, OPTION_PUNCT_TOKENS, 0),
        (G_CAUSE, RICH_OPTION_CAUSE_KEYWORDS, 0), (G_CONTEXT, RICH_OPTION_CONTEXT_KEYWORDS, 0), (G_ROLE, RICH_OPTION_ROLE_KEYWORDS, 0),
        (G_PROCESS, RICH_OPTION_PROCESS_KEYWORDS, 0), (G_UNITS, DIMENSION_UNIT_TOKENS, 0),
        (G_Q_RICH, QUESTION_RICH_MARKERS, 0), (G_Q_EVIDENCE, QUESTION_EVIDENCE_MARKERS, 0), (G_PACK_VAGUE, PACK_VAGUE_TERMS, 0)]
    for i, (left, right) in enumerate(CONTRADICTION_PAIRS):
        groups.append((G_CONTRA_L + i, (left,), 0))
        groups.append((G_CONTRA_R + i, (right,), 0))
    blobs, grp, whole = [], [], []
    for g, pats, w in groups:
        for p in pats:
            blobs.append(p.encode("utf-8")); grp.append(g); whole.append(w)
    off = np.zeros(len(blobs) + 1, np.int32)
    off[1:] = np.cumsum([len(b) for b in blobs])
    return (np.frombuffer(b"".join(blobs) + b"\0" * 16, np.uint8).copy(), off, np.asarray(grp, np.uint8), np.asarray(whole, np.uint8))


_PATTERNS = _pattern_table()


# ----------------------------------------------------------------------------- GPU feature extraction
def text_features(texts: Sequence[str], ctx=None):
    """(mask u64[n], stripped length i32[n], flags u8[n]) for each text; flags bit 0 = has a str.isdigit() char."""
    ctx = ctx or _lib.Context.get()
    n = len(texts)
    mask = np.zeros(max(n, 1), np.uint64); length = np.zeros(max(n, 1), np.int32); flags = np.zeros(max(n, 1), np.uint8)
    if n:
        data, off = pack_texts(texts)
        pb, po, pg, pw = _PATTERNS
        rc = ctx.lib.dvb_text_features(ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), n, ptr(pb), ptr(po), ptr(pg), ptr(pw), len(pg),
                                       ptr(mask), ptr(length), ptr(flags), None, None)
        ctx.check(rc, "dvb_text_features")
    return mask[:n], length[:n], flags[:n]


def contains(haystacks: Sequence[str], needles: Sequence[str], needle_hay: Sequence[int], ctx=None) -> np.ndarray:
    """bits per needle: 1 = needle in hay.strip(); 2 = needle.strip() (non-empty) in hay.strip(); 4 = stripped texts equal."""
    ctx = ctx or _lib.Context.get()
    nn = len(needles)
    out = np.zeros(max(nn, 1), np.uint8)
    if nn:
        hu, ho = pack_texts(haystacks)
        nu, no = pack_texts(needles)
        nh = np.asarray(needle_hay, np.int32)
        rc = ctx.lib.dvb_contains(ctx.handle, _lib.MEM_HOST, ptr(hu), ptr(ho), len(haystacks), ptr(nu), ptr(no), ptr(nh), nn, ptr(out))
        ctx.check(rc, "dvb_contains")
    return out[:nn]


@dataclass
class OptionFeatures:
    is_str: bool
    length: int = 0            # len(str(option or "").strip())
    mask: int = 0
    has_digit: bool = False
    raw_in_answer: bool = False      # option in answer.strip()              (web/server.py:21321)
    strip_in_answer: bool = False    # option.strip() in answer.strip()      (20327)
    equals_answer: bool = False      # option.strip() == answer.strip()      (20326)


@dataclass
class TurnFeatures:
    q_mask: int = 0
    a_mask: int = 0
    a_len: int = 0
    a_digit: bool = False
    options: List[OptionFeatures] = field(default_factory=list)


def _bit(mask: int, g: int) -> bool:
    return bool((mask >> g) & 1)


def extract_turn_features(turns: Sequence[Tuple[str, str, Sequence[Any]]], ctx=None) -> List[TurnFeatures]:
    """turns: (question, answer, options).  Two kernel launches for the whole batch."""
    texts: List[str] = []
    needles: List[str] = []
    needle_hay: List[int] = []
    layout = []
    for ti, (question, answer, options) in enumerate(turns):
        q, a = str(question or ""), str(answer)
        base = len(texts)
        texts.extend((q, a))
        opt_texts = []
        for opt in (options or []):
            is_str = isinstance(opt, str)
            t = opt if is_str else str(opt or "")
            opt_texts.append((is_str, t))
            texts.append(t)
            needles.append(t)
            needle_hay.append(ti)
        layout.append((base, opt_texts))
    mask, length, flags = text_features(texts, ctx=ctx)
    bits = contains([str(t[1]) for t in turns], needles, needle_hay, ctx=ctx) if needles else np.zeros(0, np.uint8)
    out: List[TurnFeatures] = []
    nd = 0
    for (base, opt_texts), (_q, answer, _o) in zip(layout, turns):
        tf = TurnFeatures(q_mask=int(mask[base]), a_mask=int(mask[base + 1]), a_len=int(length[base + 1]),
                          a_digit=bool(flags[base + 1] & 1))
        for j, (is_str, t) in enumerate(opt_texts):
            k = base + 2 + j
            b = int(bits[nd]); nd += 1
            tf.options.append(OptionFeatures(is_str=is_str, length=int(length[k]), mask=int(mask[k]), has_digit=bool(flags[k] & 1),
                                             raw_in_answer=bool(b & 1), strip_in_answer=bool(b & 2), equals_answer=bool(b & 4)))
        out.append(tf)
    return out


# ----------------------------------------------------------------------------- decision logic over the features
def normalize_question_answer_mode(answer_mode: Any, fallback: str = "pick_only") -> str:                # 19629
    normalized = str(answer_mode or "").strip().lower()
    if normalized in QUESTION_ANSWER_MODES:
        return normalized
    fb = str(fallback or "").strip().lower()
    return fb if fb in QUESTION_ANSWER_MODES else "pick_only"


def normalize_question_evidence_intent(evidence_intent: Any, fallback: str = "low") -> str:               # 19639
    normalized = str(evidence_intent or "").strip().lower()
    if normalized in QUESTION_EVIDENCE_INTENTS:
        return normalized
    fb = str(fallback or "").strip().lower()
    return fb if fb in QUESTION_EVIDENCE_INTENTS else "low"


def _rich_option(tf: TurnFeatures, dimension: Any) -> Tuple[bool, bool]:
    """(_score_rich_option_answer(...)["is_rich_option"], ["exact_option"])   web/server.py:20318-20367"""
    cands = [o for o in tf.options if o.length > 0]                 # option_list: stripped, non-empty
    if tf.a_len == 0 or not cands:
        return False, False
    exact = any(o.equals_answer for o in cands)
    matched = [o for o in cands if o.strip_in_answer]
    if not exact and not matched:
        return False, False
    if exact:                                                        # primary option == the stripped answer itself
        p_len, p_mask, p_digit = tf.a_len, tf.a_mask, tf.a_digit
    else:
        best = matched[0]
        for o in matched[1:]:                                        # max(..., key=len): first of the longest
            if o.length > best.length:
                best = o
        p_len, p_mask, p_digit = best.length, best.mask, best.has_digit
    score = 0
    score += 1 if exact else 0
    score += 1 if p_len >= 10 else 0
    score += 1 if p_len >= 16 else 0
    for g in (G_OPT_PUNCT, G_CAUSE, G_CONTEXT, G_ROLE, G_PROCESS):
        score += 1 if _bit(p_mask, g) else 0
    score += 1 if p_digit else 0
    score += 1 if _bit(tf.q_mask, G_Q_RICH) else 0
    if str(dimension or "").strip().lower() in {"tech_constraints", "project_constraints"} and _bit(p_mask, G_UNITS):
        score += 1
    return bool(score >= 4 or (exact and score >= 3 and p_len >= 12)), exact


_FOLLOW_UP_REASONS = {                                                                                   # 21406-21414
    "too_short": "回答过于简短，需要补充具体细节",
    "vague_expression": "回答包含模糊表述，需要明确具体要求",
    "generic_answer": "回答过于笼统，需要深入了解具体需求",
    "option_only": "当前答案方向明确，但仍需补一个关键依据以支撑后续结论",
    "no_quantification": "缺少量化指标，需要明确具体数据要求",
    "single_selection": "只选择了单一选项，需要了解是否还有其他需求",
    "contradiction_detected": "回答中存在前后冲突，需要澄清真实约束",
}
_SIGNAL_WEIGHTS = {"too_short": 0.4, "vague_expression": 0.5, "generic_answer": 0.8, "option_only": 0.3, "no_quantification": 0.2,
                   "single_selection": 0.2, "contradiction_detected": 0.6}
_NEGATIVE_PENALTY = {"too_short": 0.2, "vague_expression": 0.25, "generic_answer": 0.3, "option_only": 0.2, "no_quantification": 0.1,
                     "single_selection": 0.1, "contradiction_detected": 0.25}


def depth_from_features(tf: TurnFeatures, dimension: Any, options: Any, is_follow_up: bool = False, multi_select: bool = False,
                        answer_mode: Any = "", requires_rationale: bool = False, evidence_intent: Any = "",
                        sensitivity_table: Optional[dict] = None, hard_signals: Optional[Iterable[str]] = None) -> dict:
    """evaluate_answer_depth (web/server.py:21221-21401) with every string test replaced by a feature bit."""
    table = DIMENSION_FOLLOW_UP_SENSITIVITY if sensitivity_table is None else sensitivity_table
    hard = HARD_FOLLOW_UP_SIGNALS if hard_signals is None else hard_signals
    signals: List[str] = []
    answer_len = tf.a_len
    sensitivity = table.get(dimension, 0.5)
    mode = normalize_question_answer_mode(answer_mode, fallback="pick_with_reason" if requires_rationale else "pick_only")
    needs_rationale = bool(requires_rationale or mode == "pick_with_reason")
    intent = normalize_question_evidence_intent(evidence_intent, fallback="high" if needs_rationale else "low")
    pick_only = mode == "pick_only" and not needs_rationale
    rich, exact = _rich_option(tf, dimension)

    short_threshold = int(20 + sensitivity * 20)
    if options:
        short_threshold = min(short_threshold, (10 if multi_select else 6) if pick_only else (18 if multi_select else 12))
    if answer_len < short_threshold and not rich:
        signals.append("too_short")
    if _bit(tf.a_mask, G_VAGUE):
        signals.append("vague_expression")
    if _bit(tf.a_mask, G_GENERIC):
        signals.append("generic_answer")
    if options:
        informative = answer_len >= 10 or rich or _bit(tf.a_mask, G_INFORMATIVE)
        if not pick_only and exact and answer_len < 16 and not informative:
            signals.append("option_only")
        if rich:
            signals.append("rich_option_answer")
    has_numbers = tf.a_digit
    if dimension in ("tech_constraints", "project_constraints") and not has_numbers and answer_len < 60:
        signals.append("no_quantification")
    if multi_select and options and not _bit(tf.a_mask, G_SEMICOLON) and len(options) >= 3 and not rich:
        selected = 0
        for o in tf.options:
            if not o.is_str:
                raise TypeError("'in <string>' requires string as left operand")      # what `opt in answer` raises in the reference
            selected += 1 if o.raw_in_answer else 0
        if selected <= 1 and answer_len < 30:
            signals.append("single_selection")
    if any(_bit(tf.a_mask, G_CONTRA_L + i) and _bit(tf.a_mask, G_CONTRA_R + i) for i in range(len(CONTRADICTION_PAIRS))):
        signals.append("contradiction_detected")

    sufficient: List[str] = []
    if answer_len > 80:
        sufficient.append("detailed_answer")
    if _bit(tf.a_mask, G_SEMICOLON) and answer_len > 40:
        sufficient.append("multi_point_answer")
    if has_numbers and answer_len > 30:
        sufficient.append("quantified_answer")
    if rich:
        sufficient.append("rich_option_answer")

    follow_up_score = sum(_SIGNAL_WEIGHTS.get(s, 0.1) for s in signals)
    follow_up_score *= sensitivity
    sufficient_weights = {"detailed_answer": 0.5, "multi_point_answer": 0.3, "quantified_answer": 0.2,
                          "rich_option_answer": 0.25 if intent == "high" else 0.18}
    follow_up_score -= sum(sufficient_weights.get(s, 0) for s in sufficient)
    hard_triggered = any(s in hard for s in signals)

    reasons = [_FOLLOW_UP_REASONS.get(s, "") for s in signals if s in _FOLLOW_UP_REASONS]
    reason = reasons[0] if reasons else "需要进一步了解详细需求"
    if follow_up_score >= 0.4:
        needs, suggest = True, False
    elif follow_up_score >= 0.15 and not sufficient:
        needs, suggest = False, True
    else:
        needs, suggest, reason = False, False, None
    return {"needs_follow_up": needs, "suggest_ai_eval": suggest, "reason": reason, "signals": signals,
            "hard_triggered": hard_triggered, "has_numbers": has_numbers, "follow_up_score": round(follow_up_score, 2)}


def quality_from_features(eval_result: dict, a_len: int, a_mask: int, is_follow_up: Any, follow_up_round: Any) -> dict:
    """evaluate_answer_quality (web/server.py:20370-20426); a_len = len(answer.strip()), a_mask from the answer text."""
    signals = eval_result.get("signals", [])
    hard_triggered = eval_result.get("hard_triggered", False)
    quality_score = 0.5
    if a_len >= 120:
        quality_score += 0.2
    elif a_len >= 80:
        quality_score += 0.1
    if eval_result.get("has_numbers"):
        quality_score += 0.1
    if _bit(a_mask, G_SCENARIO):
        quality_score += 0.1
    if "rich_option_answer" in signals:
        quality_score += 0.12
    quality_score -= sum(_NEGATIVE_PENALTY.get(signal, 0.05) for signal in signals)
    if is_follow_up and follow_up_round >= 1 and not hard_triggered:
        quality_score += 0.05
    quality_score = max(0.0, min(1.0, quality_score))
    quality_signals: List[str] = []
    if quality_score >= 0.75:
        quality_signals.append("high_quality")
    if eval_result.get("has_numbers"):
        quality_signals.append("quantified")
    if a_len >= 80:
        quality_signals.append("detailed")
    if "rich_option_answer" in signals:
        quality_signals.append("rich_option")
    for signal in signals:
        if signal not in quality_signals:
            quality_signals.append(signal)
    return {"quality_score": round(quality_score, 2), "quality_signals": quality_signals, "hard_triggered": bool(hard_triggered)}


def classify_answer_evidence_class_v3(signals: Any, quality_score: Any, evidence_intent: Any) -> str:     # 22552 (no text work)
    normalized = {str(item or "").strip().lower() for item in (signals or []) if str(item or "").strip()}
    intent = normalize_question_evidence_intent(evidence_intent, fallback="low")
    try:
        q = float(quality_score)
    except Exception:
        q = 0.0
    q = max(0.0, min(1.0, q))
    if "rich_option_answer" in normalized:
        return "rich_option"
    if normalized & {"vague_expression", "generic_answer", "contradiction_detected"}:
        return "pending_follow_up" if intent in {"medium", "high"} else "weak_inferred"
    if normalized & {"too_short", "option_only", "no_quantification", "single_selection"}:
        return "pending_follow_up" if intent == "high" else "weak_inferred"
    if q < 0.35 and intent in {"medium", "high"}:
        return "weak_inferred"
    return "explicit"


# ----------------------------------------------------------------------------- reference-shaped single calls
def evaluate_answer_depth(question: str, answer: str, dimension: str, options: list = None, is_follow_up: bool = False,
                          multi_select: bool = False, answer_mode: str = "", requires_rationale: bool = False,
                          evidence_intent: str = "", *, ctx=None, sensitivity_table=None, hard_signals=None) -> dict:
    """Drop-in for web/server.py:21221 (a batch of one turn)."""
    answer.strip                                   # the reference calls answer.strip(): non-str answers raise AttributeError
    tf = extract_turn_features([(question, answer, options)], ctx=ctx)[0]
    return depth_from_features(tf, dimension, options, is_follow_up, multi_select, answer_mode, requires_rationale, evidence_intent,
                               sensitivity_table, hard_signals)


def evaluate_answer_quality(eval_result: dict, answer: str, is_follow_up: bool, follow_up_round: int, *, ctx=None) -> dict:
    """Drop-in for web/server.py:20370."""
    mask, length, _flags = text_features([str(answer or "")], ctx=ctx)
    return quality_from_features(eval_result, int(length[0]), int(mask[0]), is_follow_up, follow_up_round)


def build_question_capture_contract(*, should_follow_up: bool, hard_triggered: bool, missing_aspects: Optional[list] = None,
                                    formal_questions_count: int = 0, dimension: str = "") -> dict:       # 19669 (no text work)
    missing_count = len(list(missing_aspects or []))
    normalized_dimension = str(dimension or "").strip().lower()
    if should_follow_up or hard_triggered or missing_count > 0:
        return {"answer_mode": "pick_with_reason", "requires_rationale": True, "evidence_intent": "high"}
    if formal_questions_count >= 2 or normalized_dimension in {"tech_constraints", "project_constraints"}:
        return {"answer_mode": "pick_with_reason", "requires_rationale": True, "evidence_intent": "medium"}
    return {"answer_mode": "pick_only", "requires_rationale": False, "evidence_intent": "low"}


def _contract_from_features(log: dict, prior_dim_logs: list, tf: TurnFeatures) -> dict:
    """_infer_legacy_log_capture_contract (web/server.py:21419-21466) over the feature bits."""
    dimension = str(log.get("dimension", "") or "").strip()
    is_follow_up = bool(log.get("is_follow_up", False) or int(log.get("follow_up_round", 0) or 0) > 0)
    hard_triggered = bool(log.get("hard_triggered", False))
    formal = len([item for item in (prior_dim_logs or []) if not item.get("is_follow_up", False)])
    contract = build_question_capture_contract(should_follow_up=is_follow_up, hard_triggered=hard_triggered, missing_aspects=[],
                                               formal_questions_count=formal, dimension=dimension)
    asks_for_evidence = _bit(tf.q_mask, G_Q_EVIDENCE)
    marker_bits = (1 << G_CAUSE) | (1 << G_CONTEXT) | (1 << G_ROLE) | (1 << G_PROCESS)
    options_are_rich = any(o.length >= 10 and (o.mask & marker_bits) for o in tf.options)
    if asks_for_evidence or options_are_rich:
        contract["answer_mode"] = "pick_with_reason"
        contract["requires_rationale"] = True
        current = normalize_question_evidence_intent(contract.get("evidence_intent", ""), fallback="low")
        if is_follow_up or asks_for_evidence:
            contract["evidence_intent"] = "high"
        elif current == "low":
            contract["evidence_intent"] = "medium"
    return {"answer_mode": normalize_question_answer_mode(contract.get("answer_mode", ""), fallback="pick_only"),
            "requires_rationale": bool(contract.get("requires_rationale", False)),
            "evidence_intent": normalize_question_evidence_intent(contract.get("evidence_intent", ""), fallback="low")}


# ----------------------------------------------------------------------------- a12: backfill (bulk re-index body)
_SNAPSHOT_KEYS = ("answer_mode", "requires_rationale", "evidence_intent", "follow_up_signals", "quality_score", "quality_signals",
                  "hard_triggered", "needs_follow_up", "answer_evidence_class")
_ZERO = {"changed": False, "logs_total": 0, "logs_updated": 0, "field_updates": {}}


def _log_texts(log: dict) -> Tuple[str, str, list]:
    options = log.get("options", []) if isinstance(log.get("options", []), list) else []
    return str(log.get("question", "") or ""), str(log.get("answer", "") or ""), options


SIGNAL_NAMES = ("too_short", "vague_expression", "generic_answer", "option_only", "rich_option_answer", "no_quantification",
                "single_selection", "contradiction_detected")          # ids of dvb_annotate_turns = the order evaluate_answer_depth appends
_MODE_NAMES = ("", "pick_only", "pick_with_reason")
_INTENT_NAMES = ("", "low", "medium", "high")
_CLASS_NAMES = ("explicit", "rich_option", "weak_inferred", "pending_follow_up")
_MODE_IDS = {"pick_only": 1, "pick_with_reason": 2}
_INTENT_IDS = {"low": 1, "medium": 2, "high": 3}
_SIGNAL_LISTS = [[SIGNAL_NAMES[i] for i in range(8) if (m >> i) & 1] for m in range(256)]
_TC_PC = ("tech_constraints", "project_constraints")


_DECODED: Dict[int, tuple] = {}


def _decode_turn_bits(b: int) -> tuple:
    """out_bits of dvb_annotate_turns -> (answer_mode, requires_rationale, evidence_intent, class, signals, needs_follow_up,
    quality_signals, hard_triggered); cached per distinct value (a few hundred occur)."""
    sig = _SIGNAL_LISTS[b & 255]
    qsig = []
    if b & (1 << 21):
        qsig.append("high_quality")
    if b & 2048:
        qsig.append("quantified")
    if b & (1 << 22):
        qsig.append("detailed")
    if b & 16:
        qsig.append("rich_option")
    qsig.extend(sig)
    return (_MODE_NAMES[(b >> 13) & 3], bool(b & 4096), _INTENT_NAMES[(b >> 15) & 3], _CLASS_NAMES[(b >> 17) & 7], sig, bool(b & 256), qsig, bool(b & 1024))


def annotate_turns(q_mask, a_mask, a_len, a_flags, opt_off, opt_mask, opt_len, opt_flags, opt_bits, turn_flags, follow_up_round,
                   formal_prior, sensitivity, hard_mask: int, overwrite_contract: bool, ctx=None):
    """dvb_annotate_turns on host arrays: (bits u32[n], follow_up_score f64[n], quality_score f64[n]) — the whole decision
    stage of a11/a12 for a batch of turns in one launch (csrc/annotate.cu, k_annotate_turns)."""
    ctx = ctx or _lib.Context.get()
    n = len(turn_flags)
    bits = np.zeros(max(n, 1), np.uint32); fs = np.zeros(max(n, 1), np.float64); qs = np.zeros(max(n, 1), np.float64)
    if n:
        rc = ctx.lib.dvb_annotate_turns(ctx.handle, _lib.MEM_HOST, n, ptr(q_mask), ptr(a_mask), ptr(a_len), ptr(a_flags), ptr(opt_off),
                                        ptr(opt_mask), ptr(opt_len), ptr(opt_flags), ptr(opt_bits), ptr(turn_flags), ptr(follow_up_round),
                                        ptr(formal_prior), ptr(sensitivity), int(hard_mask), 1 if overwrite_contract else 0,
                                        ptr(bits), ptr(fs), ptr(qs))
        ctx.check(rc, "dvb_annotate_turns")
    return bits[:n], fs[:n], qs[:n]


def backfill_sessions(sessions: Sequence[Any], *, refresh_quality: bool = True, overwrite_contract: bool = False, ctx=None,
                      sensitivity_table=None, hard_signals=None) -> List[dict]:
    """Batch form of backfill_session_interview_log_evidence_annotations (web/server.py:21469-21584): three kernel launches
    for every turn of every session — text features, option containment, and the decision stage (contract inference,
    depth, quality, evidence class; dvb_annotate_turns).  Python only extracts the fields of each log and writes the nine
    annotation fields back (each session dict is mutated in place, like the reference) with the reference's change counts."""
    table = DIMENSION_FOLLOW_UP_SENSITIVITY if sensitivity_table is None else sensitivity_table
    hard = HARD_FOLLOW_UP_SIGNALS if hard_signals is None else hard_signals
    hard_mask = sum(1 << i for i, name in enumerate(SIGNAL_NAMES) if name in hard)
    texts: List[str] = []                    # question, answer of every turn, then every option text
    opt_texts: List[str] = []
    opt_str: List[int] = []
    needle_hay: List[int] = []
    opt_off: List[int] = [0]
    tflags: List[int] = []
    fround: List[int] = []
    formal: List[int] = []
    sens: List[float] = []
    spans: List[Optional[Tuple[int, int]]] = []
    turn_logs: List[dict] = []
    for session in sessions:
        logs = session.get("interview_log", []) if isinstance(session, dict) else None
        if not isinstance(logs, list):
            spans.append(None)
            continue
        start = len(turn_logs)
        prior_formal: Dict[str, int] = {}
        for log in logs:
            if not isinstance(log, dict):
                continue
            get = log.get
            ti = len(turn_logs)
            turn_logs.append(log)
            texts.append(str(get("question", "") or ""))
            texts.append(str(get("answer", "") or ""))
            options = get("options", [])
            if not isinstance(options, list):
                options = []
            for opt in options:
                is_str = isinstance(opt, str)
                opt_texts.append(opt if is_str else str(opt or ""))
                opt_str.append(2 if is_str else 0)
                needle_hay.append(ti)
            opt_off.append(len(opt_texts))
            dimension = str(get("dimension", "") or "").strip()
            is_fu = bool(get("is_follow_up", False))
            rnd = int(get("follow_up_round", 0) or 0)
            raw_mode = str(get("answer_mode", "") or "").strip().lower()
            raw_intent = str(get("evidence_intent", "") or "").strip().lower()
            f = ((1 if options else 0) | (2 if get("question_multi_select", get("multi_select", False)) else 0) | (4 if is_fu else 0)
                 | (8 if get("hard_triggered", False) else 0) | (16 if get("requires_rationale", False) else 0)
                 | (32 if dimension in _TC_PC else 0) | (64 if dimension.lower() in _TC_PC else 0)
                 | (_MODE_IDS.get(raw_mode, 0) << 8) | (_INTENT_IDS.get(raw_intent, 0) << 10) | (8192 if (is_fu or rnd > 0) else 0))
            tflags.append(f)
            fround.append(rnd)
            formal.append(prior_formal.get(dimension, 0))
            sens.append(table.get(dimension, 0.5))
            if not is_fu:
                prior_formal[dimension] = prior_formal.get(dimension, 0) + 1
        spans.append((start, len(turn_logs)))
    n = len(turn_logs)
    if n:
        mask, length, flags = text_features(texts + opt_texts, ctx=ctx)
        cbits = contains(texts[1::2], opt_texts, needle_hay, ctx=ctx) if opt_texts else np.zeros(0, np.uint8)
        n2 = 2 * n
        no = len(opt_texts)
        bits, fscore, qscore = annotate_turns(
            np.ascontiguousarray(mask[0:n2:2]), np.ascontiguousarray(mask[1:n2:2]), np.ascontiguousarray(length[1:n2:2]), np.ascontiguousarray(flags[1:n2:2]),
            np.asarray(opt_off, np.int32), np.ascontiguousarray(mask[n2:]) if no else np.zeros(1, np.uint64),
            np.ascontiguousarray(length[n2:]) if no else np.zeros(1, np.int32),
            ((flags[n2:] & 1) | np.asarray(opt_str, np.uint8)) if no else np.zeros(1, np.uint8),
            np.ascontiguousarray(cbits) if no else np.zeros(1, np.uint8),
            np.asarray(tflags, np.uint32), np.asarray(fround, np.int32), np.asarray(formal, np.int32), np.asarray(sens, np.float64),
            hard_mask, overwrite_contract, ctx=ctx)
        bits_l, q_l = bits.tolist(), qscore.tolist()
    out = []
    for session, span in zip(sessions, spans):
        if span is None:
            out.append(dict(_ZERO, field_updates={}))
            continue
        logs_updated, field_updates = 0, {}
        for ti in range(span[0], span[1]):
            log = turn_logs[ti]
            get = log.get
            b = bits_l[ti]
            if b & (1 << 20):
                raise TypeError("'in <string>' requires string as left operand")      # what `opt in answer` raises in the reference
            dec = _DECODED.get(b)
            if dec is None:
                dec = _DECODED[b] = _decode_turn_bits(b)
            mode, rr, intent, cls, sig, needs, qsig, hard_t = dec
            touched = False
            # the nine fields in the reference's snapshot order (web/server.py:21549-21559); lists are fresh objects per log
            if get("answer_mode") != mode:
                field_updates["answer_mode"] = field_updates.get("answer_mode", 0) + 1; touched = True
            log["answer_mode"] = mode
            if get("requires_rationale") != rr:
                field_updates["requires_rationale"] = field_updates.get("requires_rationale", 0) + 1; touched = True
            log["requires_rationale"] = rr
            if get("evidence_intent") != intent:
                field_updates["evidence_intent"] = field_updates.get("evidence_intent", 0) + 1; touched = True
            log["evidence_intent"] = intent
            fresh_sig = refresh_quality or "follow_up_signals" not in log
            fresh_q = refresh_quality or "quality_score" not in log
            if fresh_sig:
                if get("follow_up_signals") != sig:
                    field_updates["follow_up_signals"] = field_updates.get("follow_up_signals", 0) + 1; touched = True
                log["follow_up_signals"] = list(sig)
            if fresh_q:
                q = q_l[ti]
                if get("quality_score") != q:
                    field_updates["quality_score"] = field_updates.get("quality_score", 0) + 1; touched = True
                log["quality_score"] = q
                if get("quality_signals") != qsig:
                    field_updates["quality_signals"] = field_updates.get("quality_signals", 0) + 1; touched = True
                log["quality_signals"] = list(qsig)
                if get("hard_triggered") != hard_t:
                    field_updates["hard_triggered"] = field_updates.get("hard_triggered", 0) + 1; touched = True
                log["hard_triggered"] = hard_t
            if fresh_sig:
                if get("needs_follow_up") != needs:
                    field_updates["needs_follow_up"] = field_updates.get("needs_follow_up", 0) + 1; touched = True
                log["needs_follow_up"] = needs
            if get("answer_evidence_class") != cls:
                field_updates["answer_evidence_class"] = field_updates.get("answer_evidence_class", 0) + 1; touched = True
            log["answer_evidence_class"] = cls
            if touched:
                logs_updated += 1
        out.append({"changed": logs_updated > 0, "logs_total": span[1] - span[0], "logs_updated": logs_updated, "field_updates": field_updates})
    return out


def backfill_session_interview_log_evidence_annotations(session: dict, *, refresh_quality: bool = True, overwrite_contract: bool = False,
                                                        ctx=None) -> dict:
    """Drop-in for web/server.py:21469 (mutates session["interview_log"][i] in place, returns the summary)."""
    return backfill_sessions([session], refresh_quality=refresh_quality, overwrite_contract=overwrite_contract, ctx=ctx)[0]


# ----------------------------------------------------------------------------- a14: the bulk re-index script body
def get_utc_now() -> str:
    from datetime import datetime, timezone
    return datetime.now(timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")


def backfill_session_files(session_files: Iterable[Path], *, apply_changes: bool = False, backup_dir: Optional[Path] = None,
                           server_module=None, batch_sessions: int = 2048, ctx=None, shard: Optional[Tuple[int, int]] = None) -> dict:
    """scripts/migrate_session_evidence_annotations.py:64-114 with the per-file calls batched (`batch_sessions` files per
    pair of kernel launches).  `server_module` only supplies get_utc_now() for the write-back stamp.  `shard=(rank, world)`
    processes a contiguous block of the file list (multi-GPU: sessions are independent; summaries add up)."""
    files = [Path(f) for f in session_files]
    if shard is not None:
        from .parallel import shard_range
        lo, hi = shard_range(len(files), shard[0], shard[1])
        files = files[lo:hi]
    summary = {"sessions_total": 0, "sessions_changed": 0, "logs_total": 0, "logs_updated": 0, "field_updates": {}, "changed_sessions": []}
    now = getattr(server_module, "get_utc_now", None) or get_utc_now
    for b0 in range(0, len(files), max(1, int(batch_sessions))):
        batch = files[b0:b0 + max(1, int(batch_sessions))]
        datas = [json.loads(f.read_text(encoding="utf-8")) for f in batch]
        results = backfill_sessions(datas, refresh_quality=True, overwrite_contract=False, ctx=ctx)
        for session_file, session_data, result in zip(batch, datas, results):
            summary["sessions_total"] += 1
            summary["logs_total"] += int(result.get("logs_total", 0) or 0)
            summary["logs_updated"] += int(result.get("logs_updated", 0) or 0)
            for name, count in (result.get("field_updates", {}) or {}).items():
                summary["field_updates"][name] = summary["field_updates"].get(name, 0) + int(count or 0)
            if result.get("changed"):
                summary["sessions_changed"] += 1
                summary["changed_sessions"].append({
                    "session_id": session_data.get("session_id", session_file.stem) if isinstance(session_data, dict) else session_file.stem,
                    "topic": session_data.get("topic", ""), "logs_updated": int(result.get("logs_updated", 0) or 0),
                    "field_updates": dict(result.get("field_updates", {}) or {})})
                if apply_changes:
                    if backup_dir is not None:
                        shutil.copy2(session_file, Path(backup_dir) / session_file.name)
                    session_data["updated_at"] = now()
                    session_file.write_text(json.dumps(session_data, ensure_ascii=False, indent=2), encoding="utf-8")
    return summary


def merge_summaries(parts: Sequence[dict]) -> dict:
    """Sum per-rank summaries of backfill_session_files (counters add; changed_sessions concatenate in rank order)."""
    out = {"sessions_total": 0, "sessions_changed": 0, "logs_total": 0, "logs_updated": 0, "field_updates": {}, "changed_sessions": []}
    for p in parts:
        for k in ("sessions_total", "sessions_changed", "logs_total", "logs_updated"):
            out[k] += int(p.get(k, 0))
        for name, count in p.get("field_updates", {}).items():
            out["field_updates"][name] = out["field_updates"].get(name, 0) + int(count)
        out["changed_sessions"].extend(p.get("changed_sessions", []))
    return out


# ----------------------------------------------------------------------------- dimension saturation
# Related words per key aspect (web/server.py:19442-19459) and the depth cue words (19471-19485), in the reference's order.
ASPECT_KEYWORDS = {
    "核心痛点": ["痛点", "问题", "困难", "挑战", "困扰"], "期望价值": ["价值", "收益", "效果", "目标", "期望"],
    "使用场景": ["场景", "情况", "使用", "应用", "何时"], "用户角色": ["用户", "角色", "人员", "谁", "使用者"],
    "关键流程": ["流程", "步骤", "环节", "过程"], "角色分工": ["分工", "职责", "负责", "部门"],
    "触发事件": ["触发", "开始", "启动", "何时"], "异常处理": ["异常", "错误", "失败", "例外"],
    "部署方式": ["部署", "云", "本地", "服务器"], "系统集成": ["集成", "对接", "接口", "系统"],
    "性能要求": ["性能", "响应", "并发", "速度"], "安全合规": ["安全", "合规", "权限", "加密"],
    "预算范围": ["预算", "费用", "成本", "价格"], "时间节点": ["时间", "期限", "周期", "何时"],
    "资源限制": ["资源", "人力", "团队", "限制"], "优先级": ["优先", "重要", "紧急", "先后"],
}
_DEPTH_CUES = (("比如", "例如", "当", "如果", "场景", "情况下"), ("而不是", "优先", "相比", "更重要", "首先"),
               ("因为", "由于", "所以", "原因是"), ("；", "、"))
_ZERO_SATURATION = {"saturation_score": 0, "coverage_score": 0, "depth_score": 0, "volume_score": 0, "covered_aspects": [], "level": "low"}


def dimension_saturation_batch(items: Sequence[tuple], *, server_module, ctx=None) -> List[dict]:
    """calculate_dimension_saturation (web/server.py:19399-19515) for many (session, dimension) pairs.  Every
    `aspect in text` / `keyword in text` test of all pairs is ONE dvb_contains launch and every `any(c.isdigit())` ONE
    dvb_text_features launch; the weighted sums, rounding and thresholds stay in Python in the reference's order.
    The server module supplies get_dimension_info_for_session and get_mode_saturation_thresholds (app configuration)."""
    mod = server_module
    out: List[Optional[dict]] = [None] * len(items)
    hay: List[str] = []                   # per live pair: combined text (answers + questions), then the answers alone
    needles: List[str] = []
    needle_hay: List[int] = []
    host_bits: Dict[int, bool] = {}       # needles with white space at an end: dvb_contains searches the STRIPPED haystack
    live = []                             # (item index, key_aspects, first needle, total answer characters, session)

    def need(text: str, h: int) -> int:
        if not isinstance(text, str):
            raise TypeError("'in <string>' requires string as left operand, not %s" % type(text).__name__)
        j = len(needles)
        needles.append(text)
        needle_hay.append(h)
        if text != text.strip():
            host_bits[j] = text in hay[h]
        return j

    for i, (session, dimension) in enumerate(items):
        dim_logs = [log for log in session.get("interview_log", []) if log.get("dimension") == dimension]
        if not dim_logs:
            out[i] = copy.deepcopy(_ZERO_SATURATION)
            continue
        key_aspects = mod.get_dimension_info_for_session(session).get(dimension, {}).get("key_aspects", [])
        all_answers = " ".join([log.get("answer", "") for log in dim_logs])
        all_questions = " ".join([log.get("question", "") for log in dim_logs])
        h = len(hay)
        hay.append(all_answers + all_questions)
        hay.append(all_answers)
        first = len(needles)
        for aspect in key_aspects:
            need(aspect, h)
            for kw in ASPECT_KEYWORDS.get(aspect, []):
                need(kw, h)
        for group in _DEPTH_CUES:
            for kw in group:
                need(kw, h + 1)
        live.append((i, key_aspects, first, sum(len(log.get("answer", "")) for log in dim_logs), session))
    if not live:
        return out
    bits = contains(hay, needles, needle_hay, ctx=ctx)
    digit = text_features(hay[1::2], ctx=ctx)[2]

    def found(j: int) -> bool:
        return host_bits[j] if j in host_bits else bool(bits[j] & 1)

    for n, (i, key_aspects, j, total_chars, session) in enumerate(live):
        covered = []
        for aspect in key_aspects:
            hit = found(j)
            j += 1
            for _kw in ASPECT_KEYWORDS.get(aspect, []):
                hit = hit or found(j)
                j += 1
            if hit:
                covered.append(aspect)
        coverage_score = len(covered) / len(key_aspects) if key_aspects else 0
        depth_signals = 1 if (digit[n] & 1) else 0
        for group in _DEPTH_CUES:
            if any([found(j + g) for g in range(len(group))]):
                depth_signals += 1
            j += len(group)
        depth_score = min(1.0, depth_signals / 5)
        volume_score = min(1.0, total_chars / 300)
        saturation_score = coverage_score * 0.4 + depth_score * 0.3 + volume_score * 0.3
        thresholds = mod.get_mode_saturation_thresholds(session)
        level = "high" if saturation_score >= thresholds["high"] else ("medium" if saturation_score >= thresholds["medium"] else "low")
        out[i] = {"saturation_score": round(saturation_score, 2), "coverage_score": round(coverage_score, 2),
                  "depth_score": round(depth_score, 2), "volume_score": round(volume_score, 2),
                  "covered_aspects": covered, "level": level}
    return out


def calculate_dimension_saturation(session: dict, dimension: str, *, server_module, ctx=None) -> dict:
    """web/server.py:19399-19515 (single pair = a batch of one)."""
    return dimension_saturation_batch([(session, dimension)], server_module=server_module, ctx=ctx)[0]


def missing_aspects_batch(items: Sequence[tuple], *, server_module, ctx=None) -> List[list]:
    """get_dimension_missing_aspects (web/server.py:19583-19589) for many (session, dimension) pairs."""
    sats = dimension_saturation_batch(items, server_module=server_module, ctx=ctx)
    res = []
    for (session, dimension), sat in zip(items, sats):
        covered = set(sat.get("covered_aspects", []))
        key_aspects = server_module.get_dimension_info_for_session(session).get(dimension, {}).get("key_aspects", [])
        res.append([aspect for aspect in key_aspects if aspect and aspect not in covered])
    return res


def get_dimension_missing_aspects(session: dict, dimension: str, *, server_module, ctx=None) -> list:
    return missing_aspects_batch([(session, dimension)], server_module=server_module, ctx=ctx)[0]


# ----------------------------------------------------------------------------- a13: evidence pack
# (pair_id, index into CONTRADICTION_PAIRS, description)   web/server.py:22579-22585
_PACK_CONTRADICTIONS = (("need", 0, "需求取向冲突"), ("support", 2, "支持立场冲突"), ("available", 4, "资源现状冲突"),
                        ("must", 5, "约束优先级冲突"), ("ready", 3, "准备状态冲突"))


def build_report_evidence_pack(session: dict, *, server_module, ctx=None, device_coverage: bool = False) -> dict:
    """Drop-in for web/server.py:22568-22803.  Annotation (a12) and every `term in answer` test run on the GPU.  The
    per-dimension coverage block reads the app configuration through the server module (get_dimension_info_for_session,
    get_interview_mode_config, resolve_report_template_for_session); the missing aspects of ALL dimensions come from one
    batched device pass (missing_aspects_batch) when device_coverage is set, else from the module's
    get_dimension_missing_aspects (what a test that stubs that name expects).  The caller's session is not mutated
    (deep copy, 22570)."""
    from datetime import datetime, timezone
    mod = server_module
    work = copy.deepcopy(session) if isinstance(session, dict) else {}
    st, hs = getattr(mod, "DIMENSION_FOLLOW_UP_SENSITIVITY", None), getattr(mod, "HARD_FOLLOW_UP_SIGNALS", None)
    backfill_sessions([work], refresh_quality=True, overwrite_contract=False, ctx=ctx, sensitivity_table=st, hard_signals=hs)
    interview_log = work.get("interview_log", [])
    dim_info = mod.get_dimension_info_for_session(work)
    mode_config = mod.get_interview_mode_config(work)
    answers = [str(log.get("answer", "")) for log in interview_log]
    a_masks = [int(m) for m in text_features(answers, ctx=ctx)[0]] if answers else []

    facts, unknowns, contradictions, blindspots = [], [], [], []
    state, seen_contra, seen_unknown = {}, set(), set()
    mode_dist = {m: 0 for m in QUESTION_ANSWER_MODES}
    intent_dist = {i: 0 for i in QUESTION_EVIDENCE_INTENTS}
    class_dist = {"explicit": 0, "rich_option": 0, "weak_inferred": 0, "pending_follow_up": 0}
    for idx, log in enumerate(interview_log, 1):
        q_id = "Q%d" % idx
        dim_key = log.get("dimension", "")
        dim_name = dim_info.get(dim_key, {}).get("name", dim_key or "未分类")
        answer = answers[idx - 1].strip()
        a_mask = a_masks[idx - 1]
        signals = log.get("follow_up_signals") if isinstance(log.get("follow_up_signals"), list) else []
        try:
            quality = float(log.get("quality_score", 0) or 0)
        except Exception:
            quality = 0.0
        quality = max(0.0, min(1.0, quality))
        mode = normalize_question_answer_mode(log.get("answer_mode", ""), fallback="pick_only")
        needs_why = bool(log.get("requires_rationale", False) or mode == "pick_with_reason")
        intent = normalize_question_evidence_intent(log.get("evidence_intent", ""),
                                                    fallback="high" if needs_why else ("medium" if mode == "pick_with_reason" else "low"))
        mode_dist[mode] = mode_dist.get(mode, 0) + 1
        intent_dist[intent] = intent_dist.get(intent, 0) + 1
        klass = classify_answer_evidence_class_v3(signals, quality, intent)
        class_dist[klass] = class_dist.get(klass, 0) + 1
        facts.append({
            "q_id": q_id, "dimension": dim_key, "dimension_name": dim_name, "question": str(log.get("question", "")).strip(),
            "answer": answer, "is_follow_up": bool(log.get("is_follow_up", False)), "follow_up_round": int(log.get("follow_up_round", 0) or 0),
            "quality_score": quality,
            "quality_signals": log.get("quality_signals", []) if isinstance(log.get("quality_signals"), list) else [],
            "follow_up_signals": signals, "hard_triggered": bool(log.get("hard_triggered", False)), "answer_mode": mode,
            "requires_rationale": needs_why, "evidence_intent": intent, "answer_evidence_class": klass})

        reasons = []
        if klass != "rich_option" and any(sig in ("vague_expression", "generic_answer") for sig in signals):
            reasons.append("命中模糊回答信号")
        if _bit(a_mask, G_PACK_VAGUE):
            reasons.append("回答存在模糊表述")
        if klass != "rich_option" and 0 < quality < 0.45:
            reasons.append("回答质量偏低")
        if reasons:
            ordered = sorted(set(reasons))
            ukey = q_id + ":" + "|".join(ordered)
            if ukey not in seen_unknown:
                seen_unknown.add(ukey)
                unknowns.append({"q_id": q_id, "dimension": dim_name, "reason": "；".join(ordered), "answer_excerpt": answer[:120]})

        for pair_id, pi, description in _PACK_CONTRADICTIONS:
            positive, negative = CONTRADICTION_PAIRS[pi]
            has_pos, has_neg = _bit(a_mask, G_CONTRA_L + pi), _bit(a_mask, G_CONTRA_R + pi)
            if has_pos and has_neg:
                key = "self:%s:%s" % (pair_id, q_id)
                if key not in seen_contra:
                    seen_contra.add(key)
                    contradictions.append({"type": "same_answer_conflict", "pair_id": pair_id, "description": description, "dimension": dim_name,
                                           "evidence_refs": [q_id], "detail": "%s 同时出现「%s」与「%s」" % (q_id, positive, negative)})
            if not (has_pos or has_neg):
                continue
            now_state = "positive" if has_pos else "negative"
            prev = state.get((dim_key, pair_id))
            if prev and prev.get("state") != now_state:
                key = "cross:%s:%s:%s:%s" % (pair_id, prev.get("q_id"), q_id, dim_key)
                if key not in seen_contra:
                    seen_contra.add(key)
                    contradictions.append({"type": "cross_answer_conflict", "pair_id": pair_id, "description": description, "dimension": dim_name,
                                           "evidence_refs": [prev.get("q_id"), q_id],
                                           "detail": "%s 与 %s 对「%s/%s」存在冲突" % (prev.get("q_id"), q_id, positive, negative)})
            state[(dim_key, pair_id)] = {"state": now_state, "q_id": q_id}

    def _q(value):
        try:
            return max(0.0, min(1.0, float(value)))
        except Exception:
            return 0.0

    dimension_coverage, adjusted_values, count_values = {}, [], []
    missing_by_dim = {}
    if device_coverage and dim_info:
        keys = list(dim_info)
        missing_by_dim = dict(zip(keys, missing_aspects_batch([(session, k) for k in keys], server_module=mod, ctx=ctx)))
    for dim_key, info in dim_info.items():
        dim_logs = [log for log in interview_log if log.get("dimension") == dim_key]
        dim_facts = [f for f in facts if f.get("dimension") == dim_key]
        count_percent = int(session.get("dimensions", {}).get(dim_key, {}).get("coverage", 0) or 0)
        count_ratio = max(0, min(100, count_percent)) / 100.0
        missing = missing_by_dim[dim_key] if device_coverage else mod.get_dimension_missing_aspects(session, dim_key)
        key_aspects = info.get("key_aspects", []) if isinstance(info.get("key_aspects", []), list) else []
        if key_aspects:
            total = len(key_aspects)
            ratio = max(0.0, min(1.0, (total - min(total, len(missing))) / total))
            percent = int(round(ratio * 100))
        else:
            percent, ratio = count_percent, count_ratio
        scores = [_q(f.get("quality_score", 0.0)) for f in dim_facts]
        avg = (sum(scores) / len(scores)) if scores else 0.0
        adjusted = max(0.0, min(1.0, ratio * (0.35 + 0.65 * avg)))
        adjusted_values.append(adjusted)
        count_values.append(count_ratio)
        blindspots.extend({"dimension": info.get("name", dim_key), "aspect": aspect} for aspect in missing)
        dimension_coverage[dim_key] = {
            "name": info.get("name", dim_key), "coverage_percent": max(0, min(100, percent)), "coverage_ratio": round(ratio, 2),
            "question_count_coverage_percent": max(0, min(100, count_percent)), "question_count_coverage_ratio": round(count_ratio, 2),
            "quality_adjusted_coverage": round(adjusted, 3), "raw_average_quality_score": round(avg, 3),
            "formal_count": len([log for log in dim_logs if not log.get("is_follow_up", False)]),
            "follow_up_count": len([log for log in dim_logs if log.get("is_follow_up", False)]),
            "minimum_formal": mode_config.get("formal_questions_per_dim", 3),
            "maximum_formal": mode_config.get("max_formal_questions_per_dim", mode_config.get("formal_questions_per_dim", 3)),
            "missing_aspects": missing, "key_aspects": key_aspects}

    raw = [_q(f.get("quality_score", 0.0)) for f in facts]
    positive_only = [x for x in raw if x > 0]
    average = sum(raw) / len(raw) if raw else 0.0
    n_follow = len([f for f in facts if f.get("is_follow_up", False)])
    template = mod.resolve_report_template_for_session(session)
    report_cfg = (session.get("scenario_config", {}) or {}).get("report", {})
    if not isinstance(report_cfg, dict):
        report_cfg = {}
    schema = {}
    if template == getattr(mod, "REPORT_TEMPLATE_CUSTOM_V1", object()):
        schema, _issues = mod.normalize_custom_report_schema(report_cfg.get("schema"), fallback_sections=report_cfg.get("sections"))
    return {
        "topic": session.get("topic", "未知主题"),
        "report_type": session.get("scenario_config", {}).get("report", {}).get("type", "standard"),
        "report_template": template, "report_schema": schema, "generated_at": datetime.now(timezone.utc).isoformat(),
        "facts": facts, "contradictions": contradictions, "unknowns": unknowns, "blindspots": blindspots,
        "dimension_coverage": dimension_coverage,
        "overall_coverage": round(sum(adjusted_values) / len(adjusted_values) if adjusted_values else 0.0, 3),
        "question_count_overall_coverage": round(sum(count_values) / len(count_values) if count_values else 0.0, 3),
        "quality_snapshot": {
            "average_quality_score": round(average, 3), "raw_average_quality_score": round(average, 3),
            "positive_only_average_quality_score": round(sum(positive_only) / len(positive_only) if positive_only else 0.0, 3),
            "hard_triggered_count": len([f for f in facts if f.get("hard_triggered")]), "total_questions": len(facts),
            "total_formal_questions": len(facts) - n_follow, "total_follow_up_questions": n_follow,
            "follow_up_ratio": round(n_follow / len(facts), 3) if facts else 0.0,
            "answer_mode_distribution": mode_dist, "evidence_intent_distribution": intent_dist,
            "answer_evidence_class_distribution": class_dist}}


def build_bindings(mod, ctx=None) -> Dict[str, Any]:
    """Replacement callables for dropin.install(); DIMENSION_FOLLOW_UP_SENSITIVITY / HARD_FOLLOW_UP_SIGNALS are read from
    the server module at call time."""
    def _tables():
        return getattr(mod, "DIMENSION_FOLLOW_UP_SENSITIVITY", None), getattr(mod, "HARD_FOLLOW_UP_SIGNALS", None)

    def _depth(question, answer, dimension, options=None, is_follow_up=False, multi_select=False, answer_mode="",
               requires_rationale=False, evidence_intent=""):
        st, hs = _tables()
        return evaluate_answer_depth(question, answer, dimension, options, is_follow_up, multi_select, answer_mode, requires_rationale,
                                     evidence_intent, ctx=ctx, sensitivity_table=st, hard_signals=hs)

    def _quality(eval_result, answer, is_follow_up, follow_up_round):
        return evaluate_answer_quality(eval_result, answer, is_follow_up, follow_up_round, ctx=ctx)

    def _backfill(session, *, refresh_quality=True, overwrite_contract=False):
        st, hs = _tables()
        return backfill_sessions([session], refresh_quality=refresh_quality, overwrite_contract=overwrite_contract, ctx=ctx,
                                 sensitivity_table=st, hard_signals=hs)[0]

    def _saturation(session, dimension):
        return calculate_dimension_saturation(session, dimension, server_module=mod, ctx=ctx)

    def _missing(session, dimension):
        return get_dimension_missing_aspects(session, dimension, server_module=mod, ctx=ctx)

    # the pack takes the batched device pass for its coverage block unless somebody re-bound the two names after install
    # (the reference's own tests stub get_dimension_missing_aspects: tests/test_security_regression.py:4215-4219)
    untouched = {"calculate_dimension_saturation": (getattr(mod, "calculate_dimension_saturation", None), _saturation),
                 "get_dimension_missing_aspects": (getattr(mod, "get_dimension_missing_aspects", None), _missing)}

    def _pack(session):
        dev = all(fn is not None and getattr(mod, name, None) in (fn, ours) for name, (fn, ours) in untouched.items()) \
            and hasattr(mod, "get_mode_saturation_thresholds")
        return build_report_evidence_pack(session, server_module=mod, ctx=ctx, device_coverage=dev)

    return {"calculate_dimension_saturation": _saturation, "get_dimension_missing_aspects": _missing,
            "evaluate_answer_depth": _depth, "evaluate_answer_quality": _quality, "build_report_evidence_pack": _pack,
            "classify_answer_evidence_class_v3": classify_answer_evidence_class_v3,
            "backfill_session_interview_log_evidence_annotations": _backfill}
