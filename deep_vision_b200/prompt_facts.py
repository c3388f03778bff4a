"""Host mirror of the reference's evidence slimming for the report prompt.

`select_slimmed_facts_for_prompt` (reference web/server.py:22886-22993) keeps its signature and return value; the scoring,
the (-score, index) order, the three greedy passes and the final sort run in ONE `dvb_slim_facts` launch for any number of
evidence packs (`select_slimmed_facts_batch`).  The host keeps what is dict and string handling in the reference: the
contradiction / unknown reference sets, `_safe_float`, the de-dup signature strings (interned to dense ids per pack) and the
Q-number sort key.  There is no CPU path: without the CUDA library the call raises.
"""
from __future__ import annotations

import re
from typing import List, Sequence

import numpy as np

from . import _lib
from .packing import ptr

_Q_RE = re.compile(r"Q\d+")


def _safe_float(value: object, default: float = 0.0) -> float:
    """web/server.py:22860-22864"""
    try:
        return float(value)
    except Exception:
        return default


def _normalize_evidence_refs(raw_refs) -> list:
    """web/server.py:22510-22521"""
    refs = []
    if isinstance(raw_refs, str):
        refs.extend(_Q_RE.findall(raw_refs.upper()))
    elif isinstance(raw_refs, list):
        for item in raw_refs:
            if isinstance(item, str):
                refs.extend(_Q_RE.findall(item.upper()))
    return sorted(set(refs), key=lambda ref: int(ref[1:]) if ref[1:].isdigit() else 10 ** 9)


def _fact_sort_key_for_prompt(fact: dict) -> int:
    """web/server.py:22867-22874"""
    q_id = str(fact.get("q_id", "")).upper().strip()
    if re.fullmatch(r"Q\d+", q_id):
        try:
            return int(q_id[1:])
        except Exception:
            pass
    return 10 ** 9


def _fact_signature_for_prompt(fact: dict) -> str:
    """web/server.py:22876-22883"""
    dimension = str(fact.get("dimension", "")).strip().lower()
    question = re.sub(r"\s+", " ", str(fact.get("question", "")).strip().lower())
    answer = re.sub(r"\s+", " ", str(fact.get("answer", "")).strip().lower())
    return f"{dimension}|{question[:80]}|{answer[:120]}"


_I64_MAX = (1 << 63) - 1


def extract_fact_arrays(pack: dict, facts: list):
    """Per fact what the kernel takes (see include/dvb200.h, dvb_slim_facts): clamped quality, flag bits, dense signature and
    dimension ids, sort key — computed with the reference's own expressions (web/server.py:22903-22949)."""
    contradiction_refs = set()
    for item in pack.get("contradictions", []):
        if isinstance(item, dict):
            contradiction_refs.update(_normalize_evidence_refs(item.get("evidence_refs", [])))
    unknown_refs = set()
    for item in pack.get("unknowns", []):
        if isinstance(item, dict):
            q = str(item.get("q_id", "")).upper().strip()
            if re.fullmatch(r"Q\d+", q):
                unknown_refs.add(q)
    quality, flags, sig, dim, key = [], [], [], [], []
    sig_ids, dim_ids = {}, {}
    for fact in facts:
        q_id = str(fact.get("q_id", "")).upper().strip()
        quality.append(max(0.0, min(1.0, _safe_float(fact.get("quality_score", 0.0), 0.0))))
        flags.append((1 if bool(fact.get("hard_triggered", False)) else 0) | (2 if q_id in contradiction_refs else 0)
                     | (4 if q_id in unknown_refs else 0) | (8 if bool(fact.get("is_follow_up", False)) else 0))
        sig.append(sig_ids.setdefault(_fact_signature_for_prompt(fact), len(sig_ids)))
        dim.append(dim_ids.setdefault(str(fact.get("dimension", "") or "__unknown__").strip().lower() or "__unknown__", len(dim_ids)))
        key.append(min(_fact_sort_key_for_prompt(fact), _I64_MAX))             # Q + more than 18 digits: still sorts last
    return quality, flags, sig, dim, key


def select_slimmed_facts_batch(packs: Sequence[dict], limits: Sequence[object], *, slim_enabled: bool = True, dedup_enabled: bool = True,
                               min_quality: float = 0.2, dim_quota: int = 6, keep_hard_triggered: bool = True, ctx=None) -> List[list]:
    """One launch for all packs; element i equals select_slimmed_facts_for_prompt(packs[i], limits[i]) of the reference under
    the given REPORT_V3_EVIDENCE_* settings."""
    results: List[list] = [[] for _ in packs]
    todo = []                                                     # (pack index, facts, limit)
    for pi, (pack, facts_limit) in enumerate(zip(packs, limits)):
        raw_facts = pack.get("facts", [])
        if not isinstance(raw_facts, list):
            continue
        normalized_limit = max(1, int(facts_limit or 1))
        facts = [item for item in raw_facts if isinstance(item, dict)]
        if not facts:
            continue
        if not slim_enabled:
            results[pi] = facts[:normalized_limit]
            continue
        todo.append((pi, facts, normalized_limit))
    if not todo:
        return results
    n_rep = len(todo)
    fact_off = np.zeros(n_rep + 1, np.int64)
    out_off = np.zeros(n_rep + 1, np.int64)
    quality, flags, sig, dim, key, limit = [], [], [], [], [], np.zeros(n_rep, np.int32)
    for r, (pi, facts, normalized_limit) in enumerate(todo):
        q_, fl_, sg_, dm_, ky_ = extract_fact_arrays(packs[pi], facts)
        quality += q_; flags += fl_; sig += sg_; dim += dm_; key += ky_
        n = len(facts)
        limit[r] = min(normalized_limit, 2 ** 31 - 1)
        fact_off[r + 1] = fact_off[r] + n
        out_off[r + 1] = out_off[r] + min(normalized_limit, n if dedup_enabled else 3 * n)
    ctx = ctx or _lib.Context.get()
    a_q = np.asarray(quality, np.float64); a_fl = np.asarray(flags, np.uint8); a_sg = np.asarray(sig, np.int32)
    a_dm = np.asarray(dim, np.int32); a_key = np.asarray(key, np.int64)
    out_sel = np.zeros(max(int(out_off[-1]), 1), np.int32); out_cnt = np.zeros(n_rep, np.int32)
    rc = ctx.lib.dvb_slim_facts(ctx.handle, _lib.MEM_HOST, n_rep, ptr(fact_off), ptr(a_q), ptr(a_fl), ptr(a_sg), ptr(a_dm), ptr(a_key), ptr(limit),
                                ptr(out_off), int(max(1, int(dim_quota))), float(min_quality), int(bool(keep_hard_triggered)), int(bool(dedup_enabled)),
                                ptr(out_sel), ptr(out_cnt))
    ctx.check(rc, "dvb_slim_facts")
    for r, (pi, facts, _lim) in enumerate(todo):
        o = int(out_off[r])
        results[pi] = [facts[int(i)] for i in out_sel[o:o + int(out_cnt[r])]]
    return results


def select_slimmed_facts_for_prompt(evidence_pack: dict, facts_limit: int, **settings) -> list:
    """Drop-in for web/server.py:22886 (the REPORT_V3_EVIDENCE_* settings come as keyword arguments; dropin.install reads
    them from the server module at call time)."""
    return select_slimmed_facts_batch([evidence_pack], [facts_limit], **settings)[0]
