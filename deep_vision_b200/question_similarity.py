"""Question-similarity quality gate and solution-signature similarity (rows a8-a10 of SURVEY.md §8).

Host-side mirror of the reference interface (web/server.py):
  * `normalize_interview_question_text(text)`                     20462
  * `question_token_set(text)`                                    20472
  * `is_similar_interview_question(left, right, threshold=0.62)`  20485
  * `find_similar_interview_question_log(logs, question, threshold=0.62)`   20494
  * `clean_solution_text`, `_solution_jaccard_similarity(left, right)`      34997, 36808

All text work (strip, CPython lower(), ordered replaces, bigram sets, intersections, the fp64
threshold compare, "most recent match wins") runs on the GPU through `dvb_question_keys` /
`dvb_question_gate` / `dvb_solution_jaccard`.  The synonym dict and stop-word set are passed IN THE
CALLER'S ITERATION ORDER: the reference iterates a `set` (web/server.py:20467), whose order depends on
PYTHONHASHSEED, so `install()` reads `list(server.QUESTION_SIMILARITY_STOPWORDS)` from the live process.
"""
from __future__ import annotations

import html
import re
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from .packing import pack_texts, ptr

# reference defaults (web/server.py:20429-20459); the stop-word ORDER here is only a default for
# standalone use — a drop-in must pass the order of the host process.
DEFAULT_SYNONYMS: List[Tuple[str, str]] = [("痛点", "问题"), ("诉求", "需求"), ("最关键", "关键"), ("最大的", "大"), ("最大", "大"),
                                           ("当前", ""), ("目前", ""), ("头疼", ""), ("影响推进", "")]
DEFAULT_STOPWORDS: List[str] = ["当前", "目前", "主要", "最", "核心", "请", "你", "们", "让", "您", "贵司", "是什么", "哪些", "方面",
                                "一个", "一下", "的"]


def _cp_csr(strings: Sequence[str]):
    cps: List[int] = []
    off = [0]
    for s in strings:
        cps.extend(ord(c) for c in s)
        off.append(len(cps))
    return np.asarray(cps if cps else [0], np.uint32), np.asarray(off, np.int32)


class QuestionGate:
    """Holds the synonym / stop-word tables in caller order and exposes the reference-shaped functions."""

    def __init__(self, synonyms: Optional[Sequence[Tuple[str, str]]] = None, stopwords: Optional[Sequence[str]] = None, ctx=None):
        self.synonyms = [(str(a), str(b)) for a, b in (DEFAULT_SYNONYMS if synonyms is None else synonyms)]
        self.stopwords = [str(w) for w in (DEFAULT_STOPWORDS if stopwords is None else stopwords)]
        self._ctx = ctx
        self._ss, self._sso = _cp_csr([a for a, _ in self.synonyms])
        self._sd, self._sdo = _cp_csr([b for _, b in self.synonyms])
        self._st, self._sto = _cp_csr(self.stopwords)
        growth = 1
        for a, b in self.synonyms:
            if a and len(b) > len(a):
                growth *= (len(b) + len(a) - 1) // len(a)
        self._growth = growth

    @property
    def ctx(self):
        return self._ctx or _lib.Context.get()

    def _table_args(self):
        return (ptr(self._ss), ptr(self._sso), ptr(self._sd), ptr(self._sdo), len(self.synonyms),
                ptr(self._st), ptr(self._sto), len(self.stopwords))

    # ---- batch primitives
    def keys(self, texts: Sequence[str], want_norm: bool = False):
        """Per text: sorted distinct bigram keys (uint64) [and the normalised string]."""
        ctx = self.ctx
        texts = [str(t or "") for t in texts]
        data, off = pack_texts(texts)
        n = len(texts)
        cap = int(((off[1:] - off[:-1]) + 2).sum() * self._growth) if n else 0
        cap_off = np.zeros(n + 1, np.int64)
        keys = np.zeros(max(cap, 1), np.uint64)
        cnt = np.zeros(max(n, 1), np.int32)
        norm = np.zeros(max(cap, 1), np.uint32) if want_norm else None
        nlen = np.zeros(max(n, 1), np.int32) if want_norm else None
        if n:
            rc = ctx.lib.dvb_question_keys(ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), n, *self._table_args(),
                                           ptr(cap_off), ptr(keys), ptr(cnt), ptr(norm) if want_norm else None,
                                           ptr(nlen) if want_norm else None, cap)
            ctx.check(rc, "dvb_question_keys")
        out_keys = [keys[cap_off[i]:cap_off[i] + cnt[i]].copy() for i in range(n)]
        if not want_norm:
            return out_keys
        norms = ["".join(map(chr, norm[cap_off[i]:cap_off[i] + nlen[i]].tolist())) for i in range(n)]
        return out_keys, norms

    def match(self, candidates: Sequence[str], history: Sequence[str], threshold: float = 0.62, packed=None) -> np.ndarray:
        """For each candidate the largest history index whose question is similar, else -1."""
        ctx = self.ctx
        if packed is None:
            cu, co = pack_texts([str(c or "") for c in candidates])
            hu, ho = pack_texts([str(h or "") for h in history])
        else:
            cu, co, hu, ho = packed
        nc, nh = len(co) - 1, len(ho) - 1
        out = np.full(max(nc, 1), -1, np.int64)
        if nc:
            rc = ctx.lib.dvb_question_gate(ctx.handle, _lib.MEM_HOST, ptr(cu), ptr(co), nc, ptr(hu), ptr(ho), nh, *self._table_args(),
                                           float(threshold), ptr(out))
            ctx.check(rc, "dvb_question_gate")
        return out[:nc]

    # ---- reference-shaped functions
    def normalize_interview_question_text(self, text: object) -> str:
        return self.keys([str(text or "")], want_norm=True)[1][0]

    def question_token_set(self, text: object) -> set:
        _keys, norms = self.keys([str(text or "")], want_norm=True)
        s = norms[0]
        if not s:
            return set()
        if len(s) == 1:
            return {s}
        return {s[i:i + 2] for i in range(len(s) - 1)}           # strings re-assembled from the kernel's normalised text

    def is_similar_interview_question(self, left: object, right: object, threshold: float = 0.62) -> bool:
        return bool(self.match([str(right or "")], [str(left or "")], threshold)[0] >= 0)

    def find_similar_interview_question_log(self, logs: list, question: object, threshold: float = 0.62) -> Optional[dict]:
        candidate = str(question or "").strip()
        if not candidate:
            return None
        logs = list(logs or [])
        if not logs:
            return None
        hit = int(self.match([candidate], [str(log.get("question") or "") for log in logs], threshold)[0])
        return logs[hit] if hit >= 0 else None


# ----------------------------------------------------------------------------- a10

def clean_solution_text(value: object, max_len: int = 0) -> str:
    """web/server.py:34997-35007 (host: markdown/html stripping of a <= ~600 char string)."""
    text = html.unescape(str(value or ""))
    text = re.sub(r"<[^>]+>", " ", text)
    text = re.sub(r"\[([^\]]+)\]\([^\)]+\)", r"\1", text)
    text = re.sub(r"(?m)^\s{0,3}#{1,6}\s*", "", text)
    text = re.sub(r"(?m)^\s*[-*+]\s+", "", text)
    text = text.replace("**", "").replace("__", "").replace("`", "")
    text = re.sub(r"\s+", " ", text).strip(" \t\r\n-—|:：；;，。,.")
    if max_len and len(text) > max_len:
        text = text[:max_len].rstrip("，。,.;；：:!?？！ ")
    return text


def solution_jaccard_counts(pairs: Sequence[Tuple[str, str]], ngram: int = 2, ctx=None):
    """(inter, union, |L|, |R|) arrays for already-cleaned (left, right) pairs."""
    ctx = ctx or _lib.Context.get()
    lu, lo = pack_texts([p[0] for p in pairs])
    ru, ro = pack_texts([p[1] for p in pairs])
    n = len(pairs)
    outs = [np.zeros(max(n, 1), np.int32) for _ in range(4)]
    if n:
        rc = ctx.lib.dvb_solution_jaccard(ctx.handle, _lib.MEM_HOST, ptr(lu), ptr(lo), ptr(ru), ptr(ro), n, int(ngram), *[ptr(o) for o in outs])
        ctx.check(rc, "dvb_solution_jaccard")
    return [o[:n] for o in outs]


def _solution_jaccard_similarity(left: str, right: str, ngram: int = 2, ctx=None) -> float:
    """Drop-in for web/server.py:36808."""
    inter, union, nl, nr = solution_jaccard_counts([(clean_solution_text(left), clean_solution_text(right))], ngram, ctx=ctx)
    if nl[0] == 0 or nr[0] == 0:
        return 0.0
    return round(int(inter[0]) / max(1, int(union[0])), 3)


def max_solution_similarity(signature: str, others: Sequence[Tuple[str, str]], ngram: int = 2, ctx=None) -> Tuple[float, str]:
    """The similarity loop of build_solution_quality_signals (web/server.py:36970-36984) as ONE launch: `others` =
    (report_name, similarity text) of the recent sidecars that passed the fingerprint check, in the reference's order.
    -> (max score, name of the FIRST report that reaches it; (0.0, "") when no score is positive — the reference's strict '>')."""
    if not others:
        return 0.0, ""
    left = clean_solution_text(signature)
    inter, union, nl, nr = solution_jaccard_counts([(left, clean_solution_text(t)) for _n, t in others], ngram, ctx=ctx)
    best, name = 0.0, ""
    for i, (n, _t) in enumerate(others):
        score = 0.0 if nl[i] == 0 or nr[i] == 0 else round(int(inter[i]) / max(1, int(union[i])), 3)
        if score > best:
            best, name = score, str(n or "")
    return best, name
