"""Claim -> fact weak evidence binding on the GPU (rows a1/a2 of SURVEY.md §8).

Host-side mirror of the reference interface:
  * `_tokenize_similarity_text_v3(text)`            web/server.py:24722
  * `infer_weak_evidence_refs_v3(field, item, evidence_pack, min_score)`   web/server.py:24780
plus the batch form `bind_batch` that the drop-in and the bulk paths use.

What runs where: the O(claims x facts x text) work — decode, lower, tokenise, set overlap,
fp64 score, first-max argmax — is ONE CUDA kernel (csrc/bind.cu) reached through the C-ABI
`dvb_bind_claims`.  The host only does what the reference does once per claim outside the
fact loop: gate checks, joining the claim's text fields, the preferred-dimension lookup and
the threshold arithmetic (web/server.py:24787-24811, 24848-24859).  There is no CPU scorer.
"""
from __future__ import annotations

import ctypes as C
import re
from operator import is_ as _is
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from .packing import pack_texts, ptr

_QID = re.compile(r"Q\d+")

CLAIM_TEXT_FIELDS = {
    "risks": ("risk", "impact", "mitigation"),
    "actions": ("action", "owner", "timeline", "metric"),
    "open_questions": ("question", "reason", "impact", "suggested_follow_up"),
}


@dataclass
class WeakBindingConfig:
    """Reference knobs (web/server.py:977-979, 990); read at call time, never baked into kernels."""
    enabled: bool = True
    min_score: float = 0.46
    unknown_ratio_trigger: float = 0.65


DEFAULT_CONFIG = WeakBindingConfig()


def _as_float(value: Any, default: float = 0.0) -> float:      # web/server.py:22860
    try:
        return float(value)
    except Exception:
        return default


def _text(value: Any) -> str:
    return str(value or "")


# ----------------------------------------------------------------------------- host control logic

def infer_item_dimension_key(field: str, item: Any, evidence_pack: Any) -> str:
    """Preferred dimension of a claim (web/server.py:24735-24777): explicit `dimension` if it is a
    coverage key, else the coverage key whose vocabulary (key, name, missing aspects) hits the
    claim's lower-cased text most often; first best wins."""
    if not isinstance(item, dict) or not isinstance(evidence_pack, dict):
        return ""
    coverage = evidence_pack.get("dimension_coverage", {})
    if not isinstance(coverage, dict):
        return ""
    explicit = _text(item.get("dimension", "")).strip()
    if explicit in coverage:
        return explicit
    corpus = " ".join(_text(item.get(k, "")) for k in
                      ("title", "description", "risk", "action", "question", "reason", "impact")).lower()
    if not corpus:
        return ""
    winner, winner_hits = "", 0
    for key, meta in coverage.items():
        if not isinstance(meta, dict):
            continue
        aspects = meta.get("missing_aspects", [])
        vocab = [key, _text(meta.get("name", ""))] + (list(aspects) if isinstance(aspects, list) else [])
        hits = 0
        for word in vocab:
            needle = _text(word).strip().lower()
            if needle and needle in corpus:
                hits += 1
        if hits > winner_hits:
            winner, winner_hits = key, hits
    return winner


def is_evidence_sparse(evidence_pack: Any, config: WeakBindingConfig = DEFAULT_CONFIG) -> bool:
    """web/server.py:24102-24112"""
    if not isinstance(evidence_pack, dict):
        return False
    facts = evidence_pack.get("facts", [])
    unknowns = evidence_pack.get("unknowns", [])
    snapshot = evidence_pack.get("quality_snapshot", {}) or {}
    n_facts = len(facts) if isinstance(facts, list) else 0
    n_unknown = len(unknowns) if isinstance(unknowns, list) else 0
    ratio = (n_unknown / n_facts) if n_facts > 0 else 0.0
    avg_q = _as_float(snapshot.get("average_quality_score", 0.0), 0.0)
    return ratio >= config.unknown_ratio_trigger or avg_q <= 0.32


def action_strategy_flags(evidence_pack: Any, config: WeakBindingConfig = DEFAULT_CONFIG) -> Tuple[bool, int]:
    """(prefer_open_questions, min_actions) of web/server.py:24115-24190 — the two fields the
    binding threshold reads (24851-24854)."""
    if not isinstance(evidence_pack, dict):
        return False, 2
    facts = evidence_pack.get("facts", [])
    blindspots = evidence_pack.get("blindspots", [])
    unknowns = evidence_pack.get("unknowns", [])
    snapshot = evidence_pack.get("quality_snapshot", {}) or {}
    n_facts = len(facts) if isinstance(facts, list) else 0
    n_blind = len(blindspots) if isinstance(blindspots, list) else 0
    n_unknown = len(unknowns) if isinstance(unknowns, list) else 0
    unknown_ratio = (n_unknown / n_facts) if n_facts > 0 else 0.0
    avg_q = max(0.0, min(1.0, _as_float(snapshot.get("average_quality_score", 0.0), 0.0)))
    total_formal = max(0, int(snapshot.get("total_formal_questions", 0) or 0))
    modes = snapshot.get("answer_mode_distribution", {})
    intents = snapshot.get("evidence_intent_distribution", {})
    modes = modes if isinstance(modes, dict) else {}
    intents = intents if isinstance(intents, dict) else {}
    pick_only = (max(0, int(modes.get("pick_only", 0) or 0)) / total_formal) if total_formal > 0 else 0.0
    low_intent = (max(0, int(intents.get("low", 0) or 0)) / total_formal) if total_formal > 0 else 0.0
    low_signal = total_formal > 0 and (pick_only >= 0.55 or low_intent >= 0.55)
    min_actions = 3 if n_facts >= 10 else 2
    prefer_open = bool(is_evidence_sparse(evidence_pack, config) or n_blind >= 8 or low_signal)
    if n_facts <= 3 and n_blind >= 8 and (unknown_ratio >= 0.25 or avg_q <= 0.58 or low_signal):
        min_actions, prefer_open = 1, True
    elif n_facts <= 6 and (n_blind >= 6 or unknown_ratio >= 0.40 or avg_q <= 0.42 or low_signal):
        min_actions, prefer_open = min(min_actions, 2), True
    return prefer_open, int(min_actions)


def binding_threshold(field: str, min_score: Any, preferred_dimension: str, evidence_pack: dict,
                      config: WeakBindingConfig = DEFAULT_CONFIG) -> float:
    """web/server.py:24848-24859 (same additions in the same order)."""
    threshold = float(min_score or config.min_score)
    if field == "actions":
        prefer_open, min_actions = action_strategy_flags(evidence_pack, config)
        threshold += 0.04
        if prefer_open:
            threshold += 0.02
        if int(min_actions or 2) <= 1:
            threshold += 0.04
    if preferred_dimension:
        threshold -= 0.03
    avg_q = _as_float((evidence_pack.get("quality_snapshot", {}) or {}).get("average_quality_score", 0.0), 0.0)
    if avg_q <= 0.30:
        threshold += 0.02
    return max(0.25, min(threshold, 0.92))


# ----------------------------------------------------------------------------- packing of sessions

class _Session:
    """Facts of one evidence_pack in device order (only facts whose q_id matches Q\\d+, 24817-24819)."""

    __slots__ = ("qids", "texts", "quality", "dims", "hard", "dim_ids", "_packed")

    def __init__(self, evidence_pack: dict):
        self.qids: List[str] = []
        self.texts: List[str] = []
        self.quality: List[float] = []
        self.dims: List[int] = []
        self.hard: List[int] = []
        self.dim_ids: Dict[str, int] = {}
        self._packed = None
        facts = evidence_pack.get("facts", [])
        for fact in facts if isinstance(facts, list) else []:
            if not isinstance(fact, dict):
                continue
            qid = _text(fact.get("q_id", "")).upper().strip()
            if not _QID.fullmatch(qid):
                continue
            dim = _text(fact.get("dimension", ""))
            self.qids.append(qid)
            self.texts.append(" ".join((_text(fact.get("question", "")), _text(fact.get("answer", "")),
                                        _text(fact.get("dimension_name", "")), dim)))
            self.quality.append(max(0.0, min(1.0, _as_float(fact.get("quality_score", 0.0), 0.0))))
            self.dims.append(self.dim_ids.setdefault(dim, len(self.dim_ids)))
            self.hard.append(1 if bool(fact.get("hard_triggered", False)) else 0)

    def packed(self):
        """The five fact arrays of dvb_bind_claims for a call on this session alone (built once per session object)."""
        if self._packed is None:
            fu, fo = pack_texts(self.texts)
            self._packed = (fu, fo, np.asarray(self.quality, np.float64), np.asarray(self.dims, np.int32), np.asarray(self.hard, np.uint8))
        return self._packed


# The reference calls infer_weak_evidence_refs_v3 once per draft item against the SAME evidence pack (web/server.py:24871-24935):
# packing its facts (join, encode, numpy) is ~100 us of a ~200 us call.  A session is kept per pack object and reused while
# every value _Session reads from every fact is still the very same object (identity, not equality: `1`, `1.0` and `True`
# compare equal and print differently).  The entry holds references to those values, so an id cannot be recycled under it; only
# immutable value types are cached — a fact holding a list or a dict is packed afresh every call.
_FACT_KEYS = ("q_id", "question", "answer", "dimension_name", "dimension", "quality_score", "hard_triggered")
_IMMUTABLE = (str, int, float, bool, type(None))
_SESSION_CACHE: Dict[int, Tuple[list, _Session]] = {}
SESSION_CACHE_ENTRIES = 16          # 0 switches the cache off


def _session_for(pack: dict) -> _Session:
    if SESSION_CACHE_ENTRIES <= 0:
        return _Session(pack)
    facts = pack.get("facts", [])
    vals: list = []
    for fact in facts if isinstance(facts, list) else []:
        if type(fact) is not dict:                       # the reference skips non-dicts; dict subclasses may compute values
            return _Session(pack)
        vals.extend(map(fact.get, _FACT_KEYS))
    hit = _SESSION_CACHE.get(id(pack))
    if hit is not None and len(hit[0]) == len(vals) and all(map(_is, hit[0], vals)):
        return hit[1]
    sess = _Session(pack)
    if all(type(v) in _IMMUTABLE for v in vals):
        if len(_SESSION_CACHE) >= SESSION_CACHE_ENTRIES:
            _SESSION_CACHE.clear()
        _SESSION_CACHE[id(pack)] = (vals, sess)
    else:
        _SESSION_CACHE.pop(id(pack), None)
    return sess


def bind_packed(ctx: "_lib.Context", claim_utf8, claim_off, fact_utf8, fact_off, sess_claim_off, sess_fact_off,
                fact_quality, fact_dim, fact_hard, claim_pref_dim, claim_threshold, *, want_lens: bool = False,
                max_claim_group_bytes: int = 0):
    """Thin call of dvb_bind_claims on host numpy arrays (end-to-end path: H2D, kernel, D2H)."""
    n_sessions = len(sess_claim_off) - 1
    n_claims, n_facts = int(sess_claim_off[-1]), int(sess_fact_off[-1])
    best = np.full(max(n_claims, 1), -1, np.int32)
    score = np.zeros(max(n_claims, 1), np.float64)
    bound = np.zeros(max(n_claims, 1), np.uint8)
    lc = np.zeros(max(n_claims, 1), np.int32) if want_lens else None
    lf = np.zeros(max(n_facts, 1), np.int32) if want_lens else None
    rc = ctx.lib.dvb_bind_claims(ctx.handle, _lib.MEM_HOST, ptr(claim_utf8), ptr(claim_off), ptr(fact_utf8), ptr(fact_off),
                                 ptr(sess_claim_off), ptr(sess_fact_off), n_sessions,
                                 ptr(fact_quality), ptr(fact_dim), ptr(fact_hard), ptr(claim_pref_dim), ptr(claim_threshold),
                                 ptr(best), ptr(score), ptr(bound), ptr(lc) if want_lens else None,
                                 ptr(lf) if want_lens else None, int(max_claim_group_bytes))
    ctx.check(rc, "dvb_bind_claims")
    out = {"best_fact": best[:n_claims], "best_score": score[:n_claims], "bound": bound[:n_claims]}
    if want_lens:
        out["len_claim"] = lc[:n_claims]
        out["len_fact"] = lf[:n_facts]
    return out


def result_from_raw(field: str, min_score: Any, raw: Optional[Tuple[str, float, str]], evidence_pack: dict,
                    config: WeakBindingConfig = DEFAULT_CONFIG) -> dict:
    """The reference's last step (web/server.py:24848-24864) on a stored raw answer (q_id of the first best fact, unrounded
    fp64 score, preferred dimension): the best fact and its score do not depend on min_score, only the bound decision does."""
    if raw is None:
        return {"refs": [], "score": 0.0}
    qid, score, preferred = raw
    threshold = binding_threshold(field, min_score, preferred, evidence_pack, config)
    return {"refs": [qid] if not (score + 1e-9 < threshold) else [], "score": round(score, 3)}


def bind_batch(requests: Sequence[Tuple[str, dict, dict, Any]], *, config: WeakBindingConfig = DEFAULT_CONFIG,
               ctx: Optional["_lib.Context"] = None, dimension_resolver=None, raw: bool = False) -> List[Any]:
    """Batch form of infer_weak_evidence_refs_v3: one kernel launch for all (field, item, pack, min_score).
    Requests that share an evidence_pack object share one device session.
    raw=True returns, per request, (q_id, unrounded score, preferred dimension) or None instead of the result dict
    (see result_from_raw; the drop-in's a3 wrapper stores these for one repair pass)."""
    results: List[Any] = [None] * len(requests)
    prefs: Dict[int, str] = {}
    resolver = dimension_resolver or infer_item_dimension_key
    sessions: Dict[int, Tuple[_Session, List[Tuple[int, str, int, float]]]] = {}
    order: List[int] = []
    # the kept sessions serve the reference's call shape (a few packs, many calls); a bulk batch of more packs than the cache
    # holds would only evict itself
    few_packs = len(requests) <= SESSION_CACHE_ENTRIES or len({id(r[2]) for r in requests}) <= SESSION_CACHE_ENTRIES
    for i, (field, item, pack, min_score) in enumerate(requests):
        empty = {"refs": [], "score": 0.0}
        if not config.enabled or field not in CLAIM_TEXT_FIELDS or not isinstance(item, dict) or not isinstance(pack, dict):
            results[i] = empty
            continue
        facts = pack.get("facts", [])
        if not isinstance(facts, list) or not facts:
            results[i] = empty
            continue
        key = id(pack)
        if key not in sessions:
            sessions[key] = (_session_for(pack) if few_packs else _Session(pack), [])
            order.append(key)
        sess, claims = sessions[key]
        claim_text = " ".join(_text(item.get(k, "")) for k in CLAIM_TEXT_FIELDS[field])
        preferred = resolver(field, item, pack)
        prefs[i] = preferred
        pref_id = -1
        if preferred:
            pref_id = sess.dim_ids.get(preferred, 1 << 30)       # a dimension no fact carries never matches
        threshold = binding_threshold(field, min_score, preferred, pack, config)
        claims.append((i, claim_text, pref_id, threshold))
    order = [k for k in order if sessions[k][1] and sessions[k][0].texts]
    for k, (sess, claims) in sessions.items():
        if not sess.texts:
            for (i, _t, _p, _th) in claims:
                results[i] = {"refs": [], "score": 0.0}
    if order:
        ctx = ctx or _lib.Context.get()
        claim_texts, fact_texts, sco, sfo = [], [], [0], [0]
        fq, fd, fh, cp, ct = [], [], [], [], []
        if len(order) == 1:                               # one pack (the reference's call shape): its packed fact arrays are kept with the session
            sess, claims = sessions[order[0]]
            fu, fo, a_fq, a_fd, a_fh = sess.packed()
            cu, co = pack_texts([t for (_i, t, _p, _th) in claims])
            out = bind_packed(ctx, cu, co, fu, fo, np.asarray([0, len(claims)], np.int64), np.asarray([0, len(sess.texts)], np.int64),
                              a_fq, a_fd, a_fh, np.asarray([p for (_i, _t, p, _th) in claims], np.int32),
                              np.asarray([th for (_i, _t, _p, th) in claims], np.float64))
            order_multi = []
        else:
            order_multi = order
        for k in order_multi:
            sess, claims = sessions[k]
            claim_texts.extend(t for (_i, t, _p, _th) in claims)
            cp.extend(p for (_i, _t, p, _th) in claims)
            ct.extend(th for (_i, _t, _p, th) in claims)
            fact_texts.extend(sess.texts)
            fq.extend(sess.quality)
            fd.extend(sess.dims)
            fh.extend(sess.hard)
            sco.append(len(claim_texts))
            sfo.append(len(fact_texts))
        if order_multi:
            cu, co = pack_texts(claim_texts)
            fu, fo = pack_texts(fact_texts)
            out = bind_packed(ctx, cu, co, fu, fo, np.asarray(sco, np.int64), np.asarray(sfo, np.int64),
                              np.asarray(fq, np.float64), np.asarray(fd, np.int32), np.asarray(fh, np.uint8),
                              np.asarray(cp, np.int32), np.asarray(ct, np.float64))
        pos = 0
        for k in order:
            sess, claims = sessions[k]
            for (i, _t, _p, _th) in claims:
                bf = int(out["best_fact"][pos])
                if raw:
                    results[i] = (sess.qids[bf], float(out["best_score"][pos]), prefs[i]) if bf >= 0 else None
                elif bf < 0:
                    results[i] = {"refs": [], "score": 0.0}
                else:
                    sc = round(float(out["best_score"][pos]), 3)
                    results[i] = {"refs": [sess.qids[bf]] if out["bound"][pos] else [], "score": sc}
                pos += 1
    if raw:
        return [r if isinstance(r, tuple) else None for r in results]
    return [r if r is not None else {"refs": [], "score": 0.0} for r in results]


def infer_weak_evidence_refs_v3(field: str, item: dict, evidence_pack: dict, min_score: Any = None, *,
                                config: WeakBindingConfig = DEFAULT_CONFIG, ctx=None, dimension_resolver=None) -> dict:
    """Drop-in for web/server.py:24780.  Never raises for the inputs the reference accepts."""
    if min_score is None:
        min_score = config.min_score
    return bind_batch([(field, item, evidence_pack, min_score)], config=config, ctx=ctx,
                      dimension_resolver=dimension_resolver)[0]


# ----------------------------------------------------------------------------- claim -> uploaded-document chunk binding
# Not in the reference: it binds claims to interview FACTS only (a2) and selects chunks by substring COUNTS only (a7).
# This joins the two halves it does have — a2's overlap score over a7's chunks — and ships behind a flag
# (dropin.install(..., claim_chunk_binding=True)).  Definition (restated independently by the test checker):
#   A = tokens(claim), B = tokens(chunk text)          (_tokenize_similarity_text_v3, web/server.py:24722-24732)
#   ov = |A & B|; ov == 0 -> not a candidate
#   score = 0.62 * ov / max(3, min(|A|, 14)) + 0.18 * ov / max(4, min(|B|, 18))      (web/server.py:24835-24843 with the
#           quality, dimension and hard-trigger terms absent: a chunk carries none)
#   order: min(floor(score * 2^20), 2^31 - 1) descending, chunk index ascending; the first k <= 32 with score > 0
#          (the coverage term is not capped at 1 — ov may exceed 14 — so scores run into the hundreds for long claims)
#   documents of more than `max_exhaustive` chunks: a claim is scored only against the <= 32 chunks the reference's own
#   selector ranks first for it (a6 terms, a7 counts, (count desc, index asc); web/server.py:32073-32085).
CLAIM_CHUNK_KEY_SCALE = float(1 << 20)


def claim_chunk_scores(claim_texts: Sequence[str], chunk_texts: Sequence[str], ctx=None) -> np.ndarray:
    """fp64 [n_claims, n_chunks] overlap scores (0 where the token sets are disjoint): ONE dvb_bind_claims launch — every
    chunk is a session whose single fact is the chunk, so the per-claim best score IS the pair's score."""
    ctx = ctx or _lib.Context.get()
    nc, nk = len(claim_texts), len(chunk_texts)
    if nc == 0 or nk == 0:
        return np.zeros((nc, nk), np.float64)
    cu, co = pack_texts([_text(t) for t in claim_texts])
    nb = int(co[-1])
    claim_utf8 = np.concatenate([np.tile(cu[:nb], nk), np.zeros(64, np.uint8)])
    claim_off = (np.arange(nk, dtype=np.int64)[:, None] * nb + co[None, :-1]).reshape(-1)
    claim_off = np.concatenate([claim_off, np.asarray([nk * nb], np.int64)])
    fu, fo = pack_texts([_text(t) for t in chunk_texts])
    out = bind_packed(ctx, claim_utf8, claim_off, fu, fo, np.arange(nk + 1, dtype=np.int64) * nc, np.arange(nk + 1, dtype=np.int64),
                      np.zeros(nk, np.float64), np.zeros(nk, np.int32), np.zeros(nk, np.uint8), np.full(nk * nc, -1, np.int32),
                      np.zeros(nk * nc, np.float64))
    score = np.where(out["best_fact"] >= 0, out["best_score"], 0.0)
    return np.ascontiguousarray(score.reshape(nk, nc).T)


def bind_claims_to_chunks(claim_texts: Sequence[str], chunk_texts: Sequence[str], k: int = 4, *, ctx=None, max_exhaustive: int = 256,
                          index=None) -> List[List[dict]]:
    """Per claim the k best chunks of a document under a2's overlap score: [{"index": chunk position, "score": fp64}, ...].
    `index`: a resident reference_material.ChunkIndex over exactly these chunks (candidate stage of large documents)."""
    from . import reference_material as rm
    ctx = ctx or _lib.Context.get()
    nc, nk = len(claim_texts), len(chunk_texts)
    k = max(1, min(int(k), 32))
    if nc == 0 or nk == 0:
        return [[] for _ in range(nc)]
    subset = np.arange(nk, dtype=np.int64)
    allowed = None
    if nk > max_exhaustive:
        own = index is None
        ix = rm.ChunkIndex(chunk_texts, ctx=ctx) if own else index
        try:
            dev_terms = rm.query_terms_batch([_text(t) for t in claim_texts], ctx=ctx, device=True)
            ci, _cs, cn = (a.cpu().numpy() for a in ix.topk_device(dev_terms, nc, 32))
        finally:
            if own:
                ix.close()
        ci = ci - ix.index_base
        valid = np.arange(32)[None, :] < cn[:, None]
        subset = np.unique(ci[valid])
        allowed = np.zeros((nc, len(subset)), bool)
        rows = np.repeat(np.arange(nc), cn)
        allowed[rows, np.searchsorted(subset, ci[valid])] = True
    if len(subset) == 0:
        return [[] for _ in range(nc)]
    scores = claim_chunk_scores(claim_texts, [chunk_texts[int(i)] for i in subset], ctx=ctx)
    if allowed is not None:
        scores = np.where(allowed, scores, 0.0)
    keys = np.minimum(np.floor(scores * CLAIM_CHUNK_KEY_SCALE), 2147483647.0).astype(np.int32)
    idx, _ks, nv = rm.topk(keys, k, index=np.ascontiguousarray(np.broadcast_to(subset, (nc, len(subset)))), ctx=ctx)
    out = []
    for c in range(nc):
        row = []
        for j in range(int(nv[c])):
            g = int(idx[c, j])
            row.append({"index": g, "score": float(scores[c, int(np.searchsorted(subset, g))])})
        out.append(row)
    return out


# ----------------------------------------------------------------------------- a1: token keys

def tokenize_similarity_keys(texts: Sequence[str], ctx=None) -> List[np.ndarray]:
    """Sorted distinct 64-bit token keys per text (key spec: include/dvb200.h)."""
    ctx = ctx or _lib.Context.get()
    data, off = pack_texts(texts)
    n = len(texts)
    cap = int(off[-1]) // 2 + n + 1
    keys = np.zeros(cap, np.uint64)
    ooff = np.zeros(n + 1, np.int64)
    rc = ctx.lib.dvb_tokenize_v3(ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), n, ptr(keys), cap, ptr(ooff))
    ctx.check(rc, "dvb_tokenize_v3")
    return [keys[ooff[i]:ooff[i + 1]].copy() for i in range(n)]


def decode_token_key(key: int) -> Optional[str]:
    """Inverse of the injective key class (tag 0: CJK bigram); None for hashed tokens."""
    key = int(key)
    if key >> 62 == 0:
        return chr(((key >> 15) & 0x7FFF) - 1 + 0x4E00) + chr((key & 0x7FFF) - 1 + 0x4E00)
    return None


def _tokenize_similarity_text_v3(text: Any, ctx=None) -> set:
    """Drop-in for web/server.py:24722: the token STRINGS.  Decode, CPython lower() and character classes come from
    the GPU symboliser (`dvb_symbolize`); the host only slices runs of the resulting symbol stream into the reference's
    tokens (ASCII runs of >= 2 chars; CJK runs cut greedily into <= 12-char pieces, each piece and its bigrams)."""
    from .reference_material import symbolize
    source = str(text or "")
    if not source.strip():
        return set()
    sym = symbolize([source], ctx=ctx)[0]
    tokens: set = set()
    i, n = 0, len(sym)
    while i < n:
        s = int(sym[i])
        if s == 0:
            i += 1
            continue
        ascii_cls = s >= 0x8000
        j = i
        while j < n and int(sym[j]) != 0 and (int(sym[j]) >= 0x8000) == ascii_cls:
            j += 1
        if ascii_cls:
            if j - i >= 2:
                tokens.add("".join("_abcdefghijklmnopqrstuvwxyz0123456789_"[(int(c) & 0x7FFF)] if (int(c) & 0x7FFF) <= 36
                                   else "_" for c in sym[i:j]))
        else:
            run = "".join(chr(int(c) - 1 + 0x4E00) for c in sym[i:j])
            for p in range(0, len(run), 12):
                piece = run[p:p + 12]
                if len(piece) >= 2:
                    tokens.add(piece)
                    for q in range(len(piece) - 1):
                        tokens.add(piece[q:q + 2])
        i = j
    return tokens
