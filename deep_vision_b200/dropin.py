"""install(server_module): rebind the reference's hot-path names on a loaded `web.server` module.

The reference has no FFI; callers resolve these functions by bare name at call time and `server.py` re-pushes
its globals into `report_generation_runtime.py` on every entry (web/server.py:25927-26023), so assigning the
attributes is the whole integration — the reference's own tests patch functions the same way
(tests/test_security_regression.py:4211-4219).  Every knob is read from the module AT CALL TIME
(web/server.py:977-979, 990, 31882-31889, 858, 20429-20459, 35839-35845); nothing is baked into kernels.
"""
from __future__ import annotations

import json
import threading
from typing import Any, Dict, Optional

from . import evidence, prompt_facts, question_similarity as qs, reference_material as rm

_SAVED: Dict[int, Dict[str, Any]] = {}
_MISSING = object()

HOT_NAMES = (
    "_tokenize_similarity_text_v3", "infer_weak_evidence_refs_v3", "apply_deterministic_report_repairs_v3",
    "_build_reference_material_chunks", "_tokenize_reference_material_query", "select_reference_material_context",
    "normalize_interview_question_text", "question_token_set", "is_similar_interview_question",
    "find_similar_interview_question_log", "_solution_jaccard_similarity",
    "evaluate_answer_depth", "evaluate_answer_quality", "classify_answer_evidence_class_v3",
    "backfill_session_interview_log_evidence_annotations", "build_report_evidence_pack",
)
# single-pair calls of these two sit in the interactive interview flow (web/server.py:20904-21053,
# 30566-30670), where two kernel launches per call cost more than the Python loop they replace, so they are re-bound
# only on request (install(..., saturation=True)); build_report_evidence_pack uses the batched device pass either way.
SATURATION_NAMES = ("calculate_dimension_saturation", "get_dimension_missing_aspects")
# one evidence pack per call is ~10^2 facts — a launch costs more than the reference's Python loop, so the
# single-call name is re-bound only on request (install(..., slim_facts=True)); prompt_facts.select_slimmed_facts_batch is the
# batched entry (many packs, one launch).
SLIM_NAMES = ("select_slimmed_facts_for_prompt",)


# a3 (web/server.py:25241-25612): one repair pass scores all its candidate items in ONE launch; the per-item calls the
# reference then makes (25302, 25366, 25074) are answered from this per-thread store when the item is unchanged.
_PASS = threading.local()
_BINDABLE_FIELDS = ("risks", "actions", "open_questions")


def _item_key(field, item, pack):
    try:
        return (field, id(pack), json.dumps(item, sort_keys=True, ensure_ascii=False, default=str))
    except Exception:
        return None


def _weak_cfg(mod) -> evidence.WeakBindingConfig:
    return evidence.WeakBindingConfig(
        enabled=bool(getattr(mod, "REPORT_V3_WEAK_BINDING_ENABLED", True)),
        min_score=float(getattr(mod, "REPORT_V3_WEAK_BINDING_MIN_SCORE", 0.46)),
        unknown_ratio_trigger=float(getattr(mod, "REPORT_V3_UNKNOWN_RATIO_TRIGGER", 0.65)))


def _rm_cfg(mod) -> rm.ReferenceMaterialConfig:
    d = rm.DEFAULT_CONFIG
    return rm.ReferenceMaterialConfig(
        chunk_size=int(getattr(mod, "REFERENCE_MATERIAL_CHUNK_SIZE", d.chunk_size)),
        chunk_overlap=int(getattr(mod, "REFERENCE_MATERIAL_CHUNK_OVERLAP", d.chunk_overlap)),
        context_chunk_limit=int(getattr(mod, "REFERENCE_MATERIAL_CONTEXT_CHUNK_LIMIT", d.context_chunk_limit)),
        inline_chunk_preview_limit=int(getattr(mod, "REFERENCE_MATERIAL_INLINE_CHUNK_PREVIEW_LIMIT", d.inline_chunk_preview_limit)),
        preview_limit=int(getattr(mod, "REFERENCE_MATERIAL_PREVIEW_LIMIT", d.preview_limit)),
        max_doc_length=int(getattr(mod, "MAX_DOC_LENGTH", d.max_doc_length)))


def _gate(mod, ctx) -> qs.QuestionGate:
    """Synonym dict and stop-word SET in the live process's iteration order (the set order is PYTHONHASHSEED dependent)."""
    syn = getattr(mod, "QUESTION_SIMILARITY_SYNONYMS", None)
    stop = getattr(mod, "QUESTION_SIMILARITY_STOPWORDS", None)
    # keyed on CONTENT (and the set's iteration order), not on object identity: tests and hot-config change these objects in
    # place, and the reference iterates the live objects on every call (web/server.py:20464-20468)
    key = (tuple(syn.items()) if isinstance(syn, dict) else None, tuple(stop) if stop is not None else None)
    cached = getattr(mod, "_dvb200_gate_cache", None)
    if cached is not None and cached[0] == key:
        return cached[1]
    gate = qs.QuestionGate(list(syn.items()) if isinstance(syn, dict) else None,
                           list(stop) if stop is not None else None, ctx=ctx)
    try:
        setattr(mod, "_dvb200_gate_cache", (key, gate))
    except Exception:
        pass
    return gate


def build_bindings(mod, ctx=None, sidecar_cache: Optional["rm.ChunkIndexCache"] = None) -> Dict[str, Any]:
    """The replacement callables, keyed by the reference name (signatures of SURVEY.md §8b).
    sidecar_cache (opt-in): keep / persist a resident chunk index per uploaded document (chunks.dvbidx next
    to chunks.json, found through the module's own `_resolve_data_ref`); selection results are unchanged."""

    def _index_provider(doc, chunks):
        sha = str((doc or {}).get("full_content_sha256") or "")          # written into the doc record at web/server.py:32013, 32357
        if not sha:
            return None
        side = None
        ref = str((doc or {}).get("chunk_manifest_ref") or "").strip()
        resolve = getattr(mod, "_resolve_data_ref", None)
        if ref and resolve is not None:
            from pathlib import Path
            side = Path(resolve(ref)).with_name(rm.SIDECAR_NAME)
        return sidecar_cache.get(side, sha, chunks)

    def _tokenize_similarity_text_v3(text):
        return evidence._tokenize_similarity_text_v3(text, ctx=ctx)

    def infer_weak_evidence_refs_v3(field, item, evidence_pack, min_score=None):
        cfg = _weak_cfg(mod)
        store = getattr(_PASS, "store", None)
        if store:
            key = _item_key(field, item, evidence_pack)
            if key is not None and key in store:
                return evidence.result_from_raw(field, cfg.min_score if min_score is None else min_score, store[key], evidence_pack, cfg)
        try:
            return evidence.infer_weak_evidence_refs_v3(field, item, evidence_pack, cfg.min_score if min_score is None else min_score,
                                                        config=cfg, ctx=ctx)
        except evidence._lib.DvbError:
            raise
        except Exception:
            return {"refs": [], "score": 0.0}

    repairs_original = _SAVED.get(id(mod), {}).get("apply_deterministic_report_repairs_v3") or getattr(mod, "apply_deterministic_report_repairs_v3", None)
    if getattr(repairs_original, "_dvb200_wrapper", False):
        repairs_original = None

    def apply_deterministic_report_repairs_v3(draft, evidence_pack, issues, runtime_profile=""):
        """The reference's own repair pass (control flow untouched), with its a2 calls pre-scored in one batch."""
        store = {}
        try:
            cfg = _weak_cfg(mod)
            if cfg.enabled and isinstance(draft, dict) and isinstance(evidence_pack, dict):
                reqs, keys = [], []
                for field in _BINDABLE_FIELDS:
                    values = draft.get(field, [])
                    for item in values if isinstance(values, list) else []:
                        key = _item_key(field, item, evidence_pack) if isinstance(item, dict) else None
                        if key is not None and key not in store:
                            store[key] = None
                            reqs.append((field, item, evidence_pack, cfg.min_score)); keys.append(key)
                if reqs:
                    for key, raw in zip(keys, evidence.bind_batch(reqs, config=cfg, ctx=ctx, raw=True)):
                        store[key] = raw
        except evidence._lib.DvbError:
            raise
        except Exception:
            store = {}
        prev = getattr(_PASS, "store", None)
        _PASS.store = store
        try:
            return repairs_original(draft, evidence_pack, issues, runtime_profile)
        finally:
            _PASS.store = prev
    apply_deterministic_report_repairs_v3._dvb200_wrapper = True

    def _build_reference_material_chunks(content):
        return rm._build_reference_material_chunks(content, config=_rm_cfg(mod), ctx=ctx)

    def _tokenize_reference_material_query(query_text):
        return rm._tokenize_reference_material_query(query_text)

    def select_reference_material_context(doc, *, query_text="", max_chars=0, chunk_limit=0):
        loader = getattr(mod, "load_reference_material_chunks", None)
        return rm.select_reference_material_context(doc, query_text=query_text, max_chars=max_chars, chunk_limit=chunk_limit,
                                                    chunk_loader=loader, config=_rm_cfg(mod), ctx=ctx,
                                                    index_provider=_index_provider if sidecar_cache is not None else None)

    def normalize_interview_question_text(text):
        return _gate(mod, ctx).normalize_interview_question_text(text)

    def question_token_set(text):
        return _gate(mod, ctx).question_token_set(text)

    def is_similar_interview_question(left, right, threshold=0.62):
        return _gate(mod, ctx).is_similar_interview_question(left, right, threshold)

    def find_similar_interview_question_log(logs, question, threshold=0.62):
        return _gate(mod, ctx).find_similar_interview_question_log(logs, question, threshold)

    def _solution_jaccard_similarity(left, right):
        return qs._solution_jaccard_similarity(left, right, int(getattr(mod, "SOLUTION_SIMILARITY_NGRAM_SIZE", 2)), ctx=ctx)

    def select_slimmed_facts_for_prompt(evidence_pack, facts_limit):
        return prompt_facts.select_slimmed_facts_for_prompt(
            evidence_pack, facts_limit, ctx=ctx,
            slim_enabled=bool(getattr(mod, "REPORT_V3_EVIDENCE_SLIM_ENABLED", True)),
            dedup_enabled=bool(getattr(mod, "REPORT_V3_EVIDENCE_DEDUP_ENABLED", True)),
            min_quality=getattr(mod, "REPORT_V3_EVIDENCE_MIN_QUALITY", 0.2),
            dim_quota=getattr(mod, "REPORT_V3_EVIDENCE_DIM_QUOTA", 6),
            keep_hard_triggered=bool(getattr(mod, "REPORT_V3_EVIDENCE_KEEP_HARD_TRIGGERED", True)))

    bindings = dict(locals())
    for k in ("mod", "ctx", "sidecar_cache", "_index_provider", "repairs_original"):
        bindings.pop(k, None)
    if repairs_original is None:
        bindings.pop("apply_deterministic_report_repairs_v3", None)
    try:                                             # a11-a13 (annotation path); present once annotations.py is built
        from . import annotations
        bindings.update(annotations.build_bindings(mod, ctx))
    except ImportError:
        pass
    return {k: v for k, v in bindings.items() if k in HOT_NAMES or k in SATURATION_NAMES or k in SLIM_NAMES}


def check_unicode_tables(strict: bool = False) -> bool:
    """The kernels lower-case with tables generated from ONE CPython's Unicode database; the reference lower-cases with
    the host interpreter's.  A different Unicode version can change str.lower() for a few code points, which the goldens
    of this repository would not see: warn (or refuse, strict=True) instead of drifting silently."""
    import unicodedata
    import warnings
    compiled = evidence._lib.load().dvb_unicode_version().decode("ascii")
    if compiled == unicodedata.unidata_version:
        return True
    msg = ("libdvb200 was built with Unicode %s case tables, this interpreter has Unicode %s: regenerate "
           "include/dv_unicode_tables.h with tools/gen_unicode_tables.py and rebuild" % (compiled, unicodedata.unidata_version))
    if strict:
        raise RuntimeError(msg)
    warnings.warn(msg, RuntimeWarning, stacklevel=2)
    return False


def infer_reference_chunk_refs_v3_factory(mod, ctx=None):
    """SURVEY §8(f)3 (no counterpart in the reference; opt-in): bind one report item to the chunks of an uploaded document
    with a2's overlap score.  Same shape of call as infer_weak_evidence_refs_v3: (field, item, doc, top_k) ->
    {"chunks": [{"chunk_id", "index", "start", "end", "score"}], "score": best score rounded to 3 places}."""
    def infer_reference_chunk_refs_v3(field, item, doc, top_k=4):
        empty = {"chunks": [], "score": 0.0}
        try:
            if field not in evidence.CLAIM_TEXT_FIELDS or not isinstance(item, dict):
                return empty
            loader = getattr(mod, "load_reference_material_chunks", None)
            chunks = loader(doc) if loader is not None else []
            if not chunks:
                return empty
            claim = " ".join(evidence._text(item.get(k, "")) for k in evidence.CLAIM_TEXT_FIELDS[field])
            rows = evidence.bind_claims_to_chunks([claim], [str(c.get("text") or "") for c in chunks], top_k, ctx=ctx)[0]
        except evidence._lib.DvbError:
            raise
        except Exception:
            return empty
        picked = [{"chunk_id": chunks[r["index"]].get("chunk_id"), "index": chunks[r["index"]].get("index"), "start": chunks[r["index"]].get("start"),
                   "end": chunks[r["index"]].get("end"), "score": round(r["score"], 3)} for r in rows]
        return {"chunks": picked, "score": picked[0]["score"] if picked else 0.0}
    return infer_reference_chunk_refs_v3


def install(server_module, ctx=None, names: Optional[tuple] = None, sidecar_cache=None, saturation: bool = False,
            strict_unicode: bool = False, claim_chunk_binding: bool = False, slim_facts: bool = False) -> Dict[str, Any]:
    """Rebind the hot-path names that exist on `server_module`; returns {name: replacement}.  Idempotent.
    sidecar_cache: a reference_material.ChunkIndexCache to serve select_reference_material_context from resident /
    persisted chunk indexes (off by default: the drop-in then writes nothing into the data directory)."""
    check_unicode_tables(strict_unicode)
    bindings = build_bindings(server_module, ctx, sidecar_cache)
    saved = _SAVED.setdefault(id(server_module), {})
    done = {}
    for name, fn in bindings.items():
        if names is not None and name not in names:
            continue
        if name in SATURATION_NAMES and not saturation and (names is None or name not in names):
            continue
        if name in SLIM_NAMES and not slim_facts and (names is None or name not in names):
            continue
        current = getattr(server_module, name, _MISSING)
        if current is _MISSING:
            continue                                  # no-op for anything the module does not define
        if name not in saved:
            saved[name] = current
        setattr(server_module, name, fn)
        done[name] = fn
    if claim_chunk_binding:                                    # a NEW name: nothing of the reference is replaced by it
        fn = infer_reference_chunk_refs_v3_factory(server_module, ctx)
        saved.setdefault("infer_reference_chunk_refs_v3", getattr(server_module, "infer_reference_chunk_refs_v3", _MISSING))
        setattr(server_module, "infer_reference_chunk_refs_v3", fn)
        done["infer_reference_chunk_refs_v3"] = fn
    return done


def uninstall(server_module) -> None:
    for name, fn in _SAVED.pop(id(server_module), {}).items():
        if fn is _MISSING:
            try:
                delattr(server_module, name)
            except AttributeError:
                pass
        else:
            setattr(server_module, name, fn)
