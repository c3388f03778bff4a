"""TEST INFRASTRUCTURE ONLY — loads the *real* reference functions for parity checks.

The reference (`/root/reference/web/server.py`) cannot be imported whole here
(Flask is absent), and it does not exist on the GPU box at all.  This module
parses it with `ast`, picks the top-level definitions on the evidence-indexing
hot path and executes *those nodes from the read-only
mount* in a private namespace.  Nothing is copied into this repository.

Used by:
  * tools/make_golden.py   — generates tests/golden/*.json (committed fixtures)
  * tests/test_reference_diff.py — differential tests, skipped when the mount
    is absent (GPU box)

Never imported by the product package `deep_vision_b200`.
"""
from __future__ import annotations

import ast
import copy
import hashlib
import html
import json
import os
import re
import threading
import time
from pathlib import Path
from typing import Any, Dict, Iterable, Optional

REFERENCE_ROOT = Path(os.environ.get("DV_REFERENCE_ROOT", "/root/reference"))
SERVER_PY = REFERENCE_ROOT / "web" / "server.py"

# Config scalars the hot path reads (reference defaults; server.py:977-990, 31882-31889, 858, 35839).
CONFIG_SCALARS: Dict[str, Any] = {
    "REPORT_V3_WEAK_BINDING_ENABLED": True,
    "REPORT_V3_WEAK_BINDING_MIN_SCORE": 0.46,
    "REPORT_V3_UNKNOWN_RATIO_TRIGGER": 0.65,
    "REFERENCE_MATERIAL_CONTENT_LIMIT": 10000,
    "REFERENCE_MATERIAL_PREVIEW_LIMIT": 300,
    "REFERENCE_MATERIAL_CHUNK_SIZE": 1800,
    "REFERENCE_MATERIAL_CHUNK_OVERLAP": 160,
    "REFERENCE_MATERIAL_CONTEXT_CHUNK_LIMIT": 4,
    "REFERENCE_MATERIAL_INLINE_CHUNK_PREVIEW_LIMIT": 120,
    "MAX_DOC_LENGTH": 2000,
    "SOLUTION_SIMILARITY_NGRAM_SIZE": 2,
    "ENABLE_DEBUG_LOG": False,
}

# Top-level names executed from the reference (functions and constant assignments).
HOT_PATH_NAMES = [
    # a1 / a2
    "_safe_float", "_normalize_evidence_refs", "_tokenize_similarity_text_v3",
    "_infer_item_dimension_key_v3", "_is_evidence_sparse_v3",
    "_derive_action_generation_strategy_v3", "infer_weak_evidence_refs_v3",
    # a4 .. a7
    "_compact_reference_material_preview", "_build_reference_material_chunks",
    "_tokenize_reference_material_query", "select_reference_material_context",
    # a8 / a9
    "QUESTION_SIMILARITY_STOPWORDS", "QUESTION_SIMILARITY_SYNONYMS",
    "normalize_interview_question_text", "question_token_set",
    "is_similar_interview_question", "find_similar_interview_question_log",
    # a10
    "clean_solution_text", "_solution_signature_ngrams", "_solution_jaccard_similarity",
    # a11 / a12
    "QUESTION_ANSWER_MODES", "QUESTION_EVIDENCE_INTENTS", "DIMENSION_FOLLOW_UP_SENSITIVITY",
    "HARD_FOLLOW_UP_SIGNALS", "RICH_OPTION_CAUSE_KEYWORDS", "RICH_OPTION_CONTEXT_KEYWORDS",
    "RICH_OPTION_ROLE_KEYWORDS", "RICH_OPTION_PROCESS_KEYWORDS",
    "normalize_question_answer_mode", "normalize_question_evidence_intent",
    "build_question_capture_contract", "_score_rich_option_answer", "evaluate_answer_depth",
    "_build_follow_up_reason", "evaluate_answer_quality", "classify_answer_evidence_class_v3",
    "_infer_legacy_log_capture_contract", "backfill_session_interview_log_evidence_annotations",
]

_cache_lock = threading.Lock()
_cached_ns: Optional[Dict[str, Any]] = None


def available() -> bool:
    return SERVER_PY.is_file()


def _toplevel_nodes(tree: ast.Module, wanted: Iterable[str]):
    wanted = set(wanted)
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef)) and node.name in wanted:
            yield node.name, node
        elif isinstance(node, ast.Assign):
            for tgt in node.targets:
                if isinstance(tgt, ast.Name) and tgt.id in wanted:
                    yield tgt.id, node
        elif isinstance(node, ast.AnnAssign) and isinstance(node.target, ast.Name) and node.target.id in wanted:
            yield node.target.id, node


def load(extra_names: Iterable[str] = (), overrides: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
    """Return a namespace dict holding the reference's own hot-path functions."""
    global _cached_ns
    if not available():
        raise FileNotFoundError(f"reference not mounted at {SERVER_PY}")
    extra_names = list(extra_names)
    with _cache_lock:
        if _cached_ns is not None and not extra_names and not overrides:
            return _cached_ns
        source = SERVER_PY.read_text(encoding="utf-8")
        tree = ast.parse(source, filename=str(SERVER_PY))
        ns: Dict[str, Any] = {
            "__name__": "dv_reference_hot_path",
            "re": re, "html": html, "json": json, "copy": copy, "hashlib": hashlib,
            "time": time, "Path": Path, "Optional": Optional, "Any": Any,
        }
        import typing
        for tname in ("Dict", "List", "Tuple", "Set", "Union", "Callable", "Iterable"):
            ns[tname] = getattr(typing, tname)
        ns.update(CONFIG_SCALARS)
        if overrides:
            ns.update(overrides)
        names = HOT_PATH_NAMES + extra_names
        found = {}
        nodes = []                                     # a constant may be assigned twice (value, then its clamp): keep every node
        for name, node in _toplevel_nodes(tree, names):
            found[name] = node
            nodes.append((name, node))
        missing = [n for n in names if n not in found]
        # execute in source order so constants precede their users
        for name, node in sorted(nodes, key=lambda kv: kv[1].lineno):
            mod = ast.Module(body=[node], type_ignores=[])
            code = compile(mod, filename=str(SERVER_PY), mode="exec")
            exec(code, ns)  # noqa: S102 - executing the read-only reference on purpose
        ns["__missing__"] = missing
        if not extra_names and not overrides:
            _cached_ns = ns
        return ns


def select_context_with_chunks(ns: Dict[str, Any], chunks: list, *, query_text: str = "",
                               max_chars: int = 0, chunk_limit: int = 0, content: str = ""):
    """Run the reference's select_reference_material_context on an in-memory chunk list
    (the loader, server.py:32020-32034, is replaced by the list; its filter is re-applied)."""
    saved = ns.get("load_reference_material_chunks")
    ns["load_reference_material_chunks"] = lambda doc: [
        c for c in chunks if isinstance(c, dict) and str(c.get("text") or "").strip()
    ]
    try:
        return ns["select_reference_material_context"](
            {"content": content}, query_text=query_text, max_chars=max_chars, chunk_limit=chunk_limit)
    finally:
        if saved is not None:
            ns["load_reference_material_chunks"] = saved


if __name__ == "__main__":
    n = load()
    print("loaded", len([k for k in HOT_PATH_NAMES if k in n]), "names; missing:", n["__missing__"])
    print(n["_tokenize_similarity_text_v3"]("Hello 世界你好 ab"))
    print("stopword order:", list(n["QUESTION_SIMILARITY_STOPWORDS"]))
