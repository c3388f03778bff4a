"""The C-ABI library loads without a GPU and exports every symbol include/dvb200.h declares."""
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def header_symbols():
    src = (ROOT / "include" / "dvb200.h").read_text(encoding="utf-8")
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dvb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_header_symbols():
    from deep_vision_b200 import build, _lib
    build.build()
    lib = _lib.load()
    syms = header_symbols()
    assert "dvb_bind_claims" in syms and "dvb_tokenize_v3" in syms
    for s in syms:
        assert hasattr(lib, s), s
    assert sorted(_lib.exported_symbols()) == syms
    assert lib.dvb_version() >= 100


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from deep_vision_b200 import _lib, evidence
    with pytest.raises(_lib.DvbError):
        _lib.Context(0)
    pack = {"facts": [{"q_id": "Q1", "question": "alpha beta", "answer": ""}]}
    with pytest.raises(_lib.DvbError):
        evidence.infer_weak_evidence_refs_v3("risks", {"risk": "alpha beta"}, pack)


def test_product_never_imports_oracle():
    for p in (ROOT / "deep_vision_b200").rglob("*.py"):
        txt = p.read_text(encoding="utf-8")
        assert "oracle" not in txt.replace("# oracle", ""), p
