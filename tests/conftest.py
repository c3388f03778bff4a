import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (build container only)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import dv_oracle
    dv_oracle.build()
    return dv_oracle


@pytest.fixture(scope="session")
def golden():
    import json

    def load(name):
        return json.loads((ROOT / "tests" / "golden" / name).read_text(encoding="utf-8"))
    return load


@pytest.fixture(scope="session")
def gpu_ctx():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from deep_vision_b200 import _lib
    return _lib.Context.get(0)


def oracle_bind_packed(ctx, claim_utf8, claim_off, fact_utf8, fact_off, sess_claim_off, sess_fact_off,
                       fact_quality, fact_dim, fact_hard, claim_pref_dim, claim_threshold, *, want_lens=False,
                       max_claim_group_bytes=0):
    """Oracle-backed stand-in for evidence.bind_packed: lets the CPU suite exercise the HOST logic
    (packing, thresholds, result shaping) without a GPU.  Tests only."""
    from oracle import dv_oracle
    r = dv_oracle.bind_claims(None, None, sess_claim_off, sess_fact_off, fact_quality, fact_dim, fact_hard,
                              claim_pref_dim, claim_threshold, packed=(claim_utf8, claim_off, fact_utf8, fact_off))
    import numpy as np
    out = {"best_fact": r["best_fact"], "best_score": r["best_score"], "bound": r["bound"]}
    if want_lens:
        out["len_claim"] = np.minimum(r["len_claim"], 14)
        out["len_fact"] = np.minimum(r["len_fact"], 18)
    return out
