"""select_slimmed_facts_for_prompt (web/server.py:22886-22993).
CPU tier: the oracle restatement (oracle/dv_slim.py) and the product's host extraction against goldens the reference's own
function produced (tools/make_golden_slim.py).  GPU tier: dvb_slim_facts through the host mirror against the same goldens
and against the oracle on random packs."""
import json
import random
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
GOLD = json.loads((ROOT / "tests" / "golden" / "f4_slim_facts.json").read_text(encoding="utf-8"))


def _kw(st):
    return dict(slim_enabled=st["REPORT_V3_EVIDENCE_SLIM_ENABLED"], dedup_enabled=st["REPORT_V3_EVIDENCE_DEDUP_ENABLED"],
                min_quality=st["REPORT_V3_EVIDENCE_MIN_QUALITY"], dim_quota=st["REPORT_V3_EVIDENCE_DIM_QUOTA"],
                keep_hard_triggered=st["REPORT_V3_EVIDENCE_KEEP_HARD_TRIGGERED"])


def _dict_facts(pack):
    raw = pack.get("facts", [])
    return [f for f in raw if isinstance(f, dict)] if isinstance(raw, list) else []


def test_host_signature_and_sort_key_golden():
    from deep_vision_b200 import prompt_facts as pf
    n = 0
    for pack, sigs, keys in zip(GOLD["packs"], GOLD["signatures"], GOLD["sort_keys"]):
        raw = pack.get("facts", []) if isinstance(pack.get("facts", []), list) else []
        for f, s, k in zip(raw, sigs, keys):
            if isinstance(f, dict):
                assert pf._fact_signature_for_prompt(f) == s
                assert pf._fact_sort_key_for_prompt(f) == k
                n += 1
    assert n > 1000


def test_oracle_matches_reference_golden():
    from deep_vision_b200 import prompt_facts as pf
    from oracle import dv_slim
    checked = 0
    for case in GOLD["cases"]:
        st = GOLD["settings"][case["settings"]]
        pack = GOLD["packs"][case["pack"]]
        facts = _dict_facts(pack)
        limit = max(1, int(case["limit"] or 1))
        if not facts:
            assert case["selected"] == []
            continue
        if not st["REPORT_V3_EVIDENCE_SLIM_ENABLED"]:
            assert case["selected"] == list(range(min(limit, len(facts))))
            continue
        q, fl, sg, dm, ky = pf.extract_fact_arrays(pack, facts)
        got = dv_slim.slim_select(q, fl, sg, dm, ky, limit, dim_quota=max(1, st["REPORT_V3_EVIDENCE_DIM_QUOTA"]),
                                  min_quality=st["REPORT_V3_EVIDENCE_MIN_QUALITY"], keep_hard=st["REPORT_V3_EVIDENCE_KEEP_HARD_TRIGGERED"],
                                  dedup=st["REPORT_V3_EVIDENCE_DEDUP_ENABLED"])
        assert got == case["selected"], (case["settings"], case["pack"])
        checked += 1
    assert checked > 300


@pytest.mark.gpu
def test_gpu_slim_facts_golden():
    from deep_vision_b200 import prompt_facts as pf
    for si, st in enumerate(GOLD["settings"]):
        cases = [c for c in GOLD["cases"] if c["settings"] == si]
        packs = [GOLD["packs"][c["pack"]] for c in cases]
        got = pf.select_slimmed_facts_batch(packs, [c["limit"] for c in cases], **_kw(st))       # ONE launch per settings row
        for c, pack, g in zip(cases, packs, got):
            facts = _dict_facts(pack)
            assert [next(i for i, f in enumerate(facts) if f is o) for o in g] == c["selected"], (si, c["pack"])
    one = pf.select_slimmed_facts_for_prompt(GOLD["packs"][10], 8)                              # the single-call name
    c10 = next(c for c in GOLD["cases"] if c["settings"] == 0 and c["pack"] == 10)
    if c10["limit"] == 8:
        assert [_dict_facts(GOLD["packs"][10]).index(o) for o in one] == c10["selected"]


@pytest.mark.gpu
def test_gpu_slim_facts_random_vs_oracle_and_limits():
    from deep_vision_b200 import _lib, prompt_facts as pf
    from oracle import dv_slim
    rng = random.Random(7)
    packs, limits = [], []
    for _ in range(400):
        n = rng.choice([1, 7, 33, 100, 257, 1024])
        facts = [{"q_id": "Q%d" % rng.randint(1, 300), "dimension": "d%d" % rng.randint(0, 5), "question": "q%d" % rng.randint(0, n),
                  "answer": "a%d" % rng.randint(0, 3), "quality_score": rng.choice([rng.random(), round(rng.random(), 1)]),
                  "hard_triggered": rng.random() < 0.1, "is_follow_up": rng.random() < 0.3} for _ in range(n)]
        packs.append({"facts": facts, "contradictions": [{"evidence_refs": ["Q%d" % rng.randint(1, 300) for _ in range(3)]}],
                      "unknowns": [{"q_id": "Q%d" % rng.randint(1, 300)} for _ in range(5)]})
        limits.append(rng.choice([1, 5, 24, 60, 2000]))
    for dedup in (True, False):
        got = pf.select_slimmed_facts_batch(packs, limits, dedup_enabled=dedup, dim_quota=3, min_quality=0.35)
        for pack, lim, g in zip(packs, limits, got):
            facts = pack["facts"]
            q, fl, sg, dm, ky = pf.extract_fact_arrays(pack, facts)
            ref = dv_slim.slim_select(q, fl, sg, dm, ky, lim, dim_quota=3, min_quality=0.35, dedup=dedup)
            ids = {id(f): i for i, f in enumerate(facts)}
            assert [ids[id(o)] for o in g] == ref
    too_many = {"facts": [{"q_id": "Q1", "question": str(i)} for i in range(1025)]}
    with pytest.raises(_lib.DvbError):
        pf.select_slimmed_facts_batch([too_many], [5])


@pytest.mark.gpu
def test_gpu_max_solution_similarity_equals_pairwise_scan():
    from deep_vision_b200 import question_similarity as qs
    others = [("r%d" % i, t) for i, t in enumerate(["客户审批流程优化方案", "审批流程 优化 方案 客户", "", "ERP 接口改造", "客户审批流程优化方案"])]
    sig = "客户 审批流程优化 | 方案"
    best, name = 0.0, ""
    for n, t in others:
        s = qs._solution_jaccard_similarity(sig, t)
        if s > best:
            best, name = s, n
    assert qs.max_solution_similarity(sig, others) == (best, name)
    assert qs.max_solution_similarity(sig, []) == (0.0, "")


@pytest.mark.gpu
def test_gpu_slim_facts_device_buffers_match_host_buffers():
    """dvb_slim_facts with DVB_MEM_DEVICE (every array already on the GPU, outputs written in place) equals the host-buffer call."""
    import torch
    from deep_vision_b200 import _lib, prompt_facts as pf
    from deep_vision_b200.packing import ptr
    rng = random.Random(3)
    packs = [{"facts": [{"q_id": "Q%d" % rng.randint(1, 50), "dimension": "d%d" % rng.randint(0, 3), "question": "q%d" % rng.randint(0, 20),
                         "quality_score": rng.random(), "hard_triggered": rng.random() < 0.2} for _ in range(rng.choice([3, 40, 90]))]} for _ in range(64)]
    limits = [rng.choice([2, 10, 30]) for _ in packs]
    want = pf.select_slimmed_facts_batch(packs, limits)
    ctx = _lib.Context.get()
    fo, oo, q, fl, sg, dm, ky = [0], [0], [], [], [], [], []
    for p, lim in zip(packs, limits):
        a = pf.extract_fact_arrays(p, p["facts"])
        q += a[0]; fl += a[1]; sg += a[2]; dm += a[3]; ky += a[4]
        fo.append(len(q)); oo.append(oo[-1] + min(lim, len(p["facts"])))
    host = [np.asarray(fo, np.int64), np.asarray(q, np.float64), np.asarray(fl, np.uint8), np.asarray(sg, np.int32), np.asarray(dm, np.int32),
            np.asarray(ky, np.int64), np.asarray(limits, np.int32), np.asarray(oo, np.int64)]
    dev = [torch.from_numpy(x).cuda() for x in host]
    d_sel = torch.full((oo[-1],), -7, dtype=torch.int32, device="cuda"); d_cnt = torch.zeros(len(packs), dtype=torch.int32, device="cuda")
    rc = ctx.lib.dvb_slim_facts(ctx.handle, _lib.MEM_DEVICE, len(packs), *[t.data_ptr() for t in dev], 6, 0.2, 1, 1, d_sel.data_ptr(), d_cnt.data_ptr())
    ctx.check(rc, "dvb_slim_facts")
    torch.cuda.synchronize()
    sel, cnt = d_sel.cpu().numpy(), d_cnt.cpu().numpy()
    for r, (p, w) in enumerate(zip(packs, want)):
        got = [p["facts"][int(i)] for i in sel[oo[r]:oo[r] + int(cnt[r])]]
        assert [id(x) for x in got] == [id(x) for x in w]
