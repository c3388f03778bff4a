"""a11-a14: per-turn evidence annotation + bulk re-index.  Golden vectors come from the REAL reference functions
(tools/make_golden_annot.py); the CPU suite pins the oracle and the host decision logic, the GPU suite the kernels."""
import copy
import json
import random

import numpy as np
import pytest


@pytest.fixture(scope="module")
def gold(golden):
    return golden("a11_a13_annotations.json")


def test_oracle_backfill_matches_reference(gold):
    from oracle import dv_annot
    for case in gold["backfill"]:
        work = copy.deepcopy(case["input"])
        summary = dv_annot.backfill(work, case["refresh_quality"], case["overwrite_contract"])
        assert summary == case["summary"]
        assert (work.get("interview_log") if isinstance(work, dict) else None) == case["logs_after"]


def test_oracle_depth_quality_class_match_reference(gold):
    from oracle import dv_annot
    for c in gold["depth"]:
        ev = dv_annot.answer_depth(**c["kwargs"])
        assert ev == c["depth"], c["kwargs"]
        ql = dv_annot.answer_quality(ev, c["kwargs"]["answer"], c["kwargs"]["is_follow_up"], c["follow_up_round"])
        assert ql == c["quality"]
        assert dv_annot.evidence_class(ev["signals"], ql["quality_score"], c["kwargs"]["evidence_intent"]) == c["class"]


# ---- plain-Python stand-ins for the two kernels (tests only): they let the CPU suite check the HOST decision tree
def py_text_features(texts, ctx=None):
    from deep_vision_b200 import annotations as an
    pb, po, pg, pw = an._PATTERNS
    pats = [(bytes(pb[po[i]:po[i + 1]]).decode("utf-8"), int(pg[i]), int(pw[i])) for i in range(len(pg))]
    mask = np.zeros(len(texts), np.uint64); length = np.zeros(len(texts), np.int32); flags = np.zeros(len(texts), np.uint8)
    for i, t in enumerate(texts):
        s = t.strip()
        m = 0
        for p, g, w in pats:
            if (s == p) if w else (p in s):
                m |= 1 << g
        mask[i] = m; length[i] = len(s); flags[i] = (1 if any(c.isdigit() for c in s) else 0) | (2 if s else 0)
    return mask, length, flags


def py_contains(haystacks, needles, needle_hay, ctx=None):
    out = np.zeros(len(needles), np.uint8)
    for j, (nd, h) in enumerate(zip(needles, needle_hay)):
        hay, s = haystacks[h].strip(), nd.strip()
        out[j] = (1 if nd in hay else 0) | (2 if s and s in hay else 0) | (4 if s and s == hay else 0)
    return out


def py_annotate_turns(q_mask, a_mask, a_len, a_flags, opt_off, opt_mask, opt_len, opt_flags, opt_bits, turn_flags, follow_up_round,
                      formal_prior, sensitivity, hard_mask, overwrite_contract, ctx=None):
    """Stand-in for dvb_annotate_turns built from the package's reference-shaped decision functions (the ones the single-call
    drop-ins use): lets the CPU suite run the extraction / write-back glue of backfill_sessions.  Tests only."""
    from deep_vision_b200 import annotations as an
    n = len(turn_flags)
    bits = np.zeros(n, np.uint32); fs = np.zeros(n, np.float64); qs = np.zeros(n, np.float64)
    hard = {an.SIGNAL_NAMES[i] for i in range(8) if (hard_mask >> i) & 1}
    for t in range(n):
        f = int(turn_flags[t])
        opts = [an.OptionFeatures(is_str=bool(opt_flags[j] & 2), length=int(opt_len[j]), mask=int(opt_mask[j]), has_digit=bool(opt_flags[j] & 1),
                                  raw_in_answer=bool(opt_bits[j] & 1), strip_in_answer=bool(opt_bits[j] & 2), equals_answer=bool(opt_bits[j] & 4))
                for j in range(int(opt_off[t]), int(opt_off[t + 1]))]
        tf = an.TurnFeatures(q_mask=int(q_mask[t]), a_mask=int(a_mask[t]), a_len=int(a_len[t]), a_digit=bool(a_flags[t] & 1), options=opts)
        dim_exact = "tech_constraints" if f & 32 else ("Tech_Constraints" if f & 64 else "other")   # exact test / lower-cased test
        log = {"dimension": dim_exact, "is_follow_up": bool(f & 8192), "follow_up_round": 0, "hard_triggered": bool(f & 8)}
        inferred = an._contract_from_features(log, [{"is_follow_up": False}] * int(formal_prior[t]), tf)
        raw_mode, raw_intent = an._MODE_NAMES[(f >> 8) & 3], an._INTENT_NAMES[(f >> 10) & 3]
        explicit = bool(raw_mode) and bool(raw_intent)
        keep = explicit and not overwrite_contract
        mode = raw_mode if keep else inferred["answer_mode"]
        rr = bool(f & 16) or mode == "pick_with_reason"
        if overwrite_contract or not explicit:
            rr = bool(inferred["requires_rationale"] or mode == "pick_with_reason")
        intent = raw_intent if keep else inferred["evidence_intent"]
        options = [0] * len(opts) if (f & 1) else []
        err = 0
        try:
            ev = an.depth_from_features(tf, dim_exact, options, bool(f & 4), bool(f & 2), mode, rr, intent, {dim_exact: float(sensitivity[t])}, hard)
        except TypeError:
            err, ev = 1, {"signals": [], "needs_follow_up": False, "suggest_ai_eval": False, "hard_triggered": False, "has_numbers": False, "follow_up_score": 0.0}
        ql = an.quality_from_features(ev, tf.a_len, tf.a_mask, bool(f & 4), int(follow_up_round[t]))
        cls = an.classify_answer_evidence_class_v3(ev["signals"], ql["quality_score"], intent)
        sig = sum(1 << an.SIGNAL_NAMES.index(s) for s in ev["signals"])
        bits[t] = (sig | (256 if ev["needs_follow_up"] else 0) | (512 if ev["suggest_ai_eval"] else 0) | (1024 if ev["hard_triggered"] else 0)
                   | (2048 if ev["has_numbers"] else 0) | (4096 if rr else 0) | (an._MODE_IDS[mode] << 13) | (an._INTENT_IDS[intent] << 15)
                   | (an._CLASS_NAMES.index(cls) << 17) | (err << 20) | ((1 << 21) if "high_quality" in ql["quality_signals"] else 0)
                   | ((1 << 22) if tf.a_len >= 80 else 0))
        fs[t] = ev["follow_up_score"]; qs[t] = ql["quality_score"]
    return bits, fs, qs


def _check_product_against_golden(an, gold, ctx=None):
    for case in gold["backfill"]:
        work = copy.deepcopy(case["input"])
        summary = an.backfill_session_interview_log_evidence_annotations(work, refresh_quality=case["refresh_quality"],
                                                                         overwrite_contract=case["overwrite_contract"], ctx=ctx)
        assert summary == case["summary"]
        assert (work.get("interview_log") if isinstance(work, dict) else None) == case["logs_after"]
    # the batch form gives the same per-session results with two launches in total
    inputs = [copy.deepcopy(c["input"]) for c in gold["backfill"] if c["refresh_quality"]]
    want = [c for c in gold["backfill"] if c["refresh_quality"]]
    got = an.backfill_sessions(inputs, ctx=ctx)
    assert got == [c["summary"] for c in want]
    for c in gold["depth"]:
        ev = an.evaluate_answer_depth(**c["kwargs"], ctx=ctx)
        assert ev == c["depth"], c["kwargs"]
        ql = an.evaluate_answer_quality(ev, c["kwargs"]["answer"], c["kwargs"]["is_follow_up"], c["follow_up_round"], ctx=ctx)
        assert ql == c["quality"]
        assert an.classify_answer_evidence_class_v3(ev["signals"], ql["quality_score"], c["kwargs"]["evidence_intent"]) == c["class"]


def _check_packs(an, gold, ctx=None):
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tools"))
    from make_golden_annot import PackStubs
    for c in gold["packs"]:
        sess = copy.deepcopy(c["input"])
        pack = an.build_report_evidence_pack(sess, server_module=PackStubs, ctx=ctx)
        assert sess == c["input"]                                       # the caller's session is not mutated (web/server.py:22570)
        assert pack.pop("generated_at")
        assert pack == c["pack"]


def test_host_decision_logic_with_standin_features(gold, monkeypatch):
    from deep_vision_b200 import annotations as an
    monkeypatch.setattr(an, "text_features", py_text_features)
    monkeypatch.setattr(an, "contains", py_contains)
    monkeypatch.setattr(an, "annotate_turns", py_annotate_turns)
    _check_product_against_golden(an, gold)
    _check_packs(an, gold)


def test_script_body_dry_run_apply_and_shards(tmp_path, gold, monkeypatch):
    """scripts/migrate_session_evidence_annotations.py:64-114 semantics (summary schema, dry run, --apply + backup)."""
    from deep_vision_b200 import annotations as an
    monkeypatch.setattr(an, "text_features", py_text_features)
    monkeypatch.setattr(an, "contains", py_contains)
    monkeypatch.setattr(an, "annotate_turns", py_annotate_turns)
    sess_dir, bak = tmp_path / "sessions", tmp_path / "bak"
    sess_dir.mkdir(); bak.mkdir()
    cases = [c for c in gold["backfill"] if c["refresh_quality"] and isinstance(c["input"].get("interview_log"), list)][:12]
    files = []
    for c in cases:
        f = sess_dir / (c["input"]["session_id"] + ".json")
        f.write_text(json.dumps(c["input"], ensure_ascii=False), encoding="utf-8")
        files.append(f)
    files.sort()
    by_id = {c["input"]["session_id"]: c for c in cases}
    dry = an.backfill_session_files(files, batch_sessions=5)
    assert dry["sessions_total"] == len(files)
    assert dry["logs_total"] == sum(c["summary"]["logs_total"] for c in cases)
    assert dry["logs_updated"] == sum(c["summary"]["logs_updated"] for c in cases)
    assert dry["sessions_changed"] == sum(1 for c in cases if c["summary"]["changed"])
    assert [e["session_id"] for e in dry["changed_sessions"]] == [f.stem for f in files if by_id[f.stem]["summary"]["changed"]]
    assert json.loads(files[0].read_text(encoding="utf-8")) == by_id[files[0].stem]["input"]          # dry run: untouched
    parts = [an.backfill_session_files(files, shard=(r, 3)) for r in range(3)]                        # 3 "ranks"
    assert an.merge_summaries(parts) == dry

    class FakeServer:
        @staticmethod
        def get_utc_now():
            return "2026-01-01T00:00:00Z"
    applied = an.backfill_session_files(files, apply_changes=True, backup_dir=bak, server_module=FakeServer)
    assert applied == dry
    for f in files:
        c = by_id[f.stem]
        data = json.loads(f.read_text(encoding="utf-8"))
        if c["summary"]["changed"]:
            assert data["updated_at"] == "2026-01-01T00:00:00Z" and data["interview_log"] == c["logs_after"]
            assert json.loads((bak / f.name).read_text(encoding="utf-8")) == c["input"]
        else:
            assert data == c["input"] and not (bak / f.name).exists()


# ----------------------------------------------------------------------------- GPU
ADVERSARIAL = ["", " ", "　\n好的\t", "好的", "好的吧", "x²", "٣ items", "𝟘", "no digits here", "a", " 可选 必须 ", "😀需要不需要",
               "（a)：b:c/d-e，f", "如", "当 如果", "abc" * 50 + "；" + "影响" * 30, " x ", "É１２", "﻿可以"]


@pytest.mark.gpu
def test_gpu_text_features_and_contains_vs_python(gpu_ctx):
    from deep_vision_b200 import annotations as an
    m, ln, fl = an.text_features(ADVERSARIAL, ctx=gpu_ctx)
    wm, wl, wf = py_text_features(ADVERSARIAL)
    assert ln.tolist() == wl.tolist() and fl.tolist() == wf.tolist()
    assert m.tolist() == wm.tolist()
    hays = ADVERSARIAL
    needles, nh = [], []
    rng = random.Random(3)
    for h, hay in enumerate(hays):
        for nd in ("", " ", "好的", " 好的 ", hay, hay.strip(), hay[1:4], "需要", "x", "\n"):
            needles.append(nd); nh.append(h)
    got = an.contains(hays, needles, nh, ctx=gpu_ctx)
    assert got.tolist() == py_contains(hays, needles, nh).tolist()


@pytest.mark.gpu
def test_gpu_annotations_golden(gpu_ctx, gold):
    from deep_vision_b200 import annotations as an
    _check_product_against_golden(an, gold, ctx=gpu_ctx)
    _check_packs(an, gold, ctx=gpu_ctx)


@pytest.mark.gpu
def test_gpu_decision_stage_rounds_like_python(gpu_ctx):
    """round(x, 2) inside dvb_annotate_turns == Python's float.__round__ on every multiple of 0.005 in range, their
    neighbouring doubles and random values: one vague_expression signal (weight 0.5) and sensitivity 2x make follow_up_score = x."""
    import math
    from deep_vision_b200 import annotations as an
    xs = []
    for k in range(-280, 281):
        x = k * 0.005
        xs += [x, math.nextafter(x, 2.0), math.nextafter(x, -2.0), k / 200.0]
    rng = random.Random(3)
    xs += [rng.uniform(-1.4, 1.4) for _ in range(20000)] + [0.0, -0.0, 1e-9, -1e-9, 0.004999999, -0.004999999]
    n = len(xs)
    sens = np.asarray([2.0 * x for x in xs], np.float64)
    bits, fs, qs = an.annotate_turns(np.zeros(n, np.uint64), np.full(n, 1 << an.G_VAGUE, np.uint64), np.full(n, 79, np.int32), np.zeros(n, np.uint8),
                                     np.zeros(n + 1, np.int32), np.zeros(1, np.uint64), np.zeros(1, np.int32), np.zeros(1, np.uint8), np.zeros(1, np.uint8),
                                     np.zeros(n, np.uint32), np.zeros(n, np.int32), np.zeros(n, np.int32), sens, 0, False, ctx=gpu_ctx)
    want = np.asarray([round(0.5 * float(s), 2) for s in sens], np.float64)
    assert np.array_equal(fs.view(np.uint64), want.view(np.uint64))           # bit-identical, the sign of zero included
    assert np.all((bits & 255) == 2) and np.all(qs == 0.25)                    # quality 0.5 - 0.25


@pytest.mark.gpu
def test_gpu_backfill_vs_oracle_random(gpu_ctx):
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tools"))
    from make_golden_annot import make_log
    from deep_vision_b200 import annotations as an
    from oracle import dv_annot
    rng = random.Random(77)
    sessions = [{"session_id": "r%d" % s, "interview_log": [make_log(rng) for _ in range(rng.randint(0, 30))]} for s in range(300)]
    ref = copy.deepcopy(sessions)
    want = [dv_annot.backfill(s) for s in ref]
    got = an.backfill_sessions(sessions, ctx=gpu_ctx)
    assert got == want
    assert sessions == ref


# ---- dimension saturation / missing aspects (golden = the reference's own functions,
# tools/make_golden_saturation.py)
class _SatModule:
    """Server-module stand-in: the app configuration of the golden generator + the reference's threshold rule
    (web/server.py:19280-19302)."""
    SATURATION_THRESHOLDS = {"high": 0.8, "medium": 0.6, "low": 0.4}

    def __init__(self):
        import sys
        from pathlib import Path
        sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tools"))
        from make_golden_saturation import SatStubs
        for k in ("REPORT_TEMPLATE_CUSTOM_V1", "get_dimension_info_for_session", "get_interview_mode_config",
                  "resolve_report_template_for_session", "normalize_custom_report_schema"):
            setattr(self, k, getattr(SatStubs, k))

    def get_mode_saturation_thresholds(self, session):
        quality = self.get_interview_mode_config(session).get("quality_thresholds") or {}
        return {k: quality.get(k, self.SATURATION_THRESHOLDS[k]) for k in ("high", "medium", "low")}


def _check_saturation(an, sat, ctx=None):
    mod = _SatModule()
    items, want = [], []
    for c in sat["cases"]:
        for dim, w in c["per_dimension"].items():
            items.append((copy.deepcopy(c["input"]), dim)); want.append(w)
    got = an.dimension_saturation_batch(items, server_module=mod, ctx=ctx)            # one contains + one feature launch
    assert got == [w["saturation"] for w in want]
    assert an.missing_aspects_batch(items, server_module=mod, ctx=ctx) == [w["missing"] for w in want]
    assert [it[0] for it in items] == [c["input"] for c in sat["cases"] for _ in c["per_dimension"]]     # sessions untouched
    c = sat["cases"][3]
    for dim, w in c["per_dimension"].items():                                        # single-pair forms
        assert an.calculate_dimension_saturation(c["input"], dim, server_module=mod, ctx=ctx) == w["saturation"]
        assert an.get_dimension_missing_aspects(c["input"], dim, server_module=mod, ctx=ctx) == w["missing"]
    for c in sat["packs"]:                                                           # pack with the batched device coverage block
        sess = copy.deepcopy(c["input"])
        pack = an.build_report_evidence_pack(sess, server_module=mod, ctx=ctx, device_coverage=True)
        assert sess == c["input"]
        assert pack.pop("generated_at")
        assert pack == c["pack"]
    with pytest.raises(TypeError):                                                   # `5 in text` raises in the reference too
        bad = _SatModule()
        bad.get_dimension_info_for_session = lambda s: {"d": {"key_aspects": [5]}}
        an.dimension_saturation_batch([({"interview_log": [{"dimension": "d", "answer": "a", "question": "q"}]}, "d")], server_module=bad, ctx=ctx)


def test_saturation_host_logic_with_standin_features(golden, monkeypatch):
    from deep_vision_b200 import annotations as an
    monkeypatch.setattr(an, "text_features", py_text_features)
    monkeypatch.setattr(an, "contains", py_contains)
    monkeypatch.setattr(an, "annotate_turns", py_annotate_turns)
    _check_saturation(an, golden("f2_saturation.json"))


def test_dropin_saturation_rebinding_and_stub_respect(golden, monkeypatch):
    """install(..., saturation=True) rebinds the two names; the pack's device coverage block steps aside when a caller
    re-binds get_dimension_missing_aspects afterwards (the reference's own tests stub it)."""
    import types
    from deep_vision_b200 import annotations as an, dropin
    monkeypatch.setattr(an, "text_features", py_text_features)
    monkeypatch.setattr(an, "contains", py_contains)
    monkeypatch.setattr(an, "annotate_turns", py_annotate_turns)
    sat = golden("f2_saturation.json")
    base = _SatModule()
    mod = types.SimpleNamespace(**{k: getattr(base, k) for k in ("REPORT_TEMPLATE_CUSTOM_V1", "get_dimension_info_for_session",
                                                                  "get_interview_mode_config", "resolve_report_template_for_session",
                                                                  "normalize_custom_report_schema", "get_mode_saturation_thresholds")})
    mod.calculate_dimension_saturation = lambda s, d: {"covered_aspects": []}
    mod.get_dimension_missing_aspects = lambda s, d: ["原函数"]
    mod.build_report_evidence_pack = lambda s: {}
    done = dropin.install(mod, names=("build_report_evidence_pack",))
    assert set(done) == {"build_report_evidence_pack"}
    c = sat["packs"][0]
    pack = mod.build_report_evidence_pack(copy.deepcopy(c["input"]))
    pack.pop("generated_at")
    assert pack == c["pack"]                                        # originals untouched -> batched device coverage
    mod.get_dimension_missing_aspects = lambda s, d: ["桩"]         # a caller stubs the name after install
    pack = mod.build_report_evidence_pack(copy.deepcopy(c["input"]))
    assert all(v["missing_aspects"] == ["桩"] for v in pack["dimension_coverage"].values())
    dropin.uninstall(mod)
    done = dropin.install(mod, saturation=True)
    assert {"calculate_dimension_saturation", "get_dimension_missing_aspects"} <= set(done)
    w = sat["cases"][3]
    for dim, want in w["per_dimension"].items():
        assert mod.calculate_dimension_saturation(w["input"], dim) == want["saturation"]
        assert mod.get_dimension_missing_aspects(w["input"], dim) == want["missing"]
    dropin.uninstall(mod)


@pytest.mark.gpu
def test_gpu_saturation_golden(gpu_ctx, golden):
    from deep_vision_b200 import annotations as an
    _check_saturation(an, golden("f2_saturation.json"), ctx=gpu_ctx)
