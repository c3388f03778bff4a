#!/usr/bin/env python3
"""Golden vectors for SURVEY §8(f)4: select_slimmed_facts_for_prompt (web/server.py:22886-22993) run by the reference's own
code (oracle/ref_loader.py) on seeded random evidence packs and hand-made edge cases, under several REPORT_V3_EVIDENCE_*
settings.   python tools/make_golden_slim.py  ->  tests/golden/f4_slim_facts.json   (build container only)"""
import json
import random
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import ref_loader  # noqa: E402

NAMES = ["_fact_sort_key_for_prompt", "_fact_signature_for_prompt", "select_slimmed_facts_for_prompt"]
SETTINGS = [
    {"REPORT_V3_EVIDENCE_SLIM_ENABLED": True, "REPORT_V3_EVIDENCE_DIM_QUOTA": 6, "REPORT_V3_EVIDENCE_DEDUP_ENABLED": True,
     "REPORT_V3_EVIDENCE_MIN_QUALITY": 0.2, "REPORT_V3_EVIDENCE_KEEP_HARD_TRIGGERED": True},
    {"REPORT_V3_EVIDENCE_SLIM_ENABLED": True, "REPORT_V3_EVIDENCE_DIM_QUOTA": 2, "REPORT_V3_EVIDENCE_DEDUP_ENABLED": True,
     "REPORT_V3_EVIDENCE_MIN_QUALITY": 0.5, "REPORT_V3_EVIDENCE_KEEP_HARD_TRIGGERED": False},
    {"REPORT_V3_EVIDENCE_SLIM_ENABLED": True, "REPORT_V3_EVIDENCE_DIM_QUOTA": 1, "REPORT_V3_EVIDENCE_DEDUP_ENABLED": False,
     "REPORT_V3_EVIDENCE_MIN_QUALITY": 0.9, "REPORT_V3_EVIDENCE_KEEP_HARD_TRIGGERED": True},
    {"REPORT_V3_EVIDENCE_SLIM_ENABLED": False, "REPORT_V3_EVIDENCE_DIM_QUOTA": 6, "REPORT_V3_EVIDENCE_DEDUP_ENABLED": True,
     "REPORT_V3_EVIDENCE_MIN_QUALITY": 0.2, "REPORT_V3_EVIDENCE_KEEP_HARD_TRIGGERED": True},
]
DIMS = ["customer_needs", "business_process", "tech_constraints", "project_constraints", "", "Tech_Constraints ", None]
WORDS = ["审批", "流程", "系统", "客户", "预算", "上线", "数据", "接口", "权限", "报表", "风险", "合规", "ERP", "API", "  ", "\n"]


def rand_text(rng, n):
    return "".join(rng.choice(WORDS) for _ in range(n))


def rand_pack(rng, n):
    facts = []
    for i in range(n):
        q = rng.choice([f"Q{rng.randint(1, 40)}", f"q{rng.randint(1, 40)} ", f"Q{i + 1}", "", "X7", f"Q{rng.randint(1, 9)}a", None])
        f = {"q_id": q, "dimension": rng.choice(DIMS), "question": rand_text(rng, rng.randint(1, 6)),
             "answer": rand_text(rng, rng.randint(0, 30)),
             "quality_score": rng.choice([round(rng.random(), 2), rng.random(), 0, 1, 1.7, -0.3, "0.45", "bad", None, float("nan")]),
             "hard_triggered": rng.choice([False, False, False, True, 0, 1, "yes", None]),
             "is_follow_up": rng.choice([False, False, True, 0, 1])}
        if rng.random() < 0.15 and facts:                       # duplicate signature (different q_id / quality)
            src = rng.choice(facts)
            f["dimension"], f["question"], f["answer"] = src.get("dimension"), "  " + str(src.get("question", "")).upper() + " ", src.get("answer", "")
        if rng.random() < 0.1:
            f.pop(rng.choice(list(f)))
        facts.append(f)
    if rng.random() < 0.1:
        facts.insert(rng.randrange(len(facts) + 1), "not a dict")
    pack = {"facts": facts,
            "contradictions": [{"evidence_refs": rng.choice([[f"Q{rng.randint(1, 40)}", f"q{rng.randint(1, 40)}"], f"见 Q{rng.randint(1, 40)}、Q{rng.randint(1, 40)}", [], None, [3]])}
                               for _ in range(rng.randint(0, 3))] + (["x"] if rng.random() < 0.2 else []),
            "unknowns": [{"q_id": rng.choice([f"Q{rng.randint(1, 40)}", f" q{rng.randint(1, 40)}", "Q", None])} for _ in range(rng.randint(0, 4))]}
    if rng.random() < 0.1:
        pack.pop(rng.choice(["contradictions", "unknowns"]))
    return pack


def main():
    rng = random.Random(20260705)
    packs = [({"facts": []}, 5), ({"facts": "nope"}, 5), ({}, 3), ({"facts": [1, 2]}, 3),
             ({"facts": [{"q_id": "Q2", "quality_score": 0.1}, {"q_id": "Q1", "quality_score": 0.1}]}, 0),
             ({"facts": [{"q_id": f"Q{i}", "dimension": "d", "quality_score": 0.5, "question": f"问题{i}"} for i in range(12, 0, -1)]}, 9)]
    for _ in range(150):
        n = rng.choice([1, 2, 3, 5, 8, 13, 21, 34, 60])
        packs.append((rand_pack(rng, n), rng.choice([1, 2, 4, 8, 16, 40, 100, None, 0, "7"])))
    cases = []
    for si, st in enumerate(SETTINGS):
        ns = ref_loader.load(extra_names=NAMES, overrides=st)
        fn = ns["select_slimmed_facts_for_prompt"]
        for pi, (pack, limit) in enumerate(packs):
            facts = [f for f in pack.get("facts", [])] if isinstance(pack.get("facts", []), list) else []
            out = fn(pack, limit)
            # the reference returns the fact dicts themselves: record their positions among the dict facts of the pack
            dict_facts = [f for f in facts if isinstance(f, dict)]
            pos = []
            for o in out:
                pos.append(next(i for i, f in enumerate(dict_facts) if f is o))
            cases.append({"settings": si, "pack": pi, "limit": limit, "selected": pos})
    ns0 = ref_loader.load(extra_names=NAMES)                   # one parse of the reference, not one per fact
    sig_fn, key_fn = ns0["_fact_signature_for_prompt"], ns0["_fact_sort_key_for_prompt"]

    def per_fact(fn):
        return [[fn(f) if isinstance(f, dict) else None for f in (p.get("facts", []) if isinstance(p.get("facts", []), list) else [])]
                for p, _ in packs]
    out = {"settings": SETTINGS, "packs": [p for p, _ in packs], "cases": cases, "signatures": per_fact(sig_fn), "sort_keys": per_fact(key_fn)}
    path = ROOT / "tests" / "golden" / "f4_slim_facts.json"
    path.write_text(json.dumps(out, ensure_ascii=False, indent=None, separators=(",", ":")), encoding="utf-8")
    print(path, len(cases), "cases", path.stat().st_size, "bytes")


if __name__ == "__main__":
    main()
