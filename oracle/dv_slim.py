"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's evidence slimming.

Plain Python over the arrays the C-ABI takes (small cases only), each step citing the reference lines it follows.  Pinned by
tests/test_slim_facts.py against tests/golden/f4_slim_facts.json, which tools/make_golden_slim.py generates with the
reference's own select_slimmed_facts_for_prompt (web/server.py:22886-22993).  Never imported by the product package.
"""
from __future__ import annotations


def slim_select(quality, flags, sig, dim, key, limit, *, dim_quota=6, min_quality=0.2, keep_hard=True, dedup=True):
    """-> selected fact indices of ONE report in the reference's final order.
    quality: clamped floats; flags: bit 0 hard, 1 contradiction ref, 2 unknown ref, 3 follow-up; sig / dim: dense ids; key: Q numbers."""
    n = len(quality)
    score = []
    for i in range(n):                                            # server.py:22930-22938, float adds in the reference's order
        s = quality[i] * 100.0
        if flags[i] & 1:
            s += 24.0
        if flags[i] & 2:
            s += 22.0
        if flags[i] & 4:
            s += 12.0
        if flags[i] & 8:
            s += 8.0
        score.append(s)
    order = sorted(range(n), key=lambda i: (-score[i], i))      # server.py:22952
    selected, taken_sig, dim_counter = [], set(), {}

    def mandatory(i):                                             # server.py:22945
        return bool(flags[i] & 2) or (keep_hard and bool(flags[i] & 1))

    def try_take(i, bypass_quota, bypass_quality):                # server.py:22959-22977
        if len(selected) >= limit:
            return False
        if dedup and sig[i] in taken_sig:
            return False
        if not bypass_quality and not mandatory(i) and quality[i] < min_quality:
            return False
        if not bypass_quota and dim_counter.get(dim[i], 0) >= dim_quota:
            return False
        selected.append(i)
        taken_sig.add(sig[i])
        dim_counter[dim[i]] = dim_counter.get(dim[i], 0) + 1
        return True

    for i in order:                                               # server.py:22979-22984
        if not mandatory(i):
            continue
        if len(selected) >= limit:
            break
        try_take(i, True, True)
    for i in order:                                               # server.py:22986-22989
        if len(selected) >= limit:
            break
        try_take(i, False, False)
    if len(selected) < limit:                                     # server.py:22991-22996
        for i in order:
            if len(selected) >= limit:
                break
            try_take(i, True, True)
    return sorted(selected, key=lambda i: key[i])                 # server.py:22998-22999 (stable)
