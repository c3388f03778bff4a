"""Runs the bind kernels on the CPU SIMT emulator against the oracle for named cases (development helper).
usage: python tools/emu_check.py [deg mixed ovf long pair syn big] [--mode N] [--grid N]"""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "tests", ROOT / "tests" / "emu"):
    sys.path.insert(0, str(p))
import numpy as np
import emu, bind_cases as bc
from oracle import dv_oracle


def check(name, a, mode=0, grid=3):
    t0 = time.time(); got = emu.bind_claims(*a, grid=grid, mode=mode); dt = time.time() - t0
    cu, co, fu, fo, sco, sfo, fq, fd, fh, cp, ct = a
    ref = dv_oracle.bind_claims(None, None, sco, sfo, fq, fd, fh, cp, ct, packed=(cu, co, fu, fo))
    has_claims = np.repeat(np.diff(sco) > 0, np.diff(sfo))            # len_fact is only written for sessions with claims
    ok = True
    for k, cap in (("best_fact", None), ("bound", None), ("len_claim", 14), ("len_fact", 18)):
        r = ref[k] if cap is None else np.minimum(ref[k], cap)
        g = got[k]
        if k == "len_fact":
            r, g = r[has_claims], g[has_claims]
        if not np.array_equal(g, r):
            ok = False; bad = np.nonzero(g != r)[0]; print("  MISMATCH", k, len(bad), bad[:8], g[bad[:8]], r[bad[:8]])
    if not np.array_equal(got["best_score"].view(np.uint64), ref["best_score"].view(np.uint64)):
        ok = False; bad = np.nonzero(got["best_score"] != ref["best_score"])[0]
        print("  MISMATCH score", len(bad), bad[:8], got["best_score"][bad[:8]], ref["best_score"][bad[:8]])
    print(name, "OK" if ok else "FAIL", "%.1fs" % dt, "handed back:", int(got["fallback"].sum()), "of", len(got["fallback"]), flush=True)
    return ok


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    mode = int(sys.argv[sys.argv.index("--mode") + 1]) if "--mode" in sys.argv else 0
    grid = int(sys.argv[sys.argv.index("--grid") + 1]) if "--grid" in sys.argv else 3
    args = [a for a in args if not a.isdigit()]
    which = args or ["deg", "mixed", "ovf", "long", "pair", "syn", "big"]
    res = []
    if "deg" in which: res.append(check("degenerate", bc.degenerate(), mode, grid))
    if "mixed" in which: res.append(check("mixed", bc.mixed_real_text(), mode, grid))
    if "ovf" in which: res.append(check("tok overflow", bc.token_table_overflow(), mode, grid))
    if "long" in which: res.append(check("long facts", bc.long_facts(1), mode, grid))
    if "pair" in which: res.append(check("every pair", bc.every_pair(), mode, grid))
    if "big" in which: res.append(check("big claim groups", bc.big_claim_groups(), mode, grid))
    if "syn" in which:
        for n, t, c in [(8, 40, 24), (5, 70, 40), (3, 1, 1), (10, 7, 33)]:
            res.append(check("synthetic %d %d %d" % (n, t, c), bc.synthetic(n, t, c), mode, grid))
    sys.exit(0 if all(res) else 1)
