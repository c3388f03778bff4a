"""Adversarial and synthetic inputs of dvb_bind_claims shared by the GPU parity tests (tests/test_gpu_parity.py) and the
CPU SIMT-emulator tests (tests/test_emu_bind.py).  Every builder returns the 11 packed arrays of the call."""
import numpy as np


def adversarial_long_texts():
    import random
    rng = random.Random(2024)
    han = [chr(0x4E00 + rng.randrange(3000)) for _ in range(400)]
    def cjk(n): return "".join(rng.choice(han) for _ in range(n))
    def word(n): return "".join(rng.choice("abcdefghijklmnopqrstuvwxyz0123456789_") for _ in range(n))
    pieces = [lambda: cjk(rng.randint(1, 30)), lambda: cjk(rng.randint(40, 70)), lambda: word(rng.randint(1, 12)),
              lambda: word(rng.randint(60, 70)), lambda: rng.choice(["，", "。", "\n", " ", "  ", "é", "ß", "😀", "Ω", "…", "Ａ", "ｂ"]),
              lambda: rng.choice(["İ", "K", "ǅ", "ſ", "ı"]), lambda: cjk(12), lambda: cjk(13), lambda: cjk(24) + word(3), lambda: "A" + cjk(2) + "Z9_"]
    weights = [30, 2, 14, 1, 30, 1, 4, 4, 4, 4]
    texts = []
    for target in [383, 384, 385, 500, 1000, 2030, 2047, 2048, 2049, 2050, 2111, 2112, 2113, 3000, 4096, 4097, 6200, 9000] * 3:
        t = ""
        while len(t.encode("utf-8")) < target:
            t += rng.choices(pieces, weights)[0]()
        texts.append(t)
    # exact byte lengths around the tile / chunk boundaries (ASCII filler keeps the byte count exact)
    for nbytes in (2048, 2048 + 64, 4096, 2048 - 16, 2048 * 2 + 63):
        t = cjk(300)
        b = len(t.encode("utf-8"))
        texts.append(t + " " + "x" * 3 + " " + "y" * (nbytes - b - 5 - 6) + " " + cjk(1) + "ab")
    texts.append(cjk(2000))                         # one run, far longer than the staged history
    texts.append((word(5) + " ") * 500)             # ASCII only
    texts.append(("😀" + cjk(3)) * 300)             # 4-byte characters between short runs
    return texts, rng, han




def _pack(claims, facts, sco, sfo, fq, fd, fh, cp, ct):
    from deep_vision_b200 import packing
    cu, co = packing.pack_texts(claims)
    fu, fo = packing.pack_texts(facts)
    return (cu, co, fu, fo, np.asarray(sco, np.int64), np.asarray(sfo, np.int64), np.asarray(fq, np.float64),
            np.asarray(fd, np.int32), np.asarray(fh, np.uint8), np.asarray(cp, np.int32), np.asarray(ct, np.float64))


def synthetic(n_sessions, turns, claims, seed=None, **kw):
    from deep_vision_b200 import synth
    c = synth.make_sessions(n_sessions, turns=turns, claims_per_session=claims, seed=(11 + n_sessions) if seed is None else seed, **kw)
    return (c.claim_utf8, c.claim_off, c.fact_utf8, c.fact_off, c.sess_claim_off, c.sess_fact_off,
            c.fact_quality, c.fact_dim, c.fact_hard, c.claim_pref_dim, c.claim_threshold)


def long_facts(n_sess=3):
    """long facts: tile/chunk boundaries, long runs, exotic case mappings, long ASCII tokens, 2-/4-byte characters"""
    facts, rng, han = adversarial_long_texts()
    claims = []
    for f in facts[::2]:
        i = rng.randrange(max(1, len(f) - 60))
        claims.append(f[i:i + rng.randint(5, 60)])
    claims += ["".join(rng.choice(han) for _ in range(20)) for _ in range(6)]
    claims = claims[:32]                                   # one claim group
    facts = facts + ["\u3002".join(claims), " ".join(claims) * 2]  # > 600 distinct hit slots in one fact
    nf, nc = len(facts) * n_sess, len(claims) * n_sess
    return _pack(claims * n_sess, facts * n_sess, np.arange(n_sess + 1) * len(claims), np.arange(n_sess + 1) * len(facts),
                 np.linspace(0.05, 0.95, nf), np.arange(nf) % 4, np.arange(nf) % 3 == 0, np.arange(nc) % 5 - 1, np.full(nc, 0.46))


def every_pair():
    """one fact per session, so best_score exposes the overlap of EVERY (claim, fact) pair: piece boundaries
    (12/13/14/24/25/36/37-character runs), ASCII tokens across cells, every text alignment, 64/65-character ASCII tokens,
    U+212A / U+0130 (hand-back) and E2 80 xx punctuation"""
    facts, rng, han = adversarial_long_texts()
    def cjk(n): return "".join(rng.choice(han) for _ in range(n))
    for run in (11, 12, 13, 14, 23, 24, 25, 26, 35, 36, 37, 38, 48, 49, 50, 96, 97):
        for pre in range(0, 9):                              # shifts the run across the lane units
            facts.append("x" * pre + "\uff0c" + cjk(run) + "\u3002" + cjk(40) + " tail_token9 " + cjk(run) + "\uff0cab" + cjk(30) + "c" + cjk(12) + "z" * 63)
    facts.append(cjk(70) + " " + "q" * 64 + " " + cjk(70) + " " + "r" * 65 + cjk(5))
    facts.append(cjk(100) + "\u212a" + cjk(100))                # U+212A
    facts.append(cjk(100) + "\u0130" + cjk(100))                # U+0130
    facts.append("\u201c" + cjk(90) + "\u201d\u2014\u2014" + cjk(90) + "\u2026")    # E2 80 xx punctuation must NOT hand back
    facts.append("ab " * 100)                               # no bigram at all: distinct count by the exact path
    facts.append((cjk(2) + "\uff0c") * 80)                      # 2-character runs only
    facts.append((cjk(1) + " ") * 150)                      # single characters: no token
    facts.append(cjk(7) * 40)                               # few distinct bigrams repeated
    claims = []
    for f in facts[::3]:
        i = rng.randrange(max(1, len(f) - 60))
        claims.append(f[i:i + rng.randint(5, 60)])
    claims = claims[-26:] + [cjk(12), cjk(13), "tail_token9 ab c", "q" * 64, "r" * 65, "k i"]
    assert len(claims) == 32
    n_sess = len(facts)
    nf, nc = n_sess, len(claims) * n_sess
    return _pack(claims * n_sess, facts, np.arange(n_sess + 1) * len(claims), np.arange(n_sess + 1),
                 np.full(nf, 0.5), np.zeros(nf), np.zeros(nf), np.full(nc, -1), np.full(nc, 0.46))


def token_table_overflow():
    """> 704 distinct ASCII tokens in one claim group: the fast kernel hands the session back"""
    import random
    rng = random.Random(9)
    abc = "abcdefghijklmnopqrstuvwxyz0123456789_"
    words = sorted({"".join(rng.choice(abc) for _ in range(3)) for _ in range(3000)})
    claims = [" ".join(rng.sample(words, 60)) for _ in range(25)]
    facts = [" ".join(rng.sample(words, 300)) + " \u4e2d\u6587\u6d4b\u8bd5\u6570\u636e" for _ in range(12)]
    return _pack(claims, facts, [0, len(claims)], [0, len(facts)], np.linspace(0.1, 0.9, len(facts)), np.zeros(len(facts)),
                 np.zeros(len(facts)), np.full(len(claims), -1), np.full(len(claims), 0.46))


def handback_in_one_group():
    """Sessions of several claim groups (the unit of work of the fast kernels) in which ONE group is handed back: session 0 by
    the claim kernel (U+0130 in a claim of its third group), session 1 by the fact kernel (U+212A in a fact), session 2 not at
    all.  The whole session goes to the general kernel; the groups the fast kernels did score are rewritten by it."""
    import random
    rng = random.Random(77)
    han = [chr(0x4E00 + rng.randrange(2000)) for _ in range(300)]
    def cjk(n): return "".join(rng.choice(han) for _ in range(n))
    claims, facts, sco, sfo = [], [], [0], [0]
    for s_ in range(3):
        fs = [cjk(rng.randint(20, 200)) + " abc" + str(rng.randrange(50)) for _ in range(30)]
        cs = []
        for _ in range(100):                               # four groups: 32 + 32 + 32 + 4
            f = rng.choice(fs); i = rng.randrange(len(f) - 12)
            cs.append(f[i:i + rng.randint(4, 12)] + cjk(3))
        if s_ == 0: cs[70] = cs[70] + "\u0130"
        if s_ == 1: fs[11] = fs[11][:9] + "\u212a" + fs[11][9:]
        claims += cs; facts += fs; sco.append(len(claims)); sfo.append(len(facts))
    nf, nc = len(facts), len(claims)
    return _pack(claims, facts, sco, sfo, np.linspace(0.05, 0.95, nf), np.arange(nf) % 4, np.arange(nf) % 3 == 0,
                 np.arange(nc) % 5 - 1, np.full(nc, 0.3))


def many_tiny_sessions(n=4300):
    """More sessions than one step of k_bind_groups numbers (4096): sessions of one short claim and one fact, every 97th
    without claims, every 211th with 33 claims (two groups), so the running group count is carried across the steps."""
    import random
    rng = random.Random(1)
    han = [chr(0x4E00 + rng.randrange(500)) for _ in range(100)]
    def cjk(k): return "".join(rng.choice(han) for _ in range(k))
    claims, facts, sco, sfo = [], [], [0], [0]
    for s_ in range(n):
        nc = 0 if s_ % 97 == 5 else (33 if s_ % 211 == 0 else 1)
        f = cjk(8)
        for _ in range(nc):
            claims.append(f[rng.randrange(4):][:5] + cjk(1))
        facts.append(f); sco.append(len(claims)); sfo.append(len(facts))
    nf, nc = len(facts), len(claims)
    return _pack(claims, facts, sco, sfo, np.linspace(0.05, 0.95, nf), np.arange(nf) % 4, np.arange(nf) % 3 == 0,
                 np.arange(nc) % 5 - 1, np.full(nc, 0.3))


def degenerate():
    """sessions without claims / without facts; empty and white-space-only claim and fact texts"""
    claims = ["", "   ", "\u9884\u7b97\u6709\u9650 budget_2024", "\uff0c\u3002\uff1b", "a", "\u5ba1\u6279\u6d41\u7a0b\u592a\u6162\u9700\u8981\u4f18\u5316"]
    facts = ["", " \n ", "\u6211\u4eec\u7684\u9884\u7b97\u6709\u9650\uff0cbudget_2024 \u5df2\u7ecf\u786e\u5b9a\u3002\u5ba1\u6279\u6d41\u7a0b\u592a\u6162\u9700\u8981\u4f18\u5316", "a b c", "\u6d41\u7a0b"]
    # session 0: claims, no facts; 1: facts, no claims; 2: both; 3: nothing; 4: both again
    nf, nc = 15, 18
    return _pack(claims * 3, facts * 3, [0, 6, 6, 12, 12, 18], [0, 0, 5, 10, 10, 15], np.linspace(0.0, 1.0, nf), np.arange(nf) % 3,
                 np.arange(nf) % 2, np.arange(nc) % 4 - 1, np.full(nc, 0.3))


def mixed_real_text():
    """short question/answer facts the way a live session looks (config 1 style): short facts go through the exact
    distinct-token count, punctuation from the E2 80 block, full-width forms, Latin-1 and astral characters"""
    import random
    rng = random.Random(77)
    han = [chr(0x4E00 + rng.randrange(2500)) for _ in range(300)]
    def cjk(n): return "".join(rng.choice(han) for _ in range(n))
    facts = []
    for i in range(60):
        q = cjk(rng.randint(8, 30)) + "\uff1f"
        a = rng.choice(["\u662f", "\u5426", cjk(rng.randint(1, 60)), "%d\u4eba" % rng.randrange(5000), "API v%d.%d" % (rng.randrange(9), rng.randrange(20)),
                        cjk(5) + "\u00b7" + cjk(4) + " caf\u00e9 \U0001f600 " + cjk(3), cjk(13) + "\n" + cjk(25) + "\u2014\u2014" + cjk(2), ""])
        facts.append(q + " " + a + " \u5ba2\u6237\u9700\u6c42 customer_needs")
    claims = []
    for i in range(40):
        f = rng.choice(facts)
        j = rng.randrange(max(1, len(f) - 12))
        claims.append(cjk(rng.randint(0, 10)) + f[j:j + rng.randint(3, 25)] + rng.choice(["", " API", " v2", cjk(4)]))
    n_sess = 4
    nf, nc = 15 * n_sess, 10 * n_sess
    return _pack(claims, facts, np.arange(n_sess + 1) * 10, np.arange(n_sess + 1) * 15, np.round(np.linspace(0, 1, nf), 2), np.arange(nf) % 4,
                 np.arange(nf) % 5 == 0, np.arange(nc) % 5 - 1, np.full(nc, 0.46))


def big_claim_groups():
    """claim groups whose token ranges exceed the fast kernel's flag array -> handed back to the general kernel"""
    return synthetic(2, 6, 30, seed=5, claim_chars=4000, copy_frac=0.05)


def _bucket_mates(n_buckets_bits=11, want=48, seed=5):
    """CJK bigrams whose ids land in ONE bucket of the fast kernel's bigram table (bits 19..29 of id * 0x9E3779B1)"""
    import random
    rng = random.Random(seed)
    target, out = None, []
    while len(out) < want:
        a, b = rng.randrange(1, 20993), rng.randrange(1, 20993)
        h = (((a << 15) | b) * 0x9E3779B1) & 0xFFFFFFFF
        bk = (h >> 19) & ((1 << n_buckets_bits) - 1)
        if target is None:
            target = bk
        if bk == target:
            out.append(chr(0x4E00 + a - 1) + chr(0x4E00 + b - 1))
    return out


def crowded_bucket(n_keys):
    """n_keys claim bigrams in one table bucket: beyond four the keys live in the side list (exact), beyond 4 + 32 the
    session is handed back.  Facts hold some of them (hits through the side list), near misses and fillers."""
    mates = _bucket_mates(want=n_keys + 8)
    inb, out = mates[:n_keys], mates[n_keys:]
    claims = ["，".join(inb[i::6]) for i in range(6)] + ["，".join(inb[:5]), inb[-1] + "，" + inb[0]]
    facts = ["。".join(inb[::2]), "。".join(inb[1::2] + out), "。".join(out), "，".join(inb), inb[-1], inb[3] + " " + inb[4] + " " + out[0]]
    nc, nf = len(claims), len(facts)
    return _pack(claims * 2, facts * 2, [0, nc, 2 * nc], [0, nf, 2 * nf], np.linspace(0.2, 0.8, 2 * nf), np.arange(2 * nf) % 3,
                 np.arange(2 * nf) % 2, np.arange(2 * nc) % 4 - 1, np.full(2 * nc, 0.3))
