"""Rows a8-a10: question-similarity gate and solution Jaccard."""
import numpy as np
import pytest


def test_default_tables_match_reference_constants(golden):
    from deep_vision_b200 import question_similarity as qs
    g = golden("a8_a9_gate.json")
    assert [list(x) for x in qs.DEFAULT_SYNONYMS] == g["synonyms"]
    assert sorted(qs.DEFAULT_STOPWORDS) == sorted(g["stopwords"])


def test_clean_solution_text_host(golden):
    from deep_vision_b200 import question_similarity as qs
    for case in golden("a10_solution_jaccard.json"):
        assert qs.clean_solution_text(case["left"]) == case["left_clean"]
        assert qs.clean_solution_text(case["right"]) == case["right_clean"]


@pytest.mark.gpu
def test_gpu_gate_golden(gpu_ctx, golden):
    from deep_vision_b200 import question_similarity as qs
    g = golden("a8_a9_gate.json")
    gate = qs.QuestionGate([tuple(x) for x in g["synonyms"]], g["stopwords"], ctx=gpu_ctx)
    texts = [q for q, _ in g["normalized"]]
    keys, norms = gate.keys(texts, want_norm=True)
    for (q, want), got in zip(g["normalized"], norms):
        assert got == want, repr(q)
    for q, toks in g["token_sets"]:
        assert sorted(gate.question_token_set(q)) == toks
    assert gate.match(g["candidates"], g["history"], g["threshold"]).tolist() == g["match"]
    for a, b, want in g["pairs"]:
        assert gate.is_similar_interview_question(a, b) is want
    # reference known-answer test (tests/test_question_fast_strategy.py:1854-1872)
    assert gate.is_similar_interview_question("当前最核心的业务痛点是什么？", "目前最大的业务问题主要是什么？")
    assert gate.is_similar_interview_question("当前最让你们头疼的痛点是什么？", "当前最影响推进的痛点是什么？")
    assert not gate.is_similar_interview_question("当前最核心的业务痛点是什么？", "需要对接哪些外部系统？")
    logs = [{"question": h, "i": i} for i, h in enumerate(g["history"])]
    for c, m in list(zip(g["candidates"], g["match"]))[:40]:
        r = gate.find_similar_interview_question_log(logs, c, g["threshold"])
        assert (r is None and (m < 0 or not c.strip())) or (r is logs[m])


@pytest.mark.gpu
def test_gpu_gate_vs_oracle_random(gpu_ctx, oracle):
    from deep_vision_b200 import question_similarity as qs
    rng = np.random.default_rng(9)
    words = ["当前", "目前", "主要", "最", "核心", "请", "你", "们", "痛点", "诉求", "最关键", "最大的", "头疼", "影响推进", "业务", "系统", "方案",
             "A", "b", "？", " ", "，", "的", "一", "个", "下", "是什么", "哪些", "方面", "对接", "外部", "流程", "İ", "ΑΣ", "Σ", "　", "ǅ", "😀"]
    def mk(n):
        return ["".join(rng.choice(words, size=int(rng.integers(0, 14)))) for _ in range(n)]
    hist, cands = mk(700), mk(900)
    cands[:100] = [h + "？" for h in hist[:100]]
    stop = list(reversed(qs.DEFAULT_STOPWORDS))
    gate = qs.QuestionGate(qs.DEFAULT_SYNONYMS + [("系统", "平台系统"), ("", "x")], stop, ctx=gpu_ctx)       # growth > 1, empty source
    og = oracle.Gate(gate.synonyms, stop)
    keys, norms = gate.keys(cands, want_norm=True)
    for c, k, nm in zip(cands, keys, norms):
        assert nm == og.normalize(c), repr(c)
        assert k.tolist() == og.keys(c).tolist()
    for thr in (0.62, 0.3, 1.0, 0.0):
        assert gate.match(cands, hist, thr).tolist() == og.gate(cands, hist, thr).tolist()
    assert gate.match(cands[:5], [], 0.62).tolist() == [-1] * 5


@pytest.mark.gpu
def test_gpu_gate_rejects_overlong_question(gpu_ctx):
    """the 60000-byte limit per question is checked on the device since the capacities are computed there (k_gate_caps)"""
    from deep_vision_b200 import _lib, question_similarity as qs
    gate = qs.QuestionGate(ctx=gpu_ctx)
    with pytest.raises(_lib.DvbError):
        gate.match(["短问题", "长" * 20001], ["历史问题"], 0.62)
    assert gate.match(["历史问题？"], ["历史问题"], 0.62).tolist() == [0]          # the context is usable afterwards


@pytest.mark.gpu
def test_gpu_gate_astral_synonym_takes_32bit_scratch(gpu_ctx, oracle):
    """k_question_keys keeps BMP questions as 16-bit code points in shared memory; a synonym that WRITES an astral character
    (and any question holding one) must take the 32-bit global scratch — same keys and normalised text as the oracle."""
    from deep_vision_b200 import question_similarity as qs
    syn = [("痛点", "😀问题"), ("系统", "平台")]
    gate = qs.QuestionGate(syn, ["请", "你"], ctx=gpu_ctx)
    og = oracle.Gate(gate.synonyms, gate.stopwords)
    cands = ["请问当前系统的痛点是什么", "痛点痛点", "没有替换的问题", "😀" * 3 + "系统", "", "A痛点b", "\U00010400大写𐐀"]
    keys, norms = gate.keys(cands, want_norm=True)
    for c, k, nm in zip(cands, keys, norms):
        assert nm == og.normalize(c), repr(c)
        assert k.tolist() == og.keys(c).tolist()
    assert gate.match(cands, cands, 0.62).tolist() == og.gate(cands, cands, 0.62).tolist()


@pytest.mark.gpu
def test_gpu_gate_prefix_filter_equals_dense_counters(gpu_ctx, oracle, monkeypatch):
    """threshold > 0 runs k_gate_match_prefix (rarest-keys candidate generation + exact verification); DVB_GATE_KERNEL=dense
    forces the dense-counter kernel.  Same answers on near-duplicates, repeated history, very long questions
    (> 128 keys: flagged to the dense kernel) and thresholds from 0.01 to above 1; a sample also against the oracle."""
    from deep_vision_b200 import question_similarity as qs
    rng = np.random.default_rng(21)
    han = [chr(0x4E00 + int(x)) for x in rng.integers(0, 400, size=4000)]
    def q(n):
        i = int(rng.integers(0, len(han) - n))
        return "".join(han[i:i + n])
    hist = [q(int(rng.integers(1, 45))) for _ in range(3000)]
    hist += hist[:200]                                              # repeated history: the LAST index must win
    hist += [q(300), q(200) + "abc" * 40]
    cands = []
    for i in range(6000):
        r = rng.random()
        if r < 0.3:
            h = hist[int(rng.integers(0, len(hist)))]
            k = int(rng.integers(0, max(1, len(h))))
            cands.append(h[:k] + q(int(rng.integers(0, 4))) + h[k + int(rng.integers(0, 3)):])      # small edit of a history item
        elif r < 0.35:
            cands.append(hist[-1][5:] if r < 0.32 else hist[-2][: 250])                              # > 128 keys
        else:
            cands.append(q(int(rng.integers(0, 45))))
    gate = qs.QuestionGate(ctx=gpu_ctx)
    og = oracle.Gate(gate.synonyms, gate.stopwords)
    for thr in (0.62, 0.01, 0.2, 0.5, 0.9, 1.0, 1.5):
        fast = gate.match(cands, hist, thr)
        monkeypatch.setenv("DVB_GATE_KERNEL", "dense")
        dense = gate.match(cands, hist, thr)
        monkeypatch.delenv("DVB_GATE_KERNEL")
        assert np.array_equal(fast, dense), thr
        assert fast[:400].tolist() == og.gate(cands[:400], hist, thr).tolist()
    assert int((gate.match(cands, hist, 0.62) >= 0).sum()) > 500


@pytest.mark.gpu
def test_gpu_solution_jaccard_golden(gpu_ctx, golden, oracle):
    from deep_vision_b200 import question_similarity as qs
    cases = golden("a10_solution_jaccard.json")
    for case in cases:
        assert qs._solution_jaccard_similarity(case["left"], case["right"], ctx=gpu_ctx) == case["similarity"]
    pairs = [(c["left_clean"], c["right_clean"]) for c in cases]
    for n in (1, 2, 3, 4):
        inter, union, nl, nr = qs.solution_jaccard_counts(pairs, n, ctx=gpu_ctx)
        for (l, r), i, u in zip(pairs, inter, union):
            _j, oi, ou = oracle.solution_jaccard(l, r, n)
            assert (int(i), int(u)) == (oi, ou), (l, r, n)


@pytest.mark.gpu
def test_gpu_gate_config4_history_size_slice_vs_oracle(gpu_ctx, oracle):
    """Parity at scale: the FULL 50 k-question history of config 4, 100 k candidates (2 % of them a history question with one
    edit); a slice of the candidates against the C oracle over the whole history, and every planted near-duplicate must
    come back with a history index at least as recent as its source."""
    from deep_vision_b200 import question_similarity as qs, synth
    rng = np.random.default_rng(44)
    hd, ho = synth.make_texts(50000, 30, seed=31)
    hraw = hd.tobytes()
    hist = [hraw[ho[i]:ho[i + 1]].decode("utf-8") + "？" for i in range(50000)]
    cd, co = synth.make_texts(100000, 30, seed=1000)
    craw = cd.tobytes()
    cands = [craw[co[i]:co[i + 1]].decode("utf-8") + "？" for i in range(100000)]
    planted = {}
    for i in np.nonzero(rng.random(100000) < 0.02)[0].tolist():
        h = int(rng.integers(0, 50000))
        cands[i] = "请问" + hist[h][1:]                      # stop word in front, first character dropped
        planted[i] = h
    gate = qs.QuestionGate(ctx=gpu_ctx)
    got = gate.match(cands, hist, 0.62)
    og = oracle.Gate(gate.synonyms, gate.stopwords)
    sl = sorted(planted)[:60] + list(range(0, 90))
    want = og.gate([cands[i] for i in sl], hist, 0.62)
    assert got[sl].tolist() == want.tolist()
    hit = [i for i, h in planted.items() if got[i] >= h]
    assert len(hit) >= 0.95 * len(planted), (len(hit), len(planted))


@pytest.mark.gpu
def test_gpu_question_keys_survive_raw_bytes_that_are_not_utf8(gpu_ctx):
    """The C ABI accepts raw bytes: a question of continuation bytes only decodes (by the kernel's rule) to many more
    code points than it has lead bytes; the shared-memory scratch must be chosen by what the decoder produces."""
    from deep_vision_b200 import _lib, question_similarity as qs
    from deep_vision_b200.packing import ptr
    gate = qs.QuestionGate(ctx=gpu_ctx)
    blobs = [bytes([0x80 + (i % 64) for i in range(n)]) for n in (256, 300, 2000, 63, 64, 65, 130)] + ["正常的问题文本？".encode("utf-8")]
    data = np.frombuffer(b"".join(blobs) + b"\0" * 64, np.uint8).copy()
    off = np.zeros(len(blobs) + 1, np.int64); off[1:] = np.cumsum([len(b) for b in blobs])
    n = len(blobs)
    cap = int(((off[1:] - off[:-1]) + 2).sum() * gate._growth)
    cap_off = np.zeros(n + 1, np.int64); keys = np.zeros(cap, np.uint64); cnt = np.zeros(n, np.int32)
    for _ in range(3):
        rc = gpu_ctx.lib.dvb_question_keys(gpu_ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), n, *gate._table_args(),
                                           ptr(cap_off), ptr(keys), ptr(cnt), None, None, cap)
        gpu_ctx.check(rc, "dvb_question_keys")
    assert int(cnt[-1]) == len(gate.keys(["正常的问题文本？"])[0]) > 0           # the context is still healthy and the valid text unaffected
    assert all(int(c) >= 0 for c in cnt)
