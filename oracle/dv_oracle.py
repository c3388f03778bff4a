"""ctypes wrapper over the CPU oracle (oracle/dv_oracle.c).  TEST INFRASTRUCTURE ONLY.

Importers allowed: tests/, __graft_entry__.smoke(), bench.py (cpu_baseline / --impl reference).
The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path
from typing import List, Optional, Sequence, Tuple

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "_build" / "libdvoracle.so"
_lib = None

u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
f64p = C.POINTER(C.c_double)


def build(force: bool = False) -> Path:
    srcs = list(HERE.glob("*.c")) + [HERE / "dv_oracle.h", HERE.parent / "include" / "dv_unicode_tables.h"]
    if force or not LIB_PATH.exists() or any(s.stat().st_mtime > LIB_PATH.stat().st_mtime for s in srcs):
        subprocess.run(["make", "-C", str(HERE), "-s"], check=True)
    return LIB_PATH


class GateTables(C.Structure):
    _fields_ = [("syn_src", u32p), ("syn_src_off", i32p), ("syn_dst", u32p), ("syn_dst_off", i32p), ("n_syn", C.c_int32),
                ("stop", u32p), ("stop_off", i32p), ("n_stop", C.c_int32)]


def lib():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            build()
        _lib = C.CDLL(str(LIB_PATH))
        _lib.dvo_tokenize_v3.restype = C.c_int64
        _lib.dvo_tokenize_v3_keys.restype = C.c_int64
        _lib.dvo_token_key.restype = C.c_uint64
        _lib.dvo_chunk_build.restype = C.c_int64
        _lib.dvo_query_terms.restype = C.c_int64
        _lib.dvo_topk.restype = C.c_int64
        _lib.dvo_normalize_question.restype = C.c_int64
        _lib.dvo_question_keys.restype = C.c_int64
        _lib.dvo_solution_jaccard.restype = C.c_double
        _lib.dvo_utf8_decode.restype = C.c_int64
        _lib.dvo_lower.restype = C.c_int64
    return _lib


def enc(s: str) -> bytes:
    return str(s).encode("utf-8", "surrogatepass")


def _ptr(a: np.ndarray, t):
    return a.ctypes.data_as(t)


def pack(texts: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
    """CSR pack: (uint8 bytes, int64 offsets[n+1])."""
    bs = [enc(t) for t in texts]
    off = np.zeros(len(bs) + 1, dtype=np.int64)
    if bs:
        np.cumsum([len(b) for b in bs], out=off[1:])
    data = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if off[-1] else np.zeros(1, dtype=np.uint8)
    return data, off


def lower(text: str) -> str:
    b = np.frombuffer(enc(text), dtype=np.uint8).copy() if text else np.zeros(1, np.uint8)
    n = len(enc(text))
    cps = np.zeros(n + 1, np.uint32)
    m = lib().dvo_utf8_decode(_ptr(b, u8p), C.c_int64(n), _ptr(cps, u32p))
    out = np.zeros(3 * (m + 1), np.uint32)
    k = lib().dvo_lower(_ptr(cps, u32p), C.c_int64(m), _ptr(out, u32p), C.c_int64(len(out)))
    return "".join(map(chr, out[:k].tolist()))


def tokenize_v3(text: str) -> set:
    """a1: set of token strings (server.py:24722)."""
    b = enc(text)
    n = len(b)
    buf = np.frombuffer(b, dtype=np.uint8).copy() if n else np.zeros(1, np.uint8)
    low = np.zeros(3 * (n + 1), np.uint32)
    ts = np.zeros(2 * (n + 1), np.int32)
    tl = np.zeros(2 * (n + 1), np.int32)
    nl = C.c_int64(0)
    nt = lib().dvo_tokenize_v3(_ptr(buf, u8p), C.c_int64(n), _ptr(low, u32p), C.c_int64(len(low)), C.byref(nl),
                               _ptr(ts, i32p), _ptr(tl, i32p), C.c_int64(len(ts)))
    assert nt >= 0, nt
    lowl = low.tolist()
    return {"".join(map(chr, lowl[s:s + l])) for s, l in zip(ts[:nt].tolist(), tl[:nt].tolist())}


def tokenize_v3_keys(text: str) -> np.ndarray:
    b = enc(text)
    n = len(b)
    buf = np.frombuffer(b, dtype=np.uint8).copy() if n else np.zeros(1, np.uint8)
    keys = np.zeros(2 * (n + 1), np.uint64)
    nk = lib().dvo_tokenize_v3_keys(_ptr(buf, u8p), C.c_int64(n), _ptr(keys, u64p), C.c_int64(len(keys)))
    assert nk >= 0, nk
    return keys[:nk].copy()


def token_key(token: str) -> int:
    cps = np.array([ord(c) for c in token], np.uint32)
    return int(lib().dvo_token_key(_ptr(cps, u32p), C.c_int32(len(cps))))


def bind_claims(claims: Sequence[str], facts: Sequence[str], sess_claim_off, sess_fact_off,
                fact_quality, fact_dim, fact_hard, claim_pref_dim, claim_threshold,
                want_overlap: bool = False, n_threads: int = 1, packed=None):
    """a2 core.  Returns dict(best_fact, best_score, bound, len_claim, len_fact[, overlap, pair_off])."""
    if packed is None:
        cu, co = pack(claims)
        fu, fo = pack(facts)
    else:
        cu, co, fu, fo = packed
    sco = np.ascontiguousarray(sess_claim_off, np.int64)
    sfo = np.ascontiguousarray(sess_fact_off, np.int64)
    ns = len(sco) - 1
    ncl, nfa = int(sco[-1]), int(sfo[-1])
    fq = np.ascontiguousarray(fact_quality, np.float64)
    fd = np.ascontiguousarray(fact_dim, np.int32)
    fh = np.ascontiguousarray(fact_hard, np.uint8)
    cp = np.ascontiguousarray(claim_pref_dim, np.int32)
    ct = np.ascontiguousarray(claim_threshold, np.float64)
    best = np.full(max(ncl, 1), -1, np.int32)
    score = np.zeros(max(ncl, 1), np.float64)
    bound = np.zeros(max(ncl, 1), np.uint8)
    lc = np.zeros(max(ncl, 1), np.int32)
    lf = np.zeros(max(nfa, 1), np.int32)
    ov = po = None
    if want_overlap:
        po = np.zeros(ns + 1, np.int64)
        np.cumsum((sco[1:] - sco[:-1]) * (sfo[1:] - sfo[:-1]), out=po[1:])
        ov = np.zeros(max(int(po[-1]), 1), np.int32)
    rc = lib().dvo_bind_claims(_ptr(cu, u8p), _ptr(co, i64p), _ptr(fu, u8p), _ptr(fo, i64p),
                               _ptr(sco, i64p), _ptr(sfo, i64p), C.c_int64(ns),
                               _ptr(fq, f64p), _ptr(fd, i32p), _ptr(fh, u8p), _ptr(cp, i32p), _ptr(ct, f64p),
                               _ptr(best, i32p), _ptr(score, f64p), _ptr(bound, u8p), _ptr(lc, i32p), _ptr(lf, i32p),
                               _ptr(ov, i32p) if ov is not None else None, _ptr(po, i64p) if po is not None else None,
                               C.c_int(n_threads))
    assert rc == 0, rc
    out = dict(best_fact=best[:ncl], best_score=score[:ncl], bound=bound[:ncl], len_claim=lc[:ncl], len_fact=lf[:nfa])
    if want_overlap:
        out["overlap"] = ov
        out["pair_off"] = po
    return out


def chunk_build(text: str, chunk_size: int = 1800, overlap: int = 160) -> List[Tuple[int, int, int, int]]:
    """a4: list of (start, end, text_start, text_end) code-point offsets."""
    b = enc(text)
    n = len(b)
    buf = np.frombuffer(b, dtype=np.uint8).copy() if n else np.zeros(1, np.uint8)
    cap = n + 2
    arrs = [np.zeros(cap, np.int32) for _ in range(4)]
    cnt = lib().dvo_chunk_build(_ptr(buf, u8p), C.c_int64(n), C.c_int32(chunk_size), C.c_int32(overlap),
                                *[_ptr(a, i32p) for a in arrs], C.c_int64(cap))
    assert cnt >= 0, cnt
    return list(zip(*[a[:cnt].tolist() for a in arrs]))


def query_terms(query: str) -> List[str]:
    """a6: ordered term list (<=80)."""
    b = enc(query)
    n = len(b)
    buf = np.frombuffer(b, dtype=np.uint8).copy() if n else np.zeros(1, np.uint8)
    cp = np.zeros(80 * (3 * n + 8), np.uint32)
    off = np.zeros(82, np.int32)
    nt = lib().dvo_query_terms(_ptr(buf, u8p), C.c_int64(n), _ptr(cp, u32p), C.c_int64(len(cp)), _ptr(off, i32p), C.c_int64(80))
    assert nt >= 0, nt
    cpl = cp.tolist()
    return ["".join(map(chr, cpl[off[i]:off[i + 1]])) for i in range(nt)]


def pack_terms(terms: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
    off = np.zeros(len(terms) + 1, np.int32)
    cps: List[int] = []
    for i, t in enumerate(terms):
        cps.extend(ord(c) for c in t)
        off[i + 1] = len(cps)
    return np.array(cps if cps else [0], np.uint32), off


def chunk_scores(chunks: Sequence[str], terms: Sequence[str], n_threads: int = 1, packed=None) -> np.ndarray:
    """a7 scoring loop: score[c] = sum_t lower(chunk).count(term)."""
    cu, co = pack(chunks) if packed is None else packed
    n = len(co) - 1
    tcp, toff = pack_terms(terms)
    out = np.zeros(max(n, 1), np.int32)
    lib().dvo_chunk_scores(_ptr(cu, u8p), _ptr(co, i64p), C.c_int64(n), _ptr(tcp, u32p), _ptr(toff, i32p),
                           C.c_int64(len(terms)), _ptr(out, i32p), C.c_int(n_threads))
    return out[:n]


def topk(score: np.ndarray, k: int, index: Optional[np.ndarray] = None):
    score = np.ascontiguousarray(score, np.int32)
    idx = np.ascontiguousarray(index, np.int64) if index is not None else None
    oi = np.zeros(max(k, 1), np.int64)
    os_ = np.zeros(max(k, 1), np.int32)
    m = lib().dvo_topk(_ptr(score, i32p), _ptr(idx, i64p) if idx is not None else None, C.c_int64(len(score)),
                       C.c_int32(k), _ptr(oi, i64p), _ptr(os_, i32p))
    return oi[:m].copy(), os_[:m].copy()


class Gate:
    """a8/a9 with explicit synonym / stop-word tables in the caller's iteration order."""

    def __init__(self, synonyms: Sequence[Tuple[str, str]], stopwords: Sequence[str]):
        self.synonyms = list(synonyms)
        self.stopwords = list(stopwords)
        self._ss, self._sso = pack_terms([s for s, _ in self.synonyms])
        self._sd, self._sdo = pack_terms([d for _, d in self.synonyms])
        self._st, self._sto = pack_terms(self.stopwords)
        self.tables = GateTables(_ptr(self._ss, u32p), _ptr(self._sso, i32p), _ptr(self._sd, u32p), _ptr(self._sdo, i32p),
                                 len(self.synonyms), _ptr(self._st, u32p), _ptr(self._sto, i32p), len(self.stopwords))

    def normalize(self, text: str) -> str:
        b = enc(text)
        n = len(b)
        buf = np.frombuffer(b, dtype=np.uint8).copy() if n else np.zeros(1, np.uint8)
        out = np.zeros(12 * (n + 1) + 8, np.uint32)
        k = lib().dvo_normalize_question(_ptr(buf, u8p), C.c_int64(n), C.byref(self.tables), _ptr(out, u32p), C.c_int64(len(out)))
        assert k >= 0, k
        return "".join(map(chr, out[:k].tolist()))

    def keys(self, text: str) -> np.ndarray:
        b = enc(text)
        n = len(b)
        buf = np.frombuffer(b, dtype=np.uint8).copy() if n else np.zeros(1, np.uint8)
        out = np.zeros(12 * (n + 1) + 8, np.uint64)
        k = lib().dvo_question_keys(_ptr(buf, u8p), C.c_int64(n), C.byref(self.tables), _ptr(out, u64p), C.c_int64(len(out)))
        assert k >= 0, k
        return out[:k].copy()

    def gate(self, candidates: Sequence[str], history: Sequence[str], threshold: float = 0.62,
             n_threads: int = 1, packed=None) -> np.ndarray:
        if packed is None:
            cu, co = pack(candidates)
            hu, ho = pack(history)
        else:
            cu, co, hu, ho = packed
        nc, nh = len(co) - 1, len(ho) - 1
        out = np.full(max(nc, 1), -1, np.int64)
        rc = lib().dvo_question_gate(_ptr(cu, u8p), _ptr(co, i64p), C.c_int64(nc), _ptr(hu, u8p), _ptr(ho, i64p), C.c_int64(nh),
                                     C.byref(self.tables), C.c_double(threshold), _ptr(out, i64p), C.c_int(n_threads))
        assert rc == 0
        return out[:nc]


def solution_jaccard(left_clean: str, right_clean: str, n: int = 2) -> Tuple[float, int, int]:
    """a10 on clean_solution_text()'d inputs: (unrounded J, |L&R|, |L|R|)."""
    lb, rb = enc(left_clean), enc(right_clean)
    la = np.frombuffer(lb, dtype=np.uint8).copy() if lb else np.zeros(1, np.uint8)
    ra = np.frombuffer(rb, dtype=np.uint8).copy() if rb else np.zeros(1, np.uint8)
    i, u = C.c_int32(0), C.c_int32(0)
    j = lib().dvo_solution_jaccard(_ptr(la, u8p), C.c_int64(len(lb)), _ptr(ra, u8p), C.c_int64(len(rb)), C.c_int32(n),
                                   C.byref(i), C.byref(u))
    return float(j), i.value, u.value


def claim_chunk_topk(claims: Sequence[str], chunks: Sequence[str], k: int = 4, max_exhaustive: int = 256):
    """SURVEY 8(f)3 (not in the reference; definition in deep_vision_b200/evidence.py): per claim the k best chunks under a2's
    overlap score.  Independent route: ONE session holding every claim and every chunk as a fact, per-pair overlaps and set sizes
    from the C tokenizer, the score formula in Python floats in the reference's operation order (web/server.py:24835-24843)."""
    nc, nk = len(claims), len(chunks)
    if nc == 0 or nk == 0:
        return [[] for _ in range(nc)]
    r = bind_claims(list(claims), list(chunks), [0, nc], [0, nk], np.zeros(nk), np.zeros(nk, np.int32), np.zeros(nk, np.uint8),
                    np.full(nc, -1, np.int32), np.zeros(nc), want_overlap=True)
    ov = r["overlap"][: nc * nk].reshape(nc, nk)
    la, lb = r["len_claim"], r["len_fact"]
    out = []
    for c in range(nc):
        cand = range(nk)
        if nk > max_exhaustive:
            s = chunk_scores(list(chunks), query_terms(claims[c]))
            cand = sorted(topk(s, 32)[0].tolist())
        rows = []
        for j in cand:
            o = int(ov[c, j])
            if o <= 0:
                continue
            coverage = o / max(3, min(int(la[c]), 14))
            precision = o / max(4, min(int(lb[j]), 18))
            score = 0.62 * coverage + 0.18 * precision + 0.16 * 0.0 + 0.0 + 0.0
            key = min(int(np.floor(score * float(1 << 20))), 2147483647)
            if key > 0:
                rows.append((-key, j, score))
        rows.sort()
        out.append([{"index": j, "score": sc} for (_k, j, sc) in rows[: max(1, min(int(k), 32))]])
    return out
