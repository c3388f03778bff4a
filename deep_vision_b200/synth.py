"""Deterministic synthetic Chinese-text corpora for the BASELINE.json configs.

Generator: PCG64(seed); code points U+4E00..U+4E00+3499 drawn Zipf(s=1.0) by rank; every 17th
position is one of "，。；\\n"; ~5 % of positions start an ASCII word [a-z]{3,8} or a 1-4 digit
number followed by a space.  Everything is produced as packed CSR UTF-8 (numpy), never as
Python strings, so the 0.9 GB config-2 corpus builds in seconds.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import numpy as np

SEED = 20260101
ALPHABET = 3500
PUNCT = np.array([0xFF0C, 0x3002, 0xFF1B, 0x0A], dtype=np.uint32)
DIMENSIONS = ["customer_needs", "business_process", "tech_constraints", "project_constraints"]
DIMENSION_NAMES = ["客户需求", "业务流程", "技术约束", "项目约束"]


def _zipf_cdf(n: int) -> np.ndarray:
    w = 1.0 / np.arange(1, n + 1, dtype=np.float64)
    return np.cumsum(w / w.sum())


_CDF = _zipf_cdf(ALPHABET)
_LUT_BITS = 22
# inverse-CDF lookup table: rank for each of 2^22 equally spaced quantiles (exact Zipf up to 2^-22)
_LUT = np.minimum(np.searchsorted(_CDF, (np.arange(1 << _LUT_BITS) + 0.5) / (1 << _LUT_BITS), side="right"),
                  ALPHABET - 1).astype(np.uint32)


def code_points(rng: np.random.Generator, n: int) -> np.ndarray:
    """n code points of synthetic text (uint32)."""
    cp = _LUT[rng.integers(0, 1 << _LUT_BITS, size=n, dtype=np.uint32)]
    cp += np.uint32(0x4E00)
    pos = np.arange(16, n, 17)
    cp[pos] = PUNCT[rng.integers(0, 4, size=len(pos))]
    # ASCII words / numbers
    starts = np.nonzero(rng.random(n) < 0.05 / 5.0)[0]          # ~5 % of positions end up ASCII (avg len 5)
    if len(starts):
        is_num = rng.random(len(starts)) < 0.3
        lens = np.where(is_num, rng.integers(1, 5, size=len(starts)), rng.integers(3, 9, size=len(starts)))
        maxlen = 9
        idx = starts[:, None] + np.arange(maxlen)[None, :]
        letters = rng.integers(ord("a"), ord("z") + 1, size=idx.shape).astype(np.uint32)
        digits = rng.integers(ord("0"), ord("9") + 1, size=idx.shape).astype(np.uint32)
        chars = np.where(is_num[:, None], digits, letters)
        col = np.arange(maxlen)[None, :]
        chars = np.where(col == lens[:, None], np.uint32(0x20), chars)       # trailing space
        valid = (col <= lens[:, None]) & (idx < n)
        cp[idx[valid]] = chars[valid]
    return cp


def utf8_encode(cp: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Vectorised UTF-8 encoding of BMP code points. -> (bytes uint8 + 64 pad, byte offset of each cp [n+1])"""
    cp = cp.astype(np.uint32, copy=False)
    n = len(cp)
    nb = np.ones(n, dtype=np.int8)
    nb[cp >= 0x80] = 2
    nb[cp >= 0x800] = 3
    off = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(nb, out=off[1:])
    planes = np.empty((n, 3), dtype=np.uint8)
    planes[:, 0] = np.where(nb == 3, 0xE0 | (cp >> 12), np.where(nb == 2, 0xC0 | (cp >> 6), cp))
    planes[:, 1] = np.where(nb == 3, 0x80 | ((cp >> 6) & 0x3F), 0x80 | (cp & 0x3F))
    planes[:, 2] = 0x80 | (cp & 0x3F)
    valid = np.empty((n, 3), dtype=bool)
    valid[:, 0] = True
    valid[:, 1] = nb >= 2
    valid[:, 2] = nb == 3
    out = np.zeros(int(off[-1]) + 64, dtype=np.uint8)
    out[:int(off[-1])] = planes[valid]
    return out, off


@dataclass
class SessionCorpus:
    """Packed config-2 style workload: S sessions x T facts x C claims."""
    claim_utf8: np.ndarray
    claim_off: np.ndarray
    fact_utf8: np.ndarray
    fact_off: np.ndarray
    sess_claim_off: np.ndarray
    sess_fact_off: np.ndarray
    fact_quality: np.ndarray
    fact_dim: np.ndarray
    fact_hard: np.ndarray
    claim_pref_dim: np.ndarray
    claim_threshold: np.ndarray

    @property
    def n_sessions(self) -> int:
        return len(self.sess_claim_off) - 1

    @property
    def n_pairs(self) -> int:
        nc = np.diff(self.sess_claim_off)
        nf = np.diff(self.sess_fact_off)
        return int((nc * nf).sum())

    @property
    def text_bytes(self) -> int:
        return int(self.claim_off[-1] + self.fact_off[-1])

    @property
    def algorithmic_bytes(self) -> int:
        """every claim and fact byte read once + 16 B out per claim."""
        return self.text_bytes + 16 * int(self.sess_claim_off[-1])

    @property
    def max_claim_group_bytes(self) -> int:
        mx = 0
        co, sco = self.claim_off, self.sess_claim_off
        for s in range(self.n_sessions):
            for c in range(int(sco[s]), int(sco[s + 1]), 32):
                e = min(c + 32, int(sco[s + 1]))
                mx = max(mx, int(co[e] - co[c]))
        return mx

    def texts(self):
        """(claims, facts) as Python strings — small corpora only (tests)."""
        cb, fb = self.claim_utf8.tobytes(), self.fact_utf8.tobytes()
        claims = [cb[self.claim_off[i]:self.claim_off[i + 1]].decode("utf-8") for i in range(len(self.claim_off) - 1)]
        facts = [fb[self.fact_off[i]:self.fact_off[i + 1]].decode("utf-8") for i in range(len(self.fact_off) - 1)]
        return claims, facts


def make_sessions(n_sessions: int, turns: int = 40, claims_per_session: int = 24, fact_chars: int = 690,
                  claim_chars: int = 80, copy_frac: float = 0.4, seed: int = SEED) -> SessionCorpus:
    """Config 2: per turn one fact of ~2 KB UTF-8 (question 40 chars + '？' + answer + dimension names);
    per session `claims_per_session` claims of ~`claim_chars`, `copy_frac` of each copied from a random
    fact of the same session (non-trivial overlaps, rare ties)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n_facts = n_sessions * turns
    n_claims = n_sessions * claims_per_session
    body = code_points(rng, n_facts * fact_chars).reshape(n_facts, fact_chars)
    body[:, 40] = 0xFF1F                                                       # '？' ends the question
    dims = rng.integers(0, 4, size=n_facts).astype(np.int32)
    suffix = np.zeros((4, 32), dtype=np.uint32)
    suffix_len = np.zeros(4, dtype=np.int64)
    for d in range(4):
        s = " " + DIMENSION_NAMES[d] + " " + DIMENSIONS[d]
        suffix[d, :len(s)] = [ord(c) for c in s]
        suffix_len[d] = len(s)
    fact_cp = np.concatenate([body, suffix[dims]], axis=1)
    keep = np.concatenate([np.ones((n_facts, fact_chars), bool), np.arange(32)[None, :] < suffix_len[dims][:, None]], axis=1)
    flat = fact_cp[keep]
    fact_len_cp = keep.sum(axis=1)
    fact_utf8, cp_off = utf8_encode(flat)
    fact_cp_off = np.zeros(n_facts + 1, dtype=np.int64)
    np.cumsum(fact_len_cp, out=fact_cp_off[1:])
    fact_off = cp_off[fact_cp_off]

    n_copy = int(claim_chars * copy_frac)
    fresh = code_points(rng, n_claims * (claim_chars - n_copy)).reshape(n_claims, claim_chars - n_copy)
    src_fact = (np.arange(n_claims) // claims_per_session) * turns + rng.integers(0, turns, size=n_claims)
    src_pos = rng.integers(41, fact_chars - n_copy, size=n_claims)
    copied = body[src_fact[:, None], src_pos[:, None] + np.arange(n_copy)[None, :]]
    claim_cp = np.concatenate([fresh[:, : (claim_chars - n_copy) // 2], copied, fresh[:, (claim_chars - n_copy) // 2:]], axis=1)
    claim_utf8, ccp_off = utf8_encode(claim_cp.reshape(-1))
    claim_off = ccp_off[np.arange(n_claims + 1) * claim_chars]

    quality = np.round(rng.random(n_facts), 2)
    hard = (rng.random(n_facts) < 0.1).astype(np.uint8)
    pref = np.where(rng.random(n_claims) < 0.5, rng.integers(0, 4, size=n_claims), -1).astype(np.int32)
    thr = np.where(pref >= 0, 0.46 - 0.03, 0.46) + np.where(np.arange(n_claims) % claims_per_session // 8 == 1, 0.04, 0.0)
    return SessionCorpus(
        claim_utf8=claim_utf8, claim_off=claim_off.astype(np.int64), fact_utf8=fact_utf8, fact_off=fact_off.astype(np.int64),
        sess_claim_off=(np.arange(n_sessions + 1) * claims_per_session).astype(np.int64),
        sess_fact_off=(np.arange(n_sessions + 1) * turns).astype(np.int64),
        fact_quality=quality.astype(np.float64), fact_dim=dims, fact_hard=hard,
        claim_pref_dim=pref, claim_threshold=thr.astype(np.float64))


def make_texts(n: int, chars: int, seed: int = SEED) -> Tuple[np.ndarray, np.ndarray]:
    """n texts of `chars` code points each, packed (config 3 style chunks)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cp = code_points(rng, n * chars)
    data, cp_off = utf8_encode(cp)
    return data, cp_off[np.arange(n + 1) * chars].astype(np.int64)
