"""CSR packing of Python strings for the C-ABI (host-side plumbing, no scoring logic)."""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np


def encode(text) -> bytes:
    # surrogatepass keeps lone surrogates (legal in a Python str) as 3-byte sequences; they are
    # "other" symbols for every kernel, exactly as they are non-matching chars for the reference's regexes.
    return str(text).encode("utf-8", "surrogatepass")


def pack_texts(texts: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
    """-> (uint8 bytes padded by 64 zero bytes, int64 offsets[n+1])"""
    bs = [encode(t) for t in texts]
    off = np.zeros(len(bs) + 1, dtype=np.int64)
    if bs:
        np.cumsum(np.fromiter((len(b) for b in bs), dtype=np.int64, count=len(bs)), out=off[1:])
    total = int(off[-1])
    data = np.zeros(total + 64, dtype=np.uint8)
    if total:
        data[:total] = np.frombuffer(b"".join(bs), dtype=np.uint8)
    return data, off


def ptr(a: np.ndarray) -> int:
    return a.ctypes.data
