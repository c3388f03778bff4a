"""Uploaded-document chunking and top-k chunk recall (rows a4-a7 of SURVEY.md §8).

Host-side mirror of the reference interface (web/server.py):
  * `_build_reference_material_chunks(content)`                         31932
  * `_tokenize_reference_material_query(query_text)`                    32037
  * `select_reference_material_context(doc, *, query_text, max_chars, chunk_limit)`   32054
plus the batch primitives `chunk_scores` / `topk` they are built on.

On the GPU: the code-point index + window walk of the chunker (dvb_chunk_build), decode + lower +
non-overlapping substring counting (dvb_chunk_scores) and the top-k (dvb_topk).  On the host: what
the reference does once per call around those loops — query term extraction (a <=300-char string),
dict / prefix-string assembly and the `chunks.json` manifest I/O.
"""
from __future__ import annotations

import hashlib
import json
import re
from dataclasses import dataclass
from pathlib import Path
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from .packing import encode, pack_texts, ptr

_TERM = re.compile(r"[a-z0-9_]{2,}|[一-鿿]{2,}")
_WS = re.compile(r"\s+")


@dataclass
class ReferenceMaterialConfig:
    """web/server.py:31882-31889 and :858 (read at call time)."""
    chunk_size: int = 1800
    chunk_overlap: int = 160
    context_chunk_limit: int = 4
    inline_chunk_preview_limit: int = 120
    preview_limit: int = 300
    max_doc_length: int = 2000

    def normalized(self) -> "ReferenceMaterialConfig":
        size = max(800, int(self.chunk_size))
        overlap = max(0, int(self.chunk_overlap))
        if overlap >= size:
            overlap = max(0, size // 10)
        return ReferenceMaterialConfig(size, overlap, max(1, int(self.context_chunk_limit)),
                                       max(20, int(self.inline_chunk_preview_limit)), self.preview_limit, self.max_doc_length)


DEFAULT_CONFIG = ReferenceMaterialConfig()


# ----------------------------------------------------------------------------- a6 (host, per query)

def _tokenize_reference_material_query(query_text: str) -> List[str]:
    """web/server.py:32037-32051: ordered distinct terms; pure-CJK terms of >= 6 chars also
    contribute their 4-grams, then their 3-grams; at most 80 terms."""
    lowered = str(query_text or "").lower()
    terms: List[str] = []
    seen = set()

    def add(t: str):
        if t and t not in seen:
            seen.add(t)
            terms.append(t)

    for term in _TERM.findall(lowered):
        add(term)
        if len(term) >= 6 and "一" <= term[0] <= "鿿":
            for size in (4, 3):
                for i in range(len(term) - size + 1):
                    add(term[i:i + size])
    return terms[:80]


def terms_to_symbols(queries_terms: Sequence[Sequence[str]]) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Terms -> the kernels' symbol alphabet (include/dvb200.h): (term_sym u16, term_off i32, query_term_off i32)."""
    flat = [t for terms in queries_terms for t in terms]
    # one join + one encode for all terms: U+0000 cannot occur in a term, so it separates them
    cp = np.frombuffer("\x00".join(flat).encode("utf-32-le", "surrogatepass"), np.uint32)
    if cp.size and int(cp.max()) > 0xFFFF:
        raise ValueError("query terms may only contain [a-z0-9_] or CJK characters")
    sym_all = _SYM_LUT[cp]
    sep = cp == 0
    if flat and (int(np.count_nonzero(sep)) != len(flat) - 1 or bool(np.any((sym_all == 0) & ~sep))):
        raise ValueError("query terms may only contain [a-z0-9_] or CJK characters")
    sym = sym_all[~sep] if flat else np.zeros(0, np.uint16)
    toff = np.zeros(len(flat) + 1, np.int64)
    if flat:
        ends = np.flatnonzero(sep)
        toff[1:-1] = ends - np.arange(len(ends))                   # symbols before the k-th separator
        toff[-1] = sym.size
    qoff = np.zeros(len(queries_terms) + 1, np.int64)
    np.cumsum(np.fromiter(map(len, queries_terms), np.int64, len(queries_terms)), out=qoff[1:])
    return (sym if sym.size else np.zeros(1, np.uint16), toff.astype(np.int32), qoff.astype(np.int32))


def query_terms_batch(query_texts: Sequence[str], *, ctx=None, max_terms: int = 80, packed=None, device: bool = False):
    """a6 for a batch on the device (csrc/query_terms.cu): the term lists of _tokenize_reference_material_query as the symbol
    CSR (term_sym u16, term_off i32, query_term_off i32) that ChunkIndex.topk_device / dvb_chunk_index_topk read.
    device=False returns numpy arrays (host), device=True torch CUDA tensors (nothing but the texts crosses PCIe)."""
    ctx = ctx or _lib.Context.get()
    data, off = packed if packed is not None else pack_texts([str(t or "") for t in query_texts])
    nq = len(off) - 1
    nbytes = int(off[-1] - off[0])
    cap = 4 * nbytes + 64                       # symbols <= bytes per term occurrence; grams of a run cost < 7 symbols per character
    totals = np.zeros(2, np.int64)
    for _attempt in range(2):
        if device:
            import torch
            dev = torch.device("cuda", ctx.device)
            d_data = torch.from_numpy(np.ascontiguousarray(data)).to(dev); d_off = torch.from_numpy(np.ascontiguousarray(off)).to(dev)
            tsym = torch.empty(max(cap, 1), dtype=torch.int16, device=dev)
            toff = torch.empty(nq * max_terms + 1, dtype=torch.int32, device=dev)
            qoff = torch.empty(nq + 1, dtype=torch.int32, device=dev)
            torch.cuda.current_stream(dev).synchronize()
            rc = ctx.lib.dvb_query_terms(ctx.handle, _lib.MEM_DEVICE, d_data.data_ptr(), d_off.data_ptr(), nq, max_terms,
                                         tsym.data_ptr(), cap, toff.data_ptr(), qoff.data_ptr(), ptr(totals))
        else:
            tsym = np.zeros(max(cap, 1), np.uint16); toff = np.zeros(nq * max_terms + 1, np.int32); qoff = np.zeros(nq + 1, np.int32)
            rc = ctx.lib.dvb_query_terms(ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), nq, max_terms, ptr(tsym), cap, ptr(toff), ptr(qoff), ptr(totals))
        if rc == _lib.DVB_E_CAPACITY and _attempt == 0:
            cap = 8 * nbytes * 4 + 1024                            # generous second try (every character in four 4-grams and three 3-grams)
            continue
        ctx.check(rc, "dvb_query_terms")
        break
    n_terms, n_sym = int(totals[0]), int(totals[1])
    return tsym[:max(n_sym, 1)], toff[:n_terms + 1], qoff


def terms_from_symbols(tsym, toff, qoff) -> List[List[str]]:
    """Inverse of terms_to_symbols for host arrays (tests, debugging): symbol CSR -> per-query term strings."""
    tsym = np.asarray(tsym).astype(np.uint16); toff = np.asarray(toff); qoff = np.asarray(qoff)
    cp = np.where(tsym >= 0x8000, 0, tsym.astype(np.int64) + 0x4E00 - 1)
    a = (tsym & 0x7FFF).astype(np.int64)
    asc = np.where(a <= 26, a + 96, np.where(a <= 36, a + 21, 95))
    cp = np.where(tsym >= 0x8000, asc, cp)
    chars = "".join(map(chr, cp.tolist()))
    return [[chars[toff[t]:toff[t + 1]] for t in range(qoff[q], qoff[q + 1])] for q in range(len(qoff) - 1)]


def _make_sym_lut() -> np.ndarray:
    """code point (BMP) -> kernel symbol (include/dvb200.h): CJK 1..20992, 0x8000 | class for [a-z0-9_], 0 = not a term character."""
    lut = np.zeros(0x10000, np.uint16)
    lut[0x4E00:0xA000] = np.arange(1, 20993, dtype=np.uint16)
    lut[97:123] = 0x8000 | np.arange(1, 27, dtype=np.uint16)
    lut[48:58] = 0x8000 | np.arange(27, 37, dtype=np.uint16)
    lut[95] = 0x8000 | 37
    return lut


_SYM_LUT = _make_sym_lut()


# ----------------------------------------------------------------------------- GPU primitives

def chunk_scores(chunk_texts: Sequence[str], queries_terms: Sequence[Sequence[str]], ctx=None, packed=None) -> np.ndarray:
    """score[q, c] = sum_t lower(chunk_c).count(term_t)   (web/server.py:32075-32076)."""
    ctx = ctx or _lib.Context.get()
    data, off = packed if packed is not None else pack_texts(chunk_texts)
    n = len(off) - 1
    tsym, toff, qoff = terms_to_symbols(queries_terms)
    nq = len(queries_terms)
    out = np.zeros(max(nq * n, 1), np.int32)
    if n and nq:
        rc = ctx.lib.dvb_chunk_scores(ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), n, ptr(tsym), ptr(toff), len(toff) - 1,
                                      ptr(qoff), nq, ptr(out))
        ctx.check(rc, "dvb_chunk_scores")
    return out[:nq * n].reshape(nq, n)


def topk(score: np.ndarray, k: int, index: Optional[np.ndarray] = None, index_base: int = 0, ctx=None):
    """Per row of `score`: (idx[k], score[k], n_valid) under (score desc, index asc), positives only."""
    ctx = ctx or _lib.Context.get()
    score = np.ascontiguousarray(np.atleast_2d(score), np.int32)
    nq, n = score.shape
    idx = np.ascontiguousarray(np.atleast_2d(index), np.int64) if index is not None else None
    oi = np.full((nq, k), -1, np.int64)
    os_ = np.zeros((nq, k), np.int32)
    on = np.zeros(nq, np.int32)
    rc = ctx.lib.dvb_topk(ctx.handle, _lib.MEM_HOST, ptr(score) if n else None, ptr(idx) if idx is not None else None, n, nq,
                          int(index_base), int(k), ptr(oi), ptr(os_), ptr(on))
    ctx.check(rc, "dvb_topk")
    return oi, os_, on


def symbolize(texts: Sequence[str], ctx=None) -> List[np.ndarray]:
    ctx = ctx or _lib.Context.get()
    data, off = pack_texts(texts)
    n = len(texts)
    sym = np.zeros(int(off[-1]) + 8, np.uint16)
    cnt = np.zeros(max(n, 1), np.int32)
    if n:
        rc = ctx.lib.dvb_symbolize(ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), n, ptr(sym), ptr(cnt))
        ctx.check(rc, "dvb_symbolize")
    return [sym[off[i]:off[i] + cnt[i]].copy() for i in range(n)]


def chunk_spans(contents: Sequence[str], chunk_size: int, overlap: int, ctx=None):
    """dvb_chunk_build on a batch of documents -> per document a list of
    (start, end, tstart, tend, tbyte0, tbyte1) and the encoded bytes of each document."""
    ctx = ctx or _lib.Context.get()
    data, off = pack_texts(contents)
    n = len(contents)
    third = max(0, chunk_size // 3)
    step = max(1, min(chunk_size, max(200, chunk_size // 2) + 1) - min(overlap, third))
    caps = (off[1:] - off[:-1]) // step + 2
    cap_off = np.zeros(n + 1, np.int64)
    np.cumsum(caps, out=cap_off[1:])
    cap = int(cap_off[-1])
    count = np.zeros(max(n, 1), np.int32)
    arrs32 = [np.zeros(max(cap, 1), np.int32) for _ in range(4)]
    arrs64 = [np.zeros(max(cap, 1), np.int64) for _ in range(2)]
    if n:
        rc = ctx.lib.dvb_chunk_build(ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), n, int(chunk_size), int(overlap), ptr(cap_off),
                                     ptr(count), *[ptr(a) for a in arrs32], *[ptr(a) for a in arrs64])
        ctx.check(rc, "dvb_chunk_build")
    out = []
    for d in range(n):
        c = int(count[d])
        if c > caps[d]:
            raise _lib.DvbError("dvb_chunk_build: capacity bound violated")
        s = int(cap_off[d])
        out.append(list(zip(*[a[s:s + c].tolist() for a in arrs32 + arrs64])))
    return out, data, off


# ----------------------------------------------------------------------------- reference-shaped API

def _compact_preview(content: str, limit: int) -> str:
    """web/server.py:31892-31896"""
    text = _WS.sub(" ", str(content or "")).strip()
    if len(text) <= limit:
        return text
    return text[: max(0, limit - 3)].rstrip() + "..."


def _build_reference_material_chunks(content: str, *, config: ReferenceMaterialConfig = DEFAULT_CONFIG, ctx=None) -> List[dict]:
    """Drop-in for web/server.py:31932: list of {chunk_id, index, start, end, text, preview}."""
    text = str(content or "")
    if not text:
        return []
    cfg = config.normalized()
    overlap = min(cfg.chunk_overlap, max(0, cfg.chunk_size // 3))
    spans, data, off = chunk_spans([text], cfg.chunk_size, overlap, ctx=ctx)
    raw = data[: int(off[1])].tobytes()
    chunks = []
    for (start, end, _ts, _te, b0, b1) in spans[0]:
        chunk_text = raw[b0:b1].decode("utf-8", "surrogatepass")
        chunks.append({
            "chunk_id": f"chunk-{len(chunks) + 1:04d}",
            "index": len(chunks),
            "start": start,
            "end": end,
            "text": chunk_text,
            "preview": _compact_preview(chunk_text, cfg.inline_chunk_preview_limit),
        })
    return chunks


def build_chunk_manifest(*, doc_id: str, filename: str, content: str, ext: str,
                         config: ReferenceMaterialConfig = DEFAULT_CONFIG, ctx=None) -> dict:
    """The `chunks.json` payload of save_reference_material_fulltext_artifacts (web/server.py:31990-31999)."""
    text = str(content or "")
    cfg = config.normalized()
    return {
        "version": 1, "doc_id": doc_id, "name": filename, "type": ext,
        "content_sha256": hashlib.sha256(text.encode("utf-8")).hexdigest(),
        "chunk_size": cfg.chunk_size, "chunk_overlap": cfg.chunk_overlap,
        "chunks": _build_reference_material_chunks(text, config=config, ctx=ctx),
    }


def load_chunks_from_manifest(path) -> List[dict]:
    """web/server.py:32020-32034 (path already resolved by the caller); I/O errors -> []."""
    try:
        payload = json.loads(Path(path).read_text(encoding="utf-8"))
    except Exception:
        return []
    chunks = payload.get("chunks", []) if isinstance(payload, dict) else []
    return [c for c in chunks if isinstance(c, dict) and str(c.get("text") or "").strip()]


def select_reference_material_context(doc: dict, *, query_text: str = "", max_chars: int = 0, chunk_limit: int = 0,
                                      chunk_loader: Optional[Callable[[dict], List[dict]]] = None,
                                      config: ReferenceMaterialConfig = DEFAULT_CONFIG, ctx=None,
                                      index_provider: Optional[Callable[[dict, List[dict]], Optional["ChunkIndex"]]] = None) -> Tuple[str, dict]:
    """Drop-in for web/server.py:32054.  `chunk_loader` is the reference's load_reference_material_chunks.
    `index_provider(doc, chunks)` may return a resident ChunkIndex over exactly these chunks (ChunkIndexCache + sidecar,
    SURVEY §8(f)1): the query then walks posting lists instead of re-lowering every chunk; results are identical."""
    cfg = config.normalized()
    chunks = chunk_loader(doc) if chunk_loader is not None else []
    limit = max(1, int(max_chars or cfg.max_doc_length))
    if not chunks:
        content = str((doc or {}).get("content") or "")
        return content[:limit], {"mode": "stored_content", "selected_chunks": [], "chunk_count": 0}

    terms = _tokenize_reference_material_query(query_text)
    max_chunks = max(1, int(chunk_limit or cfg.context_chunk_limit))
    selected: List[dict] = []
    picked = []
    if terms:
        order_index = np.asarray([int(c.get("index", 0) or 0) for c in chunks], np.int64)
        pos = np.arange(len(chunks), dtype=np.int64)
        ix = None
        if index_provider is not None and max_chunks <= 32 and np.array_equal(order_index, pos):
            try:
                ix = index_provider(doc, chunks)
            except _lib.DvbError:
                raise
            except Exception:
                ix = None                                # provider trouble (I/O, bad ref) never breaks the prompt build
        if ix is not None:                           # resident posting lists: same counts, same (score desc, index asc) order
            idx, _sc, n_valid = ix.topk([terms], max_chunks)
            picked = (idx[0][: int(n_valid[0])] - ix.index_base).tolist()
        else:
            scores = chunk_scores([str(c.get("text") or "") for c in chunks], [terms], ctx=ctx)[0]
            picked = None
            if np.array_equal(order_index, pos):
                idx, _sc, n_valid = topk(scores, min(max_chunks, 32), ctx=ctx)
                picked = idx[0][: int(n_valid[0])].tolist()
                if max_chunks > 32 and int(n_valid[0]) == 32:
                    picked = None                    # rare: more than 32 chunks requested -> generic path below
            if picked is None:                       # chunk["index"] is not the list position, or k > 32:
                ranked = sorted((i for i in range(len(chunks)) if scores[i] > 0), key=lambda i: (-int(scores[i]), int(order_index[i]), i))
                picked = ranked[:max_chunks]
    if picked:
        selected = [chunks[i] for i in picked]
        selected.sort(key=lambda item: int(item.get("index", 0) or 0))
    else:
        candidates = [0]
        if len(chunks) > 2:
            candidates.append(len(chunks) // 2)
        if len(chunks) > 1:
            candidates.append(len(chunks) - 1)
        for i in candidates:
            if chunks[i] not in selected:
                selected.append(chunks[i])
            if len(selected) >= max_chunks:
                break

    parts, used, meta_sel = [], 0, []
    for chunk in selected:
        text = str(chunk.get("text") or "").strip()
        if not text:
            continue
        prefix = f"[片段 {int(chunk.get('index', 0) or 0) + 1}/{len(chunks)}]\n"
        available = limit - used - len(prefix)
        if available <= 0:
            break
        clipped = text[:available]
        parts.append(prefix + clipped)
        used += len(prefix) + len(clipped)
        meta_sel.append({"chunk_id": chunk.get("chunk_id"), "index": chunk.get("index"), "start": chunk.get("start"), "end": chunk.get("end")})
        if used >= limit:
            break
    return "\n\n".join(parts).strip(), {"mode": "chunk_selection", "selected_chunks": meta_sel, "chunk_count": len(chunks),
                                       "query_terms": terms[:12]}


# ----------------------------------------------------------------------------- bulk path: resident inverted index (configs 3/5)
SIDECAR_NAME = "chunks.dvbidx"


def _tag32(tag: bytes) -> bytes:
    tag = bytes(tag)
    if len(tag) != 32:
        raise ValueError("a sidecar tag is exactly 32 bytes")
    return tag


def sidecar_tag(content_sha256: str, chunks: Sequence[dict]) -> bytes:
    """32-byte identity of what a chunk index was built from: the document's content hash (`content_sha256` of the
    manifest == `full_content_sha256` of the doc record, web/server.py:31995, 32015) and the (index, start, end) of every
    indexed chunk — so re-chunking with other parameters, or another document, never matches an old sidecar."""
    h = hashlib.sha256()
    h.update(b"dvbidx1|" + str(content_sha256 or "").encode("utf-8"))
    for c in chunks:
        h.update(b"|%d,%d,%d" % (int(c.get("index", 0) or 0), int(c.get("start", 0) or 0), int(c.get("end", 0) or 0)))
    return h.digest()


def _indexable_chunks(chunks) -> List[dict]:
    """The chunks load_reference_material_chunks returns (web/server.py:32033-32034), in manifest order."""
    return [c for c in (chunks or []) if isinstance(c, dict) and str(c.get("text") or "").strip()]


def save_chunk_index_sidecar(manifest_path, manifest: Optional[dict] = None, ctx=None) -> Path:
    """Build the index over a chunks.json manifest and write chunks.dvbidx beside it.  Returns the sidecar path."""
    manifest_path = Path(manifest_path)
    if manifest is None:
        manifest = json.loads(manifest_path.read_text(encoding="utf-8"))
    chunks = _indexable_chunks(manifest.get("chunks", []))
    ix = ChunkIndex([str(c.get("text") or "") for c in chunks], ctx=ctx)
    try:
        out = manifest_path.with_name(SIDECAR_NAME)
        ix.save(out, sidecar_tag(manifest.get("content_sha256", ""), chunks))
    finally:
        ix.close()
    return out


class ChunkIndexCache:
    """Resident indexes of recently queried documents, keyed by sidecar tag (LRU).  get() loads chunks.dvbidx when its
    tag matches, else builds from the chunk texts and (write_back) rewrites the sidecar."""

    def __init__(self, capacity: int = 64, ctx=None, write_back: bool = True):
        from collections import OrderedDict
        self.capacity, self.ctx, self.write_back = int(capacity), ctx, write_back
        self._lru: "OrderedDict[bytes, ChunkIndex]" = OrderedDict()
        self.hits = self.loads = self.builds = 0

    def get(self, sidecar_path, content_sha256: str, chunks: Sequence[dict]) -> "ChunkIndex":
        """`chunks` = what the reference's loader returned (already filtered to non-blank text)."""
        tag = sidecar_tag(content_sha256, chunks)
        ix = self._lru.get(tag)
        if ix is not None:
            self._lru.move_to_end(tag)
            self.hits += 1
            return ix
        side = Path(sidecar_path) if sidecar_path else None
        ix = None
        if side is not None and side.exists():
            try:
                ix = ChunkIndex.load(side, tag, ctx=self.ctx)
                if ix.stats()["n_chunks"] != len(chunks):
                    ix.close()
                    ix = None
                else:
                    self.loads += 1
            except _lib.DvbError:
                ix = None                                # stale / truncated / corrupt: rebuild below
        if ix is None:
            ix = ChunkIndex([str(c.get("text") or "") for c in chunks], ctx=self.ctx)
            self.builds += 1
            if self.write_back and side is not None:
                try:
                    ix.save(side, tag)
                except _lib.DvbError:
                    pass                                 # read-only data dir: the resident copy still serves this process
        self._lru[tag] = ix
        while len(self._lru) > self.capacity:
            _, old = self._lru.popitem(last=False)
            old.close()
        return ix

    def clear(self) -> None:
        for ix in self._lru.values():
            ix.close()
        self._lru.clear()


class ChunkIndex:
    """Device-resident positional inverted index over a chunk corpus (csrc/index.cu).  Same results as
    chunk_scores() + topk() — exact non-overlapping counts, (score desc, index asc), positives only — for many
    queries against many chunks.  One index per rank holds that rank's contiguous shard; `index_base` makes the
    returned indices global so shards merge with parallel.merge_topk()."""

    def __init__(self, chunk_texts: Optional[Sequence[str]] = None, *, packed=None, index_base: int = 0, ctx=None, device_ptrs=None):
        import ctypes as C
        self.ctx = ctx or _lib.Context.get()
        self.index_base = int(index_base)
        self._h = C.c_void_p()
        if device_ptrs is not None:                      # (utf8_ptr, off_ptr, n): corpus already resident in HBM
            up, op, n = device_ptrs
            rc = self.ctx.lib.dvb_chunk_index_build(self.ctx.handle, _lib.MEM_DEVICE, up, op, int(n), C.byref(self._h))
        else:
            data, off = packed if packed is not None else pack_texts(chunk_texts or [])
            rc = self.ctx.lib.dvb_chunk_index_build(self.ctx.handle, _lib.MEM_HOST, ptr(data), ptr(off), len(off) - 1, C.byref(self._h))
        self.ctx.check(rc, "dvb_chunk_index_build")
        self.build_ms = self.ctx.last_kernel_ms

    # ---- persistent sidecar next to chunks.json
    def save(self, path, tag: Optional[bytes] = None) -> None:
        """Write the resident index to `path` (atomic rename).  `tag` = 32 bytes naming the content (see sidecar_tag)."""
        import ctypes as C
        t = (C.c_uint8 * 32).from_buffer_copy(_tag32(tag)) if tag is not None else None
        self.ctx.check(self.ctx.lib.dvb_chunk_index_save(self.ctx.handle, self._h, str(path).encode(), t), "dvb_chunk_index_save")

    @classmethod
    def load(cls, path, tag: Optional[bytes] = None, *, index_base: int = 0, ctx=None) -> "ChunkIndex":
        """Load a sidecar written by save(); raises DvbError if it is missing, stale (tag), truncated or corrupt."""
        import ctypes as C
        self = cls.__new__(cls)
        self.ctx = ctx or _lib.Context.get()
        self.index_base = int(index_base)
        self._h = C.c_void_p()
        self.build_ms = 0.0
        t = (C.c_uint8 * 32).from_buffer_copy(_tag32(tag)) if tag is not None else None
        self.ctx.check(self.ctx.lib.dvb_chunk_index_load(self.ctx.handle, str(path).encode(), t, C.byref(self._h)), "dvb_chunk_index_load")
        return self

    def stats(self) -> dict:
        out = np.zeros(6, np.int64)
        self.ctx.check(self.ctx.lib.dvb_chunk_index_stats(self._h, ptr(out)), "dvb_chunk_index_stats")
        return dict(zip(("n_chunks", "text_bytes", "postings", "pair_keys", "device_bytes", "device"), map(int, out)))

    def topk(self, queries_terms: Sequence[Sequence[str]], k: int):
        """(idx[nq, k] global indices or -1, score[nq, k], n_valid[nq])"""
        tsym, toff, qoff = terms_to_symbols(queries_terms)
        nq = len(queries_terms)
        oi = np.full((max(nq, 1), k), -1, np.int64); os_ = np.zeros((max(nq, 1), k), np.int32); on = np.zeros(max(nq, 1), np.int32)
        if nq:
            rc = self.ctx.lib.dvb_chunk_index_topk(self.ctx.handle, self._h, _lib.MEM_HOST, ptr(tsym), ptr(toff), len(toff) - 1, ptr(qoff), nq,
                                                   self.index_base, int(k), ptr(oi), ptr(os_), ptr(on))
            self.ctx.check(rc, "dvb_chunk_index_topk")
        return oi[:nq], os_[:nq], on[:nq]

    def topk_device(self, dev_terms, nq: int, k: int):
        """Same as topk() with everything resident: dev_terms = (term_sym u16, term_off i32, query_term_off i32) torch CUDA
        tensors (terms_to_symbols() moved to the device once per query set); returns (idx[nq, k] i64, score[nq, k] i32,
        n_valid[nq] i32) torch CUDA tensors — nothing crosses PCIe (config 3/5: the candidates go straight into the
        all-gather, parallel.sharded_index_topk_device)."""
        import torch
        tsym, toff, qoff = dev_terms
        dev = tsym.device
        oi = torch.empty((max(nq, 1), k), dtype=torch.int64, device=dev)
        os_ = torch.empty((max(nq, 1), k), dtype=torch.int32, device=dev)
        on = torch.empty((max(nq, 1),), dtype=torch.int32, device=dev)
        if nq:
            torch.cuda.current_stream(dev).synchronize()         # the library runs on its own stream
            rc = self.ctx.lib.dvb_chunk_index_topk(self.ctx.handle, self._h, _lib.MEM_DEVICE, tsym.data_ptr(), toff.data_ptr(), int(toff.numel()) - 1,
                                                   qoff.data_ptr(), int(nq), self.index_base, int(k), oi.data_ptr(), os_.data_ptr(), on.data_ptr())
            self.ctx.check(rc, "dvb_chunk_index_topk")
        return oi[:nq], os_[:nq], on[:nq]

    def close(self):
        if self._h:
            self.ctx.lib.dvb_chunk_index_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
