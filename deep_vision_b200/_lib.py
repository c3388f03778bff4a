"""ctypes binding of libdvb200.so (the C-ABI in include/dvb200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device is
present, every compute entry point raises.  Loading the library (symbol checks)
works without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path

HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("DVB_LIB", HERE / "libdvb200.so"))   # DVB_LIB: A/B-test another build of the same ABI

DVB_OK, DVB_E_INVALID, DVB_E_CAPACITY, DVB_E_CUDA, DVB_E_NOMEM = 0, -1, -2, -3, -4
MEM_HOST, MEM_DEVICE = 0, 1

_vp = C.c_void_p
_SIGS = {
    "dvb_create": (C.c_int32, [C.c_int32, C.POINTER(_vp)]),
    "dvb_destroy": (None, [_vp]),
    "dvb_last_error": (C.c_char_p, [_vp]),
    "dvb_version": (C.c_int32, []),
    "dvb_unicode_version": (C.c_char_p, []),
    "dvb_launch_count": (C.c_int64, [_vp]),
    "dvb_set_timing": (None, [_vp, C.c_int32]),
    "dvb_last_kernel_ms": (C.c_float, [_vp]),
    "dvb_set_stream": (None, [_vp, _vp]),
    "dvb_tokenize_v3": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, _vp, C.c_int64, _vp]),
    "dvb_query_terms": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, C.c_int32, _vp, C.c_int64, _vp, _vp, _vp]),
    "dvb_symbolize": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, _vp, _vp]),
    "dvb_chunk_scores": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, _vp, _vp, C.c_int32, _vp, C.c_int32, _vp]),
    "dvb_topk": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, C.c_int32, C.c_int64, C.c_int32, _vp, _vp, _vp]),
    "dvb_topk_static": (C.c_int32, [_vp, _vp, _vp, _vp, C.c_int64, C.c_int32, C.c_int64, C.c_int32, _vp, _vp, _vp]),
    "dvb_chunk_build": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, C.c_int32, C.c_int32, _vp] + [_vp] * 7),
    "dvb_question_keys": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp, C.c_int32, _vp, _vp, C.c_int32,
                                      _vp, _vp, _vp, _vp, _vp, C.c_int64]),
    "dvb_question_gate": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp, C.c_int32,
                                      _vp, _vp, C.c_int32, C.c_double, _vp]),
    "dvb_solution_jaccard": (C.c_int32, [_vp, C.c_int32, _vp, _vp, _vp, _vp, C.c_int64, C.c_int32, _vp, _vp, _vp, _vp]),
    "dvb_slim_facts": (C.c_int32, [_vp, C.c_int32, C.c_int64] + [_vp] * 8 + [C.c_int32, C.c_double, C.c_int32, C.c_int32, _vp, _vp]),
    "dvb_text_features": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp, C.c_int32, _vp, _vp, _vp, _vp, _vp]),
    "dvb_annotate_turns": (C.c_int32, [_vp, C.c_int32, C.c_int64] + [_vp] * 13 + [C.c_uint32, C.c_int32, _vp, _vp, _vp]),
    "dvb_contains": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, _vp, _vp, _vp, C.c_int64, _vp]),
    "dvb_chunk_index_build": (C.c_int32, [_vp, C.c_int32, _vp, _vp, C.c_int64, C.POINTER(_vp)]),
    "dvb_chunk_index_free": (None, [_vp]),
    "dvb_chunk_index_stats": (C.c_int32, [_vp, _vp]),
    "dvb_chunk_index_topk": (C.c_int32, [_vp, _vp, C.c_int32, _vp, _vp, C.c_int32, _vp, C.c_int32, C.c_int64, C.c_int32, _vp, _vp, _vp]),
    "dvb_chunk_index_save": (C.c_int32, [_vp, _vp, C.c_char_p, _vp]),
    "dvb_chunk_index_load": (C.c_int32, [_vp, C.c_char_p, _vp, C.POINTER(_vp)]),
    "dvb_bind_claims": (C.c_int32, [_vp, C.c_int32] + [_vp] * 6 + [C.c_int64] + [_vp] * 10 + [C.c_int64]),
}

_lib = None
_lock = threading.Lock()


class DvbError(RuntimeError):
    pass


def load():
    """dlopen the library and bind every symbol include/dvb200.h declares."""
    global _lib
    with _lock:
        if _lib is None:
            if not LIB_PATH.exists():
                raise DvbError(
                    f"{LIB_PATH} is missing: build it with `python -m deep_vision_b200.build` "
                    "(nvcc, sm_100a). There is no CPU fallback.")
            lib = C.CDLL(str(LIB_PATH))
            for name, (res, args) in _SIGS.items():
                fn = getattr(lib, name)          # AttributeError if the symbol is not exported
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def exported_symbols():
    return sorted(_SIGS)


class Context:
    """One dvb_ctx per (process, device)."""

    _instances = {}
    _ilock = threading.Lock()

    def __init__(self, device: int = 0):
        self.lib = load()
        h = _vp()
        rc = self.lib.dvb_create(int(device), C.byref(h))
        if rc != DVB_OK or not h.value:
            raise DvbError(f"dvb_create(device={device}) failed (rc={rc}): a CUDA device is required; "
                           "there is no CPU fallback")
        self.handle = h
        self.device = int(device)

    @classmethod
    def get(cls, device: int | None = None) -> "Context":
        if device is None:
            device = int(os.environ.get("DVB_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        with cls._ilock:
            ctx = cls._instances.get(device)
            if ctx is None:
                ctx = cls(device)
                cls._instances[device] = ctx
            return ctx

    def check(self, rc: int, what: str):
        if rc != DVB_OK:
            msg = self.lib.dvb_last_error(self.handle)
            raise DvbError(f"{what} failed (rc={rc}): {msg.decode('utf-8', 'replace') if msg else ''}")

    @property
    def launches(self) -> int:
        return int(self.lib.dvb_launch_count(self.handle))

    def set_timing(self, on: bool):
        self.lib.dvb_set_timing(self.handle, 1 if on else 0)

    @property
    def last_kernel_ms(self) -> float:
        return float(self.lib.dvb_last_kernel_ms(self.handle))

    def set_stream(self, stream_ptr):
        self.lib.dvb_set_stream(self.handle, stream_ptr)

    def close(self):
        if self.handle:
            self.lib.dvb_destroy(self.handle)
            self.handle = None
