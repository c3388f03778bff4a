"""ctypes driver of the CPU SIMT emulator build of the bind kernels (tests/emu/emu_bind.cpp).  TEST INFRASTRUCTURE:
executes the product kernel SOURCE (bind7.cuh, bind_general.cuh) on host threads so the CPU test tier can compare it
with the oracle; the product never loads it."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
LIB = HERE / "_build" / "libemu_bind.so"
_lib = None


def build(force: bool = False) -> Path:
    srcs = [HERE / "emu_bind.cpp", HERE / "simt_emu.h", HERE / "simt_emu_intrinsics.h"] + \
           [ROOT / "deep_vision_b200" / "csrc" / n for n in ("bind7.cuh", "bind_general.cuh", "bind_common.cuh", "text.cuh")]
    if force or not LIB.exists() or any(s.stat().st_mtime > LIB.stat().st_mtime for s in srcs):
        LIB.parent.mkdir(exist_ok=True)
        subprocess.run(["g++", "-std=c++20", "-O1", "-g", "-fPIC", "-shared", "-pthread", "-ffp-contract=off", "-Wno-attributes",
                        "-o", str(LIB), str(HERE / "emu_bind.cpp")], check=True)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(LIB))
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def bind_claims(claim_utf8, claim_off, fact_utf8, fact_off, sess_claim_off, sess_fact_off, fact_quality, fact_dim, fact_hard,
                claim_pref_dim, claim_threshold, grid: int = 2, mode: int = 0):
    """mode 0: fast kernel + general re-run of handed-back sessions (the product's flow); 1: fast only; 2: general only."""
    c = lambda a, t: np.ascontiguousarray(a, dtype=t)
    # 16 bytes of lead-in: the general kernel's tokenizer loads the 16-byte aligned block that holds the first byte
    cu = np.concatenate([np.zeros(16, np.uint8), c(claim_utf8, np.uint8), np.zeros(64, np.uint8)])
    fu = np.concatenate([np.zeros(16, np.uint8), c(fact_utf8, np.uint8), np.zeros(64, np.uint8)])
    co, fo = c(claim_off, np.int64), c(fact_off, np.int64)
    sco, sfo = c(sess_claim_off, np.int64), c(sess_fact_off, np.int64)
    fq, fd, fh = c(fact_quality, np.float64), c(fact_dim, np.int32), c(fact_hard, np.uint8)
    cp, ct = c(claim_pref_dim, np.int32), c(claim_threshold, np.float64)
    ns, nc, nf = len(sco) - 1, int(sco[-1]), int(sfo[-1])
    out = {"best_fact": np.full(max(nc, 1), -7, np.int32), "best_score": np.zeros(max(nc, 1), np.float64), "bound": np.zeros(max(nc, 1), np.uint8),
           "len_claim": np.zeros(max(nc, 1), np.int32), "len_fact": np.zeros(max(nf, 1), np.int32), "fallback": np.zeros(max(ns, 1), np.uint8)}
    rc = lib().emu_bind_claims(C.c_void_p(cu.ctypes.data + 16), _p(co), C.c_void_p(fu.ctypes.data + 16), _p(fo), _p(sco), _p(sfo), C.c_int64(ns),
                               _p(fq), _p(fd), _p(fh), _p(cp), _p(ct),
                               _p(out["best_fact"]), _p(out["best_score"]), _p(out["bound"]), _p(out["len_claim"]), _p(out["len_fact"]),
                               _p(out["fallback"]), C.c_int(grid), C.c_int(mode))
    out["overflow"] = rc
    for k, n in (("best_fact", nc), ("best_score", nc), ("bound", nc), ("len_claim", nc), ("len_fact", nf), ("fallback", ns)):
        out[k] = out[k][:n]
    return out
