#!/usr/bin/env python3
"""Pinned host -> device copy bandwidth on this box: one 870 MB copy, the same bytes as 24 slices on one stream, and as
two halves on two streams (the ceiling the end-to-end arm of bench.py runs against).  python tools/h2d_probe.py"""
import torch
n = 870 * 1000 * 1000
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
def timed(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
    return best
one = timed(lambda: d.copy_(h, non_blocking=True))
def sliced():
    k = n // 24
    for i in range(24):
        d[i * k:(i + 1) * k].copy_(h[i * k:(i + 1) * k], non_blocking=True)
sl = timed(sliced)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def two():
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur); s2.wait_stream(cur)
    with torch.cuda.stream(s1): d[:n // 2].copy_(h[:n // 2], non_blocking=True)
    with torch.cuda.stream(s2): d[n // 2:].copy_(h[n // 2:], non_blocking=True)
    cur.wait_stream(s1); cur.wait_stream(s2)
tw = timed(two)
print("one copy %.2f ms = %.1f GB/s | 24 slices %.2f ms = %.1f GB/s | two streams %.2f ms = %.1f GB/s" % (one, n / one / 1e6, sl, n / sl / 1e6, tw, n / tw / 1e6))
