"""dvb_topk_static — the allocation-free variant of dvb_topk (static device scratch, caller's stream) — against the
existing entry point dvb_topk(DVB_MEM_DEVICE), through the C-ABI.  Integer work: the outputs must be identical.
Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _scores(rng, nq, n, density):
    s = np.zeros((nq, n), np.int32)
    m = rng.random((nq, n)) < density
    s[m] = rng.integers(1, 9, size=int(m.sum()), dtype=np.int32)      # few distinct values: the index tie-break decides
    return s


def _both(ctx, torch, score, idx, index_base, k, stream):
    from deep_vision_b200 import _lib
    nq, n = score.shape
    d_score = torch.from_numpy(score).cuda()
    d_idx = torch.from_numpy(idx).cuda() if idx is not None else None
    outs = []
    for which in ("dvb_topk", "dvb_topk_static"):
        oi = torch.full((nq, k), -99, dtype=torch.int64, device="cuda")
        os_ = torch.full((nq, k), -99, dtype=torch.int32, device="cuda")
        on = torch.full((nq,), -99, dtype=torch.int32, device="cuda")
        ip = d_idx.data_ptr() if d_idx is not None else None
        if which == "dvb_topk":
            rc = ctx.lib.dvb_topk(ctx.handle, _lib.MEM_DEVICE, d_score.data_ptr(), ip, n, nq, index_base, k,
                                  oi.data_ptr(), os_.data_ptr(), on.data_ptr())
        else:
            rc = ctx.lib.dvb_topk_static(ctx.handle, C.c_void_p(stream.cuda_stream), d_score.data_ptr(), ip, n, nq, index_base, k,
                                         oi.data_ptr(), os_.data_ptr(), on.data_ptr())
            stream.synchronize()                                        # the static entry point returns without waiting
        ctx.check(rc, which)
        torch.cuda.synchronize()
        outs.append((oi.cpu().numpy(), os_.cpu().numpy(), on.cpu().numpy()))
    return outs


def test_topk_static_matches_topk_on_two_calls_in_a_row(gpu_ctx):
    import torch
    ctx = gpu_ctx
    rng = np.random.default_rng(77)
    stream = torch.cuda.Stream()
    launches0 = ctx.launches
    # call 1: many splits per query (dense rows, implicit indices); call 2: fewer, shorter lists with explicit indices, so
    # the scratch still holds call 1's lists beyond what call 2 writes
    cases = [(_scores(rng, 6, 200_003, 0.01), None, 1_000_000_007, 8),
             (_scores(rng, 37, 5_001, 0.3), rng.permutation(37 * 5_001).astype(np.int64).reshape(37, 5_001), 0, 32),
             (_scores(rng, 3, 70_000, 0.0005), None, 5, 8)]                # some queries have fewer than k positives
    for score, idx, base, k in cases:
        (ri, rs, rn), (si, ss, sn) = _both(ctx, torch, score, idx, base, k, stream)
        assert np.array_equal(ri, si) and np.array_equal(rs, ss) and np.array_equal(rn, sn)
        # and both equal sorted(positive, key=(-score, index))[:k]  (web/server.py:32081-32085)
        for q in range(score.shape[0]):
            ids = idx[q] if idx is not None else base + np.arange(score.shape[1], dtype=np.int64)
            pos = np.nonzero(score[q] > 0)[0]
            order = pos[np.lexsort((ids[pos], -score[q][pos]))][:k]
            assert int(sn[q]) == len(order)
            assert np.array_equal(si[q, :len(order)], ids[order]) and np.array_equal(ss[q, :len(order)], score[q][order])
    assert ctx.launches - launches0 == 4 * len(cases)                    # two kernels per call, both entry points


def test_topk_static_rejects_sizes_beyond_the_scratch(gpu_ctx):
    import torch
    from deep_vision_b200 import _lib
    ctx = gpu_ctx
    nq, k, n = 349_525 // 32 + 1, 32, 4                                   # nq * k * 12 bytes > 4 MiB
    d_score = torch.zeros((nq, n), dtype=torch.int32, device="cuda")
    oi = torch.empty((nq, k), dtype=torch.int64, device="cuda"); os_ = torch.empty((nq, k), dtype=torch.int32, device="cuda")
    on = torch.empty((nq,), dtype=torch.int32, device="cuda")
    rc = ctx.lib.dvb_topk_static(ctx.handle, None, d_score.data_ptr(), None, n, nq, 0, k, oi.data_ptr(), os_.data_ptr(), on.data_ptr())
    assert rc == _lib.DVB_E_CAPACITY
    assert b"static scratch" in ctx.lib.dvb_last_error(ctx.handle)
    torch.cuda.synchronize()
