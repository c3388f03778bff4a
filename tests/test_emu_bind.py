"""CPU tier: the PRODUCT kernel sources of dvb_bind_claims (csrc/bind7.cuh fast kernel + csrc/bind_general.cuh re-run
kernel) executed by the SIMT emulator under tests/emu (g++, fibers) and compared bit for bit with the oracle.
The emulator is test infrastructure; the product path has no CPU fallback (tests/test_abi.py checks that)."""
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent))
sys.path.insert(0, str(Path(__file__).resolve().parent / "emu"))

import bind_cases as bc  # noqa: E402


@pytest.fixture(scope="module")
def emu():
    import emu as e
    e.build()
    return e


def _check(emu, oracle, a, mode=0, grid=3, expect_handback=None):
    got = emu.bind_claims(*a, grid=grid, mode=mode)
    cu, co, fu, fo, sco, sfo, fq, fd, fh, cp, ct = a
    ref = oracle.bind_claims(None, None, sco, sfo, fq, fd, fh, cp, ct, packed=(cu, co, fu, fo))
    assert np.array_equal(got["best_fact"], ref["best_fact"])
    assert np.array_equal(got["best_score"].view(np.uint64), ref["best_score"].view(np.uint64))      # bit-exact fp64
    assert np.array_equal(got["bound"], ref["bound"])
    assert np.array_equal(got["len_claim"], np.minimum(ref["len_claim"], 14))
    has_claims = np.repeat(np.diff(sco) > 0, np.diff(sfo))
    assert np.array_equal(got["len_fact"][has_claims], np.minimum(ref["len_fact"], 18)[has_claims])
    if expect_handback is not None:
        assert int(got["fallback"].sum()) == expect_handback
    return got


def test_emu_degenerate_and_short_facts(emu, oracle):
    _check(emu, oracle, bc.degenerate(), expect_handback=0)
    _check(emu, oracle, bc.mixed_real_text(), expect_handback=0)


@pytest.mark.parametrize("n_sessions,turns,claims", [(6, 40, 24), (3, 1, 1), (4, 7, 33)])
def test_emu_synthetic_bit_exact(emu, oracle, n_sessions, turns, claims):
    _check(emu, oracle, bc.synthetic(n_sessions, turns, claims), expect_handback=0)


def test_emu_every_pair_of_long_facts(emu, oracle):
    got = _check(emu, oracle, bc.every_pair())
    assert 0 < int(got["fallback"].sum()) < len(got["fallback"]) // 4        # only the sessions with exotic characters


def test_emu_handbacks(emu, oracle):
    _check(emu, oracle, bc.token_table_overflow(), expect_handback=1)        # > 704 ASCII tokens in a claim group
    _check(emu, oracle, bc.big_claim_groups(), expect_handback=2)            # token ranges beyond the flag array
    _check(emu, oracle, bc.long_facts(1), expect_handback=1)                 # U+0130 / U+212A inside
    _check(emu, oracle, bc.handback_in_one_group(), expect_handback=2, grid=3)   # one group of four, from the claim and from the fact side


def test_emu_group_numbering_across_steps(emu, oracle):
    _check(emu, oracle, bc.many_tiny_sessions(), expect_handback=0, grid=7)  # k_bind_groups: 4300 sessions > one 4096-session step


@pytest.mark.parametrize("parts", [2, 3, 8])
def test_emu_split_fact_kernel(emu, oracle, monkeypatch, parts):
    """k_bind_facts_split (what dvb_bind_claims launches when a call has few claim groups): the facts of every group over
    `parts` CTAs, the last one to finish reduces the per-part bests — same answers, ties on the lower fact index included."""
    monkeypatch.setenv("EMU_BIND_PARTS", str(parts))
    _check(emu, oracle, bc.degenerate(), expect_handback=0)                  # sessions with fewer facts than parts, without facts, without claims
    _check(emu, oracle, bc.mixed_real_text(), expect_handback=0)
    _check(emu, oracle, bc.synthetic(4, 7, 33), expect_handback=0)
    _check(emu, oracle, bc.handback_in_one_group(), expect_handback=2, grid=5)
    _check(emu, oracle, bc.long_facts(1), expect_handback=1)


def test_emu_crowded_bucket_side_list(emu, oracle):
    _check(emu, oracle, bc.crowded_bucket(5), expect_handback=0)             # one key in the side list
    _check(emu, oracle, bc.crowded_bucket(30), expect_handback=0)            # 26 keys in the side list
    _check(emu, oracle, bc.crowded_bucket(40), expect_handback=2)            # more than the side list holds


def test_emu_general_kernel_alone(emu, oracle):
    _check(emu, oracle, bc.synthetic(3, 12, 40), mode=2)
    _check(emu, oracle, bc.mixed_real_text(), mode=2)


def test_emu_grid_independence(emu, oracle):
    a = bc.synthetic(5, 10, 24)
    r1 = _check(emu, oracle, a, grid=1)
    r5 = _check(emu, oracle, a, grid=5)
    assert np.array_equal(r1["best_score"], r5["best_score"])
