"""Philox4x32-10: Random123 known-answer vectors for the oracle, and (gpu) the kernel's stream
against the oracle bit for bit."""
import numpy as np
import pytest

from oracle import philox

KAT = [
    ([0, 0, 0, 0], [0, 0], "6627e8d5 e169c58d bc57ac4c 9b00dbd8"),
    ([0xffffffff] * 4, [0xffffffff] * 2, "408f276d 41c83b0e a20bc7c6 6d5451fd"),
    ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
     "d16cfe09 94fdcceb 5001e420 24126ea1"),
]


@pytest.mark.parametrize("ctr,key,want", KAT)
def test_random123_known_answers(ctr, key, want):
    out = philox.philox4x32_10(np.array(ctr, dtype=np.uint32), np.array(key, dtype=np.uint32))
    assert " ".join("%08x" % x for x in out) == want


def test_uniform_range_and_moments():
    u = philox.uniform_table(1, 0, np.arange(4096), 7, 74)
    assert u.dtype == np.float32 and u.min() >= 0.0 and u.max() < 1.0
    assert abs(u.mean() - 0.5) < 5e-3 and abs(u.var() - 1 / 12) < 2e-3
    # independence across env / step / slot: distinct counters give distinct words
    a = philox.raw_table(1, 0, np.arange(64), 7, 8)
    b = philox.raw_table(1, 0, np.arange(64), 8, 8)
    c = philox.raw_table(1, 1, np.arange(64), 7, 8)
    assert (a != b).mean() > 0.99 and (a != c).mean() > 0.99
    assert len(np.unique(a)) == a.size


def test_normal_moments():
    z = philox.normal_table(1, 1, np.arange(8192), 3, 12)
    assert abs(z.mean()) < 1e-2 and abs(z.std() - 1.0) < 1e-2 and np.isfinite(z).all()


@pytest.mark.gpu
def test_kernel_stream_matches_oracle_bit_exact():
    import torch
    from unitree_rl_gym_b200 import _lib
    n, slots, seed, off, step = 1000, 74, (5 << 32) | 1234567, 4096, (1 << 33) + 17
    out = torch.empty(n, slots, device="cuda")
    _lib.check(_lib.lib().urlg_philox_uniform(_lib.ptr(out), n, slots, seed & 0xffffffff, seed >> 32, 0, off, step,
                                              _lib.cur_stream()))
    want = philox.uniform_table(seed, 0, np.arange(n) + off, step, slots)
    assert np.array_equal(out.cpu().numpy(), want)
