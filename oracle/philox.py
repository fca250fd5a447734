"""ORACLE (test infrastructure, never on the product path).

Counter-based Philox4x32-10 restated in numpy, plus the uniform / normal
conversions the CUDA kernels use.  The reference draws from torch's stateful
generator (`torch_rand_float` at legged_gym/envs/base/legged_robot.py:530-535,
549,566,590 and `torch.rand_like` at :811); BASELINE.json's north_star replaces
that by counter-based Philox, so random *values* are defined here and injected
into the reference run by oracle/refharness.

Algorithm: Salmon et al., "Parallel random numbers: as easy as 1, 2, 3"
(Random123 philox4x32, 10 rounds).  Known-answer vectors: tests/test_philox.py.

Stream layout shared with unitree-rl-gym_b200/csrc/philox.cuh:
    key     = (seed_lo, seed_hi ^ stream_id)
    counter = (global_env_id, step_lo, step_hi, block)
    slot s  -> block = s // 4, lane = s % 4
    uniform = (x >> 8) * 2**-24            in [0, 1)
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = np.uint32(0x9E3779B9)
W1 = np.uint32(0xBB67AE85)
MASK = np.uint64(0xFFFFFFFF)

STREAM_ENV = 0      # env-side draws (commands, resets, pushes, obs noise)
STREAM_POLICY = 1   # action sampling noise


def philox4x32_10(ctr, key):
    """ctr: uint32 [..., 4], key: uint32 [..., 2] (broadcastable) -> uint32 [..., 4]."""
    ctr = np.asarray(ctr, dtype=np.uint32)
    key = np.asarray(key, dtype=np.uint32)
    c0, c1, c2, c3 = (ctr[..., i].astype(np.uint64) for i in range(4))
    k0 = np.broadcast_to(key[..., 0], c0.shape).astype(np.uint32)
    k1 = np.broadcast_to(key[..., 1], c0.shape).astype(np.uint32)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = M0 * c0
            p1 = M1 * c2
            hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
            hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
            n0 = hi1 ^ c1 ^ k0.astype(np.uint64)
            n2 = hi0 ^ c3 ^ k1.astype(np.uint64)
            c0, c1, c2, c3 = n0, lo1, n2, lo0
            k0 = (k0 + W0).astype(np.uint32)
            k1 = (k1 + W1).astype(np.uint32)
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def raw_table(seed, stream, env_ids, step, num_slots):
    """uint32 [len(env_ids), num_slots] of raw Philox words for one step."""
    env_ids = np.asarray(env_ids, dtype=np.uint32)
    nblk = (num_slots + 3) // 4
    ctr = np.zeros((env_ids.shape[0], nblk, 4), dtype=np.uint32)
    ctr[..., 0] = env_ids[:, None]
    ctr[..., 1] = np.uint32(step & 0xFFFFFFFF)
    ctr[..., 2] = np.uint32((step >> 32) & 0xFFFFFFFF)
    ctr[..., 3] = np.arange(nblk, dtype=np.uint32)[None, :]
    key = np.array([seed & 0xFFFFFFFF, ((seed >> 32) & 0xFFFFFFFF) ^ stream], dtype=np.uint32)
    out = philox4x32_10(ctr, key)
    return out.reshape(env_ids.shape[0], nblk * 4)[:, :num_slots]


def u01(x):
    """uint32 -> float32 uniform in [0,1): top 24 bits."""
    return ((np.asarray(x, dtype=np.uint32) >> np.uint32(8)).astype(np.float32)
            * np.float32(1.0 / 16777216.0))


def uniform_table(seed, stream, env_ids, step, num_slots):
    return u01(raw_table(seed, stream, env_ids, step, num_slots))


def normal_table(seed, stream, env_ids, step, num_normals):
    """Box-Muller on consecutive slot pairs (2j, 2j+1) -> normal j.

    z = sqrt(-2 ln(1-u0)) * cos(2 pi u1), computed in float32 like the kernel.
    (1-u0 is in (0,1], so the log is finite.)
    """
    u = uniform_table(seed, stream, env_ids, step, 2 * num_normals)
    u0 = u[:, 0::2]
    u1 = u[:, 1::2]
    r = np.sqrt(np.float32(-2.0) * np.log(np.float32(1.0) - u0)).astype(np.float32)
    return (r * np.cos(np.float32(6.283185307179586) * u1)).astype(np.float32)
