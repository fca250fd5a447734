"""Host-side logic of the multi-GPU path on CPU: world_size-2 `gloo` groups.

What is checked (none of it needs a GPU; the device kernels are covered by the -m gpu tests):
  * env sharding: rank r's Philox table for local envs with env_offset = r*N equals rows [r*N, (r+1)*N) of the
    single-process table (draws independent of the GPU count), and EnvSpec carries the offset;
  * advantage normalisation: all-reducing (sum, sumsq, count) reproduces the single-process mean / unbiased std
    of the oracle's RolloutStorage.compute_returns over the concatenated env set;
  * gradient exchange: all-reduce(sum) * 1/world of per-rank mean-loss gradients equals the gradient of the
    oracle's loss over the union minibatch, and the KL sum / (rows * world) equals the global KL mean, so every
    rank takes the same adaptive-LR branch;
  * bench.py --impl reference under a multi-rank launch: only rank 0 works and prints.
"""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        res = fn(rank, world)
        torch.save(res, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def run_world(fn, tmp_path, world=2):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, fn, str(tmp_path)), nprocs=world, join=True)
    return [torch.load(os.path.join(str(tmp_path), f"r{r}.pt"), weights_only=False) for r in range(world)]


# ------------------------------------------------------------------ env sharding / Philox counters
def _philox_shard(rank, world):
    from oracle import philox
    n_loc, slots, step, seed = 37, 23, 5, 1
    ids = np.arange(n_loc, dtype=np.int64) + rank * n_loc
    return philox.uniform_table(seed, philox.STREAM_ENV, ids, step, slots)


def test_philox_tables_are_shard_invariant(tmp_path):
    from oracle import philox
    parts = run_world(_philox_shard, tmp_path)
    whole = philox.uniform_table(1, philox.STREAM_ENV, np.arange(74, dtype=np.int64), 5, 23)
    assert np.array_equal(np.concatenate(parts, axis=0), whole)


def test_env_spec_carries_shard_offset():
    from tests import helpers as H
    from unitree_rl_gym_b200.envspec import build_env_spec
    from unitree_rl_gym_b200.robots import ROBOTS
    cfg = H.make_cfg("go2", 64, True)
    for rank in (0, 1, 7):
        spec = build_env_spec("go2", cfg, ROBOTS["go2"], cfg.sim.dt, seed=1, env_offset=rank * 64)
        assert spec.env_offset == rank * 64


# ------------------------------------------------------------------ GAE moments (C2)
def _storage(n, t, seed):
    from oracle.rsl_rl_port import RolloutStorage
    g = torch.Generator().manual_seed(seed)
    s = RolloutStorage(n, t, [4], [None], [2], "cpu")
    s.rewards.copy_(torch.randn(t, n, 1, generator=g))
    s.values.copy_(torch.randn(t, n, 1, generator=g))
    s.dones.copy_((torch.rand(t, n, 1, generator=g) < 0.1).byte())
    last = torch.randn(n, 1, generator=g)
    return s, last


def _gae_rank(rank, world):
    """What the product does per rank: scan locally (phase 1), all-reduce the moments, normalise (phase 2)."""
    s, last = _storage(16, 24, 100 + rank)
    adv, nv = torch.zeros(16, 1), last
    for t in reversed(range(24)):                       # oracle formula, un-normalised
        nt = 1.0 - s.dones[t].float()
        delta = s.rewards[t] + nt * 0.99 * nv - s.values[t]
        adv = delta + nt * 0.99 * 0.95 * adv
        s.returns[t] = adv + s.values[t]
        nv = s.values[t]
    a = (s.returns - s.values).double()
    m = torch.tensor([a.sum(), (a * a).sum(), float(a.numel())], dtype=torch.float64)
    dist.all_reduce(m)                                  # PPO._reduce_moments
    mean = m[0] / m[2]
    var = (m[1] - m[2] * mean * mean) / (m[2] - 1.0)
    return ((a - mean) / (var.sqrt() + 1e-8)).float()


def test_global_advantage_normalisation(tmp_path):
    from oracle.rsl_rl_port import RolloutStorage
    parts = run_world(_gae_rank, tmp_path)
    s0, l0 = _storage(16, 24, 100)
    s1, l1 = _storage(16, 24, 101)
    whole = RolloutStorage(32, 24, [4], [None], [2], "cpu")
    for name in ("rewards", "values", "dones"):
        getattr(whole, name).copy_(torch.cat([getattr(s0, name), getattr(s1, name)], dim=1))
    whole.compute_returns(torch.cat([l0, l1], dim=0), 0.99, 0.95)
    got = torch.cat(parts, dim=1)
    np.testing.assert_allclose(got.numpy(), whole.advantages.numpy(), rtol=1e-4, atol=1e-5)


# ------------------------------------------------------------------ gradient exchange (C1) + KL scalar
def _policy():
    from oracle.rsl_rl_port import ActorCritic
    torch.manual_seed(3)
    return ActorCritic(12, 12, 3, [16, 8], [16, 8], "elu", init_noise_std=0.8)


def _batch(rank, rows=40):
    g = torch.Generator().manual_seed(7 + rank)
    r = lambda *s: torch.randn(*s, generator=g)
    return dict(obs=r(rows, 12), act=r(rows, 3), adv=r(rows), ret=r(rows, 1), tv=r(rows, 1), olp=r(rows) * 0.1 - 3.0,
                omu=r(rows, 3) * 0.1, osig=torch.full((rows, 3), 0.8))


def _loss_and_kl(ac, b):
    ac.act(b["obs"])
    lp = ac.get_actions_log_prob(b["act"])
    val = ac.evaluate(b["obs"])
    mu, sig = ac.action_mean, ac.action_std
    kl = torch.sum(torch.log(sig / b["osig"] + 1e-5) + (b["osig"] ** 2 + (b["omu"] - mu) ** 2) / (2.0 * sig ** 2) - 0.5, dim=-1)
    ratio = torch.exp(lp - b["olp"])
    surr = torch.max(-b["adv"] * ratio, -b["adv"] * torch.clamp(ratio, 0.8, 1.2)).mean()
    vc = b["tv"] + (val - b["tv"]).clamp(-0.2, 0.2)
    vl = torch.max((val - b["ret"]).pow(2), (vc - b["ret"]).pow(2)).mean()
    return surr + vl - 0.01 * ac.entropy.mean(), kl.detach()


def _grad_rank(rank, world):
    ac = _policy()
    loss, kl = _loss_and_kl(ac, _batch(rank))
    loss.backward()
    flat = torch.cat([p.grad.reshape(-1) for p in ac.parameters()])
    dist.all_reduce(flat)                               # PPO.update: all_reduce(eng.grads), then grad_scale = 1/world
    flat *= 1.0 / world
    klsum = kl.sum().reshape(1)
    dist.all_reduce(klsum)                              # stats[KL_SUM]; kl_mean = sum / (rows * kl_count_scale)
    return flat, float(klsum) / (kl.numel() * world)


def test_gradient_and_kl_exchange_match_union_batch(tmp_path):
    parts = run_world(_grad_rank, tmp_path)
    b0, b1 = _batch(0), _batch(1)
    union = {k: torch.cat([b0[k], b1[k]], dim=0) for k in b0}
    ac = _policy()
    loss, kl = _loss_and_kl(ac, union)
    loss.backward()
    want = torch.cat([p.grad.reshape(-1) for p in ac.parameters()])
    for flat, kl_mean in parts:
        np.testing.assert_allclose(flat.numpy(), want.numpy(), rtol=1e-5, atol=1e-7)
        assert abs(kl_mean - float(kl.mean())) < 1e-6
    assert torch.equal(parts[0][0], parts[1][0])        # identical update on every rank


# ------------------------------------------------------------------ bench.py reference arm under a multi-rank launch
def test_reference_arm_only_rank0_works():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29999")
    out = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--gpus", "2",
                          "--steps", "1", "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == ""


@pytest.mark.timeout(300)
def test_reference_arm_line_shape():
    out = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--envs", "64",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=280)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "env-steps/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
