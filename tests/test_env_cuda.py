"""Parity of the fused CUDA env path (through the C ABI, via the drop-in LeggedRobot classes)
against fixtures produced by the unmodified reference and against the oracle port."""
import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu


def make_env(case, fx=None, frames=None, seed=1, pool_seam=False, packed=False):
    from unitree_rl_gym_b200.legged_gym.utils.helpers import SimParams
    from unitree_rl_gym_b200.physics import SyntheticGym
    from unitree_rl_gym_b200.robots import ROBOTS
    task, n, pos = H.ENV_CASES[case]
    cfg = H.make_cfg(task, n, pos, H.TWEAKS.get(case))
    cfg.seed = seed
    sp = SimParams()
    sp.dt = cfg.sim.dt
    gym = SyntheticGym(ROBOTS[task], n, cfg.control.decimation, "cuda:0",
                       frames=frames if frames is not None else H.frames_of(fx), pool_seam=pool_seam, packed=packed)
    cls = H.env_class(task)
    return cls(cfg=cfg, sim_params=sp, physics_engine=gym, sim_device="cuda:0", headless=True)


def env_snapshot(e):
    names = ["obs_buf", "rew_buf", "reset_buf", "time_out_buf", "episode_length_buf", "torques", "commands",
             "actions", "last_actions", "last_dof_vel", "last_root_vel", "feet_air_time", "last_contacts", "rpy",
             "base_lin_vel", "base_ang_vel", "projected_gravity", "root_states", "dof_state"]
    out = {k: getattr(e, k) for k in names}
    if e.privileged_obs_buf is not None:
        out["privileged_obs_buf"] = e.privileged_obs_buf
    for k, v in e.episode_sums.items():
        out["episode_sums/" + k] = v
    for k, v in e.extras.get("episode", {}).items():
        out["extras_episode/" + k] = v.reshape(1)
    if "time_outs" in e.extras:
        out["extras_time_outs"] = e.extras["time_outs"]
    torch.cuda.synchronize()
    return {k: v.detach().cpu().numpy() for k, v in out.items()}


@pytest.mark.parametrize("case", list(H.ENV_CASES))
def test_env_matches_reference_fixture(case):
    fx = H.load_case(case)
    env = make_env(case, fx)
    hs = env.spec.family == 2
    cmp = lambda k, a, b, w: (H.compare_handstand_obs if hs and k in ("obs_buf", "privileged_obs_buf") else H.compare)(k, a, b, w)
    env.reset()
    snap = env_snapshot(env)
    for k in H.snapshot_keys(fx, "reset"):
        cmp(k, snap[k], fx["reset/" + k], "reset/")
    env.episode_length_buf = torch.from_numpy(fx["planted_episode_length"]).cuda()
    for t in range(int(fx["const/scalars"][4])):
        env.step(torch.from_numpy(fx[f"actions/{t}"]).cuda())
        snap = env_snapshot(env)
        for k in H.snapshot_keys(fx, f"step{t}"):
            cmp(k, snap[k], fx[f"step{t}/{k}"], f"step{t}/")


def _oracle_pair(task, n, frames_n=3, seed=5, tweak=None, pool_seam=False, packed=False):
    """(cuda env, oracle port, spec, frames, physics callback) on the same seeded synthetic frames."""
    from unitree_rl_gym_b200 import synthetic
    from unitree_rl_gym_b200.robots import ROBOTS
    case = "tmp"
    H.ENV_CASES[case] = (task, n, False)
    if tweak:
        H.TWEAKS[case] = tweak
    try:
        spec = H.make_spec(case)
        frames = synthetic.make_frames(ROBOTS[task], None, n, frames_n, spec.decimation, seed=seed, feet_idx=spec.feet_idx,
                                       pen_idx=spec.pen_idx, term_idx=spec.term_idx, init_z=float(spec.base_init_state[2]),
                                       default_dof_pos=torch.from_numpy(spec.default_dof_pos))
        env = make_env(case, frames=frames, pool_seam=pool_seam, packed=packed)
    finally:
        del H.ENV_CASES[case]
        H.TWEAKS.pop(case, None)
    port = H.make_port(spec)
    calls = {"n": 0}

    def physics(sub):
        if sub is not None:
            calls["n"] += 1
        port.load_frame(frames, ((calls["n"] - 1) // spec.decimation) % frames_n, sub)
    return env, port, spec, frames, physics


def _compare_step(env, port, where, keys=("obs_buf", "rew_buf", "torques", "commands", "last_root_vel", "feet_air_time",
                                         "last_actions", "last_dof_vel")):
    snap = env_snapshot(env)
    H.compare("reset_buf", snap["reset_buf"], port.reset_buf.numpy(), where)
    H.compare("time_out_buf", snap["time_out_buf"], port.time_out_buf.numpy(), where)
    H.compare("episode_length_buf", snap["episode_length_buf"], port.episode_length_buf.numpy(), where)
    for k in keys:
        H.compare(k, snap[k], getattr(port, k).numpy(), where)
    if port.privileged_obs_buf is not None:
        H.compare("privileged_obs_buf", snap["privileged_obs_buf"], port.privileged_obs_buf.numpy(), where)
    for k, v in port.episode_sums.items():
        H.compare("episode_sums/" + k, snap["episode_sums/" + k], v.numpy(), where)
    for k, v in port.extras["episode"].items():
        H.compare("extras_episode/" + k, snap["extras_episode/" + k], v.reshape(1).numpy(), where)
    return snap


# BASELINE.json configs[1..4] at their own sizes (Go2 / H1 / G1 at 4096 envs, H1_2 at 16384), plus ragged tails
@pytest.mark.parametrize("task,n", [("go2", 4096), ("h1", 4096), ("g1", 4096), ("h1_2", 16384), ("go2_handstand", 4096),
                                    ("g1", 1000), ("h1_2", 2049), ("h1", 37), ("go2_handstand", 95)])
def test_env_matches_oracle_port_larger(task, n):
    """Same seeded inputs on both sides at the BASELINE sizes; n % 32 != 0 exercises the tail block."""
    env, port, spec, frames, physics = _oracle_pair(task, n)
    env.reset()
    port.reset(physics)
    g = torch.Generator().manual_seed(3)
    ep = torch.randint(0, 1001, (n,), generator=g)
    env.episode_length_buf = ep.cuda()
    port.episode_length_buf = ep.clone()
    resets = 0
    for t in range(4):
        a = torch.randn(n, spec.num_dof, generator=g)
        env.step(a.cuda())
        port.step(a, physics)
        _compare_step(env, port, f"step{t}/")
        resets += int(port.reset_buf.sum())
    assert resets > 0 or n < 64


def test_pool_seam_matches_copy_seam():
    """urlg_env_bind_frame_pool: reading the synthetic frames in place gives bit-identical env outputs to copying them
    into the acquired tensors first (the pool seam only removes the stand-in copies from a timed region) -- with one dense
    array per tensor and with packed frames (urlg_frame_pool.frame_stride: root | dof x decimation | contact | rigid body of a
    frame in one row, what a host-side engine uploads with a single copy)."""
    out = []
    for pool, packed in ((False, False), (True, False), (True, True)):
        env, port, spec, frames, physics = _oracle_pair("g1", 2048, pool_seam=pool, packed=packed)
        env.reset()
        g = torch.Generator().manual_seed(3)
        env.episode_length_buf = torch.randint(0, 1001, (2048,), generator=g).cuda()
        snaps = []
        for t in range(5):
            env.step(torch.randn(2048, spec.num_dof, generator=g).cuda())
            snap = env_snapshot(env)
            for k in ("root_states", "dof_state"):       # in pool mode only reset / pushed rows of these are current
                snap.pop(k)
            snaps.append(snap)
        out.append(snaps)
    for other in out[1:]:
        for t, (a, b) in enumerate(zip(out[0], other)):
            assert a.keys() == b.keys()
            for k in a:
                if k.startswith("extras_episode/"):      # means over reset envs: float atomics, order not defined
                    np.testing.assert_allclose(a[k], b[k], rtol=1e-5, atol=1e-7, err_msg=f"step{t}/{k}")
                else:
                    assert np.array_equal(a[k], b[k]), f"step{t}/{k}"


def test_indexed_setter_lists_match_masks():
    """the compacted int32 id lists + device counts handed to set_dof_state_tensor_indexed /
    set_actor_root_state_tensor_indexed equal reset_buf.nonzero() and (reset | push-due).nonzero()
    (legged_robot.py:552-555, 569-572, 591-594)."""
    n = 4096
    env, port, spec, frames, physics = _oracle_pair("go2", n)
    env.reset()
    g = torch.Generator().manual_seed(9)
    ep = torch.randint(0, 1001, (n,), generator=g)
    ep[:7] = spec.push_interval - 1                     # push-due without a reset
    env.episode_length_buf = ep.cuda()
    calls0 = env.gym.indexed_calls
    seen_push_only = 0
    for t in range(3):
        env.step(torch.randn(n, spec.num_dof, generator=g).cuda())
        torch.cuda.synchronize()
        ids, cnt = env.gym.last_dof_set
        want = env.reset_buf.nonzero().flatten().to(torch.int32)
        got = ids[:int(cnt)].sort().values
        assert torch.equal(got, want), t
        rids, rcnt = env.gym.last_root_set
        due = (env.episode_length_buf % spec.push_interval == 0)
        want_root = (env.reset_buf | due).nonzero().flatten().to(torch.int32)
        assert torch.equal(rids[:int(rcnt)].sort().values, want_root), t
        seen_push_only += int((due & ~env.reset_buf).sum())
    assert env.gym.indexed_calls == calls0 + 6 and seen_push_only >= 7


def test_resets_and_pushes_reach_a_stateful_simulator():
    """With a simulator whose state lives apart from the exposed tensors (physics.StatefulGym), an in-step reset / push
    only survives the next refresh if the env commits it through the indexed setters."""
    from unitree_rl_gym_b200.legged_gym.utils.helpers import SimParams
    from unitree_rl_gym_b200.physics import StatefulGym
    from unitree_rl_gym_b200.robots import ROBOTS
    n = 512
    cfg = H.make_cfg("go2", n, True)
    cfg.noise.add_noise = False
    sp = SimParams()
    sp.dt = cfg.sim.dt
    gym = StatefulGym(ROBOTS["go2"], n, cfg.control.decimation, "cuda:0")
    env = H.env_class("go2")(cfg=cfg, sim_params=sp, physics_engine=gym, sim_device="cuda:0", headless=True)
    env.reset()
    ep = torch.zeros(n, dtype=torch.long)
    ep[:5] = 1000                                       # time out at the next step -> reset
    ep[5:9] = env.spec.push_interval - 1                # push-due, no reset
    ep[9:] = 3
    env.episode_length_buf = ep.cuda()
    before_root = gym.h_root.clone()
    env.step(torch.zeros(n, 12).cuda())
    torch.cuda.synchronize()
    assert env.reset_buf[:5].all() and not env.reset_buf[5:].any()
    # reset envs: the hidden state took the reset DOF positions (default * U(0.5, 1.5), velocity 0) and root state
    h_dof = gym.h_dof.view(n, 12, 2)
    assert torch.equal(h_dof[:5], env.dof_state.view(n, 12, 2)[:5]) and (h_dof[:5, :, 1] == 0).all()
    assert torch.equal(gym.h_root[:9], env.root_states[:9])
    assert (gym.h_root[:5, 2] - 0.42).abs().max() < 1e-6
    # pushed envs: xy velocity redrawn in the hidden state; everything the env did not list is untouched
    assert (gym.h_root[5:9, 7:9] != before_root[5:9, 7:9]).all()
    untouched = gym.h_root[9:].clone()
    untouched[:, 0] -= cfg.control.decimation * gym.drift
    assert torch.allclose(untouched, before_root[9:], atol=1e-6)
    # and the next step starts from the committed state: a reset env's dof_pos observation is its reset pose + drift
    env.step(torch.zeros(n, 12).cuda())
    torch.cuda.synchronize()
    want = h_dof[:5, :, 0] - env.default_dof_pos
    assert torch.allclose(env.obs_buf[:5, 12:24], want, atol=1e-5)


def test_unknown_controller_raises_nameerror():
    from unitree_rl_gym_b200.legged_gym.envs import LeggedRobot
    from unitree_rl_gym_b200.legged_gym.utils.helpers import SimParams
    cfg = H.make_cfg("go2", 32, True)
    cfg.control.control_type = "X"
    sp = SimParams()
    sp.dt = cfg.sim.dt
    with pytest.raises(NameError):
        LeggedRobot(cfg=cfg, sim_params=sp, physics_engine=None, sim_device="cuda:0", headless=True)


# ------------------------------------------------------------------ size-independent properties
def _sliced_frames(frames, lo, hi, A, B, n):
    """Rows [lo, hi) of every env-major frame tensor (dof / contact / rigid_body are flattened over env)."""
    out = {"root": frames["root"][:, lo:hi].contiguous()}
    P, D = frames["dof"].shape[0], frames["dof"].shape[1]
    out["dof"] = frames["dof"].view(P, D, n, A, 2)[:, :, lo:hi].reshape(P, D, (hi - lo) * A, 2).contiguous()
    out["contact"] = frames["contact"].view(P, n, B, 3)[:, lo:hi].reshape(P, (hi - lo) * B, 3).contiguous()
    out["rigid_body"] = frames["rigid_body"].view(P, n, B, 13)[:, lo:hi].reshape(P, (hi - lo) * B, 13).contiguous()
    return out


@pytest.mark.parametrize("task,n,cut", [("go2", 4096, 2048), ("g1", 600, 250)])
def test_env_shards_reproduce_the_whole(task, n, cut):
    """Multi-GPU sharding property at the BASELINE size: stepping envs [0,cut) and [cut,n) as two env
    objects with env_offset 0 / cut gives bit-identical tensors to stepping all n envs as one object — every Philox
    draw is keyed by the global env id, and nothing in K1/K2 couples envs except the any-push flag and the extras means,
    which are compared after re-deriving them from the shards."""
    from unitree_rl_gym_b200 import synthetic
    from unitree_rl_gym_b200.legged_gym.utils.helpers import SimParams
    from unitree_rl_gym_b200.physics import SyntheticGym
    from unitree_rl_gym_b200.robots import ROBOTS
    H.ENV_CASES["tmp"] = (task, n, False)
    try:
        spec = H.make_spec("tmp")
    finally:
        del H.ENV_CASES["tmp"]
    robot = ROBOTS[task]
    frames = synthetic.make_frames(robot, None, n, 3, spec.decimation, seed=11, feet_idx=spec.feet_idx, pen_idx=spec.pen_idx,
                                   term_idx=spec.term_idx, init_z=float(spec.base_init_state[2]),
                                   default_dof_pos=torch.from_numpy(spec.default_dof_pos))

    def build(lo, hi):
        cfg = H.make_cfg(task, hi - lo, False)
        cfg.seed = 1
        cfg.env.env_offset = lo
        sp = SimParams()
        sp.dt = cfg.sim.dt
        fr = frames if (lo, hi) == (0, n) else _sliced_frames(frames, lo, hi, robot.num_dof, robot.num_bodies, n)
        gym = SyntheticGym(robot, hi - lo, cfg.control.decimation, "cuda:0", frames=fr)
        cls = H.env_class(task)
        return cls(cfg=cfg, sim_params=sp, physics_engine=gym, sim_device="cuda:0", headless=True)
    whole, parts = build(0, n), [build(0, cut), build(cut, n)]
    g = torch.Generator().manual_seed(4)
    ep = torch.randint(0, 1001, (n,), generator=g)
    for e, (lo, hi) in zip([whole] + parts, [(0, n), (0, cut), (cut, n)]):
        e.reset()
        e.episode_length_buf = ep[lo:hi].cuda()
    names = ["obs_buf", "rew_buf", "reset_buf", "time_out_buf", "episode_length_buf", "torques", "commands", "actions",
             "last_actions", "last_dof_vel", "feet_air_time", "last_contacts", "rpy", "base_lin_vel", "dof_state"]
    for t in range(4):
        a = torch.randn(n, spec.num_dof, generator=g).cuda()
        whole.step(a)
        parts[0].step(a[:cut].contiguous())
        parts[1].step(a[cut:].contiguous())
        torch.cuda.synchronize()
        for k in names + (["privileged_obs_buf"] if whole.privileged_obs_buf is not None else []):
            w = getattr(whole, k)
            got = torch.cat([getattr(parts[0], k).reshape(cut, -1), getattr(parts[1], k).reshape(n - cut, -1)]).reshape(w.shape)
            assert torch.equal(w, got), f"step{t}/{k}"
        for k in whole.episode_sums:
            got = torch.cat([parts[0].episode_sums[k], parts[1].episode_sums[k]])
            assert torch.equal(whole.episode_sums[k], got), f"step{t}/episode_sums/{k}"


def test_masks_at_thresholds():
    """Adversarial inputs within +-2 ulp of every mask threshold: termination contact norm 1.0, penalised
    contact 0.1, feet F_z 1.0, |roll| 0.8, |pitch| 1.0, command norm 0.2.  Contact / command thresholds are exact IEEE
    sqrt / compare and must match bit for bit; roll / pitch go through atan2f / asinf (CUDA libm vs the CPU's SLEEF), so
    an env whose oracle angle lies within 2 ulp of its threshold is excluded from the reset comparison."""
    from oracle.env_port import EnvPort
    from unitree_rl_gym_b200 import synthetic
    from unitree_rl_gym_b200.robots import ROBOTS
    task, n = "go2", 512
    H.ENV_CASES["tmp"] = (task, n, False)
    try:
        spec = H.make_spec("tmp")
        frames = synthetic.make_frames(ROBOTS[task], None, n, 2, spec.decimation, seed=21, feet_idx=spec.feet_idx,
                                       pen_idx=spec.pen_idx, term_idx=spec.term_idx, init_z=float(spec.base_init_state[2]),
                                       default_dof_pos=torch.from_numpy(spec.default_dof_pos))
        B = ROBOTS[task].num_bodies

        def near(x, i):          # x shifted by (i % 5) - 2 ulps
            v = np.float32(x)
            for _ in range(abs(i % 5 - 2)):
                v = np.nextafter(v, np.float32(np.inf if i % 5 > 2 else -np.inf))
            return float(v)
        con = frames["contact"].view(2, n, B, 3)
        root = frames["root"]
        for f in range(2):
            for e in range(n):
                kind = e % 6
                if kind == 0:      # termination body norm ~ 1.0 along one axis (exact sqrt of a square)
                    con[f, e, spec.term_idx[0]] = torch.tensor([0.0, near(1.0, e // 6), 0.0])
                elif kind == 1:    # penalised body norm ~ 0.1
                    con[f, e, spec.pen_idx[e % len(spec.pen_idx)]] = torch.tensor([near(0.1, e // 6), 0.0, 0.0])
                elif kind == 2:    # foot F_z ~ 1.0
                    con[f, e, spec.feet_idx[e % len(spec.feet_idx)]] = torch.tensor([0.0, 0.0, near(1.0, e // 6)])
                elif kind == 3:    # pure roll ~ 0.8
                    th = near(0.8, e // 6) * (1 if e % 12 < 6 else -1)
                    root[f, e, 3:7] = torch.tensor([np.sin(th / 2), 0.0, 0.0, np.cos(th / 2)], dtype=torch.float32)
                elif kind == 4:    # pure pitch ~ 1.0
                    th = near(1.0, e // 6) * (1 if e % 12 < 6 else -1)
                    root[f, e, 3:7] = torch.tensor([0.0, np.sin(th / 2), 0.0, np.cos(th / 2)], dtype=torch.float32)
        env = make_env("tmp", frames=frames)
    finally:
        H.ENV_CASES.pop("tmp", None)
    port = EnvPort(spec)
    calls = {"n": 0}

    def physics(sub):
        if sub is not None:
            calls["n"] += 1
        port.load_frame(frames, ((calls["n"] - 1) // spec.decimation) % 2, sub)
    env.reset()
    port.reset(physics)
    g = torch.Generator().manual_seed(8)
    checked = 0
    for t in range(3):
        a = torch.randn(n, spec.num_dof, generator=g)
        env.step(a.cuda())
        port.step(a, physics)
        snap = env_snapshot(env)
        rpy = port.rpy.numpy()
        band = lambda v, th: np.abs(np.abs(v) - np.float32(th)) <= 2 * np.spacing(np.float32(th))
        skip = band(rpy[:, 0], 0.8) | band(rpy[:, 1], 1.0)
        keep = ~skip
        checked += int(keep.sum())
        assert np.array_equal(snap["reset_buf"][keep], port.reset_buf.numpy()[keep]), f"step{t}"
        H.compare("time_out_buf", snap["time_out_buf"], port.time_out_buf.numpy(), f"step{t}/")
        H.compare("last_contacts", snap["last_contacts"], port.last_contacts.numpy(), f"step{t}/")
        # collision / feet_air_time terms flip by a whole unit when their mask flips: they must agree everywhere
        for k in ("collision", "feet_air_time"):
            H.compare("episode_sums/" + k, snap["episode_sums/" + k][keep], port.episode_sums[k].numpy()[keep], f"step{t}/")
        # an excluded env may legitimately reset on one side only: re-align the port with the device for the next step
        # (every per-env state the next step reads) and keep going -- all three steps are checked
        m = torch.from_numpy(skip)
        if m.any():
            port.episode_length_buf[m] = env.episode_length_buf.cpu()[m]
            port.reset_buf[m] = env.reset_buf.cpu()[m]
            for k in ("commands", "actions", "last_actions", "last_dof_vel", "feet_air_time", "last_contacts"):
                getattr(port, k)[m] = getattr(env, k).cpu()[m]
            for k in port.episode_sums:
                port.episode_sums[k][m] = env.episode_sums[k].cpu()[m]
    assert checked >= 3 * n * 0.6
