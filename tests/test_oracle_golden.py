"""The oracle's env port (oracle/env_port.py) against fixtures produced by the unmodified
reference classes (tests/golden/gen_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from tests import helpers as H
from oracle.env_port import EnvPort


def port_snapshot(p):
    out = {"obs_buf": p.obs_buf, "rew_buf": p.rew_buf, "reset_buf": p.reset_buf, "time_out_buf": p.time_out_buf,
           "episode_length_buf": p.episode_length_buf, "torques": p.torques, "commands": p.commands,
           "actions": p.actions, "last_actions": p.last_actions, "last_dof_vel": p.last_dof_vel,
           "last_root_vel": p.last_root_vel, "feet_air_time": p.feet_air_time, "last_contacts": p.last_contacts,
           "rpy": p.rpy, "base_lin_vel": p.base_lin_vel, "base_ang_vel": p.base_ang_vel,
           "projected_gravity": p.projected_gravity, "root_states": p.root_states, "dof_state": p.dof_state}
    if p.privileged_obs_buf is not None:
        out["privileged_obs_buf"] = p.privileged_obs_buf
    for k, v in p.episode_sums.items():
        out["episode_sums/" + k] = v
    for k, v in p.extras.get("episode", {}).items():
        out["extras_episode/" + k] = v.reshape(1)
    if "time_outs" in p.extras:
        out["extras_time_outs"] = p.extras["time_outs"]
    return {k: v.numpy() for k, v in out.items()}


def run_port(case, fx):
    spec = H.make_spec(case)
    frames = H.frames_of(fx)
    port = H.make_port(spec)
    calls = {"n": 0}
    P = frames["root"].shape[0]

    def physics(sub):
        if sub is not None:
            calls["n"] += 1
        f = ((calls["n"] - 1) // spec.decimation) % P
        port.load_frame(frames, f, sub if sub is not None else None)
    return spec, port, physics


@pytest.mark.parametrize("case", list(H.ENV_CASES))
def test_constants_match_reference(case):
    fx = H.load_case(case)
    s = H.make_spec(case)
    np.testing.assert_array_equal(s.p_gains, fx["const/p_gains"])
    np.testing.assert_array_equal(s.d_gains, fx["const/d_gains"])
    np.testing.assert_array_equal(s.default_dof_pos.reshape(1, -1), fx["const/default_dof_pos"])
    np.testing.assert_array_equal(s.torque_limits, fx["const/torque_limits"])
    np.testing.assert_array_equal(s.dof_vel_limits, fx["const/dof_vel_limits"])
    np.testing.assert_array_equal(s.dof_pos_limits, fx["const/dof_pos_limits"])
    np.testing.assert_array_equal(s.noise_scale_vec, fx["const/noise_scale_vec"])
    assert s.feet_idx == fx["const/feet_indices"].tolist()
    assert s.pen_idx == fx["const/penalised_contact_indices"].tolist()
    assert s.term_idx == fx["const/termination_contact_indices"].tolist()
    np.testing.assert_array_equal(s.env_origins, fx["const/env_origins"])
    np.testing.assert_array_equal(s.base_init_state, fx["const/base_init_state"])
    assert s.reward_names == fx["const/reward_names"].tolist()
    np.testing.assert_array_equal(np.array(s.reward_scales), fx["const/reward_scales"])
    dt, max_len, push, resample = fx["const/scalars"][:4]
    assert (s.dt, s.max_episode_length, s.push_interval, s.resample_steps) == (dt, max_len, int(push), int(resample))


@pytest.mark.parametrize("case", list(H.ENV_CASES))
def test_env_port_matches_reference_fixture(case):
    fx = H.load_case(case)
    spec, port, physics = run_port(case, fx)
    hs = spec.family == 2
    cmp = lambda k, a, b, w: (H.compare_handstand_obs if hs and k in ("obs_buf", "privileged_obs_buf") else H.compare)(k, a, b, w)
    port.reset(physics)
    snap = port_snapshot(port)
    for k in H.snapshot_keys(fx, "reset"):
        cmp(k, snap[k], fx["reset/" + k], "reset/")
    port.episode_length_buf = torch.from_numpy(fx["planted_episode_length"]).clone()
    steps = int(fx["const/scalars"][4])
    for t in range(steps):
        port.step(torch.from_numpy(fx[f"actions/{t}"]), physics)
        snap = port_snapshot(port)
        for k in H.snapshot_keys(fx, f"step{t}"):
            cmp(k, snap[k], fx[f"step{t}/{k}"], f"step{t}/")


def test_handstand_reference_fixture_frame_transforms():
    """SURVEY.md 8(f) row 1 groundwork: the fixture recorded from the UNMODIFIED GO2HandstandEnv (46-dim obs, 17 active
    reward terms + termination) is well-formed, and the oracle's frame-transform restatements (the part of the path the
    handstand task shares, E4) reproduce its rpy / base_ang_vel / projected_gravity on every env that was not reset
    in the step (a reset rewrites root_states after the transforms were taken)."""
    from oracle.env_port import quat_rotate_inverse, euler_xyz
    fx = np.load(os.path.join(H.GOLDEN, "env_go2_handstand_n64.npz"))
    names = [str(x) for x in fx["const/reward_names"]]
    for n in ("front_feet_contact", "front_feet_together", "front_hip_neutral", "hind_feet_no_contact", "pose",
              "stability", "stay_still", "orientation", "base_height"):
        assert n in names
    steps = int(fx["const/scalars"][4])
    assert fx["step0/obs_buf"].shape == (64, 46)
    checked = 0
    for t in range(steps):
        keep = ~fx[f"step{t}/reset_buf"].astype(bool)
        assert keep.sum() > 32 and (~keep).sum() > 0          # resets fire in every recorded step
        root = torch.from_numpy(fx[f"step{t}/root_states"])[torch.from_numpy(keep)]
        q = root[:, 3:7]
        grav = torch.tensor([[0.0, 0.0, -1.0]]).repeat(q.shape[0], 1)
        np.testing.assert_allclose(quat_rotate_inverse(q, grav).numpy(), fx[f"step{t}/projected_gravity"][keep],
                                   rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(quat_rotate_inverse(q, root[:, 10:13]).numpy(), fx[f"step{t}/base_ang_vel"][keep],
                                   rtol=1e-5, atol=1e-6)
        rpy = euler_xyz(q)
        rpy = torch.stack(list(rpy), dim=1) if isinstance(rpy, (tuple, list)) else rpy
        np.testing.assert_allclose(rpy.numpy(), fx[f"step{t}/rpy"][keep], rtol=1e-5, atol=1e-6)
        assert np.isfinite(fx[f"step{t}/rew_buf"]).all() and np.isfinite(fx[f"step{t}/obs_buf"]).all()
        checked += int(keep.sum())
    assert checked > 300


def test_handstand_termination_restatement_is_bit_exact():
    """oracle/handstand_port.handstand_termination against the reference's reset / time-out masks, step by step.  The
    snapshot's episode_length_buf is already zeroed for reset envs, so the pre-reset counter is rebuilt from the planted
    lengths (+1 per step, 0 after a reset) exactly as post_physics_step does (legged_robot.py:681)."""
    from oracle.handstand_port import handstand_termination
    fx = np.load(os.path.join(H.GOLDEN, "env_go2_handstand_n64.npz"))
    steps, P = int(fx["const/scalars"][4]), int(fx["const/scalars"][5])
    max_len = int(fx["const/scalars"][1])
    n = fx["step0/obs_buf"].shape[0]
    term_idx = torch.from_numpy(fx["const/termination_contact_indices"]).long()
    contact = torch.from_numpy(fx["frames/contact"]).reshape(P, n, -1, 3)
    ep = torch.from_numpy(fx["planted_episode_length"]).long()
    fired = 0
    for t in range(steps):
        ep = ep + 1
        reset, time_out = handstand_termination(contact[(t + 1) % P], term_idx,   # frame 0 went to reset()
                                                torch.from_numpy(fx[f"step{t}/projected_gravity"]), ep, max_len)
        assert np.array_equal(reset.numpy(), fx[f"step{t}/reset_buf"].astype(bool)), t
        assert np.array_equal(time_out.numpy(), fx[f"step{t}/time_out_buf"].astype(bool)), t
        fired += int(reset.sum())
        ep = torch.where(reset, torch.zeros_like(ep), ep)
        assert np.array_equal(ep.numpy(), fx[f"step{t}/episode_length_buf"])
    assert fired >= 10


def test_handstand_pd_restatement_matches_reference_torques():
    """oracle/handstand_port.handstand_pd_torques against the reference's `torques` buffer (the last decimation substep
    of each recorded step: dof frame (t+1) % P, substep decimation-1; frame 0 went to reset()) to 1e-5."""
    from oracle.handstand_port import handstand_pd_torques
    fx = np.load(os.path.join(H.GOLDEN, "env_go2_handstand_n64.npz"))
    steps, P = int(fx["const/scalars"][4]), int(fx["const/scalars"][5])
    n, A = fx["step0/torques"].shape
    scale, clip = fx["const/handstand/action_scale_clip"]
    t_ = torch.from_numpy
    idx = t_(fx["const/handstand/rear_dof_idx"]).long()
    assert idx.tolist() == [6, 7, 8, 9, 10, 11]                   # RL_*, RR_* in the go2 DOF order of SURVEY 8(a).4
    clamped = 0
    for t in range(steps):
        a = t_(fx[f"actions/{t}"]).clamp(-clip, clip)
        d = t_(fx["frames/dof"][(t + 1) % P, 3]).view(n, A, 2)
        tq = handstand_pd_torques(a, d[..., 0], d[..., 1], t_(fx["const/p_gains"]), t_(fx["const/d_gains"]),
                                  t_(fx["const/default_dof_pos"]), float(scale), idx,
                                  t_(fx["const/handstand/rear_center"]), float(fx["const/handstand/rear_limit"]))
        np.testing.assert_allclose(tq.numpy(), fx[f"step{t}/torques"], rtol=1e-5, atol=1e-5)
        lim = t_(fx["const/torque_limits"])
        clamped += int((tq.abs() > lim).sum())
    assert clamped > 0            # the override really leaves torques beyond the URDF effort limits unclipped


def test_handstand_observation_restatement_matches_reference():
    """oracle/handstand_port.handstand_observations against the reference obs_buf: the 42 continuous columns (incl. the
    noise, drawn from the shared Philox table at step t+2 — reset() took steps 0 and 1) to 1e-5; the four contact flags
    are exact up to the reference's 1 % Bernoulli flips, which come from torch's global generator (not replayable)."""
    from oracle import philox
    from oracle.handstand_port import handstand_observations
    fx = np.load(os.path.join(H.GOLDEN, "env_go2_handstand_n64.npz"))
    steps, P, seed = int(fx["const/scalars"][4]), int(fx["const/scalars"][5]), int(fx["const/scalars"][6])
    n, A = fx["step0/torques"].shape
    O = fx["step0/obs_buf"].shape[1]
    t_ = torch.from_numpy
    feet = t_(fx["const/feet_indices"]).long()
    nsv = t_(fx["const/noise_scale_vec"])
    flipped = total = 0
    for t in range(steps):
        d = t_(fx[f"step{t}/dof_state"]).view(n, A, 2)
        u = philox.uniform_table(seed, philox.STREAM_ENV, np.arange(n), t + 2, 14 + A + O)[:, 14 + A:14 + A + O]
        obs = handstand_observations(t_(fx[f"step{t}/base_ang_vel"]), t_(fx[f"step{t}/projected_gravity"]), d[..., 0],
                                     d[..., 1], t_(fx["const/default_dof_pos"]), t_(fx[f"step{t}/actions"]),
                                     t_(fx["frames/contact"][(t + 1) % P]).view(n, -1, 3), feet[[0, 1]], feet[[2, 3]],
                                     0.25, 1.0, 0.05, t_(u.astype(np.float32)), nsv)
        ref = fx[f"step{t}/obs_buf"]
        np.testing.assert_allclose(obs[:, :-4].numpy(), ref[:, :-4], rtol=1e-5, atol=1e-6)
        assert set(np.unique(ref[:, -4:])) <= {0.0, 1.0}
        flipped += int((obs[:, -4:].numpy() != ref[:, -4:]).sum())
        total += 4 * n
        np.testing.assert_array_equal(fx[f"step{t}/privileged_obs_buf"], ref)
    assert flipped <= 0.04 * total          # p = 0.01 flips; 1536 draws


def test_handstand_reward_terms_match_reference_episode_sums():
    """Ten handstand reward terms (oracle/handstand_port.reward_*) against the reference: on envs that were not reset
    in the step, episode_sums[name] grows by term * (scale * dt) (legged_robot.py:778-781; const/reward_scales already
    carry the dt factor, :824-826)."""
    from oracle import handstand_port as HP
    fx = np.load(os.path.join(H.GOLDEN, "env_go2_handstand_n64.npz"))
    steps, P = int(fx["const/scalars"][4]), int(fx["const/scalars"][5])
    n, A = fx["step0/torques"].shape
    t_ = torch.from_numpy
    feet = t_(fx["const/feet_indices"]).long()
    scales = dict(zip([str(x) for x in fx["const/reward_names"]], fx["const/reward_scales"]))
    for t in range(steps):
        keep = ~fx[f"step{t}/reset_buf"].astype(bool)
        con = t_(fx["frames/contact"][(t + 1) % P]).view(n, -1, 3)
        lin, ang = t_(fx[f"step{t}/base_lin_vel"]), t_(fx[f"step{t}/base_ang_vel"])
        q = t_(fx[f"step{t}/dof_state"]).view(n, A, 2)[..., 0]          # unchanged by the step on non-reset envs
        terms = {"orientation": HP.reward_orientation(t_(fx[f"step{t}/projected_gravity"])),
                 "front_feet_contact": HP.reward_front_feet_contact(con, feet[[0, 1]]),
                 "hind_feet_no_contact": HP.reward_hind_feet_no_contact(con, feet[[2, 3]]),
                 "stability": HP.reward_stability(lin, ang),
                 "front_hip_neutral": HP.reward_front_hip_neutral(q, 0, 3),     # FL_hip, FR_hip in the go2 DOF order
                 "stay_still": HP.reward_stay_still(lin, ang),
                 "lin_vel_xy": HP.reward_lin_vel_xy(lin),
                 "base_height": HP.reward_base_height(t_(fx[f"step{t}/root_states"])[:, 2],
                                                      float(fx["const/handstand/base_height_target"])),
                 "pose": HP.reward_pose(q, t_(fx["const/handstand/pose_targets"])),
                 "front_feet_together": HP.reward_front_feet_together(
                     t_(fx["frames/rigid_body"][(t + 1) % P]).view(n, -1, 13)[:, feet, :3])}
        prev = "reset" if t == 0 else f"step{t - 1}"
        for name, r in terms.items():
            delta = fx[f"step{t}/episode_sums/{name}"] - fx[f"{prev}/episode_sums/{name}"]
            want = (r.numpy() * np.float32(scales[name]))
            np.testing.assert_allclose(delta[keep], want[keep], rtol=2e-4, atol=2e-6, err_msg=f"{name} step {t}")
