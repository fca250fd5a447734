#!/usr/bin/env python
"""bench.py — env-steps/s (obs + reward + PD + reset + policy forward) and PPO-iteration ms for the
B200-native hot path of unitree-rl-gym, on the configuration BASELINE.json quotes the metric on:
Go2 flat terrain, 4096 envs per GPU, decimation 4, 24-step rollout, MLP actor/critic [512,256,128].

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one PPO iteration = 24 env steps of all envs (policy forward, PD x decimation, fused
post-physics, transition store) + GAE + the 5x4 minibatch update.  `value` is the rollout-phase
env-steps/s (the metric's first half), `ppo_iter_ms` the whole iteration (its second half).
Physics is the synthetic tensor seam (BASELINE.json north_star): a device-resident pool of frames.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

TASK = "go2"
ENVS_PER_GPU = 4096
ROLLOUT = 24
POOL_FRAMES = 8
METRIC = "env-steps/sec (obs+reward+PD+reset+policy fwd) @4096 envs/GPU; PPO iter ms"


# ------------------------------------------------------------------ helpers
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = sorted(int(r[1]) for r in self.rows if len(r) > 2 and r[1].isdigit())
        mx = [int(r[2]) for r in self.rows if len(r) > 2 and r[2].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = set()
        for r in self.rows:
            for i, n in enumerate(names):
                if len(r) > 5 + i and r[5 + i].lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def measured_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


def make_frames(task, n, frames, seed=1):
    import torch
    from unitree_rl_gym_b200 import synthetic
    from unitree_rl_gym_b200.robots import ROBOTS
    from tests import helpers as H
    H.ENV_CASES["_bench"] = (task, n, True)
    try:
        spec = H.make_spec("_bench")
    finally:
        del H.ENV_CASES["_bench"]
    fr = synthetic.make_frames(ROBOTS[task], None, n, frames, spec.decimation, seed=seed, feet_idx=spec.feet_idx,
                               pen_idx=spec.pen_idx, term_idx=spec.term_idx, init_z=float(spec.base_init_state[2]),
                               default_dof_pos=torch.from_numpy(spec.default_dof_pos))
    return spec, fr


def make_env(task, n, frames, device, env_offset=0):
    from unitree_rl_gym_b200.legged_gym.envs import LeggedRobot, H1Robot, H1_2Robot, G1Robot
    from unitree_rl_gym_b200.legged_gym.utils.helpers import SimParams
    from unitree_rl_gym_b200.physics import SyntheticGym
    from unitree_rl_gym_b200.robots import ROBOTS
    from tests import helpers as H
    cfg = H.make_cfg(task, n, True)
    cfg.seed = 1
    cfg.env.env_offset = env_offset
    sp = SimParams()
    sp.dt = cfg.sim.dt
    gym = SyntheticGym(ROBOTS[task], n, cfg.control.decimation, device, frames=frames)
    cls = {"go2": LeggedRobot, "h1": H1Robot, "h1_2": H1_2Robot, "g1": G1Robot}[task]
    return cls(cfg=cfg, sim_params=sp, physics_engine=gym, sim_device=device, headless=True), cfg
