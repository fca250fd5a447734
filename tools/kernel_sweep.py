#!/usr/bin/env python
"""Per-kernel timing sweep of the env half (K1 PD torques, K2 fused post-physics) over num_envs,
CUDA events on the launching stream.  At 4096 envs the working set is L2-resident and below launch
latency, so the HBM fraction is only meaningful at the large sizes; all are printed.

    python tools/kernel_sweep.py [--task go2] [--sizes 4096,65536,1048576] [--iters 50]
"""
import argparse
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

# algorithmic bytes per env: K1 per substep, K2 per env-step
K1_BYTES = {"go2": 192, "g1": 192, "h1_2": 192, "h1": 160}
K2_BYTES = {"go2": 1102, "g1": 1338, "h1_2": 1338, "h1": 1194}


def main():
    import torch
    import bench
    from unitree_rl_gym_b200 import _lib
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="go2")
    ap.add_argument("--sizes", default="4096,65536,1048576")
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    hbm, _, _, src = bench.measured_peaks()
    L = _lib.lib()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for n in [int(s) for s in a.sizes.split(",")]:
        spec, frames = bench.make_frames(a.task, n, 2)
        env, cfg = bench.make_env(a.task, n, frames, "cuda:0")
        env.reset()
        env.episode_length_buf = torch.randint(0, 1000, (n,), device="cuda")
        act = torch.randn(n, env.num_actions, device="cuda")
        for _ in range(3):
            env.step(act)
        st = _lib.cur_stream()
        res = {"task": a.task, "num_envs": n}
        for name in ("pd", "post"):
            ts = []
            for _ in range(a.iters):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                if name == "pd":
                    _lib.check(L.urlg_pd_torques(env._handle, _lib.ptr(act), st))
                else:
                    env.common_step_counter += 1
                    env._obs_slot ^= 1
                    _lib.check(L.urlg_post_physics(env._handle, _lib.ptr(env._step_base), env.common_step_counter,
                                                   env._obs_slot, st))
                e1.record()
                e1.synchronize()
                ts.append(e0.elapsed_time(e1) * 1e3)
            ts.sort()
            us = ts[len(ts) // 2]
            b = (K1_BYTES if name == "pd" else K2_BYTES)[a.task] * n
            res[name] = {"us_median": round(us, 2), "us_min": round(ts[0], 2), "alg_bytes": b,
                         "gbs": round(b / us / 1e3, 1), "frac_hbm": round(b / us / 1e3 / hbm, 4)}
        res["peak_hbm_gbs"] = hbm
        res["peak_source"] = src
        res["l2"] = "flushed (256 MiB write) before every timed launch"
        print(json.dumps(res), flush=True)
        del env


if __name__ == "__main__":
    main()
