#!/usr/bin/env python
"""K2 (+K2b) and K1 timed alone: `reps` back-to-back launches captured in one CUDA graph, CUDA events on the replaying
stream (how bench.py's roofline_env / roofline_env_large legs time them), over num_envs and tasks.

    python tools/k2_bench.py [--tasks go2,g1] [--sizes 4096,262144] [--reps 20] [--once]   (--once: a few eager launches, for ncu)
"""
import argparse
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    import torch
    import bench
    from unitree_rl_gym_b200 import _lib
    ap = argparse.ArgumentParser()
    ap.add_argument("--tasks", default="go2")
    ap.add_argument("--sizes", default="4096,262144")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--once", action="store_true")
    a = ap.parse_args()
    hbm = bench.measured_peaks()[0]
    L = _lib.lib()
    for task in a.tasks.split(","):
        for n in [int(s) for s in a.sizes.split(",")]:
            spec, frames = bench.make_frames(task, n, 2)
            env, cfg = bench.make_env(task, n, frames, "cuda:0")
            env.reset()
            env.episode_length_buf = torch.randint(0, 1000, (n,), device="cuda")
            act = torch.randn(n, env.num_actions, device="cuda")
            for _ in range(3):
                env.step(act)

            def post():
                env.common_step_counter += 1
                env._obs_slot ^= 1
                env._launch_post()

            def pd():
                env._launch_pd(act)
            if a.once:
                for _ in range(3):
                    post()
                    pd()
                torch.cuda.synchronize()
                continue
            res = {"task": task, "num_envs": n, "resets_last_step": int(env.reset_buf.sum())}
            for name, fn, by in (("k2_post_physics_plus_finalize", post, bench.K2_BYTES[task]), ("k1_pd", pd, bench.K1_BYTES[task])):
                us = bench.time_graph(fn, a.reps)
                res[name] = {"us_per_launch": round(us, 2), "alg_bytes": by * n, "gbs": round(by * n / us / 1e3, 1),
                             "frac_hbm": round(by * n / us / 1e3 / hbm, 4)}
            print(json.dumps(res), flush=True)
            del env


if __name__ == "__main__":
    main()
