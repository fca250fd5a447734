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
                                          "-lms", "20", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
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


def ncu_traffic_per_launch(path=None):
    """dram__bytes_read.sum + dram__bytes_write.sum per GEMM launch, averaged over the 22 launches of one update minibatch,
    from the committed `ncu --set full` capture (tools/ncu_summary.py output); None when the file is absent."""
    try:
        import csv
        if path is None:
            cands = [os.path.join(REPO, "profiles", f) for f in ("r2_tensor_minibatch_ncu.csv", "r1_gemm_minibatch_ncu.csv")]
            path = next(c for c in cands if os.path.exists(c))
        rows = {r[0]: r for r in csv.reader(open(path))}
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot, n = 0.0, 0
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            r = rows[key]
            vals = [float(x) * scale[r[1]] for x in r[2:]]
            tot += sum(vals)
            n = len(vals)
        return tot / n
    except Exception:
        return None


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


def make_env(task, n, frames, device, env_offset=0, pool_seam=False):
    from unitree_rl_gym_b200.legged_gym.utils.helpers import SimParams
    from unitree_rl_gym_b200.physics import SyntheticGym
    from unitree_rl_gym_b200.robots import ROBOTS
    from tests import helpers as H
    cfg = H.make_cfg(task, n, True)
    cfg.seed = 1
    cfg.env.env_offset = env_offset
    sp = SimParams()
    sp.dt = cfg.sim.dt
    # (SyntheticGym(packed=True) would hold a frame's tensors in one row for single-copy uploads; measured on B200 it does not
    # shorten the one-graph e2e rollout and the eager e2e loop got slower, so the bench keeps one dense array per tensor)
    gym = SyntheticGym(ROBOTS[task], n, cfg.control.decimation, device, frames=frames, pool_seam=pool_seam)
    cls = H.env_class(task)
    return cls(cfg=cfg, sim_params=sp, physics_engine=gym, sim_device=device, headless=True), cfg


def train_cfg_dict(task, cuda_graph=False):
    from unitree_rl_gym_b200.legged_gym.envs import robot_configs as rc
    from unitree_rl_gym_b200.legged_gym.utils.helpers import class_to_dict
    t = {"go2": rc.GO2RoughCfgPPO, "h1": rc.H1RoughCfgPPO, "g1": rc.G1RoughCfgPPO, "h1_2": rc.H1_2RoughCfgPPO,
         "go2_handstand": rc.GO2HandstandCfgPPO}[task]()
    d = class_to_dict(t)
    d["runner"]["cuda_graph"] = bool(cuda_graph)
    return d


def config_dict(task, n, world):
    """The workload (identical in both arms: the driver compares it key by key)."""
    pol = "MLP actor/critic [512,256,128]" if task in ("go2", "go2_handstand") else "ActorCriticRecurrent LSTM(64)+[32]"
    return {"workload": f"{task} flat terrain, {n} envs/GPU, decimation {DECIM[task]}, {ROLLOUT}-step rollout, {pol}; "
                        "synthetic root/dof/contact state frames (PhysX is external to the path)",
            "task": task, "envs_per_gpu": n, "rollout_steps": ROLLOUT, "epochs": 5, "mini_batches": 4,
            "parallelism": f"dp{world} (env shards, flat-gradient all-reduce)" if world > 1 else "single GPU"}


def run_dict(graph, pool_seam):
    """How this arm ran the workload (not part of the workload itself)."""
    return {"rollout_launch": "CUDA graph replay (24 env steps incl. policy forward per graph; the 20 minibatch steps of "
                              "the update as a second graph, N>1: with its per-minibatch gradient exchange -- the one-shot "
                              "peer-memory all-reduce urlg_peer_allreduce, or NCCL when URLG_PEER_XCHG=0 -- as graph nodes)"
                              if graph else "eager",
            "physics_seam": ("pool-indexed: kernels read the current synthetic frame in place (urlg_env_bind_frame_pool), no "
                             "stand-in copies inside the timed region" if pool_seam else
                             "copied: one urlg_synth_refresh launch per physics tensor per refresh inside the timed region"),
            "l2_policy": "working set per iteration (35 MB rollout storage + 8-frame physics pool) is cycled; every timed "
                         "iteration additionally starts after a 256 MiB L2 flush write outside the event bracket"}


DECIM = {"go2": 4, "h1": 4, "g1": 4, "h1_2": 8, "go2_handstand": 5}
# algorithmic bytes per env: K2 post-physics per env-step, K1 per substep
K2_BYTES = {"go2": 1102, "g1": 1338, "h1_2": 1338, "h1": 1194, "go2_handstand": 1102 + 2 * 184 - 192 + 32}
K1_BYTES = {"go2": 192, "g1": 192, "h1_2": 192, "h1": 160, "go2_handstand": 192}
MLP_FLOP_PER_ENV_STEP = {"go2": 756992, "h1": 117952, "g1": 124224, "h1_2": 124224, "go2_handstand": 2 * (46 * 512 + 512 * 256 + 256 * 128) + 2 * 128 * 13}


def time_graph(fn, reps):
    """us per call: `reps` back-to-back launches captured in one CUDA graph (no host launch cost in the bracket),
    CUDA events on the replaying stream."""
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(reps):
            fn()
    gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    gr.replay()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


# ------------------------------------------------------------------ reference arm / cpu baseline (oracle port on host cores)
def cpu_run(task, n, iterations, warmup):
    import contextlib
    import torch
    from oracle import cpu_bench
    spec, frames = make_frames(task, n, POOL_FRAMES)
    with contextlib.redirect_stdout(sys.stderr):       # the port prints its module summary like rsl_rl does
        out = cpu_bench.run(spec, frames, train_cfg_dict(task), iterations, warmup, threads=os.cpu_count())
    roll = sum(out["rollout_s"]) / len(out["rollout_s"])
    upd = sum(out["update_s"]) / len(out["update_s"])
    return {"env_steps_per_s": n * ROLLOUT / roll, "rollout_ms": roll * 1e3, "update_ms": upd * 1e3,
            "iter_ms": (roll + upd) * 1e3, "threads": out["threads"], "cores": os.cpu_count()}


def reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if a.device != "cpu":            # context number only: the same eager torch code on the GPU
        r = eager_gpu_run(a.task, a.envs, max(a.steps, 1), max(a.warmup, 1))
        print(json.dumps({"impl": "reference", "device": a.device, "metric": METRIC, "value": r["value"], "unit": r["unit"],
                          "n_gpus": 1, "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["ppo_iter_ms"],
                          "higher_is_better": True, "config": config_dict(a.task, a.envs, 1), "eager_gpu": r}), flush=True)
        return
    r = cpu_run(a.task, a.envs, max(a.steps, 1), a.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": r["env_steps_per_s"], "unit": "env-steps/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["iter_ms"],
            "ppo_iter_ms": r["iter_ms"], "rollout_ms": r["rollout_ms"], "update_ms": r["update_ms"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(a.task, a.envs, a.gpus),
            "cpu_baseline": {"value": r["env_steps_per_s"], "unit": "env-steps/s", "cores": r["threads"], "kind": "port",
                             "sample": f"{a.steps} full PPO iteration(s) ({a.envs} envs x {ROLLOUT} steps + 5x4 minibatch "
                                       f"update) after {a.warmup} warm-up, torch CPU with {r['threads']} threads of "
                                       f"{r['cores']} cores; the unmodified reference cannot run on this box (imports "
                                       "isaacgym; /root/reference absent), so the oracle port pinned to it by "
                                       "tests/golden stands in"},
            "e2e": {"value": r["env_steps_per_s"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def quick_config(task, n, steps, dev, world, rank, pool_seam=True):
    """One of BASELINE.json's other configs (H1 / G1 at 4096 envs, H1_2 at 16384 envs per GPU) measured the same way as the
    headline: rollout and update replayed as CUDA graphs, CUDA events per phase, max over ranks, L2 flush between
    iterations.  Returns the sub-object stored under `configs` in the JSON line."""
    import torch
    import torch.distributed as dist
    from unitree_rl_gym_b200.rsl_rl.runners import OnPolicyRunner
    from unitree_rl_gym_b200.rsl_rl.runners.rollout_graph import RolloutGraph
    spec, frames = make_frames(task, n, POOL_FRAMES, seed=1 + rank)
    env, cfg = make_env(task, n, frames, dev, env_offset=rank * n, pool_seam=pool_seam)
    runner = OnPolicyRunner(env, train_cfg_dict(task), log_dir=None, device=dev)
    alg = runner.alg
    env.episode_length_buf = torch.randint_like(env.episode_length_buf, high=int(env.max_episode_length))
    alg.cuda_graph_update = True
    T = ROLLOUT

    def eager():
        obs = env.get_observations()
        priv = env.get_privileged_observations()
        cobs = priv if priv is not None else obs
        with torch.inference_mode():
            for _ in range(T):
                act = alg.act(obs, cobs)
                obs, priv, rew, dones, infos = env.step(act)
                cobs = priv if priv is not None else obs
                alg.process_env_step(rew, dones, infos)
        return obs, cobs

    def finish(cobs):
        with torch.inference_mode():
            alg.compute_returns(cobs)
        return alg.update()
    _, cobs = eager()
    finish(cobs)
    graph = RolloutGraph(runner).capture()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(3):
        _, cobs = graph.replay()
        finish(cobs)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    marks = []
    for _ in range(steps):
        flush.zero_()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        _, cobs = graph.replay()
        e[1].record()
        finish(cobs)
        e[2].record()
        marks.append(e)
    torch.cuda.synchronize()
    t = torch.tensor([sum(a.elapsed_time(b) for a, b, _ in marks) / steps, sum(b.elapsed_time(c) for _, b, c in marks) / steps],
                     dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    roll_ms, upd_ms = t.tolist()
    out = {"config": config_dict(task, n, world), "value": n * T * world / (roll_ms * 1e-3), "unit": "env-steps/s",
           "rollout_ms": roll_ms, "update_ms": upd_ms, "ppo_iter_ms": roll_ms + upd_ms,
           "train_env_steps_per_s": n * T * world / ((roll_ms + upd_ms) * 1e-3), "steps": steps, "n_gpus": world,
           "update_tensor_frac": MLP_FLOP_PER_ENV_STEP[task] * 3 * n * T * 5 / (upd_ms * 1e-3) / 1e12 / measured_peaks()[2]}
    if world > 1:
        alg.release_graphs()
    del graph, runner, env, alg
    torch.cuda.empty_cache()
    return out


def eager_gpu_run(task, n, iterations=1, warmup=1):
    """The oracle's eager torch code (what the reference runs: ~300 launches and 3 host syncs per env step) on the GPU --
    the context number SURVEY.md 0.4 / BASELINE.md 4 name.  Not the comparison anchor (that is the CPU arm)."""
    import contextlib
    from oracle import cpu_bench
    spec, frames = make_frames(task, n, POOL_FRAMES)
    with contextlib.redirect_stdout(sys.stderr):
        out = cpu_bench.run(spec, frames, train_cfg_dict(task), iterations, warmup, device="cuda")
    roll = sum(out["rollout_s"]) / len(out["rollout_s"])
    it = sum(out["iter_wall_s"]) / len(out["iter_wall_s"])
    return {"value": n * ROLLOUT / roll, "unit": "env-steps/s", "rollout_ms": roll * 1e3, "ppo_iter_ms": it * 1e3,
            "kind": "port", "device": "cuda", "sample": f"{iterations} PPO iteration(s) after {warmup} warm-up, eager torch "
            "(oracle port of LeggedRobot.step + rsl_rl v1.0.2) on the same B200"}


# ------------------------------------------------------------------ own arm
def own_arm(a):
    import torch
    import torch.distributed as dist
    from unitree_rl_gym_b200 import _lib
    from unitree_rl_gym_b200.rsl_rl.runners import OnPolicyRunner
    from unitree_rl_gym_b200.rsl_rl.runners.rollout_graph import RolloutGraph

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device (the product path has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    L = _lib.lib()
    task, n, T = a.task, a.envs, ROLLOUT
    hbm, tf_burst, tf_sus, peak_src = measured_peaks()

    spec, frames = make_frames(task, n, POOL_FRAMES, seed=1 + rank)
    env, cfg = make_env(task, n, frames, dev, env_offset=rank * n, pool_seam=not a.copy_seam)
    runner = OnPolicyRunner(env, train_cfg_dict(task), log_dir=None, device=dev)
    alg = runner.alg
    env.episode_length_buf = torch.randint_like(env.episode_length_buf, high=int(env.max_episode_length))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    use_graph = (not a.no_graph) and RolloutGraph.supported(runner)
    alg.cuda_graph_update = use_graph      # PPO.update replays its 20 minibatch steps as one graph (N>1: NCCL all-reduces captured too unless URLG_GRAPH_NCCL=0)

    def eager_rollout():
        obs = env.get_observations()
        priv = env.get_privileged_observations()
        cobs = priv if priv is not None else obs
        with torch.inference_mode():
            for _ in range(T):
                act = alg.act(obs, cobs)
                obs, priv, rew, dones, infos = env.step(act)
                cobs = priv if priv is not None else obs
                alg.process_env_step(rew, dones, infos)
        return obs, cobs

    def finish(cobs):
        with torch.inference_mode():
            alg.compute_returns(cobs)
        return alg.update()

    # one eager iteration first (allocations, lazy module loads), then capture
    _, cobs = eager_rollout()
    finish(cobs)
    graph = RolloutGraph(runner).capture() if use_graph else None
    rollout = (lambda: graph.replay()) if graph is not None else eager_rollout

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        _, cobs = rollout()
        finish(cobs)
    barrier()

    # ---- timed region: exactly K iterations; device events per phase; wall clock across (barrier + sync both sides)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    marks = []
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n_launch0 = L.urlg_launch_count()
    barrier()
    w0 = ev(); w0.record()
    t_wall0 = time.perf_counter()
    for _ in range(a.steps):
        flush.zero_()                      # L2 flush, outside the per-phase brackets (inside the wall bracket)
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        _, cobs = rollout()
        e1.record()
        finish(cobs)
        e2.record()
        marks.append((e0, e1, e2))
    w1 = ev(); w1.record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    eager_launches = L.urlg_launch_count() - n_launch0
    clocks = sampler.stop() if rank == 0 else None
    roll_ms = sum(e0.elapsed_time(e1) for e0, e1, _ in marks) / a.steps
    upd_ms = sum(e1.elapsed_time(e2) for _, e1, e2 in marks) / a.steps
    total_ms = w0.elapsed_time(w1) / a.steps
    t = torch.tensor([roll_ms, upd_ms, total_ms, t_wall * 1e3 / a.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    roll_ms, upd_ms, total_ms, wall_ms = t.tolist()
    launches = eager_launches + (graph.launches * a.steps if graph is not None else 0)
    if alg.cuda_graph_update and alg._update_graph is not None:
        launches += alg.update_graph_launches * a.steps

    # ---- e2e: the same rollout through the public API (eager), physics frames arriving from pinned HOST memory
    #      every env step (H2D inside the timed region) and rewards / dones read back to the host every step
    if env.gym.packed:          # one pinned host row per frame, one copy per upload
        pin = {"frame": env.gym.frame_buf.cpu().pin_memory()}
        pool = {"frame": env.gym.frame_buf}
    else:
        pin = {k: v.pin_memory() for k, v in frames.items() if v is not None and (k != "rigid_body" or spec.family == 1)}
        pool = env.gym.frames
    D = spec.decimation
    rew_h = torch.empty(n, dtype=torch.float, pin_memory=True)
    done_h = torch.empty(n, dtype=torch.bool, pin_memory=True)
    h2d = sum(v[0].numel() * 4 for v in pin.values())
    d2h = n * 4 + n
    # the host link alone: the same pinned frames copied back to back (floor of the e2e rollout = 24 frames / this rate)
    torch.cuda.synchronize()
    c0, c1 = ev(), ev()
    c0.record()
    for r in range(2 * POOL_FRAMES):
        for k, v in pin.items():
            pool[k][r % POOL_FRAMES].copy_(v[r % POOL_FRAMES], non_blocking=True)
    c1.record()
    c1.synchronize()
    h2d_gbs = h2d * 2 * POOL_FRAMES / (c0.elapsed_time(c1) * 1e-3) / 1e9

    def e2e_rollout():
        obs = env.get_observations()
        priv = env.get_privileged_observations()
        cobs = priv if priv is not None else obs
        with torch.inference_mode():
            for i in range(T):
                f = ((env.gym.sim_calls + base_host) // D) % POOL_FRAMES
                for k, v in pin.items():
                    pool[k][f].copy_(v[f], non_blocking=True)
                act = alg.act(obs, cobs)
                obs, priv, rew, dones, infos = env.step(act)
                cobs = priv if priv is not None else obs
                alg.process_env_step(rew, dones, infos)
                rew_h.copy_(rew, non_blocking=True)
                done_h.copy_(dones, non_blocking=True)
        return obs, cobs

    env.gym.begin_relative()
    base_host = int(env.gym.calls_base.item())      # device part of the simulate() counter (advanced by the graph replays)
    e2e_ms = []
    for i in range(2 + min(a.steps, 5)):
        barrier()
        e0, e1 = ev(), ev()
        e0.record()
        _, cobs = e2e_rollout()
        e1.record()
        e1.synchronize()
        if i >= 2:
            e2e_ms.append(e0.elapsed_time(e1))
        finish(cobs)
    t = torch.tensor([sum(e2e_ms) / len(e2e_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_eager_ms = t.item()
    e2e_roll_ms, e2e_mode = e2e_eager_ms, "eager"

    # ---- e2e, per-step graphs (public API: rsl_rl.runners.rollout_graph.StepGraphs): one graph launch per env step; the
    #      frame of step k+1 is copied from pinned host memory on a copy stream while step k computes; rewards / dones
    #      of every step are read back to pinned host memory.  Same bytes per step as the eager loop.
    if use_graph:
        from unitree_rl_gym_b200.rsl_rl.runners.rollout_graph import StepGraphs
        sg = StepGraphs(runner).capture()
        copy_stream = torch.cuda.Stream(device=dev)
        cur = torch.cuda.current_stream()
        slot_free = [torch.cuda.Event() for _ in range(POOL_FRAMES)]
        for e_ in slot_free:
            e_.record(cur)
        frame_ready = [torch.cuda.Event() for _ in range(T)]
        # rewards / dones of env step i are read back from their rollout-storage rows (written by that step's graph and
        # not touched again during the rollout) on a third stream, so the read-back never stalls step i + 1
        d2h_stream = torch.cuda.Stream(device=dev)
        step_done = [torch.cuda.Event() for _ in range(T)]
        st_ = alg.storage
        rew_rows = torch.empty(T, n, dtype=torch.float, pin_memory=True)
        done_rows = torch.empty(T, n, dtype=st_.dones.dtype, pin_memory=True)
        step_no = {"k": (int(env.gym.calls_base.item()) + env.gym.sim_calls) // D}

        h2d_batch = [_lib.CopyBatch([(pool[key][f], v[f]) for key, v in pin.items()]) for f in range(POOL_FRAMES)]
        d2h_batch = [_lib.CopyBatch([(rew_rows[i], st_.rewards[i]), (done_rows[i], st_.dones[i])]) for i in range(T)]

        def stage_frame(k, idx):
            f = k % POOL_FRAMES
            copy_stream.wait_event(slot_free[f])              # the last step that read this pool slot has finished
            h2d_batch[f].enqueue(copy_stream)                 # one C call: root, dof x decimation, contact (, rigid body)
            frame_ready[idx].record(copy_stream)

        def e2e_graph_rollout():
            k0 = step_no["k"]
            cur.wait_stream(d2h_stream)                       # last rollout's read-backs are done with the storage rows
            stage_frame(k0, 0)
            with torch.inference_mode():
                for i in range(T):
                    if i + 1 < T:
                        stage_frame(k0 + i + 1, i + 1)        # overlaps with the compute of step i
                    cur.wait_event(frame_ready[i])
                    obs, cobs, rew, dones = sg.replay(i)
                    slot_free[(k0 + i) % POOL_FRAMES].record(cur)
                    step_done[i].record(cur)
                    d2h_stream.wait_event(step_done[i])
                    d2h_batch[i].enqueue(d2h_stream)
            cur.wait_stream(d2h_stream)                       # the timed region ends after the last read-back
            step_no["k"] = k0 + T
            return obs, cobs

        def timed(fn):
            ms = []
            for i in range(2 + min(a.steps, 5)):
                barrier()
                e0, e1 = ev(), ev()
                e0.record()
                _, cobs = fn()
                e1.record()
                e1.synchronize()
                if i >= 2:
                    ms.append(e0.elapsed_time(e1))
                finish(cobs)
            t = torch.tensor([sum(ms) / len(ms)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()

        e2e_steps_ms = timed(e2e_graph_rollout)
        e2e_roll_ms = e2e_steps_ms
        e2e_mode = ("one CUDA graph per env step (StepGraphs), next frame's H2D overlapped on a copy stream, per-step "
                    "reward/done read-back on a third stream")

        # ---- e2e, whole rollout as ONE graph with the host traffic inside it (RolloutGraph.capture hooks): the upload
        #      of frame t + 1 is a memcpy node on a forked branch beside the kernels of step t, the read-back of step t's
        #      rewards / dones a node on a second branch.  Needs the frame ring to divide the rollout (24 % 8 == 0), so
        #      that the slot a captured step reads is the same on every replay.
        e2e_one_ms = None
        if T % POOL_FRAMES == 0:
            k0 = step_no["k"]
            fr, sd = [torch.cuda.Event() for _ in range(T)], [torch.cuda.Event() for _ in range(T)]

            def pre_step(i):
                c = torch.cuda.current_stream()
                if i == 0:
                    copy_stream.wait_stream(c)                # fork
                    d2h_stream.wait_stream(c)
                    h2d_batch[k0 % POOL_FRAMES].enqueue(copy_stream)
                    fr[0].record(copy_stream)
                if i + 1 < T:
                    if i + 1 >= POOL_FRAMES:                  # ring slot (i + 1) % 8 was last read by step i + 1 - 8:
                        copy_stream.wait_event(sd[i + 1 - POOL_FRAMES])   # uploads run up to a ring ahead of the compute
                    h2d_batch[(k0 + i + 1) % POOL_FRAMES].enqueue(copy_stream)
                    fr[i + 1].record(copy_stream)
                c.wait_event(fr[i])

            def post_step(i):
                c = torch.cuda.current_stream()
                sd[i].record(c)
                d2h_stream.wait_event(sd[i])
                d2h_batch[i].enqueue(d2h_stream)

            def join():
                c = torch.cuda.current_stream()
                c.wait_stream(copy_stream)
                c.wait_stream(d2h_stream)

            g1 = RolloutGraph(runner).capture(pre_step, post_step, join)
            e2e_one_ms = timed(g1.replay)
            if e2e_one_ms < e2e_roll_ms:
                e2e_roll_ms = e2e_one_ms
                e2e_mode = ("ONE CUDA graph per 24-step rollout (RolloutGraph with capture hooks): per-step H2D frame "
                            "uploads and reward/done read-backs are memcpy nodes on forked branches of the same graph")

    # ---- roofline of the dominant kernels (K3 / K7 are tensor-bound, 1.12 TFLOP per Go2 iteration).
    #      Every tensor-core launch of ONE eager iteration (the fused multi-layer forward of the rollout and of the update,
    #      the dgrad / wgrad GEMMs) is bracketed by CUDA events on its stream by the library hook urlg_gemm_timing (one
    #      launch per bracket, in situ); achieved = sum of algorithmic flops (2*M*N*K once, although split-bf16 issues three
    #      MMAs per product) / sum of durations, against the sustained bf16 peak.  The HBM view of the same launches and
    #      the path's own algorithmic bytes stand beside it.  The fused env kernels (K2 + K2b, K1) are timed alone under
    #      graph replay against the HBM peak: `roofline_env*`.
    roof = None
    extra = {}
    import ctypes as C

    def gemm_timed(fn):
        _lib.check(L.urlg_gemm_timing(1))
        out = fn()
        torch.cuda.synchronize()
        ms, cnt, fl, by = C.c_double(), C.c_int64(), C.c_double(), C.c_double()
        _lib.check(L.urlg_gemm_timing_read(C.byref(ms), C.byref(cnt), C.byref(fl), C.byref(by)))
        _lib.check(L.urlg_gemm_timing(0))
        return out, ms.value, cnt.value, (fl.value, by.value)
    # every rank runs this iteration (the update holds collectives); rank 0 reports its own times
    alg.cuda_graph_update = False          # the hook needs real launches, not graph replays
    (_, cobs), ms_r, n_r, fl_r = gemm_timed(eager_rollout)
    _, ms_u, n_u, fl_u = gemm_timed(lambda: finish(cobs))
    barrier()
    if rank == 0:
        (fl_r, by_r), (fl_u, by_u) = fl_r, fl_u
        sec = max((ms_r + ms_u) * 1e-3, 1e-12)
        tf, gbs = (fl_r + fl_u) / sec / 1e12, (by_r + by_u) / sec / 1e9
        obs_w, act_w = spec.num_obs, spec.num_dof
        crit_w = spec.num_priv if spec.num_priv else spec.num_obs
        # path-level algorithmic HBM bytes of an iteration's update: every minibatch row's gathered
        # columns once per epoch + parameters, gradients and both Adam moments once per minibatch step
        n_params = sum(p_.numel() for p_ in alg.actor_critic.parameters())
        row_bytes = 4 * (obs_w + (crit_w if spec.num_priv else 0) + 3 * act_w + 4)
        path_alg = 5 * n * T * row_bytes + 20 * n_params * 4 * 7
        ncu_traffic = ncu_traffic_per_launch()
        sub = lambda ms_, n_, fl_, by_: {"launches": int(n_), "ms": ms_, "tflops": fl_ / max(ms_ * 1e-3, 1e-12) / 1e12,
                                         "gbs": by_ / max(ms_ * 1e-3, 1e-12) / 1e9}
        roof = {"kernel": "tcgen05 kernels of the step: fused::mlp_fused_fwd_kernel (all four layers of actor + critic per "
                          "128-row tile: rollout policy forward + update forward) and tc::gemm_bf16_kernel (dgrad / wgrad)",
                "bound": "tensor", "achieved": tf, "peak": tf_sus, "unit": "TFLOP/s", "frac": tf / tf_sus,
                "traffic": ncu_traffic,
                "traffic_scope": "ncu --set full capture of the tensor-core launches of one update minibatch "
                                 "(profiles/): DRAM read + write per launch",
                "alg_flops_per_launch": (fl_r + fl_u) / max(n_r + n_u, 1),
                "alg_bytes_per_launch": (by_r + by_u) / max(n_r + n_u, 1),
                "launches_per_step": int(n_r + n_u), "us_per_launch": sec * 1e6 / max(n_r + n_u, 1),
                "hbm_view": {"achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                             "note": "per-launch operand + output bytes of the same launches over the same time"},
                "path_alg_bytes": path_alg,
                "path_alg_bytes_note": "update phase per iteration: gathered minibatch rows (5 epochs) + parameters, "
                                       "gradients, Adam moments per minibatch step; launch-level bytes above it are "
                                       "activations saved by the forward and re-read by the backward",
                "launch_bytes_over_path_bytes": by_u / max(path_alg, 1),
                "arithmetic_intensity_flop_per_byte": (fl_r + fl_u) / max(by_r + by_u, 1),
                "ridge_flop_per_byte": tf_sus * 1e12 / (hbm * 1e9), "peak_source": peak_src,
                "rollout": sub(ms_r, n_r, fl_r, by_r), "update": sub(ms_u, n_u, fl_u, by_u),
                "note": "one eager PPO iteration, one launch per event bracket; flops counted once (3 MMAs per product "
                        "in split-bf16), so frac is the fraction of the bf16 peak an fp32-faithful result reaches"}
        iters = 200

        time_call = time_graph

        def post():
            env.common_step_counter += 1
            env._obs_slot ^= 1
            env._launch_post()
        raw_act = torch.randn(n, env.num_actions, device=dev)
        us_post = time_call(post, iters)
        us_pd = time_call(lambda: env._launch_pd(raw_act), iters)
        b2 = K2_BYTES[task] * n
        extra["roofline_env"] = {
            "kernel": "post_physics_kernel (+finalize_step_kernel), K2 of SURVEY.md 8(a).7", "bound": "hbm",
            "achieved": b2 / us_post / 1e3, "peak": hbm, "unit": "GB/s", "frac": b2 / us_post / 1e3 / hbm,
            "traffic": None, "alg_bytes_per_launch": b2, "us_per_launch": us_post, "peak_source": peak_src,
            "note": f"{n} envs x {K2_BYTES[task]} B = {b2 / 1e6:.1f} MB is L2-resident and below launch latency "
                    "(SURVEY.md 7): HBM fraction at this size measures latency, not bandwidth; see roofline_env_large"}
        extra["k1_pd"] = {"us_per_launch": us_pd, "alg_bytes": K1_BYTES[task] * n,
                          "gbs": K1_BYTES[task] * n / us_pd / 1e3}
        # tensor side: update phase flops / time
        flop_upd = MLP_FLOP_PER_ENV_STEP[task] * 3 * n * T * 5
        extra["update_tensor"] = {"bound": "tensor", "achieved": flop_upd / (upd_ms * 1e-3) / 1e12, "peak": tf_sus,
                                  "unit": "TFLOP/s", "frac": flop_upd / (upd_ms * 1e-3) / 1e12 / tf_sus,
                                  "note": "whole update phase (20 minibatches: gather, fwd, head, bwd, Adam; split-bf16 = "
                                          "3 MMAs per product, counted once) vs sustained bf16 peak"}
        flop_fwd = MLP_FLOP_PER_ENV_STEP[task] * n * T
        extra["rollout_policy_flops_frac"] = flop_fwd / (roll_ms * 1e-3) / 1e12 / tf_sus
        if not a.no_large:
            del raw_act
            big = a.large_envs
            spec_b, frames_b = make_frames(task, big, 2)
            env_b, _ = make_env(task, big, frames_b, dev)
            env_b.reset()
            env_b.episode_length_buf = torch.randint(0, 1000, (big,), device=dev)
            act_b = torch.randn(big, env_b.num_actions, device=dev)
            for _ in range(2):
                env_b.step(act_b)

            def post_b():
                env_b.common_step_counter += 1
                env_b._obs_slot ^= 1
                env_b._launch_post()
            us_b = time_call(post_b, 20)
            us_pd_b = time_call(lambda: env_b._launch_pd(act_b), 20)
            bb = K2_BYTES[task] * big
            extra["roofline_env_large"] = {"kernel": "post_physics_kernel (+finalize)", "envs": big, "bound": "hbm",
                                       "achieved": bb / us_b / 1e3, "peak": hbm, "unit": "GB/s",
                                       "frac": bb / us_b / 1e3 / hbm, "us_per_launch": us_b,
                                       "alg_bytes_per_launch": bb,
                                       "k1_pd": {"us_per_launch": us_pd_b, "gbs": K1_BYTES[task] * big / us_pd_b / 1e3,
                                                 "frac": K1_BYTES[task] * big / us_pd_b / 1e3 / hbm},
                                       "note": "working set >> 126 MB L2, back-to-back launches"}
            del env_b, frames_b, act_b
            torch.cuda.empty_cache()

    # ---- BASELINE.json configs[2..4]: H1 and G1 at 4096 envs/GPU, H1_2 at 16384 envs/GPU (every rank takes part: the
    #      update holds the gradient all-reduce), so the driver's single run records them at every N
    configs = None
    if not a.no_configs and task == TASK:
        configs = {}
        for name, tk, ne in (("h1_4096", "h1", 4096), ("g1_4096", "g1", 4096), ("h1_2_16384", "h1_2", 16384)):
            try:
                configs[name] = quick_config(tk, ne, max(3, min(a.steps, 5)), dev, world, rank, pool_seam=not a.copy_seam)
            except Exception as ex:              # a sub-benchmark must not take the headline line down with it
                configs[name] = {"error": repr(ex)[:300]}
        barrier()

    # ---- eager torch on the same GPU (context, rank 0 at N=1 only)
    eager_gpu = None
    if rank == 0 and world == 1 and not a.no_eager_gpu:
        try:
            eager_gpu = eager_gpu_run(task, n, 1, 1)
        except Exception as ex:
            eager_gpu = {"error": repr(ex)[:300]}

    # ---- CPU baseline (rank 0, N=1 only): the oracle port on the host cores, bounded sample
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu:
        r = cpu_run(task, n, 1, 1)
        cpu = {"value": r["env_steps_per_s"], "unit": "env-steps/s", "cores": r["threads"], "kind": "port",
               "ppo_iter_ms": r["iter_ms"], "rollout_ms": r["rollout_ms"], "update_ms": r["update_ms"],
               "sample": f"1 full PPO iteration ({n} envs x {T} steps + 5x4 minibatch update) after 1 warm-up iteration, "
                         f"torch CPU {r['threads']} threads / {r['cores']} cores (oracle port; see --impl reference)"}

    if rank == 0:
        value = n * T * world / (roll_ms * 1e-3)
        line = {"metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": a.steps,
                "warmup": max(a.warmup, 3), "ms_per_step": total_ms, "ppo_iter_ms": total_ms, "rollout_ms": roll_ms,
                "update_ms": upd_ms, "wall_ms_per_step": wall_ms, "train_env_steps_per_s": n * T * world / (total_ms * 1e-3),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32 env path; split-bf16 (hi+lo) tensor-core operands, fp32 accumulate, policy/PPO",
                "data": "synthetic", "config": config_dict(task, n, world), "run": dict(run_dict(graph is not None, not a.copy_seam), gradient_exchange=alg.exchange),
                "clocks": clocks, "gpu_launches": int(launches),
                "e2e": {"value": n * T * world / (e2e_roll_ms * 1e-3), "unit": "env-steps/s",
                        "h2d_bytes_per_step": int(h2d * T), "d2h_bytes_per_step": int(d2h * T), "rollout_ms": e2e_roll_ms,
                        "launch": e2e_mode, "step_graphs_ms": e2e_steps_ms if use_graph else None,
                        "one_graph_ms": e2e_one_ms if use_graph else None, "h2d_link_gbs": h2d_gbs, "h2d_floor_ms": h2d * T / h2d_gbs / 1e6,
                        "eager": {"value": n * T * world / (e2e_eager_ms * 1e-3), "rollout_ms": e2e_eager_ms,
                                  "launch": "eager PPO.act / LeggedRobot.step / PPO.process_env_step, ~40 launches per env step"},
                        "note": "public API with HOST physics frames: per env step the frame (root, dof x decimation, contact) "
                                "is copied from pinned host memory and rewards + dones are read back to pinned host memory; "
                                "bytes are per 24-step rollout (one bench step)"},
                "roofline": roof, "cpu_baseline": cpu, "eager_gpu": eager_gpu, "configs": configs}
        line.update(extra)
        print(json.dumps(line), flush=True)
    if world > 1:
        alg.release_graphs()               # the update graph holds NCCL kernels: destroy it before the communicator
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--task", default=TASK)
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU)
    ap.add_argument("--no-graph", action="store_true", help="eager rollout instead of CUDA-graph replay")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-large", action="store_true", help="skip the large-N roofline point")
    ap.add_argument("--no-configs", action="store_true", help="skip the H1 / G1 / H1_2 sub-benchmarks")
    ap.add_argument("--no-eager-gpu", action="store_true", help="skip the eager-torch-on-GPU context number")
    ap.add_argument("--copy-seam", action="store_true",
                    help="serve synthetic frames by copying them into the physics tensors (urlg_synth_refresh launches "
                         "inside the timed region) instead of the pool-indexed seam")
    ap.add_argument("--device", default="cpu", help="--impl reference only: cpu (the arm the driver compares) or cuda "
                                                    "(the same eager torch code on the GPU, context number)")
    ap.add_argument("--large-envs", type=int, default=1 << 18)
    a = ap.parse_args()
    if a.impl == "reference":
        reference_arm(a)
    else:
        own_arm(a)


if __name__ == "__main__":
    main()
