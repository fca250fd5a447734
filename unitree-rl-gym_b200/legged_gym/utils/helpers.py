"""Host glue the drop-in classes need (legged_gym/utils/helpers.py:11-47,73-95,150-189).
Argument parsing goes through argparse instead of isaacgym.gymutil."""
import argparse
import copy
import os
import random

import numpy as np
import torch


def class_to_dict(obj) -> dict:
    """Recursive namespace -> dict in dir() (alphabetical) order: this ordering is what fixes
    the reward summation order (helpers.py:15)."""
    if not hasattr(obj, "__dict__"):
        return obj
    out = {}
    for key in dir(obj):
        if key.startswith("_"):
            continue
        val = getattr(obj, key)
        out[key] = [class_to_dict(v) for v in val] if isinstance(val, list) else class_to_dict(val)
    return out


def update_class_from_dict(obj, d):
    for key, val in d.items():
        attr = getattr(obj, key, None)
        if isinstance(attr, type):
            update_class_from_dict(attr, val)
        else:
            setattr(obj, key, val)


def set_seed(seed):
    if seed == -1:
        seed = np.random.randint(0, 10000)
    print("Setting seed: {}".format(seed))
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    os.environ['PYTHONHASHSEED'] = str(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


class SimParams:
    """Plain stand-in for gymapi.SimParams (only dt / use_gpu_pipeline are read on the hot path)."""
    def __init__(self):
        self.dt = 1. / 60.
        self.substeps = 2
        self.up_axis = 1
        self.gravity = [0., 0., -9.81]
        self.use_gpu_pipeline = True
        self.physx = type("physx", (), {})()


def parse_sim_params(args, cfg):
    sp = SimParams()
    sp.use_gpu_pipeline = getattr(args, "use_gpu_pipeline", True)

    def fill(dst, src):
        for k, v in src.items():
            if isinstance(v, dict):
                sub = getattr(dst, k, None)
                if sub is None:
                    sub = type(k, (), {})()
                    setattr(dst, k, sub)
                fill(sub, v)
            else:
                setattr(dst, k, v)
    if "sim" in cfg:
        fill(sp, cfg["sim"])
    return sp


def get_load_path(root, load_run=-1, checkpoint=-1):
    try:
        runs = sorted(os.listdir(root))
        if 'exported' in runs:
            runs.remove('exported')
        last_run = os.path.join(root, runs[-1])
    except Exception:
        raise ValueError("No runs in this directory: " + root)
    load_run = last_run if load_run == -1 else os.path.join(root, load_run)
    if checkpoint == -1:
        models = [f for f in os.listdir(load_run) if 'model' in f]
        models.sort(key=lambda m: '{0:0>15}'.format(m))
        model = models[-1]
    else:
        model = "model_{}.pt".format(checkpoint)
    return os.path.join(load_run, model)


def update_cfg_from_args(env_cfg, cfg_train, args):
    if env_cfg is not None and getattr(args, "num_envs", None) is not None:
        env_cfg.env.num_envs = args.num_envs
    if cfg_train is not None:
        if getattr(args, "seed", None) is not None:
            cfg_train.seed = args.seed
        for name in ("max_iterations", "experiment_name", "run_name", "load_run", "checkpoint"):
            if getattr(args, name, None) is not None:
                setattr(cfg_train.runner, name, getattr(args, name))
        if getattr(args, "resume", False):
            cfg_train.runner.resume = args.resume
    return env_cfg, cfg_train


def get_args(argv=None):
    p = argparse.ArgumentParser(description="RL Policy")
    p.add_argument("--task", type=str, default="go2")
    p.add_argument("--resume", action="store_true", default=False)
    p.add_argument("--experiment_name", type=str)
    p.add_argument("--run_name", type=str)
    p.add_argument("--load_run", type=str)
    p.add_argument("--checkpoint", type=int)
    p.add_argument("--headless", action="store_true", default=False)
    p.add_argument("--horovod", action="store_true", default=False)
    p.add_argument("--rl_device", type=str, default="cuda:0")
    p.add_argument("--sim_device", type=str, default="cuda:0")
    p.add_argument("--num_envs", type=int)
    p.add_argument("--seed", type=int)
    p.add_argument("--max_iterations", type=int)
    args = p.parse_args(argv)
    args.physics_engine = "synthetic"
    args.use_gpu_pipeline = args.sim_device.startswith("cuda")
    kind, _, idx = args.sim_device.partition(":")
    args.sim_device_type, args.sim_device_id = kind, int(idx or 0)
    return args


def export_policy_as_jit(actor_critic, path):
    if hasattr(actor_critic, 'memory_a'):
        PolicyExporterLSTM(actor_critic).export(path)
    else:
        os.makedirs(path, exist_ok=True)
        model = copy.deepcopy(actor_critic.actor).to('cpu')
        torch.jit.script(model).save(os.path.join(path, 'policy_1.pt'))


class PolicyExporterLSTM(torch.nn.Module):
    """Stateful LSTM + actor head as one scriptable module; same buffer / method names as the
    shipped deploy/pre_train/*/motion.pt (hidden_state, cell_state, reset_memory)."""
    def __init__(self, actor_critic):
        super().__init__()
        self.actor = copy.deepcopy(actor_critic.actor)
        self.is_recurrent = actor_critic.is_recurrent
        self.memory = copy.deepcopy(actor_critic.memory_a.rnn)
        self.memory.cpu()
        self.register_buffer('hidden_state', torch.zeros(self.memory.num_layers, 1, self.memory.hidden_size))
        self.register_buffer('cell_state', torch.zeros(self.memory.num_layers, 1, self.memory.hidden_size))

    def forward(self, x):
        out, (h, c) = self.memory(x.unsqueeze(0), (self.hidden_state, self.cell_state))
        self.hidden_state[:] = h
        self.cell_state[:] = c
        return self.actor(out.squeeze(0))

    @torch.jit.export
    def reset_memory(self):
        self.hidden_state[:] = 0.
        self.cell_state[:] = 0.

    def export(self, path):
        os.makedirs(path, exist_ok=True)
        self.to('cpu')
        torch.jit.script(self).save(os.path.join(path, 'policy_lstm_1.pt'))
