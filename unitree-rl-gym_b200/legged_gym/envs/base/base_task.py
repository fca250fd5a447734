"""BaseTask: buffer allocation, getters and reset() of the vectorised env
(mirrors legged_gym/envs/base/base_task.py:11-86; viewer / render are graphics and out of scope)."""
import torch

from ....physics import SyntheticGym


class BaseTask:
    def __init__(self, cfg, sim_params, physics_engine, sim_device, headless):
        self.sim_params = sim_params
        self.physics_engine = physics_engine
        self.sim_device = sim_device
        self.headless = headless
        dev = torch.device(sim_device)
        if dev.type != "cuda":
            raise RuntimeError("unitree_rl_gym_b200 runs the env hot path on CUDA only (sim_device=%r); "
                               "there is no CPU fallback" % (sim_device,))
        self.sim_device_id = dev.index or 0
        self.device = str(dev) if dev.index is not None else "cuda:0"
        self.graphics_device_id = -1

        self.num_envs = cfg.env.num_envs
        self.num_obs = cfg.env.num_observations
        self.num_privileged_obs = cfg.env.num_privileged_obs
        self.num_actions = cfg.env.num_actions

        f = dict(device=self.device, dtype=torch.float)
        # obs (and privileged obs) are ping-ponged: the reference re-binds obs_buf to a fresh tensor
        # every step (legged_robot.py:800,644) and stock rsl_rl keeps a reference to the previous one.
        self._obs_slots = [torch.zeros(self.num_envs, self.num_obs, **f) for _ in range(2)]
        self._priv_slots = ([torch.zeros(self.num_envs, self.num_privileged_obs, **f) for _ in range(2)]
                            if self.num_privileged_obs is not None else None)
        self._obs_slot = 0
        self.obs_buf = self._obs_slots[0]
        self.privileged_obs_buf = self._priv_slots[0] if self._priv_slots else None
        self.rew_buf = torch.zeros(self.num_envs, **f)
        self.reset_buf = torch.ones(self.num_envs, device=self.device, dtype=torch.bool)
        self._episode_length_buf = torch.zeros(self.num_envs, device=self.device, dtype=torch.long)
        self.time_out_buf = torch.zeros(self.num_envs, device=self.device, dtype=torch.bool)
        self.extras = {}

        self.create_sim()
        self.enable_viewer_sync = True
        self.viewer = None

    # rsl_rl's runner ASSIGNS env.episode_length_buf (randint_like, [UPSTREAM] OnPolicyRunner.learn);
    # the kernels hold the raw pointer, so assignment copies into the bound buffer.
    @property
    def episode_length_buf(self):
        return self._episode_length_buf

    @episode_length_buf.setter
    def episode_length_buf(self, value):
        self._episode_length_buf.copy_(value)

    def create_sim(self):
        raise NotImplementedError

    def get_observations(self):
        return self.obs_buf

    def get_privileged_observations(self):
        return self.privileged_obs_buf

    def reset_idx(self, env_ids):
        raise NotImplementedError

    def reset(self):
        self.reset_idx(torch.arange(self.num_envs, device=self.device))
        obs, privileged_obs, _, _, _ = self.step(torch.zeros(self.num_envs, self.num_actions, device=self.device,
                                                             requires_grad=False))
        return obs, privileged_obs

    def step(self, actions):
        raise NotImplementedError

    def render(self, sync_frame_time=True):
        return None   # headless only: no graphics stack on this platform
