"""RolloutStorage ([UPSTREAM rsl_rl v1.0.2] storage/rollout_storage.py).

Same constructor, attributes (time-major [T, N, ...] tensors, dones uint8) and methods as the reference.
compute_returns is the GAE reverse-scan kernel (csrc/storage_kernels.cu); on the PPO hot path the
transition is written into the [t] slabs directly by urlg_policy_act / urlg_process_env_step and the
minibatch gather happens inside urlg_ppo_minibatch_grads, so add_transitions and the generator methods
are kept only for callers of the public API.
"""
import ctypes as C

import torch

from ... import _lib
from ..utils import split_and_pad_trajectories


class RolloutStorage:
    class Transition:
        def __init__(self):
            self.observations = None
            self.critic_observations = None
            self.actions = None
            self.rewards = None
            self.dones = None
            self.values = None
            self.actions_log_prob = None
            self.action_mean = None
            self.action_sigma = None
            self.hidden_states = None

        def clear(self):
            self.__init__()

    def __init__(self, num_envs, num_transitions_per_env, obs_shape, privileged_obs_shape, actions_shape, device="cpu"):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("unitree_rl_gym_b200 RolloutStorage lives on CUDA (got %r); no CPU fallback" % (device,))
        self.device = device
        self.obs_shape = obs_shape
        self.privileged_obs_shape = privileged_obs_shape
        self.actions_shape = actions_shape
        T, N = num_transitions_per_env, num_envs

        def z(*s, dtype=torch.float):
            return torch.zeros(T, N, *s, device=self.device, dtype=dtype)
        self.observations = z(*obs_shape)
        self.privileged_observations = z(*privileged_obs_shape) if privileged_obs_shape[0] is not None else None
        self.rewards = z(1)
        self.actions = z(*actions_shape)
        self.dones = z(1, dtype=torch.uint8)
        self.actions_log_prob = z(1)
        self.values = z(1)
        self.returns = z(1)
        self.advantages = z(1)
        self.mu = z(*actions_shape)
        self.sigma = z(*actions_shape)
        self.num_transitions_per_env = T
        self.num_envs = N
        self.saved_hidden_states_a = None
        self.saved_hidden_states_c = None
        self.step = 0
        self._moments = torch.zeros(3, dtype=torch.float64, device=self.device)
        self._desc = None

    # ---- C-ABI view of the tensors
    def ensure_hidden(self, hidden_size):
        """Allocate saved_hidden_states_{a,c} = [h, c], each [T, 1, N, H] (the reference does it lazily in
        _save_hidden_states)."""
        if self.saved_hidden_states_a is None:
            T, N = self.num_transitions_per_env, self.num_envs
            mk = lambda: [torch.zeros(T, 1, N, hidden_size, device=self.device) for _ in range(2)]
            self.saved_hidden_states_a, self.saved_hidden_states_c = mk(), mk()
            self._desc = None

    def desc(self):
        """urlg_storage; rebuilt when a public tensor attribute was re-bound (e.g. the reference-style
        `self.advantages = ...`)."""
        key = tuple(t.data_ptr() if t is not None else 0 for t in (
            self.observations, self.privileged_observations, self.actions, self.mu, self.sigma, self.rewards,
            self.values, self.returns, self.advantages, self.actions_log_prob, self.dones))
        if self._desc is None or self._desc[0] != key:
            d = _lib.Storage()
            d.T, d.N = self.num_transitions_per_env, self.num_envs
            d.num_obs = self.obs_shape[0]
            d.num_priv = self.privileged_obs_shape[0] or 0
            d.num_actions = self.actions_shape[0]
            p = _lib.ptr
            for name in ("observations", "privileged_observations", "actions", "mu", "sigma", "rewards", "values",
                         "returns", "advantages", "actions_log_prob", "dones"):
                t = getattr(self, name)
                if t is not None and not t.is_contiguous():
                    raise RuntimeError(f"RolloutStorage.{name} must stay contiguous")
                setattr(d, name, p(t))
            if self.saved_hidden_states_a is not None:
                d.rnn_hidden = self.saved_hidden_states_a[0].shape[-1]
                d.hid_a_h, d.hid_a_c = p(self.saved_hidden_states_a[0]), p(self.saved_hidden_states_a[1])
                d.hid_c_h, d.hid_c_c = p(self.saved_hidden_states_c[0]), p(self.saved_hidden_states_c[1])
            self._desc = (key, d)
        return self._desc[1]

    # ---- public API
    def add_transitions(self, transition: "RolloutStorage.Transition"):
        if self.step >= self.num_transitions_per_env:
            raise AssertionError("Rollout buffer overflow")
        s = self.step
        self.observations[s].copy_(transition.observations)
        if self.privileged_observations is not None:
            self.privileged_observations[s].copy_(transition.critic_observations)
        self.actions[s].copy_(transition.actions)
        self.rewards[s].copy_(transition.rewards.view(-1, 1))
        self.dones[s].copy_(transition.dones.view(-1, 1))
        self.values[s].copy_(transition.values)
        self.actions_log_prob[s].copy_(transition.actions_log_prob.view(-1, 1))
        self.mu[s].copy_(transition.action_mean)
        self.sigma[s].copy_(transition.action_sigma)
        self._save_hidden_states(transition.hidden_states)
        self.step += 1

    def _save_hidden_states(self, hidden_states):
        if hidden_states is None or hidden_states == (None, None):
            return
        hid_a = hidden_states[0] if isinstance(hidden_states[0], tuple) else (hidden_states[0],)
        hid_c = hidden_states[1] if isinstance(hidden_states[1], tuple) else (hidden_states[1],)
        self.ensure_hidden(hid_a[0].shape[-1])
        for i in range(len(hid_a)):
            self.saved_hidden_states_a[i][self.step].copy_(hid_a[i])
            self.saved_hidden_states_c[i][self.step].copy_(hid_c[i])

    def clear(self):
        self.step = 0

    def compute_returns(self, last_values, gamma, lam, reduce_moments=None):
        """GAE + advantage normalisation.  reduce_moments(tensor[3] float64: sum, sumsq, count), if given, runs
        between the scan and the normalisation (multi-GPU: all-reduce for a global mean/std)."""
        L, st, d = _lib.lib(), _lib.cur_stream(), self.desc()
        lv = last_values.to(device=self.device, dtype=torch.float).contiguous()
        if lv.numel() != self.num_envs:
            raise ValueError("last_values must have one value per env")
        if reduce_moments is None:
            _lib.check(L.urlg_gae(C.byref(d), _lib.ptr(lv), gamma, lam, _lib.ptr(self._moments), 3, st))
        else:
            _lib.check(L.urlg_gae(C.byref(d), _lib.ptr(lv), gamma, lam, _lib.ptr(self._moments), 1, st))
            reduce_moments(self._moments)
            _lib.check(L.urlg_gae(C.byref(d), _lib.ptr(lv), gamma, lam, _lib.ptr(self._moments), 2, st))

    def get_statistics(self):
        done = self.dones.clone()
        done[-1] = 1
        flat = done.permute(1, 0, 2).reshape(-1, 1)
        ends = torch.cat((flat.new_tensor([-1], dtype=torch.int64), flat.nonzero(as_tuple=False)[:, 0]))
        lengths = ends[1:] - ends[:-1]
        return lengths.float().mean(), self.rewards.mean()

    def mini_batch_generator(self, num_mini_batches, num_epochs=8):
        batch_size = self.num_envs * self.num_transitions_per_env
        mini_batch_size = batch_size // num_mini_batches
        indices = torch.randperm(num_mini_batches * mini_batch_size, requires_grad=False, device=self.device)
        flat = lambda t: t.flatten(0, 1)
        obs = flat(self.observations)
        cobs = flat(self.privileged_observations) if self.privileged_observations is not None else obs
        cols = [flat(t) for t in (self.actions, self.values, self.advantages, self.returns, self.actions_log_prob,
                                  self.mu, self.sigma)]
        for _ in range(num_epochs):
            for i in range(num_mini_batches):
                b = indices[i * mini_batch_size:(i + 1) * mini_batch_size]
                yield (obs[b], cobs[b], *[c[b] for c in cols], (None, None), None)

    def reccurent_mini_batch_generator(self, num_mini_batches, num_epochs=8):
        padded_obs, masks = split_and_pad_trajectories(self.observations, self.dones)
        if self.privileged_observations is not None:
            padded_cobs, _ = split_and_pad_trajectories(self.privileged_observations, self.dones)
        else:
            padded_cobs = padded_obs
        mb = self.num_envs // num_mini_batches
        dones = self.dones.squeeze(-1)
        starts = torch.zeros_like(dones, dtype=torch.bool)
        starts[1:] = dones[:-1]
        starts[0] = True
        per_env_first = starts.permute(1, 0)
        for _ in range(num_epochs):
            first = 0
            for i in range(num_mini_batches):
                lo, hi = i * mb, (i + 1) * mb
                last = first + int(torch.sum(starts[:, lo:hi]))
                pick = lambda hs: [h.permute(2, 0, 1, 3)[per_env_first][first:last].transpose(1, 0).contiguous() for h in hs]
                hid_a, hid_c = pick(self.saved_hidden_states_a), pick(self.saved_hidden_states_c)
                hid_a = hid_a[0] if len(hid_a) == 1 else hid_a
                hid_c = hid_c[0] if len(hid_c) == 1 else hid_c
                yield (padded_obs[:, first:last], padded_cobs[:, first:last], self.actions[:, lo:hi],
                       self.values[:, lo:hi], self.advantages[:, lo:hi], self.returns[:, lo:hi],
                       self.actions_log_prob[:, lo:hi], self.mu[:, lo:hi], self.sigma[:, lo:hi], (hid_a, hid_c),
                       masks[:, first:last])
                first = last
