"""ActorCritic ([UPSTREAM rsl_rl v1.0.2] modules/actor_critic.py; constructed by
OnPolicyRunner from legged_robot_config.py:247-255, attribute use utils/helpers.py:150-160).

`actor` / `critic` stay ordinary scriptable nn.Sequential stacks (state_dict keys actor.0.weight ...,
export_policy_as_jit works unchanged), but act / evaluate / act_inference run through the fused
sm_100a path (csrc/policy_kernels.cu) once the module is on a CUDA device; training goes through
PPO.update -> urlg_ppo_minibatch_grads, not autograd.
"""
import torch
import torch.nn as nn
from torch.distributions import Normal

from ..engine import PolicyEngine


def get_activation(act_name):
    if act_name == "elu":
        return nn.ELU()
    if act_name in ("selu", "relu", "lrelu", "tanh", "sigmoid", "crelu"):
        raise NotImplementedError(f"activation {act_name!r}: the fused sm_100a MLP implements 'elu' "
                                  "(the reference's setting, legged_robot_config.py:251)")
    print("invalid activation function!")
    return None


def _stack(n_in, hidden, n_out, activation):
    widths = [n_in] + list(hidden)
    layers = []
    for a, b in zip(widths[:-1], widths[1:]):
        layers += [nn.Linear(a, b), activation]
    layers.append(nn.Linear(widths[-1], n_out))
    return nn.Sequential(*layers)


class ActorCritic(nn.Module):
    is_recurrent = False

    def __init__(self, num_actor_obs, num_critic_obs, num_actions, actor_hidden_dims=[256, 256, 256],
                 critic_hidden_dims=[256, 256, 256], activation="elu", init_noise_std=1.0, **kwargs):
        if kwargs:
            print("ActorCritic.__init__ got unexpected arguments, which will be ignored: "
                  + str([key for key in kwargs.keys()]))
        super().__init__()
        activation = get_activation(activation)
        self.actor = _stack(num_actor_obs, actor_hidden_dims, num_actions, activation)
        self.critic = _stack(num_critic_obs, critic_hidden_dims, 1, activation)
        print(f"Actor MLP: {self.actor}")
        print(f"Critic MLP: {self.critic}")
        self.std = nn.Parameter(init_noise_std * torch.ones(num_actions))
        self.distribution = None
        Normal.set_default_validate_args = False
        self._engine = None
        # precision 1 = split bf16 (hi+lo operands, 3 MMAs: ~fp32 operand precision, what the 1e-3 update parity
        # needs); 0 = plain bf16 operands.  URLG_PRECISION=bf16 selects 0.
        import os
        self._engine_args = dict(max_rows=4096, max_seq=1, seed=1, env_offset=0,
                                 precision=0 if os.environ.get("URLG_PRECISION", "bf16x3") == "bf16" else 1)

    # ---- binding to the CUDA policy handle
    def engine(self, **kw):
        """The PolicyEngine bound to this module; (re)built when shapes grow or the parameters moved."""
        grow = any(kw.get(k, 0) > self._engine_args[k] for k in ("max_rows", "max_seq"))
        changed = any(k in kw and kw[k] != self._engine_args[k] for k in ("seed", "env_offset", "precision"))
        if self._engine is None or grow or changed or not self._engine.owns(self):
            for k, v in kw.items():
                self._engine_args[k] = max(v, self._engine_args[k]) if k in ("max_rows", "max_seq") else v
            old = self._engine
            # normal (not inference) tensors even when the first call comes from inside torch.inference_mode() (the
            # runner's rollout): the optimizer updates them in place later, outside inference mode
            with torch.inference_mode(False):
                self._engine = PolicyEngine(self, **self._engine_args)
                if old is not None and old.num_params == self._engine.num_params:
                    self._engine.exp_avg.copy_(old.exp_avg)
                    self._engine.exp_avg_sq.copy_(old.exp_avg_sq)
                    self._engine.stats.copy_(old.stats)
        return self._engine

    def load_state_dict(self, state_dict, strict=True):
        out = super().load_state_dict(state_dict, strict=strict)
        if self._engine is not None and self._engine.owns(self):
            self._engine.sync_weights()
        return out

    @staticmethod
    def init_weights(sequential, scales):
        [torch.nn.init.orthogonal_(module.weight, gain=scales[idx]) for idx, module in
         enumerate(mod for mod in sequential if isinstance(mod, nn.Linear))]

    def reset(self, dones=None):
        pass

    def forward(self):
        raise NotImplementedError

    @property
    def action_mean(self):
        return self.distribution.mean

    @property
    def action_std(self):
        return self.distribution.stddev

    @property
    def entropy(self):
        return self.distribution.entropy().sum(dim=-1)

    def _actor_forward(self, observations):
        return self.engine(max_rows=observations.reshape(-1, observations.shape[-1]).shape[0]).forward(0, observations)

    def _critic_forward(self, critic_observations):
        return self.engine(max_rows=critic_observations.reshape(-1, critic_observations.shape[-1]).shape[0]).forward(
            1, critic_observations)

    def update_distribution(self, observations):
        mean = self._actor_forward(observations)
        self.distribution = Normal(mean, mean * 0. + self.std.detach())

    def act(self, observations, **kwargs):
        self.update_distribution(observations)
        return self.distribution.sample()

    def get_actions_log_prob(self, actions):
        return self.distribution.log_prob(actions).sum(dim=-1)

    def act_inference(self, observations):
        return self._actor_forward(observations)

    def evaluate(self, critic_observations, **kwargs):
        return self._critic_forward(critic_observations)
