"""ActorCriticRecurrent + Memory ([UPSTREAM rsl_rl v1.0.2] modules/actor_critic_recurrent.py; SURVEY.md
§8(a).6 R6; configs h1_config.py:103-115, g1_config.py:91-105; attribute use utils/helpers.py:163-189:
`memory_a.rnn` must be an nn.LSTM so PolicyExporterLSTM can copy it).

Rollout-mode calls (no masks) step the LSTM through csrc/lstm_kernels.cu with the (h, c) state kept in
one device buffer [4, N, H] = (actor h, actor c, critic h, critic c); `memory_*.hidden_states` are views of
it.  Update-mode calls (masks + hidden_states) belong to PPO.update, which runs the fused masked time
loop instead of padded trajectories.
"""
import torch
import torch.nn as nn

from .actor_critic import ActorCritic


class Memory(nn.Module):
    def __init__(self, input_size, type="lstm", num_layers=1, hidden_size=256):
        super().__init__()
        if type.lower() != "lstm" or num_layers != 1:
            raise NotImplementedError("Memory: the fused sm_100a path implements a single-layer LSTM "
                                      "(h1_config.py:110-113); got type=%r num_layers=%r" % (type, num_layers))
        self.rnn = nn.LSTM(input_size=input_size, hidden_size=hidden_size, num_layers=num_layers)
        self.hidden_states = None

    def forward(self, input, masks=None, hidden_states=None):
        raise NotImplementedError("Memory.forward runs inside ActorCriticRecurrent (fused LSTM step); "
                                  "batched update mode is PPO.update")

    def reset(self, dones=None):
        if self.hidden_states is None:
            return
        for hidden_state in self.hidden_states:
            hidden_state[..., dones, :] = 0.0


class ActorCriticRecurrent(ActorCritic):
    is_recurrent = True

    def __init__(self, num_actor_obs, num_critic_obs, num_actions, actor_hidden_dims=[256, 256, 256],
                 critic_hidden_dims=[256, 256, 256], activation="elu", rnn_type="lstm", rnn_hidden_size=256,
                 rnn_num_layers=1, init_noise_std=1.0, **kwargs):
        if kwargs:
            print("ActorCriticRecurrent.__init__ got unexpected arguments, which will be ignored: " + str(kwargs.keys()))
        super().__init__(num_actor_obs=rnn_hidden_size, num_critic_obs=rnn_hidden_size, num_actions=num_actions,
                         actor_hidden_dims=actor_hidden_dims, critic_hidden_dims=critic_hidden_dims,
                         activation=activation, init_noise_std=init_noise_std)
        self.memory_a = Memory(num_actor_obs, type=rnn_type, num_layers=rnn_num_layers, hidden_size=rnn_hidden_size)
        self.memory_c = Memory(num_critic_obs, type=rnn_type, num_layers=rnn_num_layers, hidden_size=rnn_hidden_size)
        print(f"Actor RNN: {self.memory_a}")
        print(f"Critic RNN: {self.memory_c}")
        self._hidden = None

    # ---- rollout hidden state [4, N, H]
    def hidden_buffer(self, num_rows):
        H = self.memory_a.rnn.hidden_size
        dev = self.std.device
        if self._hidden is None or self._hidden.shape[1] != num_rows or self._hidden.device != dev:
            # a normal tensor even when the first call comes from inside torch.inference_mode() (the runner's rollout):
            # reset(dones) is also called from plain code (play.py-style loops) and must be able to write in place
            with torch.inference_mode(False):
                self._hidden = torch.zeros(4, num_rows, H, device=dev)
                h = self._hidden
                self.memory_a.hidden_states = (h[0].unsqueeze(0), h[1].unsqueeze(0))
                self.memory_c.hidden_states = (h[2].unsqueeze(0), h[3].unsqueeze(0))
        return self._hidden

    def reset(self, dones=None):
        self.memory_a.reset(dones)
        self.memory_c.reset(dones)

    def _rollout_forward(self, which, x):
        if x.dim() != 2:
            raise ValueError("rollout-mode recurrent forward takes [num_envs, num_obs]")
        h = self.hidden_buffer(x.shape[0])
        return self.engine(max_rows=x.shape[0]).forward(which, x, h[2 * which], h[2 * which + 1])

    def _actor_forward(self, observations):
        return self._rollout_forward(0, observations)

    def _critic_forward(self, critic_observations):
        return self._rollout_forward(1, critic_observations)

    def act(self, observations, masks=None, hidden_states=None):
        if masks is not None:
            raise NotImplementedError("batched (masks=...) mode is fused into PPO.update on this backend")
        return super().act(observations)

    def act_inference(self, observations):
        return self._actor_forward(observations)

    def evaluate(self, critic_observations, masks=None, hidden_states=None):
        if masks is not None:
            raise NotImplementedError("batched (masks=...) mode is fused into PPO.update on this backend")
        return self._critic_forward(critic_observations)

    def get_hidden_states(self):
        return self.memory_a.hidden_states, self.memory_c.hidden_states
