"""ORACLE: rsl_rl v1.0.2 algorithms/ppo.py restated."""
import torch
import torch.nn as nn
import torch.optim as optim
from .storage import RolloutStorage


class PPO:
    def __init__(self, actor_critic, num_learning_epochs=1, num_mini_batches=1, clip_param=0.2, gamma=0.998,
                 lam=0.95, value_loss_coef=1.0, entropy_coef=0.0, learning_rate=1e-3, max_grad_norm=1.0,
                 use_clipped_value_loss=True, schedule="fixed", desired_kl=0.01, device="cpu"):
        self.device = device
        self.desired_kl = desired_kl
        self.schedule = schedule
        self.learning_rate = learning_rate
        self.actor_critic = actor_critic
        self.actor_critic.to(self.device)
        self.storage = None
        self.optimizer = optim.Adam(self.actor_critic.parameters(), lr=learning_rate)
        self.transition = RolloutStorage.Transition()
        self.clip_param = clip_param
        self.num_learning_epochs = num_learning_epochs
        self.num_mini_batches = num_mini_batches
        self.value_loss_coef = value_loss_coef
        self.entropy_coef = entropy_coef
        self.gamma = gamma
        self.lam = lam
        self.max_grad_norm = max_grad_norm
        self.use_clipped_value_loss = use_clipped_value_loss
        self.lr_trace = []          # oracle-only: LR decision per minibatch step
        self.sample_noise = None    # oracle-only: injected N(0,1) for act()
        self.perm = None            # oracle-only: injected minibatch permutation

    def init_storage(self, num_envs, num_transitions_per_env, actor_obs_shape, critic_obs_shape, action_shape):
        self.storage = RolloutStorage(num_envs, num_transitions_per_env, actor_obs_shape, critic_obs_shape,
                                      action_shape, self.device)

    def test_mode(self):
        self.actor_critic.test()

    def train_mode(self):
        self.actor_critic.train()

    def act(self, obs, critic_obs):
        ac = self.actor_critic
        if ac.is_recurrent:
            self.transition.hidden_states = ac.get_hidden_states()
        sampled = ac.act(obs).detach()
        if self.sample_noise is not None:   # mean + std * eps with shared eps (parity runs)
            sampled = (ac.action_mean + ac.action_std * self.sample_noise).detach()
        self.transition.actions = sampled
        self.transition.values = ac.evaluate(critic_obs).detach()
        self.transition.actions_log_prob = ac.get_actions_log_prob(self.transition.actions).detach()
        self.transition.action_mean = ac.action_mean.detach()
        self.transition.action_sigma = ac.action_std.detach()
        self.transition.observations = obs
        self.transition.critic_observations = critic_obs
        return self.transition.actions

    def process_env_step(self, rewards, dones, infos):
        self.transition.rewards = rewards.clone()
        self.transition.dones = dones
        if "time_outs" in infos:
            self.transition.rewards += self.gamma * torch.squeeze(
                self.transition.values * infos["time_outs"].unsqueeze(1).to(self.device), 1)
        self.storage.add_transitions(self.transition)
        self.transition.clear()
        self.actor_critic.reset(dones)

    def compute_returns(self, last_critic_obs):
        last_values = self.actor_critic.evaluate(last_critic_obs).detach()
        self.storage.compute_returns(last_values, self.gamma, self.lam)

    def update(self):
        mean_value_loss = 0
        mean_surrogate_loss = 0
        ac = self.actor_critic
        if ac.is_recurrent:
            generator = self.storage.reccurent_mini_batch_generator(self.num_mini_batches, self.num_learning_epochs)
        elif self.perm is not None:
            generator = self.storage.mini_batch_generator(self.num_mini_batches, self.num_learning_epochs, self.perm)
        else:
            generator = self.storage.mini_batch_generator(self.num_mini_batches, self.num_learning_epochs)
        for (obs_b, critic_obs_b, actions_b, target_values_b, advantages_b, returns_b, old_lp_b,
             old_mu_b, old_sigma_b, hid_b, masks_b) in generator:
            ac.act(obs_b, masks=masks_b, hidden_states=hid_b[0])
            lp_b = ac.get_actions_log_prob(actions_b)
            value_b = ac.evaluate(critic_obs_b, masks=masks_b, hidden_states=hid_b[1])
            mu_b = ac.action_mean
            sigma_b = ac.action_std
            entropy_b = ac.entropy

            if self.desired_kl is not None and self.schedule == "adaptive":
                with torch.inference_mode():
                    kl = torch.sum(torch.log(sigma_b / old_sigma_b + 1.e-5)
                                   + (torch.square(old_sigma_b) + torch.square(old_mu_b - mu_b))
                                   / (2.0 * torch.square(sigma_b)) - 0.5, axis=-1)
                    kl_mean = torch.mean(kl)
                    if kl_mean > self.desired_kl * 2.0:
                        self.learning_rate = max(1e-5, self.learning_rate / 1.5)
                    elif kl_mean < self.desired_kl / 2.0 and kl_mean > 0.0:
                        self.learning_rate = min(1e-2, self.learning_rate * 1.5)
                    for g in self.optimizer.param_groups:
                        g["lr"] = self.learning_rate
                    self.lr_trace.append((float(kl_mean), self.learning_rate))

            ratio = torch.exp(lp_b - torch.squeeze(old_lp_b))
            surrogate = -torch.squeeze(advantages_b) * ratio
            surrogate_clipped = -torch.squeeze(advantages_b) * torch.clamp(ratio, 1.0 - self.clip_param,
                                                                           1.0 + self.clip_param)
            surrogate_loss = torch.max(surrogate, surrogate_clipped).mean()

            if self.use_clipped_value_loss:
                value_clipped = target_values_b + (value_b - target_values_b).clamp(-self.clip_param, self.clip_param)
                value_losses = (value_b - returns_b).pow(2)
                value_losses_clipped = (value_clipped - returns_b).pow(2)
                value_loss = torch.max(value_losses, value_losses_clipped).mean()
            else:
                value_loss = (returns_b - value_b).pow(2).mean()

            loss = surrogate_loss + self.value_loss_coef * value_loss - self.entropy_coef * entropy_b.mean()
            self.optimizer.zero_grad()
            loss.backward()
            nn.utils.clip_grad_norm_(ac.parameters(), self.max_grad_norm)
            self.optimizer.step()
            mean_value_loss += value_loss.item()
            mean_surrogate_loss += surrogate_loss.item()

        num_updates = self.num_learning_epochs * self.num_mini_batches
        mean_value_loss /= num_updates
        mean_surrogate_loss /= num_updates
        self.storage.clear()
        return mean_value_loss, mean_surrogate_loss
