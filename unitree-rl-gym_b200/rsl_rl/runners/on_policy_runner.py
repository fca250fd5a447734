"""OnPolicyRunner ([UPSTREAM rsl_rl v1.0.2] runners/on_policy_runner.py; SURVEY.md §3.2, §8(a).6 R2-R4;
call sites legged_gym/utils/task_registry.py:118-126, scripts/train.py:14, scripts/play.py:34-39).

Same constructor / learn / save / load / get_inference_policy.  Differences by design: per-step logging
bookkeeping (reward / episode-length sums, finished-episode statistics, infos["episode"]) is accumulated on the
device and read back once per iteration instead of a `.cpu()` per env step (_EpisodeBook), which also lets the
rollout CUDA graph stay on while logging; TensorBoard is optional (used when importable and log_dir is given).
"""
import os
import statistics
import time
from collections import deque

import torch

from ..algorithms import PPO  # noqa: F401  (resolved by name from the config, like the reference's eval())
from ..modules import ActorCritic, ActorCriticRecurrent  # noqa: F401

_CLASSES = {"PPO": PPO, "ActorCritic": ActorCritic, "ActorCriticRecurrent": ActorCriticRecurrent}


class _EpisodeBook:
    """The logging bookkeeping of the reference's learn() loop ([UPSTREAM rsl_rl v1.0.2] on_policy_runner.py: ep_infos,
    cur_reward_sum, cur_episode_length, rewbuffer / lenbuffer) as stream-ordered device work: `step` launches a handful
    of elementwise kernels and reads nothing, so it can be recorded into the rollout CUDA graph; `read` is the one
    device->host transfer of the iteration."""

    def __init__(self, T, N, device):
        self.T, self.N, self.device = T, N, device
        self.cur_reward_sum = torch.zeros(N, dtype=torch.float, device=device)
        self.cur_episode_length = torch.zeros(N, dtype=torch.float, device=device)
        # (reward sum, length) of the episodes that finished at (step, env); NaN elsewhere
        self.fin = torch.full((2, T, N), float("nan"), device=device)
        self._nan = torch.full((), float("nan"), device=device)
        self.keys = None            # names of the tensor-valued entries of infos["episode"]
        self.consts = {}            # host-valued entries (e.g. max_command_x): logged as they are
        self.ep = None              # [T, K]
        self.present = [False] * T

    def step(self, t, rewards, dones, infos):
        ep = infos.get("episode") if isinstance(infos, dict) else None
        if ep is not None:
            if self.keys is None:
                self.keys = [k for k, v in ep.items() if torch.is_tensor(v)]
                self.ep = torch.zeros(self.T, max(len(self.keys), 1), device=self.device)
            if self.keys:
                vals = [ep[k] if ep[k].dim() == 0 else ep[k].float().mean() for k in self.keys]
                torch.stack(vals, out=self.ep[t, :len(vals)])
            self.consts = {k: float(v) for k, v in ep.items() if not torch.is_tensor(v)}
            self.present[t] = True
        self.cur_reward_sum += rewards
        self.cur_episode_length += 1
        d = dones > 0
        torch.where(d, self.cur_reward_sum, self._nan, out=self.fin[0, t])
        torch.where(d, self.cur_episode_length, self._nan, out=self.fin[1, t])
        self.cur_reward_sum.masked_fill_(d, 0.0)
        self.cur_episode_length.masked_fill_(d, 0.0)

    def read(self):
        """-> (finished reward sums, finished lengths, {episode key: mean over the iteration's steps})."""
        n_fin = self.fin.numel()
        parts = [self.fin.flatten()] + ([self.ep.flatten()] if self.ep is not None else [])
        host = torch.cat(parts).cpu()                      # the iteration's single logging read-back
        fin = host[:n_fin].view(2, -1)
        keep = ~torch.isnan(fin[0])
        means = {}
        if self.ep is not None:
            ep = host[n_fin:].view(self.T, -1).double()
            rows = [t for t in range(self.T) if self.present[t]]
            if rows:
                m = ep[rows].mean(dim=0)
                means = {k: float(m[i]) for i, k in enumerate(self.keys)}
                means.update(self.consts)
        return fin[0][keep].tolist(), fin[1][keep].tolist(), means


class OnPolicyRunner:
    def __init__(self, env, train_cfg, log_dir=None, device="cpu"):
        self.cfg = train_cfg["runner"]
        self.alg_cfg = train_cfg["algorithm"]
        self.policy_cfg = train_cfg["policy"]
        self.device = device
        self.env = env
        num_critic_obs = self.env.num_privileged_obs if self.env.num_privileged_obs is not None else self.env.num_obs
        try:
            actor_critic_class = _CLASSES[self.cfg["policy_class_name"]]
            alg_class = _CLASSES[self.cfg["algorithm_class_name"]]
        except KeyError as e:
            raise NameError(f"name {e.args[0]!r} is not defined")      # what the reference's eval() raises
        actor_critic = actor_critic_class(self.env.num_obs, num_critic_obs, self.env.num_actions,
                                          **self.policy_cfg).to(self.device)
        self.alg = alg_class(actor_critic, device=self.device, **self.alg_cfg)
        self.alg.cuda_graph_update = bool(self.cfg.get("cuda_graph", False))
        seed = train_cfg.get("seed", 1)
        self.alg.seed = seed if seed is not None and seed >= 0 else 1
        self.num_steps_per_env = self.cfg["num_steps_per_env"]
        self.save_interval = self.cfg["save_interval"]
        self.alg.init_storage(self.env.num_envs, self.num_steps_per_env, [self.env.num_obs],
                              [self.env.num_privileged_obs], [self.env.num_actions])
        self.log_dir = log_dir
        self.writer = None
        self.tot_timesteps = 0
        self.tot_time = 0
        self.current_learning_iteration = 0
        self.last_stats = {}
        _, _ = self.env.reset()

    def learn(self, num_learning_iterations, init_at_random_ep_len=False):
        if self.log_dir is not None and self.writer is None:
            try:
                from torch.utils.tensorboard import SummaryWriter
                self.writer = SummaryWriter(log_dir=self.log_dir, flush_secs=10)
            except Exception:
                self.writer = None
        env, alg = self.env, self.alg
        if init_at_random_ep_len:
            env.episode_length_buf = torch.randint_like(env.episode_length_buf, high=int(env.max_episode_length))
        obs = env.get_observations()
        privileged_obs = env.get_privileged_observations()
        critic_obs = privileged_obs if privileged_obs is not None else obs
        obs, critic_obs = obs.to(self.device), critic_obs.to(self.device)
        alg.actor_critic.train()
        rewbuffer, lenbuffer = deque(maxlen=100), deque(maxlen=100)
        N, T = env.num_envs, self.num_steps_per_env
        book = _EpisodeBook(T, N, self.device) if self.log_dir is not None else None
        tot_iter = self.current_learning_iteration + num_learning_iterations
        self._tot_iter = tot_iter
        # opt-in (runner cfg key `cuda_graph`): the whole rollout replayed as one CUDA graph (rollout_graph.py).  The
        # logging bookkeeping is stream-ordered device work (no host read per step), so it is recorded into the same
        # graph through the post_step hook and the graph stays on when log_dir is set.
        graph = None
        if self.cfg.get("cuda_graph", False):
            from .rollout_graph import RolloutGraph
            if RolloutGraph.supported(self):
                graph = self._rollout_graph = RolloutGraph(self)
                graph.capture(post_step=None if book is None else (lambda t: book.step(t, *graph.last)))
        for it in range(self.current_learning_iteration, tot_iter):
            start = time.time()
            with torch.inference_mode():
                if graph is not None:
                    obs, critic_obs = graph.replay()
                for i in range(T if graph is None else 0):
                    actions = alg.act(obs, critic_obs)
                    obs, privileged_obs, rewards, dones, infos = env.step(actions)
                    critic_obs = privileged_obs if privileged_obs is not None else obs
                    obs, critic_obs, rewards, dones = (obs.to(self.device), critic_obs.to(self.device),
                                                       rewards.to(self.device), dones.to(self.device))
                    alg.process_env_step(rewards, dones, infos)
                    if book is not None:
                        book.step(i, rewards, dones, infos)
                if book is not None:
                    torch.cuda.current_stream().synchronize()     # collection_time covers the rollout, like the reference
                stop = time.time()
                collection_time = stop - start
                start = stop
                alg.compute_returns(critic_obs)
            mean_value_loss, mean_surrogate_loss = alg.update()
            ep_means = {}
            if book is not None:              # ONE read-back per iteration, in the reference's (step, env) order
                fin_r, fin_l, ep_means = book.read()
                rewbuffer.extend(fin_r)
                lenbuffer.extend(fin_l)
            stop = time.time()
            learn_time = stop - start
            self.tot_timesteps += T * N
            self.tot_time += collection_time + learn_time
            self.last_stats = {
                "it": it, "collection_time": collection_time, "learn_time": learn_time,
                "fps": int(T * N / (collection_time + learn_time)),
                "mean_value_loss": mean_value_loss, "mean_surrogate_loss": mean_surrogate_loss,
                "learning_rate": alg.learning_rate,
                "mean_noise_std": float(alg.actor_critic.std.detach().mean()),
                "mean_reward": statistics.mean(rewbuffer) if len(rewbuffer) > 0 else None,
                "mean_episode_length": statistics.mean(lenbuffer) if len(lenbuffer) > 0 else None,
            }
            if self.log_dir is not None:
                self.log(dict(self.last_stats), ep_means)
                if it % self.save_interval == 0:
                    self.save(os.path.join(self.log_dir, "model_{}.pt".format(it)))
        self.current_learning_iteration += num_learning_iterations
        if self.log_dir is not None:
            self.save(os.path.join(self.log_dir, "model_{}.pt".format(self.current_learning_iteration)))
            if self.writer is not None:
                self.writer.flush()

    def log(self, st, ep_means, width=80, pad=35):
        """Same scalar names as the reference's TensorBoard output ([UPSTREAM rsl_rl v1.0.2]
        runners/on_policy_runner.py log(); SURVEY.md §5 metrics row).  `ep_means`: {key: mean over the iteration's
        steps of infos["episode"][key]}, already on the host."""
        w, it = self.writer, st["it"]
        if w is not None:
            for key, v in ep_means.items():
                w.add_scalar("Episode/" + key, v, it)
            w.add_scalar("Loss/value_function", st["mean_value_loss"], it)
            w.add_scalar("Loss/surrogate", st["mean_surrogate_loss"], it)
            w.add_scalar("Loss/learning_rate", st["learning_rate"], it)
            w.add_scalar("Policy/mean_noise_std", st["mean_noise_std"], it)
            w.add_scalar("Perf/total_fps", st["fps"], it)
            w.add_scalar("Perf/collection time", st["collection_time"], it)
            w.add_scalar("Perf/learning_time", st["learn_time"], it)
            if st["mean_reward"] is not None:
                w.add_scalar("Train/mean_reward", st["mean_reward"], it)
                w.add_scalar("Train/mean_episode_length", st["mean_episode_length"], it)
                w.add_scalar("Train/mean_reward/time", st["mean_reward"], self.tot_time)
                w.add_scalar("Train/mean_episode_length/time", st["mean_episode_length"], self.tot_time)
        tot_iter = getattr(self, "_tot_iter", self.current_learning_iteration)
        head = f" \033[1m Learning iteration {it}/{tot_iter} \033[0m "
        lines = [f"{'#' * width}", head.center(width, " "), "",
                 f"{'Computation:':>{pad}} {st['fps']:.0f} steps/s (collection: {st['collection_time']:.3f}s, "
                 f"learning {st['learn_time']:.3f}s)",
                 f"{'Value function loss:':>{pad}} {st['mean_value_loss']:.4f}",
                 f"{'Surrogate loss:':>{pad}} {st['mean_surrogate_loss']:.4f}",
                 f"{'Mean action noise std:':>{pad}} {st['mean_noise_std']:.2f}"]
        if st["mean_reward"] is not None:
            lines += [f"{'Mean reward:':>{pad}} {st['mean_reward']:.2f}",
                      f"{'Mean episode length:':>{pad}} {st['mean_episode_length']:.2f}"]
        lines += [f"{'Mean episode ' + k + ':':>{pad}} {v:.4f}" for k, v in ep_means.items()]
        iters_done = it + 1 - self.current_learning_iteration       # advanced only after the loop
        lines += ["-" * width,
                  f"{'Total timesteps:':>{pad}} {self.tot_timesteps}",
                  f"{'Iteration time:':>{pad}} {st['collection_time'] + st['learn_time']:.2f}s",
                  f"{'Total time:':>{pad}} {self.tot_time:.2f}s",
                  f"{'ETA:':>{pad}} {self.tot_time / max(iters_done, 1) * (tot_iter - it - 1):.1f}s"]
        print("\n".join(lines))

    def save(self, path, infos=None):
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        torch.save({"model_state_dict": self.alg.actor_critic.state_dict(),
                    "optimizer_state_dict": self.alg.optimizer.state_dict(),
                    "iter": self.current_learning_iteration, "infos": infos}, path)

    def load(self, path, load_optimizer=True):
        loaded = torch.load(path, map_location=self.device, weights_only=False)
        self.alg.actor_critic.load_state_dict(loaded["model_state_dict"])
        self.alg._check_engine().sync_weights()
        if load_optimizer:
            self.alg.optimizer.load_state_dict(loaded["optimizer_state_dict"])
        self.current_learning_iteration = loaded["iter"]
        return loaded["infos"]

    def get_inference_policy(self, device=None):
        self.alg.actor_critic.eval()
        if device is not None:
            self.alg.actor_critic.to(device)
        return self.alg.actor_critic.act_inference
