"""PPO ([UPSTREAM rsl_rl v1.0.2] algorithms/ppo.py; hyper-parameters
legged_robot_config.py:257-270).  Same constructor and methods; what each does on this backend:

  act                 urlg_policy_act: actor+critic forward on tcgen05, Philox sampling, log-prob, and the
                      transition written straight into the storage slabs (no separate add_transitions pass)
  process_env_step    urlg_process_env_step: timeout bootstrap, reward/done store, LSTM state reset
  compute_returns     critic forward + urlg_gae (reverse scan + normalisation)
  update              num_learning_epochs x num_mini_batches of urlg_ppo_minibatch_grads (+ NCCL all-reduce of the
                      flat gradient when torch.distributed is initialised) + urlg_ppo_apply; the adaptive-KL
                      learning-rate rule runs on the device, so there is ONE host read per update (the returned
                      mean losses) instead of the reference's 40.
"""
import atexit
import ctypes as C
import gc
import os
import weakref

import torch
import torch.distributed as dist

from ... import _lib
from ..storage import RolloutStorage


class _FlatAdam(torch.optim.Optimizer):
    """torch.optim.Adam-compatible shell (param_groups / state_dict / load_state_dict) around the flat
    exp_avg / exp_avg_sq buffers the fused Adam kernel updates; checkpoints interchange with the reference's
    `optim.Adam(actor_critic.parameters(), lr)` ([UPSTREAM] ppo.py)."""

    def __init__(self, ppo, params, lr):
        self._ppo = ppo
        super().__init__(list(params), dict(lr=lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False,
                                            maximize=False, foreach=None, capturable=False, differentiable=False,
                                            fused=None))

    def _views(self):
        eng = self._ppo.engine
        out, off = [], 0
        for p in self.param_groups[0]["params"]:
            k = p.numel()
            out.append((p, eng.exp_avg[off:off + k].view(p.shape), eng.exp_avg_sq[off:off + k].view(p.shape)))
            off += k
        return out

    def state_dict(self):
        eng = self._ppo.engine
        if eng is None:
            return super().state_dict()
        step = eng.stats[_lib.STAT_STEP].detach().clone().cpu()
        self.param_groups[0]["lr"] = self._ppo.learning_rate
        if float(step) > 0:
            for p, m, v in self._views():
                self.state[p] = {"step": step.clone(), "exp_avg": m, "exp_avg_sq": v}
        return super().state_dict()

    def load_state_dict(self, state_dict):
        eng = self._ppo.engine
        params = self.param_groups[0]["params"]
        ids = state_dict["param_groups"][0]["params"]
        step = 0.0
        for (p, m, v), i in zip(self._views(), ids):
            st = state_dict["state"].get(i)
            if st is None:
                continue
            m.copy_(st["exp_avg"])
            v.copy_(st["exp_avg_sq"])
            step = float(st["step"])
        eng.stats[_lib.STAT_STEP] = step
        lr = state_dict["param_groups"][0]["lr"]
        self.param_groups[0]["lr"] = lr
        self._ppo.learning_rate = lr
        assert len(params) == len(ids)

    def zero_grad(self, set_to_none=False):
        self._ppo.engine.grads.zero_()

    def step(self, closure=None):
        raise NotImplementedError("the optimiser step is fused into PPO.update (urlg_ppo_apply)")


class PPO:
    actor_critic = None

    def __init__(self, actor_critic, num_learning_epochs=1, num_mini_batches=1, clip_param=0.2, gamma=0.998,
                 lam=0.95, value_loss_coef=1.0, entropy_coef=0.0, learning_rate=1e-3, max_grad_norm=1.0,
                 use_clipped_value_loss=True, schedule="fixed", desired_kl=0.01, device="cpu"):
        self.device = device
        if torch.device(device).type != "cuda":
            raise RuntimeError("unitree_rl_gym_b200 PPO runs on CUDA only (device=%r); no CPU fallback" % (device,))
        self.desired_kl = desired_kl
        self.schedule = schedule
        self._lr_host = learning_rate
        self.actor_critic = actor_critic
        self.actor_critic.to(self.device)
        self.storage = None
        self.engine = None
        self.optimizer = _FlatAdam(self, self.actor_critic.parameters(), learning_rate)
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
        self.seed = 1
        self.env_offset = 0
        self._act_calls = 0
        self._act_base = None
        self.sample_noise = None     # parity hook: N(0,1) [N, A] used instead of Philox for the next act()
        self.perm = None             # parity hook: minibatch permutation [T*N] used instead of randperm
        self.lr_trace = None         # parity hook: a list -> (kl_mean, lr) read back after every minibatch (eager update)
        self.cuda_graph_update = False
        self.update_graph_launches = 0
        self._idx_buf = None
        self._update_graph = None
        self._atexit_hooked = False
        self._update_warm = None
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        self._graph_nccl = os.environ.get("URLG_GRAPH_NCCL", "1") != "0" and (
            self.world > 1 and dist.get_backend() == "nccl")
        # gradient exchange: one-shot all-reduce over peer memory (urlg_peer_allreduce) when the ranks can map each other's
        # memory; URLG_PEER_XCHG=0 or a failed symmetric-memory rendezvous keeps the NCCL all-reduce
        self._peer, self._peer_fused = None, False
        self.exchange = "none" if self.world == 1 else "nccl"


    # ---- learning rate lives on the device (adaptive schedule); the attribute stays readable / writable
    @property
    def learning_rate(self):
        return self._lr_host          # host copy: set here and refreshed by update()'s one read-back (no device sync)

    @learning_rate.setter
    def learning_rate(self, v):
        self._lr_host = float(v)
        if self.engine is not None:
            self.engine.stats[_lib.STAT_LR] = float(v)

    def init_storage(self, num_envs, num_transitions_per_env, actor_obs_shape, critic_obs_shape, action_shape):
        self.storage = RolloutStorage(num_envs, num_transitions_per_env, actor_obs_shape, critic_obs_shape,
                                      action_shape, self.device)
        ac = self.actor_critic
        T, N = num_transitions_per_env, num_envs
        if ac.is_recurrent:
            self.storage.ensure_hidden(ac.memory_a.rnn.hidden_size)
            mb_rows = T * (N // self.num_mini_batches)
        else:
            mb_rows = (N * T) // self.num_mini_batches
        if self.world > 1:
            self.env_offset = self.rank * N
        self.engine = ac.engine(max_rows=max(N, mb_rows), max_seq=T, seed=self.seed, env_offset=self.env_offset)
        self.engine.stats[_lib.STAT_LR] = self._lr_host
        if self.world > 1:       # identical initial parameters on every rank
            dist.broadcast(self.engine.flat, 0)
            self.engine.sync_weights()
            self._setup_peer_exchange()
        dev = self.engine.device
        self._actions = torch.zeros(N, action_shape[0], device=dev)
        self._act_base = torch.zeros(1, dtype=torch.int64, device=dev)
        self._last_values = torch.zeros(N, 1, device=dev)

    def _setup_peer_exchange(self):
        self._peer, self.exchange = None, "nccl"
        # measured on B200 (bench.py, Go2 4096 envs/GPU, update as one graph): 2 ranks 7.30 ms fused peer exchange vs 7.60 ms
        # NCCL; 8 ranks 8.28 ms (one- and two-shot alike: flag + remote-read round trips, not bytes) vs 8.00 ms NCCL.  So the
        # default is the peer exchange for 2 ranks and NCCL beyond; URLG_PEER_XCHG=1 forces it at any world size, 0 disables it
        want = os.environ.get("URLG_PEER_XCHG", "auto")
        if want == "0" or (want == "auto" and self.world > 2) or dist.get_backend() != "nccl" or self.world > _lib.PEER_MAX_WORLD:
            return
        ok = torch.ones(1, device=self.engine.device)
        try:
            from ..peer_exchange import PeerExchange
            peer = PeerExchange(self.engine.grads_and_sums.numel(), self.engine.device)
        except Exception as ex:            # no peer access / symmetric memory on this box: every rank must agree
            peer, ok[0] = None, 0.0
            self._peer_error = repr(ex)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok) > 0:
            self._peer = peer
            self._bind_peer()

    def _bind_peer(self):
        """Fused form: the gradient-finalising kernel publishes, the gradient-norm kernel reduces (no extra launch per
        minibatch step); URLG_PEER_FUSED=0 keeps the exchange as the two stand-alone launches of urlg_peer_allreduce."""
        self._peer_fused = False
        self.exchange = "peer-memory one-shot (urlg_peer_allreduce)"
        if os.environ.get("URLG_PEER_FUSED", "1") != "0":
            _lib.check(_lib.lib().urlg_policy_bind_peer_exchange(self.engine.handle, C.byref(self._peer.x)))
            self._peer_fused = True
            self.exchange = "peer-memory one-shot, fused into the gradient finalize / norm kernels (urlg_policy_bind_peer_exchange)"

    def _check_engine(self):
        if self.engine is None:
            raise RuntimeError("PPO.init_storage must be called first")
        if not self.engine.owns(self.actor_critic):
            self.engine = self.actor_critic.engine()
            self.engine.stats[_lib.STAT_LR] = self._lr_host
            if self._peer is not None:
                self._bind_peer()
        return self.engine

    # ---- CUDA-graph support: the Philox step of act() is (device base + host offset), like the env's counters
    def begin_relative(self):
        """Fold the host act() counter into the device base (call before capturing a rollout graph)."""
        self._act_base += self._act_calls
        self._act_calls = 0

    def advance_base(self):
        """Captured as a node of the graph: base += act() calls made inside the graph."""
        self._act_base += self._act_calls

    def end_relative(self):
        self._act_calls = 0

    def test_mode(self):
        self.actor_critic.test()

    def train_mode(self):
        self.actor_critic.train()

    def _hparams(self):
        h = _lib.PpoHparams()
        h.clip_param, h.value_loss_coef, h.entropy_coef = self.clip_param, self.value_loss_coef, self.entropy_coef
        h.max_grad_norm = self.max_grad_norm if self.max_grad_norm is not None else 0.0
        h.desired_kl = self.desired_kl if self.desired_kl is not None else 0.0
        h.beta1, h.beta2, h.eps = 0.9, 0.999, 1e-8
        h.lr_min, h.lr_max = 1e-5, 1e-2
        h.use_clipped_value_loss = int(bool(self.use_clipped_value_loss))
        h.adaptive_lr = int(self.desired_kl is not None and self.schedule == "adaptive")
        h.grad_scale = 1.0 / self.world
        h.kl_count_scale = float(self.world)
        return h

    # ---- rollout
    def act(self, obs, critic_obs):
        eng, s = self._check_engine(), self.storage
        ac = self.actor_critic
        obs = obs.to(device=eng.device, dtype=torch.float).contiguous()
        critic_obs = critic_obs.to(device=eng.device, dtype=torch.float).contiguous()
        hidden = ac.hidden_buffer(s.num_envs) if ac.is_recurrent else None
        if ac.is_recurrent:
            self.transition.hidden_states = ac.get_hidden_states()
        noise = self.sample_noise
        if noise is not None:
            noise = noise.to(device=eng.device, dtype=torch.float).contiguous()
        self._act_calls += 1
        _lib.check(_lib.lib().urlg_policy_act(eng.handle, _lib.ptr(obs), _lib.ptr(critic_obs), C.byref(s.desc()), s.step,
                                              _lib.ptr(hidden), _lib.ptr(noise), _lib.ptr(self._actions),
                                              _lib.ptr(self._act_base), self._act_calls, _lib.cur_stream()))
        t = s.step
        tr = self.transition
        tr.actions, tr.values, tr.actions_log_prob = s.actions[t], s.values[t], s.actions_log_prob[t].view(-1)
        tr.action_mean, tr.action_sigma = s.mu[t], s.sigma[t]
        tr.observations, tr.critic_observations = obs, critic_obs
        return self._actions

    def process_env_step(self, rewards, dones, infos):
        eng, s = self._check_engine(), self.storage
        ac = self.actor_critic
        rewards = rewards.to(device=eng.device, dtype=torch.float).contiguous()
        dones = dones.to(device=eng.device)
        d8 = (dones if dones.dtype in (torch.bool, torch.uint8) else (dones != 0)).contiguous()
        to = infos.get("time_outs") if isinstance(infos, dict) else None
        if to is not None:
            to = to.to(device=eng.device)
            to = (to if to.dtype in (torch.bool, torch.uint8) else (to != 0)).contiguous()
        hidden = ac.hidden_buffer(s.num_envs) if ac.is_recurrent else None
        _lib.check(_lib.lib().urlg_process_env_step(C.byref(s.desc()), s.step, _lib.ptr(rewards),
                                                    _lib.ptr(d8.view(torch.uint8)),
                                                    _lib.ptr(to.view(torch.uint8)) if to is not None else None,
                                                    self.gamma, _lib.ptr(hidden), _lib.cur_stream()))
        s.step += 1
        self.transition.clear()

    def _reduce_moments(self, m):
        dist.all_reduce(m)

    def compute_returns(self, last_critic_obs):
        self._check_engine()
        last_values = self.actor_critic.evaluate(last_critic_obs).detach()
        self.storage.compute_returns(last_values, self.gamma, self.lam,
                                     reduce_moments=self._reduce_moments if self.world > 1 else None)

    # ---- update
    def _minibatch_loop(self, idx):
        """num_learning_epochs x num_mini_batches of (gather + forward + loss + backward) -> (all-reduce) -> (LR rule,
        clip, Adam).  Pure stream-ordered device work: capturable as one CUDA graph."""
        eng, s = self.engine, self.storage
        L, st = _lib.lib(), _lib.cur_stream()
        h = self._hparams()
        T, N = s.num_transitions_per_env, s.num_envs
        eng.stats[_lib.STAT_ACC_VLOSS:_lib.STAT_ACC_N + 1] = 0.0
        sd = s.desc()
        if self.actor_critic.is_recurrent:
            mb = N // self.num_mini_batches
            batches = [(None, 0, i * mb, mb) for i in range(self.num_mini_batches)]
        else:
            mini = (N * T) // self.num_mini_batches
            batches = [(idx[i * mini:(i + 1) * mini], mini, 0, 0) for i in range(self.num_mini_batches)]
        for _ in range(self.num_learning_epochs):
            for ind, rows, e0, ec in batches:
                _lib.check(L.urlg_ppo_minibatch_grads(eng.handle, C.byref(sd), _lib.ptr(ind), rows, e0, ec,
                                                      C.byref(h), st))
                if self._peer is not None:
                    if not self._peer_fused:                           # fused: minibatch_grads published, apply reduces
                        self._peer.all_reduce(eng.grads_and_sums, st)  # flat gradients + (KL, surrogate, value, entropy) sums
                elif self.world > 1:
                    dist.all_reduce(eng.grads_and_sums)
                _lib.check(L.urlg_ppo_apply(eng.handle, C.byref(h), st))
                if self.lr_trace is not None:
                    v = eng.stats.tolist()
                    self.lr_trace.append((v[_lib.STAT_KL_MEAN], v[_lib.STAT_LR]))

    def release_graphs(self):
        """Destroy the captured update graph.  With several ranks it holds NCCL kernels: NCCL keeps the communicator
        alive until every graph that captured its collectives is destroyed, so this must run before
        `dist.destroy_process_group()` (also registered with atexit when such a graph is captured)."""
        if self._update_graph is not None:
            torch.cuda.synchronize()
            self._update_graph = None
            gc.collect()
            torch.cuda.synchronize()

    def update(self):
        eng, s = self._check_engine(), self.storage
        ac = self.actor_critic
        T, N = s.num_transitions_per_env, s.num_envs
        idx = None
        if not ac.is_recurrent:
            mini = (N * T) // self.num_mini_batches
            if self.perm is not None:
                idx = self.perm.to(device=eng.device, dtype=torch.int64).contiguous()
            else:
                idx = torch.randperm(self.num_mini_batches * mini, device=eng.device)
            if self._idx_buf is None or self._idx_buf.numel() != idx.numel():
                self._idx_buf = torch.empty_like(idx)
                self._update_graph = None
            self._idx_buf.copy_(idx)           # static buffer: the captured graph gathers through it
            idx = self._indices = self._idx_buf
        # opt-in (`cuda_graph_update`, set by the runner's `cuda_graph` cfg key): the 20 minibatch steps (~800 launches
        # for the MLP, ~8000 for the LSTM's per-step BPTT) replayed as one graph, invalidated when the engine / storage
        # it captured is replaced.  With several ranks the per-minibatch NCCL all-reduces (flat gradients + KL / loss
        # sums) are captured as nodes of the same graph (NCCL >= 2.9.6 enqueues capturable kernels; every rank captures
        # and replays the same sequence; measured clean at 2, 4 and 8 ranks); URLG_GRAPH_NCCL=0 keeps the multi-rank update eager.
        key = (id(eng), s.desc().observations, self.clip_param, self.value_loss_coef, self.entropy_coef, self.max_grad_norm,
               self.desired_kl, self.schedule, self.num_learning_epochs, self.num_mini_batches, self.use_clipped_value_loss)
        if self.cuda_graph_update and self.lr_trace is None and (self.world == 1 or self._graph_nccl):
            if self._update_graph is None or self._update_graph[0] != key:
                if self._update_warm != key:
                    self._minibatch_loop(idx)          # first update runs eagerly (lazy init, warm-up), capture on the next
                    self._update_warm = key
                else:
                    torch.cuda.synchronize()
                    g = torch.cuda.CUDAGraph()
                    n0 = _lib.lib().urlg_launch_count()
                    # thread-local capture mode: the NCCL watchdog thread may poll events of earlier collectives
                    with torch.cuda.graph(g, capture_error_mode="thread_local" if self.world > 1 else "global"):
                        self._minibatch_loop(idx)
                    self.update_graph_launches = _lib.lib().urlg_launch_count() - n0
                    self._update_graph = (key, g)
                    if self.world > 1 and not self._atexit_hooked:
                        self._atexit_hooked = True
                        ref = weakref.ref(self)
                        def _at_exit():
                            alg = ref()
                            if alg is not None:
                                try:
                                    alg.release_graphs()
                                except Exception:      # interpreter / CUDA context already going down
                                    pass
                        atexit.register(_at_exit)
                    g.replay()
            else:
                self._update_graph[1].replay()
        else:
            self._minibatch_loop(idx)
        stats = eng.stats.tolist()        # the one host read of the update
        n = max(stats[_lib.STAT_ACC_N], 1.0)
        self._lr_host = stats[_lib.STAT_LR]
        self.optimizer.param_groups[0]["lr"] = self._lr_host
        s.clear()
        return stats[_lib.STAT_ACC_VLOSS] / n, stats[_lib.STAT_ACC_SURR] / n
