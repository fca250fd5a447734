"""PolicyEngine: binds an ActorCritic / ActorCriticRecurrent nn.Module to the CUDA policy handle
(include/urlg_b200.h urlg_policy_*).

The module's parameters are re-seated as views of ONE flat fp32 buffer (same for .grad and the Adam
moments), in `parameters()` order, so state_dict / checkpoints / TorchScript export keep working on the
ordinary nn.Module while the kernels (and the multi-GPU all-reduce) see one contiguous array.
"""
import ctypes as C

import torch

from .. import _lib


def _linears(seq):
    return [(i, m) for i, m in enumerate(seq) if isinstance(m, torch.nn.Linear)]


class PolicyEngine:
    def __init__(self, actor_critic, max_rows, max_seq=1, seed=1, env_offset=0, precision=1):
        self.ac = actor_critic
        params = list(actor_critic.parameters())
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("unitree_rl_gym_b200.rsl_rl runs on CUDA only (policy is on %s); no CPU fallback" % dev)
        self.device = dev
        for _, m in actor_critic.named_modules():
            if isinstance(m, (torch.nn.SELU, torch.nn.ReLU, torch.nn.LeakyReLU, torch.nn.Tanh, torch.nn.Sigmoid)):
                raise NotImplementedError("the fused sm_100a MLP implements activation='elu' (the reference's setting, "
                                          "legged_robot_config.py:251); got %s" % type(m).__name__)
        self.max_rows, self.max_seq = int(max_rows), int(max_seq)
        n = sum(p.numel() for p in params)
        self.num_params = n
        self.flat = torch.zeros(n, device=dev)
        # flat gradients and the stats block share one allocation: [grads (n) | stats (STAT_COUNT)], so that the
        # data-parallel update all-reduces the gradients and the four per-minibatch sums (stats[0:4]) in one collective
        self._grads_stats = torch.zeros(n + _lib.STAT_COUNT, device=dev)
        self.grads = self._grads_stats[:n]
        self.exp_avg = torch.zeros(n, device=dev)
        self.exp_avg_sq = torch.zeros(n, device=dev)
        self.stats = self._grads_stats[n:]
        self.grads_and_sums = self._grads_stats[:n + _lib.STAT_ENT_SUM + 1]
        self.offsets = {}
        off = 0
        with torch.no_grad():
            for name, p in actor_critic.named_parameters():
                k = p.numel()
                self.flat[off:off + k].copy_(p.detach().reshape(-1))
                p.data = self.flat[off:off + k].view(p.shape)
                p.grad = self.grads[off:off + k].view(p.shape)
                self.offsets[name] = off
                off += k
        d = _lib.PolicyDesc()
        rec = bool(getattr(actor_critic, "is_recurrent", False))
        d.recurrent = int(rec)
        d.num_actions = actor_critic.std.numel()
        self._fill_mlp(d.actor, actor_critic.actor, "actor")
        self._fill_mlp(d.critic, actor_critic.critic, "critic")
        d.std_off = self.offsets["std"]
        if rec:
            ra, rc = actor_critic.memory_a.rnn, actor_critic.memory_c.rnn
            if not isinstance(ra, torch.nn.LSTM) or ra.num_layers != 1:
                raise NotImplementedError("recurrent policy: single-layer LSTM only (h1_config.py:110-113)")
            d.rnn_hidden = ra.hidden_size
            d.num_actor_obs, d.num_critic_obs = ra.input_size, rc.input_size
            for i, nm in enumerate(("weight_ih_l0", "weight_hh_l0", "bias_ih_l0", "bias_hh_l0")):
                d.lstm_a_off[i] = self.offsets["memory_a.rnn." + nm]
                d.lstm_c_off[i] = self.offsets["memory_c.rnn." + nm]
        else:
            d.num_actor_obs, d.num_critic_obs = d.actor.dims[0], d.critic.dims[0]
        d.num_params = n
        d.max_rows, d.max_seq = self.max_rows, self.max_seq
        d.seed_lo, d.seed_hi = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
        d.env_offset = env_offset
        d.precision = int(precision)
        self.precision = int(precision)
        self.desc = d
        L = _lib.lib()
        nbytes = L.urlg_policy_workspace_bytes(C.byref(d))
        if nbytes < 0:
            _lib.check(int(nbytes))
        self.workspace = torch.zeros(int(nbytes) + 1024, dtype=torch.uint8, device=dev)
        base = (self.workspace.data_ptr() + 1023) & ~1023
        h = C.c_void_p()
        _lib.check(L.urlg_policy_create(C.byref(d), _lib.ptr(self.flat), _lib.ptr(self.grads), _lib.ptr(self.exp_avg),
                                        _lib.ptr(self.exp_avg_sq), _lib.ptr(self.stats), C.c_void_p(base), int(nbytes),
                                        C.byref(h)))
        self.handle = h
        self.sync_weights()

    def _fill_mlp(self, md, seq, prefix):
        lin = _linears(seq)
        if len(lin) > _lib.MAX_LAYERS:
            raise ValueError("MLP deeper than URLG_MAX_LAYERS")
        md.num_layers = len(lin)
        for l, (i, m) in enumerate(lin):
            md.dims[l] = m.in_features
            md.dims[l + 1] = m.out_features
            md.w_off[l] = self.offsets[f"{prefix}.{i}.weight"]
            md.b_off[l] = self.offsets[f"{prefix}.{i}.bias"]

    def owns(self, actor_critic):
        """True while the module's parameters are still the views this engine created (a later .to() /
        load into new tensors would silently detach them)."""
        p = next(actor_critic.parameters())
        return p.data_ptr() == self.flat.data_ptr() + 4 * self.offsets[next(iter(self.offsets))]

    def sync_weights(self):
        _lib.check(_lib.lib().urlg_policy_sync_weights(self.handle, _lib.cur_stream()))

    def forward(self, which, x, hidden_h=None, hidden_c=None):
        """actor (which=0) / critic (which=1) forward on [rows, in] -> [rows, out]."""
        x = x.to(device=self.device, dtype=torch.float).contiguous()
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        md = self.desc.actor if which == 0 else self.desc.critic
        out = torch.empty(x2.shape[0], md.dims[md.num_layers], device=self.device)
        _lib.check(_lib.lib().urlg_policy_forward(self.handle, which, _lib.ptr(x2), x2.shape[0], _lib.ptr(hidden_h),
                                                  _lib.ptr(hidden_c), _lib.ptr(out), _lib.cur_stream()))
        return out.reshape(*lead, out.shape[-1])

    def __del__(self):
        h = getattr(self, "handle", None)
        if h is not None:
            try:
                _lib.lib().urlg_policy_destroy(h)
            except Exception:
                pass
            self.handle = None
