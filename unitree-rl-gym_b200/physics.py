"""The physics seam.  The reference consumes PhysX through Isaac Gym's tensor API
(legged_robot.py:83-119 acquire/wrap, :629-639 actuate/simulate/refresh, :678-679 refresh,
:553,570,592 indexed setters; h1_env.py:37-49 rigid-body tensor).  PhysX itself is external
to the hot path (BASELINE.json north_star); `SyntheticGym` implements the same calls over a
device-resident pool of synthetic frames so the env runs, is testable and is benchable
without Isaac Gym.  A real-PhysX binding would expose the same method names.

Frame addressing is graph-safe: kernels read (device `calls_base` + host offset), and a captured
graph ends with `advance_base()` so replays keep cycling through the pool.

Two ways of serving a frame: `pool_seam=False` copies it into the acquired tensors on every refresh_* (one
urlg_synth_refresh launch per tensor: the stand-in for PhysX filling its buffers); `pool_seam=True` hands the
kernels the pool itself (urlg_env_bind_frame_pool) so refresh_* launches nothing and a timed region holds only the
path.  Both give identical env outputs.  `StatefulGym` is the protocol a real simulator follows: its state lives
apart from the exposed tensors and only the indexed setters write back into it.
"""
import torch

from . import _lib


class SyntheticGym:
    def __init__(self, robot, num_envs, decimation, device, frames=None, pool_seam=False, packed=False):
        self.pool_seam = bool(pool_seam)
        # packed: frame f of root / dof (all substeps) / contact (/ rigid body) are consecutive sections of ONE row of
        # `frame_buf` [P, frame_floats], so a host-side engine uploads a frame with a single contiguous copy
        # (urlg_frame_pool.frame_stride); the per-tensor `frames[k]` stay available as views of it.  Pool seam only.
        self.packed = bool(packed)
        self.frame_buf = None
        if self.packed and not self.pool_seam:
            raise ValueError("SyntheticGym(packed=True) needs pool_seam=True")
        self.robot = robot
        self.num_envs = num_envs
        self.decimation = decimation
        self.device = torch.device(device)
        A, B = robot.num_dof, robot.num_bodies
        dev = self.device
        self.root_states = torch.zeros(num_envs, 13, device=dev)
        self.root_states[:, 6] = 1.0
        self.dof_state = torch.zeros(num_envs * A, 2, device=dev)
        self.net_contact_forces = torch.zeros(num_envs * B, 3, device=dev)
        self.rigid_body_states = torch.zeros(num_envs * B, 13, device=dev)
        self.calls_base = torch.zeros(1, dtype=torch.int64, device=dev)   # device part of the simulate() counter
        self.sim_calls = 0                                                # host part (offset)
        self.frames = None
        self.last_torques = None
        self.indexed_calls = 0
        self.last_dof_set = None      # (ids int32 [N], count int32 [1]) of the last set_dof_state_tensor_indexed
        self.last_root_set = None
        if frames is not None:
            self.set_frames(frames)

    def set_frames(self, frames):
        """frames: dict from synthetic.make_frames (CPU or CUDA tensors)."""
        if self.packed:
            keys = [k for k in ("root", "dof", "contact", "rigid_body") if frames.get(k) is not None]
            P = frames["root"].shape[0]
            sizes = [frames[k][0].numel() for k in keys]
            if any(n % 4 for n in sizes):
                raise ValueError("packed frames: every section must be a multiple of 4 floats")
            self.frame_buf = torch.empty(P, sum(sizes), device=self.device)
            self.frames, off = {k: None for k in frames}, 0
            for k, n in zip(keys, sizes):
                view = self.frame_buf[:, off:off + n].view((P,) + tuple(frames[k].shape[1:]))
                view.copy_(frames[k])
                self.frames[k] = view
                off += n
            self.num_frames = P
            return
        self.frames = {k: (v.to(self.device).contiguous() if v is not None else None) for k, v in frames.items()}
        self.num_frames = self.frames["root"].shape[0]

    # ---- construction-time asset queries (answered from robots.RobotTable) ----
    def get_asset_dof_names(self):
        return list(self.robot.dof_names)

    def get_asset_rigid_body_names(self):
        return list(self.robot.body_names)

    def find_actor_rigid_body_handle(self, name):
        return self.robot.body_names.index(name)

    # ---- tensor API ----
    def acquire_actor_root_state_tensor(self, sim=None):
        return self.root_states

    def acquire_dof_state_tensor(self, sim=None):
        return self.dof_state

    def acquire_net_contact_force_tensor(self, sim=None):
        return self.net_contact_forces

    def acquire_rigid_body_state_tensor(self, sim=None):
        return self.rigid_body_states

    def simulate(self, sim=None):
        self.sim_calls += 1

    def fetch_results(self, *a):
        pass

    def frame_pool_desc(self):
        """urlg_frame_pool over the device-resident frames (pool_seam=True)."""
        if self.frames is None:
            raise RuntimeError("SyntheticGym(pool_seam=True) needs frames before the env binds")
        d = _lib.FramePool()
        d.root, d.dof, d.contact = _lib.ptr(self.frames["root"]), _lib.ptr(self.frames["dof"]), _lib.ptr(self.frames["contact"])
        d.rigid_body = _lib.ptr(self.frames.get("rigid_body"))
        d.num_frames, d.substeps = self.num_frames, self.decimation
        d.calls_base = _lib.ptr(self.calls_base)
        d.frame_stride = self.frame_buf.shape[1] if self.packed else 0
        return d

    def _refresh(self, dst, key, per_substep):
        if self.pool_seam or self.frames is None or self.frames.get(key) is None:
            return
        pool = self.frames[key]
        elems = dst.numel()
        _lib.check(_lib.lib().urlg_synth_refresh(_lib.ptr(dst), _lib.ptr(pool), elems, self.num_frames,
                                                 self.decimation, 1 if per_substep else 0,
                                                 _lib.ptr(self.calls_base), self.sim_calls, _lib.cur_stream()))

    def refresh_dof_state_tensor(self, sim=None):
        self._refresh(self.dof_state, "dof", True)

    def refresh_actor_root_state_tensor(self, sim=None):
        self._refresh(self.root_states, "root", False)

    def refresh_net_contact_force_tensor(self, sim=None):
        self._refresh(self.net_contact_forces, "contact", False)

    def refresh_rigid_body_state_tensor(self, sim=None):
        self._refresh(self.rigid_body_states, "rigid_body", False)

    def set_dof_actuation_force_tensor(self, sim, torques):
        self.last_torques = torques

    # Indexed setters (legged_robot.py:552-555, 569-572, 591-594).  The drop-in env passes (state tensor, int32 id
    # tensor [N], int32 count tensor [1]) — ids and count stay on the device.  A pool of synthetic frames has no state
    # to write back into, so these only record what they were given.
    def set_dof_state_tensor_indexed(self, sim, state, ids, count):
        self.indexed_calls += 1
        self.last_dof_set = (ids, count)

    def set_actor_root_state_tensor_indexed(self, sim, state, ids, count):
        self.indexed_calls += 1
        self.last_root_set = (ids, count)

    # ---- graph support ----
    def begin_relative(self):
        """Fold the host counter into the device base (call before capturing a graph)."""
        self.calls_base += self.sim_calls
        self.sim_calls = 0

    def advance_base(self):
        """Captured as the last node of a graph: base += calls made inside the graph."""
        self.calls_base += self.sim_calls


class StatefulGym(SyntheticGym):
    """A simulator whose state lives APART from the tensors it exposes — how PhysX behaves: `simulate` advances the
    hidden state (here: a fixed drift per call), `refresh_*` copies hidden -> exposed, and ONLY the indexed setters
    copy exposed rows back into the hidden state.  An env that rewrites reset / pushed rows in the exposed tensors
    without committing them through the setters loses those writes at the next refresh (tests/test_env_cuda.py)."""

    def __init__(self, robot, num_envs, decimation, device, seed=0):
        super().__init__(robot, num_envs, decimation, device, frames=None, pool_seam=False)
        g = torch.Generator().manual_seed(seed)
        A, B = robot.num_dof, robot.num_bodies
        dev = self.device
        self.h_root = torch.zeros(num_envs, 13)
        self.h_root[:, 2] = 0.4
        self.h_root[:, 6] = 1.0
        self.h_root[:, 7:13] = 0.1 * torch.randn(num_envs, 6, generator=g)
        self.h_root = self.h_root.to(dev)
        self.h_dof = (0.2 * torch.randn(num_envs * A, 2, generator=g)).to(dev)
        self.h_contact = torch.zeros(num_envs * B, 3, device=dev)
        self.drift = 1e-3

    def simulate(self, sim=None):
        self.sim_calls += 1
        self.h_dof[:, 0] += self.drift
        self.h_root[:, 0] += self.drift

    def refresh_dof_state_tensor(self, sim=None):
        self.dof_state.copy_(self.h_dof)

    def refresh_actor_root_state_tensor(self, sim=None):
        self.root_states.copy_(self.h_root)

    def refresh_net_contact_force_tensor(self, sim=None):
        self.net_contact_forces.copy_(self.h_contact)

    def refresh_rigid_body_state_tensor(self, sim=None):
        pass

    def _scatter(self, dst, src, ids, count, row_floats):
        _lib.check(_lib.lib().urlg_scatter_rows(_lib.ptr(dst), _lib.ptr(src), _lib.ptr(ids), _lib.ptr(count), row_floats,
                                                self.num_envs, _lib.cur_stream()))

    def set_dof_state_tensor_indexed(self, sim, state, ids, count):
        super().set_dof_state_tensor_indexed(sim, state, ids, count)
        self._scatter(self.h_dof, state, ids, count, 2 * self.robot.num_dof)

    def set_actor_root_state_tensor_indexed(self, sim, state, ids, count):
        super().set_actor_root_state_tensor_indexed(sim, state, ids, count)
        self._scatter(self.h_root, state, ids, count, 13)
