"""The physics seam.  The reference consumes PhysX through Isaac Gym's tensor API
(legged_robot.py:83-119 acquire/wrap, :629-639 actuate/simulate/refresh, :678-679 refresh,
:553,570,592 indexed setters; h1_env.py:37-49 rigid-body tensor).  PhysX itself is external
to the hot path (BASELINE.json north_star); `SyntheticGym` implements the same calls over a
device-resident pool of synthetic frames so the env runs, is testable and is benchable
without Isaac Gym.  A real-PhysX binding would expose the same method names.

Frame addressing is graph-safe: kernels read (device `calls_base` + host offset), and a captured
graph ends with `advance_base()` so replays keep cycling through the pool.
"""
import torch

from . import _lib


class SyntheticGym:
    def __init__(self, robot, num_envs, decimation, device, frames=None):
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
        if frames is not None:
            self.set_frames(frames)

    def set_frames(self, frames):
        """frames: dict from synthetic.make_frames (CPU or CUDA tensors)."""
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

    def _refresh(self, dst, key, per_substep):
        if self.frames is None or self.frames.get(key) is None:
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

    def set_dof_state_tensor_indexed(self, *a):
        self.indexed_calls += 1

    def set_actor_root_state_tensor_indexed(self, *a):
        self.indexed_calls += 1

    # ---- graph support ----
    def begin_relative(self):
        """Fold the host counter into the device base (call before capturing a graph)."""
        self.calls_base += self.sim_calls
        self.sim_calls = 0

    def advance_base(self):
        """Captured as the last node of a graph: base += calls made inside the graph."""
        self.calls_base += self.sim_calls
