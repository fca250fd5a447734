"""Stub gymapi: a fake `gym` whose tensor API serves synthetic physics frames.

Seam replaced (reference call sites): acquire_*_tensor legged_robot.py:83-88,
h1_env.py:37; refresh_* :90-92,639,678-679, h1_env.py:49; simulate :630;
set_dof_actuation_force_tensor :629; set_*_tensor_indexed :553,570,592.
"""
import types
import numpy as np
import torch

SIM_PHYSX = 1
SIM_FLEX = 0
KEY_ESCAPE = 0
KEY_V = 1


class _NS(types.SimpleNamespace):
    pass


def Vec3(*a):
    return _NS(x=a[0] if len(a) > 0 else 0., y=a[1] if len(a) > 1 else 0., z=a[2] if len(a) > 2 else 0.)


def PlaneParams():
    return _NS()


def Transform():
    return _NS(p=None, r=None)


def AssetOptions():
    return _NS()


def CameraProperties():
    return _NS()


def SimParams():
    return _NS(dt=1. / 60., substeps=2, up_axis=1, gravity=None, use_gpu_pipeline=False,
               physx=_NS(use_gpu=False, num_subscenes=0, num_threads=0))


class FakeGym:
    """Serves frames from `self.frames` (set by the harness):
       frames = dict(root [P,N,13], dof [P,D,N*A,2], contact [P,N*B,3], rigid_body [P,N*B,13] or None)
       frame index = (simulate calls - 1) // decimation, substep = (calls - 1) % decimation."""
    def __init__(self):
        self.robot = None      # RobotTable set by the harness before env construction
        self.frames = None
        self.decimation = 1
        self.sim_calls = 0
        self.N = 0
        self.t_root = self.t_dof = self.t_contact = self.t_rb = None
        self.last_torques = None
        self.indexed_calls = []

    # ---- construction-time API (answered from the URDF-derived table) ----
    def create_sim(self, *a): return "sim"
    def add_ground(self, *a): pass
    def prepare_sim(self, *a): pass
    def load_asset(self, *a): return "asset"
    def get_asset_dof_count(self, a): return len(self.robot.dof_names)
    def get_asset_rigid_body_count(self, a): return len(self.robot.body_names)
    def get_asset_rigid_body_names(self, a): return list(self.robot.body_names)
    def get_asset_dof_names(self, a): return list(self.robot.dof_names)
    def get_asset_dof_properties(self, a):
        n = len(self.robot.dof_names)
        p = np.zeros(n, dtype=[('lower', 'f4'), ('upper', 'f4'), ('velocity', 'f4'), ('effort', 'f4')])
        p['lower'] = self.robot.dof_lower; p['upper'] = self.robot.dof_upper
        p['velocity'] = self.robot.dof_velocity; p['effort'] = self.robot.dof_effort
        return p
    def get_asset_rigid_shape_properties(self, a): return []
    def set_asset_rigid_shape_properties(self, *a): pass
    def create_env(self, *a): self.N += 1; return self.N - 1
    def create_actor(self, *a): return 0
    def set_actor_dof_properties(self, *a): pass
    def get_actor_rigid_body_properties(self, *a): return [_NS(mass=1.0)]
    def set_actor_rigid_body_properties(self, *a, **k): pass
    def find_actor_rigid_body_handle(self, env, actor, name): return self.robot.body_names.index(name)

    # ---- tensor API ----
    def _alloc(self):
        if self.t_root is None:
            N, A, B = self.N, len(self.robot.dof_names), len(self.robot.body_names)
            self.t_root = torch.zeros(N, 13)
            self.t_root[:, 6] = 1.0
            self.t_dof = torch.zeros(N * A, 2)
            self.t_contact = torch.zeros(N * B, 3)
            self.t_rb = torch.zeros(N * B, 13)
    def acquire_actor_root_state_tensor(self, sim): self._alloc(); return self.t_root
    def acquire_dof_state_tensor(self, sim): self._alloc(); return self.t_dof
    def acquire_net_contact_force_tensor(self, sim): self._alloc(); return self.t_contact
    def acquire_rigid_body_state_tensor(self, sim): self._alloc(); return self.t_rb

    def _idx(self):
        c = self.sim_calls - 1
        P = self.frames['root'].shape[0]
        return (c // self.decimation) % P, c % self.decimation
    def simulate(self, sim): self.sim_calls += 1
    def fetch_results(self, *a): pass
    def refresh_dof_state_tensor(self, sim):
        if self.frames is None or self.sim_calls == 0: return
        f, s = self._idx(); self.t_dof.copy_(self.frames['dof'][f, s])
    def refresh_actor_root_state_tensor(self, sim):
        if self.frames is None or self.sim_calls == 0: return
        f, _ = self._idx(); self.t_root.copy_(self.frames['root'][f])
    def refresh_net_contact_force_tensor(self, sim):
        if self.frames is None or self.sim_calls == 0: return
        f, _ = self._idx(); self.t_contact.copy_(self.frames['contact'][f])
    def refresh_rigid_body_state_tensor(self, sim):
        if self.frames is None or self.sim_calls == 0 or self.frames.get('rigid_body') is None: return
        f, _ = self._idx(); self.t_rb.copy_(self.frames['rigid_body'][f])
    def set_dof_actuation_force_tensor(self, sim, t): self.last_torques = t
    def set_dof_state_tensor_indexed(self, sim, t, ids, n): self.indexed_calls.append(('dof', ids.clone(), n))
    def set_actor_root_state_tensor_indexed(self, sim, t, ids, n): self.indexed_calls.append(('root', ids.clone(), n))
    def get_elapsed_time(self, sim): return 0.
    def get_sim_time(self, sim): return 0.


_GYM = None


def acquire_gym():
    global _GYM
    if _GYM is None:
        _GYM = FakeGym()
    return _GYM


def _new_gym():
    global _GYM
    _GYM = FakeGym()
    return _GYM
