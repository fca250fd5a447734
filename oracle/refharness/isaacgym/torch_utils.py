"""[UPSTREAM restatement] isaacgym.torch_utils: the six helpers the reference
pulls in with `from isaacgym.torch_utils import *` (legged_robot.py:7).
Quaternions are xyzw.  `torch_rand_float` can be redirected to a shared
counter-based uniform table (see RNG below) so the reference, the oracle port
and the CUDA kernels consume identical random numbers.
"""
import numpy as np
import torch


def to_torch(x, dtype=torch.float, device='cuda:0', requires_grad=False):
    return torch.tensor(x, dtype=dtype, device=device, requires_grad=requires_grad)


def get_axis_params(value, axis_idx, x_value=0., dtype=float, n_dims=3):
    zs = np.zeros((n_dims,))
    assert axis_idx < n_dims, "the axis dim should be within the vector dimensions"
    zs[axis_idx] = 1.
    params = np.where(zs == 1., value, zs)
    params[0] = x_value
    return list(params.astype(dtype))


def normalize(x, eps: float = 1e-9):
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)


def quat_apply(a, b):
    shape = b.shape
    a = a.reshape(-1, 4)
    b = b.reshape(-1, 3)
    xyz = a[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return (b + a[:, 3:] * t + xyz.cross(t, dim=-1)).view(shape)


def quat_rotate_inverse(q, v):
    shape = q.shape
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(shape[0], 1, 3), v.view(shape[0], 3, 1)).squeeze(-1) * 2.0
    return a - b + c


class _RngContext:
    """When `table_fn` is set, torch_rand_float reads uniforms from
    table_fn(step)[env_ids, slot:slot+cols] instead of torch.rand."""
    def __init__(self):
        self.table_fn = None
        self.step = 0
        self.env_ids = None
        self.slot = 0

    def active(self):
        return self.table_fn is not None and self.env_ids is not None

    def take(self, rows, cols):
        tab = self.table_fn(self.step)
        ids = self.env_ids
        assert rows == len(ids), f"rng ctx rows {rows} != env_ids {len(ids)}"
        u = tab[ids, self.slot:self.slot + cols]
        assert u.shape[1] == cols, "rng table too narrow"
        self.slot += cols
        return u


RNG = _RngContext()


def torch_rand_float(lower, upper, shape, device):
    if RNG.active():
        u = RNG.take(shape[0], shape[1]).to(device)
    else:
        u = torch.rand(*shape, device=device)
    return (upper - lower) * u + lower
