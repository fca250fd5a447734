"""Host-side math helpers kept for API parity (legged_gym/utils/math.py:8-26).  The
hot path computes these inside the fused CUDA kernels; these torch versions serve
user code and subclasses that call them directly."""
import math
import torch


def quat_apply(q, v):
    """[UPSTREAM isaacgym.torch_utils] rotate v by xyzw quaternion q."""
    shape = v.shape
    q = q.reshape(-1, 4)
    v = v.reshape(-1, 3)
    u = q[:, :3]
    t = 2.0 * torch.cross(u, v, dim=-1)
    return (v + q[:, 3:4] * t + torch.cross(u, t, dim=-1)).view(shape)


def quat_rotate_inverse(q, v):
    """[UPSTREAM isaacgym.torch_utils] rotate v by the inverse of xyzw quaternion q."""
    w = q[:, 3:4]
    u = q[:, :3]
    return v * (2.0 * w * w - 1.0) - 2.0 * w * torch.cross(u, v, dim=-1) + 2.0 * u * (u * v).sum(-1, keepdim=True)


def normalize(x, eps: float = 1e-9):
    return x / x.norm(p=2, dim=-1).clamp(min=eps).unsqueeze(-1)


def quat_apply_yaw(quat, vec):
    yaw_only = quat.clone().view(-1, 4)
    yaw_only[:, :2] = 0.
    return quat_apply(normalize(yaw_only), vec)


def wrap_to_pi(angles):
    """In-place wrap to (-pi, pi] with Python-sign modulo, as math.py:15-18."""
    angles %= 2 * math.pi
    angles -= 2 * math.pi * (angles > math.pi)
    return angles


def torch_rand_sqrt_float(lower, upper, shape, device):
    r = 2 * torch.rand(*shape, device=device) - 1
    r = torch.sign(r) * torch.sqrt(r.abs())
    return (upper - lower) * (r + 1.) / 2. + lower
