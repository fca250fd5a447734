"""Euler angles from xyzw quaternions in (-pi, pi] (legged_gym/utils/isaacgym_utils.py:11-30);
check_termination's abs() relies on this range (legged_robot.py:718)."""
import math
import torch


def copysign(a: float, b):
    return torch.full_like(b, abs(a)) * torch.sign(b)


def get_euler_xyz(q):
    x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    roll = torch.atan2(2.0 * (w * x + y * z), w * w - x * x - y * y + z * z)
    sinp = 2.0 * (w * y - z * x)
    pitch = torch.where(sinp.abs() >= 1, copysign(math.pi / 2.0, sinp), torch.asin(sinp))
    yaw = torch.atan2(2.0 * (w * z + x * y), w * w + x * x - y * y - z * z)
    return torch.stack((roll, pitch, yaw), dim=-1)
