"""Synthetic physics frames: the stand-in for PhysX at the tensor seam.

The reference reads root_states / dof_state / net_contact_forces (and, for the
H1 family, rigid_body_states) as zero-copy views of PhysX buffers
(legged_robot.py:83-119, h1_env.py:37-39).  BASELINE.json's north_star keeps
that seam and drives it with synthetic tensors; distributions follow SURVEY.md
§8(d) so that every branch (terminations, collisions, limits, contacts) fires at
realistic rates.  Frames are generated on CPU from a seeded generator so that the
reference harness, the oracle port and the CUDA path see identical bits.
"""
import math
import torch


def make_frames(robot, cfg_like, num_envs, num_frames, decimation, seed=1, feet_idx=None,
                pen_idx=None, term_idx=None, init_z=1.0, default_dof_pos=None, env_spacing=3.0):
    """Returns dict of CPU float32 tensors:
         root [P,N,13], dof [P,D,N*A,2], contact [P,N*B,3], rigid_body [P,N*B,13]"""
    g = torch.Generator().manual_seed(seed)
    P, N, D = num_frames, num_envs, decimation
    A, B = robot.num_dof, robot.num_bodies
    feet_idx = list(feet_idx or [])
    pen_idx = list(pen_idx or [])
    term_idx = list(term_idx or [])

    def U(*shape, lo=0., hi=1.):
        return torch.rand(*shape, generator=g) * (hi - lo) + lo

    def G(*shape, std=1.):
        return torch.randn(*shape, generator=g) * std

    root = torch.zeros(P, N, 13)
    cols = math.floor(math.sqrt(N))
    rows = math.ceil(N / cols)
    xx, yy = torch.meshgrid(torch.arange(rows), torch.arange(cols), indexing="ij")
    root[:, :, 0] = env_spacing * xx.flatten()[:N].float()
    root[:, :, 1] = env_spacing * yy.flatten()[:N].float()
    root[:, :, 2] = U(P, N, lo=0.8, hi=1.2) * init_z
    axis = G(P, N, 3)
    axis = axis / axis.norm(dim=-1, keepdim=True).clamp(min=1e-6)
    ang = G(P, N, std=0.35).abs()
    root[:, :, 3:6] = axis * torch.sin(0.5 * ang).unsqueeze(-1)
    root[:, :, 6] = torch.cos(0.5 * ang)
    root[:, :, 7:10] = G(P, N, 3, std=0.5)
    root[:, :, 10:13] = G(P, N, 3, std=0.7)

    if default_dof_pos is None:
        default_dof_pos = torch.zeros(A)
    dof = torch.zeros(P, D, N, A, 2)
    dof[..., 0] = default_dof_pos.view(1, 1, 1, A) + U(P, D, N, A, lo=-0.6, hi=0.6)
    dof[..., 1] = G(P, D, N, A, std=3.0)
    dof = dof.reshape(P, D, N * A, 2).contiguous()

    contact = torch.zeros(P, N, B, 3)
    for f in feet_idx:
        on = (U(P, N) >= 0.45).float()
        fz = U(P, N, lo=0., hi=250.) * on
        contact[:, :, f, 2] = fz
        contact[:, :, f, 0:2] = G(P, N, 2, std=10.) * on.unsqueeze(-1)
    for b in pen_idx:
        hit = (U(P, N) < 0.03).float()
        d = G(P, N, 3)
        d = d / d.norm(dim=-1, keepdim=True).clamp(min=1e-6)
        contact[:, :, b, :] = d * (U(P, N, lo=0., hi=50.) * hit).unsqueeze(-1)
    for b in term_idx:
        hit = (U(P, N) < 0.005).float()
        d = G(P, N, 3)
        d = d / d.norm(dim=-1, keepdim=True).clamp(min=1e-6)
        contact[:, :, b, :] = d * (U(P, N, lo=1.5, hi=80.) * hit + U(P, N, lo=0., hi=0.5) * (1 - hit)).unsqueeze(-1)
    contact = contact.reshape(P, N * B, 3).contiguous()

    rb = torch.zeros(P, N, B, 13)
    rb[..., 6] = 1.0
    for f in feet_idx:
        rb[:, :, f, 0:2] = root[:, :, 0:2] + G(P, N, 2, std=0.2)
        rb[:, :, f, 2] = U(P, N, lo=0., hi=0.25)
        rb[:, :, f, 7:10] = G(P, N, 3, std=1.0)
    rb = rb.reshape(P, N * B, 13).contiguous()
    return {"root": root.contiguous(), "dof": dof, "contact": contact, "rigid_body": rb}
