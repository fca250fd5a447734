"""Trajectory (un)padding helpers of the public rsl_rl API ([UPSTREAM rsl_rl v1.0.2] utils/utils.py; SURVEY.md
§8(a).6 R9).  The fused recurrent update does NOT use them: it runs the equivalent masked time loop on the
device (csrc/lstm_kernels.cu) with no host sync; they stay for callers of the public generator API."""
import torch


def split_and_pad_trajectories(tensor, dones):
    """[T,N,...] + dones [T,N,1] -> (padded [T', n_traj, ...], masks [T, n_traj]); one trajectory per stretch
    between dones, and the window end closes every env's last one."""
    dones = dones.clone()
    dones[-1] = 1
    flat = dones.transpose(1, 0).reshape(-1, 1)
    ends = torch.cat((flat.new_tensor([-1], dtype=torch.int64), flat.nonzero()[:, 0]))
    lengths = ends[1:] - ends[:-1]
    pieces = torch.split(tensor.transpose(1, 0).flatten(0, 1), lengths.tolist())
    padded = torch.nn.utils.rnn.pad_sequence(pieces)
    masks = lengths > torch.arange(0, tensor.shape[0], device=tensor.device).unsqueeze(1)
    return padded, masks


def unpad_trajectories(trajectories, masks):
    return (trajectories.transpose(1, 0)[masks.transpose(1, 0)]
            .view(-1, trajectories.shape[0], trajectories.shape[-1]).transpose(1, 0))
