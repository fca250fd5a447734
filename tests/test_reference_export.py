"""the reference's UNMODIFIED exporter (`/root/reference/legged_gym/utils/helpers.py:150-189`,
`export_policy_as_jit` / `PolicyExporterLSTM`) is handed this repo's ActorCritic / ActorCriticRecurrent and must produce
the TorchScript modules that deploy_mujoco.py:77 / deploy_real.py:29 load: same attribute surface (`.actor` scriptable,
`.memory_a.rnn` an nn.LSTM, `.is_recurrent`), same numbers as the module's own layers.

Container-only (`refonly`: imports the reference under oracle/refharness); CPU: only the nn.Module surface is exercised.
"""
import os

import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.refonly


def _exporter():
    from oracle.refharness.run_reference import import_reference
    import_reference()
    from legged_gym.utils.helpers import export_policy_as_jit
    import legged_gym.utils.helpers as ref_helpers
    assert ref_helpers.__file__.startswith("/root/reference")
    return export_policy_as_jit


def test_reference_exporter_scripts_the_mlp_policy(tmp_path):
    from unitree_rl_gym_b200.rsl_rl.modules import ActorCritic
    torch.manual_seed(3)
    ac = ActorCritic(48, 48, 12, [512, 256, 128], [512, 256, 128], "elu", init_noise_std=1.0)
    _exporter()(ac, str(tmp_path))
    m = torch.jit.load(os.path.join(str(tmp_path), "policy_1.pt"))
    x = torch.randn(7, 48)
    with torch.no_grad():
        want = ac.actor(x)              # the plain nn.Sequential the CUDA path mirrors (act_inference = actor(obs))
    np.testing.assert_allclose(m(x).detach().numpy(), want.numpy(), rtol=1e-6, atol=1e-6)


def test_reference_exporter_scripts_the_lstm_policy_and_reproduces_motion_pt(tmp_path):
    """The exported module of a B200 ActorCriticRecurrent loaded with the weights of deploy/pre_train/g1/motion.pt (itself a
    PolicyExporterLSTM export of a trained rsl_rl model) reproduces that file's recorded 8-step sequence."""
    from unitree_rl_gym_b200.rsl_rl.modules import ActorCriticRecurrent
    z = np.load(os.path.join(H.GOLDEN, "motion_pt.npz"))
    robot = "g1"
    sd = {k[len(robot) + 4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith(robot + "/sd/")}
    n_in, n_act = sd["memory.weight_ih_l0"].shape[1], sd["actor.2.weight"].shape[0]
    ac = ActorCriticRecurrent(n_in, n_in, n_act, [32], [32], "elu", rnn_type="lstm", rnn_hidden_size=64, rnn_num_layers=1)
    own = ac.state_dict()
    for k in ("weight_ih_l0", "weight_hh_l0", "bias_ih_l0", "bias_hh_l0"):
        own["memory_a.rnn." + k] = sd["memory." + k]
    for k in ("actor.0.weight", "actor.0.bias", "actor.2.weight", "actor.2.bias"):
        own[k] = sd[k]
    ac.load_state_dict(own)
    assert ac.is_recurrent and isinstance(ac.memory_a.rnn, torch.nn.LSTM)
    _exporter()(ac, str(tmp_path))
    m = torch.jit.load(os.path.join(str(tmp_path), "policy_lstm_1.pt"))
    assert {"hidden_state", "cell_state"} <= {k for k, _ in m.named_buffers()}
    m.reset_memory()
    seq_in, seq_out = torch.from_numpy(z[robot + "/seq_in"]), z[robot + "/seq_out"]
    got = np.stack([m(seq_in[t]).detach().numpy() for t in range(seq_in.shape[0])])
    np.testing.assert_allclose(got, seq_out, rtol=1e-5, atol=2e-6)
