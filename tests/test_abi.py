"""The C-ABI library loads on a CPU-only box and exports every symbol include/urlg_b200.h declares."""
import ctypes
import os
import re

from tests.helpers import GOLDEN  # noqa: F401

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(REPO, "include", "urlg_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(urlg_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from unitree_rl_gym_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/urlg_b200.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature in _lib.py"
    assert L.urlg_version() == _lib.ABI_VERSION == 3


def test_struct_sizes_match_header():
    """sizeof(EnvDesc/EnvBuffers) in ctypes == what nvcc compiled (checked through a tiny C probe)."""
    import subprocess
    import tempfile
    from unitree_rl_gym_b200 import _lib
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "p.c")
        open(c, "w").write('#include <stdio.h>\n#include "urlg_b200.h"\nint main(){printf("%zu %zu\\n",'
                           'sizeof(urlg_env_desc),sizeof(urlg_env_buffers));return 0;}\n')
        exe = os.path.join(d, "p")
        subprocess.check_call(["gcc", "-I", os.path.join(REPO, "include"), c, "-o", exe])
        a, b = map(int, subprocess.check_output([exe]).split())
    assert ctypes.sizeof(_lib.EnvDesc) == a
    assert ctypes.sizeof(_lib.EnvBuffers) == b


def test_no_cpu_fallback():
    import pytest
    import torch
    from unitree_rl_gym_b200 import _lib
    with pytest.raises(RuntimeError):
        _lib.ptr(torch.zeros(4))


def _probe(body):
    import subprocess
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "p.c")
        open(c, "w").write('#include <stdio.h>\n#include "urlg_b200.h"\nint main(){' + body + 'return 0;}\n')
        exe = os.path.join(d, "p")
        subprocess.check_call(["gcc", "-I", os.path.join(REPO, "include"), c, "-o", exe])
        return subprocess.check_output([exe]).split()


def test_policy_side_struct_sizes_match_header():
    """The ctypes mirrors of the policy / storage / PPO structs have the size gcc gives the header's."""
    from unitree_rl_gym_b200 import _lib
    names = [("urlg_gemm_desc", _lib.GemmDesc), ("urlg_mlp_desc", _lib.MlpDesc), ("urlg_policy_desc", _lib.PolicyDesc),
             ("urlg_storage", _lib.Storage), ("urlg_ppo_hparams", _lib.PpoHparams), ("urlg_peer_exchange", _lib.PeerExchange)]
    body = 'printf("' + " ".join("%zu" for _ in names) + '\\n",' + ",".join(f"sizeof({n})" for n, _ in names) + ");"
    sizes = list(map(int, _probe(body)))
    for (n, cls), sz in zip(names, sizes):
        assert ctypes.sizeof(cls) == sz, n


def test_stat_slots_match_header_and_lead_with_the_minibatch_sums():
    """_lib.STAT_* equal the header's enum; the four per-minibatch sums are slots 0..3 so that a data-parallel caller
    all-reduces [flat gradients | sums] as one contiguous range."""
    from unitree_rl_gym_b200 import _lib
    names = ["KL_SUM", "SURR_SUM", "VLOSS_SUM", "ENT_SUM", "LR", "STEP", "GNORM_SQ", "CLIP_COEF", "ACC_VLOSS", "ACC_SURR",
             "ACC_N", "KL_MEAN", "BC1", "BC2", "COUNT"]
    body = 'printf("' + " ".join("%d" for _ in names) + '\\n",' + ",".join("URLG_STAT_" + n for n in names) + ");"
    vals = list(map(int, _probe(body)))
    for n, v in zip(names, vals):
        assert getattr(_lib, "STAT_" + n) == v, n
    assert [_lib.STAT_KL_SUM, _lib.STAT_SURR_SUM, _lib.STAT_VLOSS_SUM, _lib.STAT_ENT_SUM] == [0, 1, 2, 3]


def test_peer_exchange_constants_and_layout_match_header():
    """The Python mirror of the peer-exchange constants and the buffer layout rsl_rl/peer_exchange.py allocates
    ([2 halves | result area | publish flags + slice flags]) agree with include/urlg_b200.h; header compiles as plain C."""
    from unitree_rl_gym_b200 import _lib
    from unitree_rl_gym_b200.rsl_rl.peer_exchange import _Exchange, _round4
    w, ver = map(int, _probe('printf("%d %d\\n", URLG_PEER_MAX_WORLD, URLG_ABI_VERSION);'))
    assert w == _lib.PEER_MAX_WORLD and ver == _lib.ABI_VERSION
    for n in (1, 4, 380_000, 380_005):
        c = _round4(n)
        assert c >= n and c % 4 == 0 and c - n < 4
        assert _Exchange.floats_needed(c) == 3 * c + 2 * w
    offs = dict(_lib.PeerExchange._fields_)
    assert list(offs) == ["world", "rank", "bufs", "flags", "count", "epoch", "ticket", "mode", "reserved"]
