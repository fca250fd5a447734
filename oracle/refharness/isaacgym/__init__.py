"""ORACLE (test infrastructure).  Stub of NVIDIA Isaac Gym, just enough for the
UNMODIFIED reference `LeggedRobot` / `H1Robot` / `H1_2Robot` / `G1Robot`
(/root/reference/legged_gym/envs) to be imported and stepped on CPU with
synthetic physics tensors.  Isaac Gym itself is a proprietary download
(doc/setup_en.md:59-85) and is not in the container; see SURVEY.md §8(c).
Nothing here ships: only tests/golden/gen_golden.py and tests/ use it, and only
in the build container where /root/reference exists.
"""
from . import gymapi, gymtorch, gymutil, torch_utils, terrain_utils  # noqa: F401
