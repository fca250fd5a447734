"""Stub: legged_gym/utils/terrain.py:5 imports this name; terrain is dead code in this fork."""
