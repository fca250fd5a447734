"""Nested-class config namespaces (mirrors legged_gym/envs/base/base_config.py:3-25:
instantiating a config replaces every member *class* by an *instance*, recursively)."""
import inspect


class BaseConfig:
    def __init__(self) -> None:
        _instantiate_members(self)

    @staticmethod
    def init_member_classes(obj):
        _instantiate_members(obj)


def _instantiate_members(obj):
    for name in dir(obj):
        if name == "__class__":
            continue
        member = getattr(obj, name)
        if inspect.isclass(member):
            inst = member()
            setattr(obj, name, inst)
            _instantiate_members(inst)
