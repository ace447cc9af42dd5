import importlib

registry = {}


class EnvSpec:
    def __init__(self, id, entry_point, kwargs):
        self.id, self.entry_point, self.kwargs = id, entry_point, dict(kwargs or {})


def register(id, entry_point=None, kwargs=None, **_ignored):
    registry[id] = EnvSpec(id, entry_point, kwargs)


def make(id, **overrides):
    spec = registry[id]
    mod, cls = spec.entry_point.split(":")
    kw = dict(spec.kwargs)
    kw.update(overrides)
    env = getattr(importlib.import_module(mod), cls)(**kw)
    env.spec = spec
    return env
