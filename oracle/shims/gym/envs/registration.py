import importlib

_REGISTRY = {}


def register(id, entry_point=None, kwargs=None, **_ignored):
    _REGISTRY[id] = (entry_point, dict(kwargs or {}))


def make(id, **kw):
    entry_point, kwargs = _REGISTRY[id]
    mod, cls = entry_point.split(':')
    kwargs = dict(kwargs, **kw)
    return getattr(importlib.import_module(mod), cls)(**kwargs)
