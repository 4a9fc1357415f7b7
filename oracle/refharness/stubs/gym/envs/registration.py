"""Minimal gym registry stand-in (test infrastructure only)."""
import importlib

registry = {}


def register(id, entry_point=None, **kwargs):
    registry[id] = (entry_point, kwargs)


def make(id, **kwargs):
    entry_point, defaults = registry[id]
    mod_name, cls_name = entry_point.split(":")
    cls = getattr(importlib.import_module(mod_name), cls_name)
    args = dict(defaults.get("kwargs", {}))
    args.update(kwargs)
    return cls(**args)
