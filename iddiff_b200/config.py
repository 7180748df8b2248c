"""Minimal stand-in for ml_collections.ConfigDict (not installed offline): attribute + item access, .get,
nested dicts wrapped on the fly.  The drop-in entry points accept ANY object with the same attribute
surface (a real ml_collections.ConfigDict, a pickled one, or this class)."""
from __future__ import annotations


class ConfigDict(dict):
    def __init__(self, *a, **k):
        super().__init__()
        for key, v in dict(*a, **k).items():
            self[key] = v

    def __setitem__(self, key, v):
        if isinstance(v, dict) and not isinstance(v, ConfigDict):
            v = ConfigDict(v)
        super().__setitem__(key, v)

    def __getattr__(self, key):
        try:
            return self[key]
        except KeyError as e:
            raise AttributeError(key) from e

    def __setattr__(self, key, v):
        self[key] = v


def lookup(config, dotted: str, default=None):
    """Walks `a.b.c` through attributes/items (the reference probes dotted keys with hasattr,
    dim_reduction.py:144-147)."""
    cur = config
    for part in dotted.split("."):
        try:
            cur = getattr(cur, part)
        except AttributeError:
            try:
                cur = cur[part]
            except (KeyError, TypeError, IndexError):
                return default
    return cur
