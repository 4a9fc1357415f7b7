"""Import shim for the unmodified reference package (TEST INFRASTRUCTURE ONLY).

Puts oracle/refharness/stubs (gym + matplotlib stand-ins) and /root/reference on
sys.path and applies the `numpy.int = int` shim the reference's RMCSA env needs
(SURVEY.md F11).  Nothing in the product package may import this module; only
oracle/refharness generator scripts (run in the build container, where
/root/reference exists) use it.
"""
import os
import pickle
import sys

import numpy as np

REFERENCE_ROOT = os.environ.get("ORLG_REFERENCE_ROOT", "/root/reference")
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "optical_rl_gym"))


def install():
    if not available():
        raise RuntimeError(f"reference checkout not found at {REFERENCE_ROOT}")
    if not hasattr(np, "int"):
        np.int = int  # rmcsa_env.py:139,472 use the removed alias
    for p in (_STUBS, REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    import gym  # noqa: F401  (the stub)
    import optical_rl_gym  # noqa: F401  (registers env ids)

    return gym


def load_topology(name: str):
    """Unpickle one of the reference's topology files (examples/topologies/*.h5)."""
    install()
    path = os.path.join(REFERENCE_ROOT, "examples", "topologies", name)
    with open(path, "rb") as f:
        return pickle.load(f)
