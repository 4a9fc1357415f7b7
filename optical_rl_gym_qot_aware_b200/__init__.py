"""B200-native batched stepping engine for the QoT-aware Optical RL-Gym environments.

Host side (Python, mirroring the reference's gym surface) over a C-ABI shared library
(``include/orlg.h`` -> ``liborlg.so``, hand-written sm_100a CUDA).  There is no CPU fallback:
constructing an env without the built library or without a CUDA device raises.
"""
from .spec import ENV_IDS, EnvSpec, make_spec  # noqa: F401
from .tables import TopologyTables, tables_from_topology  # noqa: F401

from .topology_builder import build_from_file, build_tables  # noqa: F401

__all__ = ["ENV_IDS", "EnvSpec", "make_spec", "TopologyTables", "tables_from_topology", "build_from_file", "build_tables",
           "make", "register_all"]


def make(env_id, **kwargs):
    from .envs import make as _make

    return _make(env_id, **kwargs)


def register_all():
    """Ids this package serves (optical_rl_gym/__init__.py:3-31)."""
    return sorted(ENV_IDS)
