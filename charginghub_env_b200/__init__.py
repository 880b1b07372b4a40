"""charginghub_env_b200 -- B200-native batched stepping engine for the charging-hub environment.

The hot path (reset()/step(action) of XJTU-Power-AI/charginghub-env's `charging-hub-v6`) runs as
hand-written sm_100a CUDA behind the C ABI of include/chub.h; this package binds that ABI and keeps
the reference's gym surface.  There is no CPU fallback: importing the engine without a built
libchub.so, or creating one without a CUDA device, raises.
"""
from .engine import ChubError, HubEngine, load_tables, make_hub_class  # noqa: F401

__all__ = ["HubEngine", "ChubError", "load_tables", "make_hub_class"]
