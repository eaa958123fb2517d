"""g1_locomotion_b200 -- B200-native batched SRBD-MPC hot path (drop-in for g1_mpc's MPC.update)."""
from .presets import PRESETS, get_preset  # noqa: F401


def __getattr__(name):
    if name in ("MPC", "BatchedMPC"):
        from . import mpc
        return getattr(mpc, name)
    raise AttributeError(name)
