"""B200-native batched Minigrid engine (step + partial observation + Bellman sweeps on sm_100a)."""
from .constants import Actions  # noqa: F401

__version__ = "0.1.0"
