"""B200-native batched-forward engine behind the ray-dynamic-batching scheduler interface.

Import as ``rdb_b200`` (see ``rdb_b200.py`` at the repo root: the directory name carries a hyphen).
"""
from . import lib  # noqa: F401

__all__ = ["lib"]
