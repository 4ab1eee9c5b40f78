"""totalsegmentator2d_b200: the TS2D inference hot path on hand-written sm_100a CUDA kernels.

``from totalsegmentator2d_b200 import TS2D`` (or ``from ts2d import TS2D`` through the alias package)
mirrors ``from ts2d import TS2D`` of the reference (reference ts2d/__init__.py:1).
"""
__version__ = "0.1.0"


def __getattr__(name):
    if name == "TS2D":
        from .tool import TS2D
        return TS2D
    raise AttributeError(name)
