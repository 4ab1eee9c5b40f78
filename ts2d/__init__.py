"""Alias package: ``from ts2d import TS2D`` keeps working after switching to the B200-native implementation."""


def __getattr__(name):
    if name == "TS2D":
        from totalsegmentator2d_b200.tool import TS2D
        return TS2D
    raise AttributeError(name)
