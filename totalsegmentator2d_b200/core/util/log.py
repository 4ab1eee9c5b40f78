"""print-based logging (reference ts2d/core/util/log.py:12-36)."""
import sys

_SILENT = False
_WARNED = set()


def log_silent(silent: bool = True):
    global _SILENT
    _SILENT = bool(silent)


def log(*args, **kwargs):
    if not _SILENT:
        print(*args, **kwargs)


def warn(msg, once: bool = False):
    if once:
        if msg in _WARNED:
            return
        _WARNED.add(msg)
    print(msg, file=sys.stderr)
