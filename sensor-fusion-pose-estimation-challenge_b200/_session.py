"""Process-wide projector cache for the drop-in (single-sweep, host in / host out) entries.

The reference's functions are free functions with import-time constants; the drop-ins
keep that shape by holding one LidarProjector per parameter set on the default device
(env SFPE_DEVICE, default 0)."""
import os
import threading

from .projector import LidarProjector

_lock = threading.Lock()
_cache = {}


def default_device():
    return int(os.environ.get("SFPE_DEVICE", "0"))


def projector(**kw):
    key = (default_device(),) + tuple(sorted((k, tuple(v) if isinstance(v, (list, tuple)) else v)
                                             for k, v in kw.items()))
    with _lock:
        pr = _cache.get(key)
        if pr is None:
            pr = LidarProjector(device=default_device(), **kw)
            _cache[key] = pr
        return pr


def clear():
    with _lock:
        for pr in _cache.values():
            pr.close()
        _cache.clear()
