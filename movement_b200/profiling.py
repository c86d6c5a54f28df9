"""Profiling ranges at the sites the reference instruments with ``PETSc.Log.EventDecorator``
(movement/monge_ampere.py:186,225,312,373,420,452,487,564,609,648,703,729; laplacian_smoothing.py:36,54,79;
spring.py:71,93,117,148,183,246; tangling.py:39).

The reference's decorator opens a PETSc log event named ``<module>.<qualname>``; here the same name opens an NVTX
range (visible in Nsight Systems / ncu ``--nvtx``) and feeds a host-side table of call counts and wall time, the
analogue of PETSc's ``-log_view``:

    import movement_b200.profiling as prof
    prof.enable()            # or MV_PROFILE=1 in the environment
    ...
    print(prof.log_view())

NVTX ranges are pushed whenever CUDA is initialised and ``MV_NVTX`` is not 0; the table costs two
``perf_counter`` calls per event and is off by default.
"""
from __future__ import annotations

import functools
import os
import time

_enabled = os.environ.get("MV_PROFILE", "0") not in ("0", "")
_nvtx = os.environ.get("MV_NVTX", "1") not in ("0", "")
_table = {}   # name -> [calls, seconds]
_stack = []


def enable(on=True):
    global _enabled
    _enabled = bool(on)


def reset():
    _table.clear()


def _nvtx_mod():
    if not _nvtx:
        return None
    try:
        import torch

        if torch.cuda.is_available() and torch.cuda.is_initialized():
            return torch.cuda.nvtx
    except Exception:   # pragma: no cover - torch without CUDA support
        pass
    return None


class event:
    """Decorator / context manager: ``@event("movement.monge_ampere.MongeAmpereMover_Relaxation.move")``."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        nv = _nvtx_mod()
        if nv is not None:
            nv.range_push(self.name)
        _stack.append((nv, time.perf_counter() if _enabled else 0.0))
        return self

    def __exit__(self, *exc):
        nv, t0 = _stack.pop()
        if nv is not None:
            nv.range_pop()
        if _enabled:
            rec = _table.setdefault(self.name, [0, 0.0])
            rec[0] += 1
            rec[1] += time.perf_counter() - t0
        return False

    def __call__(self, fn):
        name = self.name

        @functools.wraps(fn)
        def wrapper(*args, **kwargs):
            with event(name):
                return fn(*args, **kwargs)

        return wrapper


def events():
    """{name: (calls, seconds)} recorded since the last reset (host wall time, inclusive of nested events)."""
    return {k: (v[0], v[1]) for k, v in _table.items()}


def log_view():
    rows = sorted(_table.items(), key=lambda kv: -kv[1][1])
    w = max([len(k) for k, _ in rows] + [5])
    out = [f"{'Event':<{w}}  {'Count':>7}  {'Time (s)':>10}"]
    out += [f"{k:<{w}}  {v[0]:>7d}  {v[1]:>10.4f}" for k, v in rows]
    return "\n".join(out)
