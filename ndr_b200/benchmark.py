"""benchmark_* hooks of pyVoxelFEM (VoxelFEM.cc:245-255; MeshFEM GlobalBenchmark.hh): named wall-clock
sections.  Device work is asynchronous, so a section synchronises the device when it stops.

The reference times its own calls under fixed names ("applyK 128x64x64" TPS.hh:907, "MG Solver" MG.hh:449, "CG Iterations"
MG.hh:693, "setVars" / "evaluateObjective..." Problem.hh:43-132, "OC step" OptimalityCriterion.hh:42).  The same names are
recorded here around the corresponding library calls (`timed`), but only after `benchmark_reset()` has switched the
sections on: a section costs a device synchronisation, which an unprofiled run must not pay.  The sections the reference
opens INSIDE a solve ("V Cycle ...", "smoothingMulticoloredGS ...", "Restriction") have no host-side boundary here — one
preconditioner pass is a single CUDA-graph launch — and are reported per level by tools/solve_profile.py instead."""
import functools
import time

from ._lib import lib

_sections = {}
_open = {}
_enabled = False


def benchmark_reset():
    global _enabled
    _sections.clear()
    _open.clear()
    _enabled = True


def timed(name):
    """decorator for the bound methods: records a section under the reference's name (a string, or a function of self)"""
    def deco(fn):
        @functools.wraps(fn)
        def wrapper(self, *a, **k):
            if not _enabled:
                return fn(self, *a, **k)
            n = name(self) if callable(name) else name
            benchmark_start_timer_section(n)
            try:
                return fn(self, *a, **k)
            finally:
                benchmark_stop_timer_section(n)
        return wrapper
    return deco


def benchmark_start_timer_section(name):
    _open[name] = time.perf_counter()


def benchmark_stop_timer_section(name):
    lib().vfem_device_synchronize()
    t0 = _open.pop(name, None)
    if t0 is not None:
        tot, cnt = _sections.get(name, (0.0, 0))
        _sections[name] = (tot + time.perf_counter() - t0, cnt + 1)


benchmark_start_timer = benchmark_start_timer_section
benchmark_stop_timer = benchmark_stop_timer_section


def benchmark_report(include_messages=False):
    for name, (tot, cnt) in _sections.items():
        print("%-40s %10.6f s  (%d calls)" % (name, tot, cnt))


def to_dict():
    return {k: {"time": v[0], "calls": v[1]} for k, v in _sections.items()}
