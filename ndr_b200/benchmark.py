"""benchmark_* hooks of pyVoxelFEM (VoxelFEM.cc:245-255; MeshFEM GlobalBenchmark.hh): named wall-clock
sections.  Device work is asynchronous, so a section synchronises the device when it stops."""
import time

from ._lib import lib

_sections = {}
_open = {}


def benchmark_reset():
    _sections.clear()
    _open.clear()


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
