"""ndr_b200 — B200-native (sm_100a) implementation of the VoxelFEM elasticity solve behind the
reference's pyVoxelFEM interface.  `import ndr_b200.pyVoxelFEM as pyVoxelFEM`, or put
`ndr_b200/shims` on PYTHONPATH to make `import pyVoxelFEM` (and the sibling modules the
reference scripts import) resolve to this package.

The submodules are imported on demand so that `ndr_b200.build` (which creates libvfem_b200.so) can be
imported before the library exists; `ndr_b200.pyVoxelFEM` itself fails at import when the library is missing."""
import importlib as _importlib


def __getattr__(name):
    if name in ("pyVoxelFEM", "build", "slab", "benchmark", "voxelfem", "voxelfem_mg", "voxelfem_top", "_lib"):
        return _importlib.import_module("." + name, __name__)
    raise AttributeError(name)
