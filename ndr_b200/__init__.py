"""ndr_b200 — B200-native (sm_100a) implementation of the VoxelFEM elasticity solve behind the
reference's pyVoxelFEM interface.  `import ndr_b200.pyVoxelFEM as pyVoxelFEM`, or put
`ndr_b200/shims` on PYTHONPATH to make `import pyVoxelFEM` (and the sibling modules the
reference scripts import) resolve to this package."""
from . import pyVoxelFEM  # noqa: F401
