from ndr_b200.pyVoxelFEM import *  # noqa: F401,F403
from ndr_b200.pyVoxelFEM import detail  # noqa: F401
