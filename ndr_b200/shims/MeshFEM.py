"""Import shim: the reference's pyVoxelFEM imports the MeshFEM python module (VoxelFEM.cc:222)."""
