"""Drop-in module with the names of the reference's pybind11 module `pyVoxelFEM`
(VoxelFEM/python_bindings/VoxelFEM.cc:218-289)."""
import types as _types

from ._lib import lib as _lib
from .voxelfem import TensorProductSimulator
from .voxelfem_mg import ComplianceObjective, MultigridComplianceObjective, MultigridSolver
from .voxelfem_top import (LangelaarFilter, OCOptimizer, ProjectionFilter, PythonFilter, SmoothingFilter,
                           TopologyOptimizationProblem, TotalVolumeConstraint, applyFilter)
from .benchmark import (benchmark_report, benchmark_reset, benchmark_start_timer, benchmark_start_timer_section,
                        benchmark_stop_timer, benchmark_stop_timer_section)

_lib()  # fail at import time when the CUDA library is missing: there is no CPU fallback

# name-mangled classes of the reference live in pyVoxelFEM.detail (VoxelFEM.cc:20-33)
detail = _types.ModuleType("pyVoxelFEM.detail")
for _suffix in ("1_1", "1_1_1"):
    setattr(detail, "TensorProductSimulator" + _suffix, TensorProductSimulator)
    setattr(detail, "MultigridSolver" + _suffix, MultigridSolver)
    setattr(detail, "TopologyOptimizationProblem" + _suffix, TopologyOptimizationProblem)
    setattr(detail, "MultigridComplianceObjective" + _suffix, MultigridComplianceObjective)
    setattr(detail, "OCOptimizer" + _suffix, OCOptimizer)
