"""Torch side of the neural-design loop: `VoxelFEMFunction`, the autograd bridge of the reference (fem.py:109-134),
on the device-resident solver.

The reference moves the MLP's densities to numpy (`densities.numpy().astype(np.float64)`), runs the CPU solve, and
hands the per-voxel gradient back through numpy as float32.  Here a CUDA tensor is used in place: the FP64 view of the
densities is passed to `setVars` as a device pointer (`__cuda_array_interface__`, the same memory DLPack exports), the
compliance-gradient kernel writes straight into a torch-owned FP64 buffer, and only the scalar objective crosses to the
host.  Conventions kept exactly: the loss is 2*J (fem.py:122), the gradient handed to autograd is dJ/drho
WITHOUT the factor 2 (fem.py:125), both cast to float32.  CPU tensors take the reference's numpy path unchanged.
"""
import numpy as np
import torch
from torch import autograd


class VoxelFEMFunction(autograd.Function):
    @staticmethod
    def forward(ctx, densities, top):  # type: ignore
        """densities: predicted physical densities (any float dtype, CPU or CUDA); top: pyVoxelFEM.TopologyOptimizationProblem"""
        x = densities.detach()
        if x.is_cuda:
            # stream contract: the library works on the legacy default stream, which orders against torch only
            # while torch's current stream is its default stream
            if torch.cuda.current_stream(x.device) != torch.cuda.default_stream(x.device):
                raise RuntimeError("VoxelFEMFunction: call it on torch's default CUDA stream (the solver enqueues on stream 0)")
            x64 = x.reshape(-1).to(torch.float64).contiguous()      # the FP32 -> FP64 cast is the only pass over the data
            top.setVars(x64)                                        # device pointer, no host copy
            objective = 2.0 * top.evaluateObjective()
            grad64 = torch.empty_like(x64)
            top.evaluateObjectiveGradient(out=grad64)               # written in place by the gradient chain on the GPU
            output_gradient = grad64.to(torch.float32).reshape(densities.shape)
        else:
            top.setVars(x.numpy().astype(np.float64).ravel())
            objective = 2.0 * top.evaluateObjective()
            output_gradient = torch.from_numpy(top.evaluateObjectiveGradient().astype(np.float32)).reshape(densities.shape)
        ctx.save_for_backward(output_gradient)
        return torch.tensor(objective, device=densities.device).float()

    @staticmethod
    def backward(ctx, grad_output):
        output_gradient = ctx.saved_tensors[0]
        return (output_gradient * grad_output), None
