"""Shim of MeshFEM's `parallelism` module (src/python_bindings/parallelism.cc): TBB thread counts have
no meaning on the device path, so the setters are accepted and ignored."""


def set_max_num_tbb_threads(n):
    pass


def set_gradient_assembly_num_threads(n):
    pass


def set_hessian_assembly_num_threads(n):
    pass
