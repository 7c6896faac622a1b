"""Import shim for `cyipopt` (helpers/ipopt_helpers.py:19 imports it even on the OC path, fem.py:47).

If a real cyipopt is installed anywhere else on sys.path it is loaded and re-exported unchanged.  Otherwise `problem`
records its arguments and options so `initializeIpoptProblem(top)` works, and `solve` fails with a clear message: Ipopt
itself (the LBFGS branch of fem.py:89-95) is outside the VoxelFEM hot path."""
import importlib.util
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
_real = None
for _p in sys.path:
    if os.path.abspath(_p or ".") == _here:
        continue
    _spec = importlib.util.spec_from_file_location("cyipopt", os.path.join(_p, "cyipopt", "__init__.py")) \
        if os.path.isfile(os.path.join(_p or ".", "cyipopt", "__init__.py")) else None
    if _spec is not None:
        _real = importlib.util.module_from_spec(_spec)
        sys.modules["cyipopt"] = _real
        _spec.loader.exec_module(_real)
        break

if _real is not None:
    globals().update({k: v for k, v in vars(_real).items() if not k.startswith("__")})
else:
    class problem:  # noqa: N801  (the reference calls ipopt.problem(...))
        def __init__(self, n, m, problem_obj=None, lb=None, ub=None, cl=None, cu=None):
            self.n, self.m, self.problem_obj, self.lb, self.ub, self.cl, self.cu = n, m, problem_obj, lb, ub, cl, cu
            self.options = {}

        def addOption(self, key, value):  # noqa: N802
            self.options[key] = value

        add_option = addOption

        def solve(self, x0, lagrange=None, zl=None, zu=None):
            raise RuntimeError("cyipopt (Ipopt) is not installed: the LBFGS branch of fem.py needs the real package; "
                               "the OC optimizer and the neural loop do not use it")

        def close(self):
            pass

    Problem = problem
