from ndr_b200.benchmark import *  # noqa: F401,F403
from ndr_b200.benchmark import benchmark_report as report, benchmark_reset as reset, to_dict  # noqa: F401
