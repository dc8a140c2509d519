"""racing-slam_b200 -- B200-native (sm_100a) ORB descriptor matching for Racing-SLAM.

Only the reference's data-parallel hot path lives here: the CUDA kernels
and C ABI (csrc/, libslamgpu.so), the ctypes binding (binding), the host-side mirror of the
reference's matcher interface (matcher), the multi-GPU sharding of the batched workloads
(sharding) and synthetic workload generators (synthetic).  There is no CPU compute path.

The directory name carries a hyphen; `import racing_slam_b200` works through the loader shim
at the repository root (racing_slam_b200.py).
"""
from . import binding, matcher, sharding, synthetic  # noqa: F401
from .binding import Context, ProjParams, SlamGpuError, pinned_empty  # noqa: F401
from .matcher import MapMatcher, match_features  # noqa: F401
