"""orbslam3lib_b200 — B200 (sm_100a) ORB front-end: ORBextractor + Hamming kNN behind the reference's call signatures.

The product is liborbfe.so (CUDA kernels + C ABI, include/orbfe_c.h); the modules here are the host-side mirror of
the reference interface used by tests and bench.py.  Nothing in this package imports oracle/."""
from . import scenes  # noqa: F401
from ._lib import KP_DTYPE, OrbfeError, build  # noqa: F401
from .extractor import ORBextractor  # noqa: F401
from .matcher import BFMatcher, ORBmatcher  # noqa: F401
from .sharded import ShardedMatcher, shard_range  # noqa: F401
from .wire import LynxHardwareAccelerator  # noqa: F401
