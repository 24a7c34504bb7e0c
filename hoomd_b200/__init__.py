"""hoomd_b200 -- B200-native (sm_100a) implementation of the HOOMD-blue MD inner loop behind the reference's
plugin / job-script API.  See DESIGN.md.  The compute path is the CUDA C-ABI library hoomd_b200/libhoomd_b200.so;
there is no CPU fallback."""
from .engine import System  # noqa: F401
from . import workloads  # noqa: F401

__version__ = "0.1.0"
from . import domain  # noqa: F401,E402
