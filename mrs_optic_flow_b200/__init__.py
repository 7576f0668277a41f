"""B200-native (sm_100a) backend for mrs_optic_flow's per-section image-shift estimation.

The compute lives in libmof_b200.so (hand-written CUDA behind the C ABI of include/mof.h);
`api` mirrors the reference's operator interface on top of it.  No CPU fallback exists: using the
package without the built extension or without a CUDA device raises.
"""
from . import synth  # noqa: F401  (numpy only)

__all__ = ["api", "sharding", "synth", "_lib"]
__version__ = "0.1"
