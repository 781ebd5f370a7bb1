"""cornerstone_octree_b200 — B200-native (sm_100a) octree/domain hot path of cornerstone-octree.

The product is libcstone_b200.so (hand-written CUDA behind the C ABI in include/cstone_b200.h); this package is the
thin host-side mirror of the reference's free functions used by tests and benchmarks.  Importing it fails loudly if
the shared library has not been built; there is no CPU fallback."""
from ._lib import HILBERT, MORTON, OPEN, PERIODIC, FIXED, CstoneError, launch_count, LIB_PATH  # noqa: F401
from . import ops  # noqa: F401
