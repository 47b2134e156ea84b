"""b200-native pull-SpMV / PageRank engine behind cuGraph's API.

All compute lives in libcugraph_b200.so (hand-written sm_100a CUDA, C ABI in
include/b200_cugraph.h and include/cugraph_c/*.h).  Importing the package does not load the
library; the first call does, and raises if it has not been built."""
from ._lib import LIB_PATH, CugraphError  # noqa: F401

__all__ = ["LIB_PATH", "CugraphError"]
