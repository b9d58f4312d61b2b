"""walkai-nos_b200 — B200-native GPU-partitioning planner hot path (libnosplan.so + thin bindings).

The product is the C-ABI shared library built from csrc/ (see include/nosplan.h).  This package
holds only what the path needs: the CUDA sources, a ctypes binding used by the tests and the bench,
the synthetic-cluster generator of SURVEY §8(d), and the host-side mirror of the reference's
planner interface.
"""
from . import dist, host, nosplan, synth  # noqa: F401
from .nosplan import Planner, NosplanError, build_library, library_path  # noqa: F401
