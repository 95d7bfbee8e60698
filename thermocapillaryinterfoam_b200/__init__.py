"""B200-native interface hot path of thermocapillaryInterFoam (capillary-stress surface force +
MULES phase-fraction advection).  Compute lives in libtcif.so (hand-written CUDA, sm_100a) behind
the C ABI of include/tcif.h; this package holds the thin host-side binding and the case harness."""
from . import abi  # noqa: F401
