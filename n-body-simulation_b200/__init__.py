"""n-body-simulation_b200 — B200-native all-pairs gravity + Velocity-Verlet/Euler step.

The product is the C-ABI shared library `libnbody_b200.so` next to this file (sources in csrc/, interface in
include/nbody_b200.h) and the C++17 host drop-in under host/. This Python module only locates and loads the library;
there is no Python or CPU implementation of the path."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnbody_b200.so")


def load() -> ctypes.CDLL:
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} not built: run `make lib` (requires nvcc; there is no CPU fallback)")
    return ctypes.CDLL(LIB_PATH)
