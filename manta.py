"""`from manta import *` -- the module name the reference's scenes import (source/pwrapper/registry.cpp registers the
embedded module as `manta`).  It is the host-side mirror of that API for the device-resident grid path:
everything comes from mantaflow_b200.manta (ctypes -> libmantacuda.so -> CUDA kernels)."""
from mantaflow_b200.manta import *            # noqa: F401,F403
from mantaflow_b200 import manta as _m

__all__ = [n for n in dir(_m) if not n.startswith("_")]
