"""mantaflow_b200 -- B200 (sm_100a) implementation of Mantaflow's Eulerian grid hot path.

  mantaflow_b200.manta   : drop-in mirror of the reference's Python scene API for the path
  mantaflow_b200._lib    : ctypes binding of libmantacuda.so (include/mantacuda.h)
  mantaflow_b200/csrc    : CUDA kernels + C-ABI
"""
__version__ = "0.1.0"
