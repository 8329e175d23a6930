"""svtrek_b200 -- B200-native (sm_100a) drop-in for SVTrek's `audt` breakpoint-refinement hot path.

csrc/      hand-written CUDA kernels + the C ABI of include/svtk.h (libsvtk_b200.so, built in-tree by build.py)
api.py     ctypes mirror of the C ABI, named after the reference's own functions
simreads.py  seeded synthetic long-read / VCF generator for BASELINE.json's configs (inputs only)
"""
__version__ = "0.1.0"
