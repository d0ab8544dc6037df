"""B200-native scan-to-map registration hot path (libpointmatcher ICP chain + open3d_slam voxel-map crop/insert).

The package holds only what the path needs: csrc/ (CUDA kernels + the C ABI, built into libb200reg.so) and the
Python-side binding of that ABI (capi).  There is no CPU fallback.
"""
__version__ = "0.1.0"
