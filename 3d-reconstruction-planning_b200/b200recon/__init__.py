"""b200recon -- B200-native drop-in for the Open3D subset that the reference's stitcher.py uses.

    import b200recon as o3d      # instead of `import open3d as o3d` (stitcher.py:1)

Every operation runs as a hand-written sm_100a CUDA kernel of libb200recon.so, reached through the
C ABI in include/b200recon.h.  There is no CPU fallback: without the built library or without a B200
the first device call raises.
"""
from . import camera, geometry, cloud_ops, io, pipelines, utility, visualization  # noqa: F401
from ._lib import B2RError  # noqa: F401

__version__ = "0.1.0"
