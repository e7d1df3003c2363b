"""B200-native stereo bundle adjuster — drop-in for Victor-YG/stereo_orb_slam's BundleAdjuster.

The compute lives in lib/libsba.so (hand-written sm_100a CUDA behind the C-ABI of include/sba.h).
Importing this package does not load the library; the first Solver / BundleAdjuster does, and it
raises if the library has not been built (no CPU fallback).
"""
from ._capi import Options, Summary, SbaError, default_options, LIB_PATH  # noqa: F401
from .solver import Solver  # noqa: F401
from .bundle_adjuster import BundleAdjuster, pack_window, unpack_window, matrix_to_pose, pose_to_matrix  # noqa: F401
from .device_map import ResidentBundleAdjuster, replay  # noqa: F401
from .pose_graph import optimize_pose_graph, odometry_graph  # noqa: F401
from .ransac import RansacOption, StereoRansac, CalcTransformation, GlibcRand, triangulate_stereo  # noqa: F401
from . import synthetic  # noqa: F401
from . import dump_io  # noqa: F401
