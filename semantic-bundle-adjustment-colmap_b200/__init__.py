"""B200-native bundle-adjustment hot path behind COLMAP's BundleAdjuster API.

The product is `libsba_b200.so` (CUDA, sm_100a; C ABI in include/sba_b200.h).  This package
is the host-side mirror used by tests and bench.py: ctypes binding (`_capi`), flattened problem
+ handle wrapper (`problem`), the BundleAdjuster / BundleAdjustmentConfig /
BundleAdjustmentOptions mirror (`bundle_adjustment`), the synthetic scene generator (`scene`) and the on-disk formats
either side of the path (`model_io`: sparse model binary / text, float32 TIFF maps, cylinders.txt).
"""
from . import _capi as capi  # noqa: F401
from .problem import FlatProblem, ResidentMaps, Solver, comm_unique_id  # noqa: F401
from .bundle_adjustment import (  # noqa: F401
    BundleAdjuster, BundleAdjustmentConfig, BundleAdjustmentOptions, Camera, Image, Point2D, Point3D,
    Reconstruction, Track, TrackElement)
from . import scene  # noqa: F401
from . import model_io  # noqa: F401
from . import filters  # noqa: F401
from . import bundle_adjustment  # noqa: F401
from . import semantic_adjusters  # noqa: F401  (SemanticBundleAdjuster / GeometricSemanticBundleAdjuster as solves)
from . import estimators  # noqa: F401  (RefineAbsolutePose through the device solve)
from . import rig_adjuster  # noqa: F401  (RigBundleAdjuster over sba_rig_solve)
