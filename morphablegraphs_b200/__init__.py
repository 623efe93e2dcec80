"""morphablegraphs_b200 -- B200-native engine for the sample-and-evaluate hot
path of dfki-asr/morphablegraphs, behind the reference's own Python interface.

Host side: Python (this package) over a C-ABI CUDA library (csrc/ ->
libmgb200.so, header include/mgb200.h).  There is no CPU fallback: every
compute call fails loudly if the library or a GPU is missing.
"""
from .skeleton import Skeleton, SkeletonNode  # noqa: F401
from .motion_model import MotionPrimitive, MotionPrimitiveModelWrapper, MotionStateGraphNode, MotionSpline  # noqa: F401
from .constraints import (MotionPrimitiveConstraints, GlobalTransformConstraint, RelativeTransformConstraint,  # noqa: F401
                          Direction2DConstraint, PoseConstraint, TwoHandConstraintSet, FeetConstraint,
                          LookAtConstraint, GlobalTransformCAConstraint, CONSTRAINT_FACTORY)
from .motion_primitive_generator import MotionPrimitiveGenerator  # noqa: F401
from .objective_functions import obj_spatial_error_sum, obj_spatial_error_sum_and_magnitude  # noqa: F401

__version__ = "0.1.0"
