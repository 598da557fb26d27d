"""scflow_b200 -- B200-native (sm_100a) implementation of SCFlow's recurrent shape-constrained
flow refinement hot path behind the reference's mmflow-style module surface."""
from .registry import MODELS, install  # noqa: F401
from .modules import ConvGRU, ConvModule, CorrelationPyramid, CorrLookup, MotionEncoder, XHead  # noqa: F401
from .pose_head import MultiClassPoseHead  # noqa: F401
from .decoder import SCFlowDecoder  # noqa: F401
from .pose import (PointSet, cal_3d_2d_corr, get_flow_from_delta_pose_and_points,  # noqa: F401
                   get_pose_from_delta_pose)

__all__ = ["MODELS", "install", "SCFlowDecoder", "CorrelationPyramid", "CorrLookup", "ConvGRU", "MotionEncoder",
           "XHead", "MultiClassPoseHead", "ConvModule", "PointSet", "cal_3d_2d_corr",
           "get_flow_from_delta_pose_and_points", "get_pose_from_delta_pose"]
