"""B200-native (sm_100a) implementation of the phase1_lifting ``LinearModel`` 2D->3D lifter hot path of
RHnejad/3D_PoseEstimation.  The directory name starts with a digit: import it with
``importlib.import_module("3d_poseestimation_b200")`` or put this directory on ``sys.path`` and
``from baselineModel import LinearModel`` exactly as the reference's scripts do (train_1.py:15-17).
"""
from .baselineModel import Attention, Linear, LinearModel, MyViT, MyViTBlock, weight_init  # noqa: F401

__all__ = ["LinearModel", "Linear", "weight_init", "MyViT", "MyViTBlock", "Attention"]
