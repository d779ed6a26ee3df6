"""Drop-in for the reference's ``phase5_loop/losses.py`` import name (``from losses import loss_MPJPE, TriangleLoss,
TriangleLoss_sep``, train_5.py:14, backup_june8.py:14) -- put this directory on ``sys.path`` in place of
``phase5_loop/``'s own module.  ``flip_pose`` (imported from ``utils`` by the reference, train_5.py:19) and the
(de)normalisation / fused pose-loss helpers are re-exported here too.  Implementation: ``pose_losses.py``."""
try:
    from . import pose_losses as _m
except ImportError:                      # imported as top-level ``losses`` through sys.path
    import importlib as _il
    import os as _os
    import sys as _sys
    _here = _os.path.dirname(_os.path.abspath(__file__))
    if _os.path.dirname(_here) not in _sys.path:
        _sys.path.insert(0, _os.path.dirname(_here))
    _m = _il.import_module(_os.path.basename(_here) + ".pose_losses")

loss_MPJPE, TriangleLoss, TriangleLoss_sep = _m.loss_MPJPE, _m.TriangleLoss, _m.TriangleLoss_sep
flip_pose, flip_average, epoch_mpjpe_mm = _m.flip_pose, _m.flip_average, _m.epoch_mpjpe_mm
normalise, denormalise = _m.normalise, _m.denormalise
normalise_minmax, denormalise_minmax = _m.normalise_minmax, _m.denormalise_minmax
PoseLoss, l1_loss, coco2h36m = _m.PoseLoss, _m.l1_loss, _m.coco2h36m
keypoint_stats, process_data = _m.keypoint_stats, _m.process_data
keypoints_to_heatmap_3D = _m.keypoints_to_heatmap_3D          # phase3_direct/my_HybrIK/H36_dataset.py:196-202
H36M_EDGES, H36M_INTRINSICS, LEFT_JOINTS, RIGHT_JOINTS = _m.H36M_EDGES, _m.H36M_INTRINSICS, _m.LEFT_JOINTS, _m.RIGHT_JOINTS

__all__ = ["loss_MPJPE", "TriangleLoss", "TriangleLoss_sep", "flip_pose", "flip_average", "epoch_mpjpe_mm",
           "normalise", "denormalise", "normalise_minmax", "denormalise_minmax", "PoseLoss", "l1_loss", "coco2h36m", "keypoint_stats", "process_data", "keypoints_to_heatmap_3D",
           "H36M_EDGES", "H36M_INTRINSICS", "LEFT_JOINTS", "RIGHT_JOINTS"]
