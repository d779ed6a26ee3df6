"""CPU restatement of the loss / metric / normalisation math around the lifter.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  PyTorch fp32 on the CPU.

Pinned against the imported reference (``tests/golden/losses.npz``):
* ``mpjpe_per_joint_sum``  <- ``phase5_loop/losses.py:3-7`` (= ``phase1_lifting/train_1.py:19-23``)
* ``mpjpe_mean``           <- ``phase1_lifting/two2three.py:387-390``
* ``epoch_mpjpe_mm``       <- ``phase1_lifting/train_1.py:100-104``
* ``mse_mean`` / ``l1_mean`` <- ``torch.nn.MSELoss/L1Loss`` as used at ``train_1.py:37``, ``losses.py:21``
* ``triangle_loss``        <- ``losses.py:24-53``; ``triangle_loss_sep`` <- ``losses.py:81-110``
  (including the batch-dimension "centring" quirk: ``t[1:] -= t[0]`` subtracts SAMPLE 0)
* ``flip_pose``            <- ``phase3_direct/my_HybrIK/utils.py:372-396``
* ``flip_average``         <- ``phase1_lifting/train_1.py:128-134`` / ``phase5_loop/train_5.py:153-168``
* ``normalise`` / ``denormalise`` <- ``phase1_lifting/main.py:334,337`` / ``main.py:473-479``,
  ``phase5_loop/train_5.py:42-47``

PARITY UNPINNED (no implementation in the reference; BASELINE.json's north_star asks for them,
SURVEY.md §0 D3 / §8c): ``pinhole_reprojection`` + ``reprojection_l1`` and ``bone_lengths`` +
``bone_length_l1``.  Their math follows the nearest reference material: the intrinsics table
``phase3_direct/my_HybrIK/utils.py:131-172``, the pixel/1000 convention of the 2D data and the
``sk_points`` edge list ``utils.py:37``.
"""
from __future__ import annotations

import numpy as np
import torch

LEFT_JOINTS = [4, 5, 6, 11, 12, 13]      # utils.py:382
RIGHT_JOINTS = [1, 2, 3, 14, 15, 16]     # utils.py:383

# utils.py:37 lists [5,6] twice; the 16 unique parent->child edges of the H36M-17 skeleton:
H36M_EDGES = [(0, 1), (1, 2), (2, 3), (0, 4), (4, 5), (5, 6), (0, 7), (7, 8), (8, 9), (9, 10),
              (8, 11), (11, 12), (12, 13), (8, 14), (14, 15), (15, 16)]

# (fx, fy, cx, cy) in pixels, utils.py:131-172 (cameras 54138969, 55011271, 58860488, 60457274)
H36M_INTRINSICS = [
    (1145.0494384765625, 1143.7811279296875, 512.54150390625, 515.4514770507812),
    (1149.6756591796875, 1147.5916748046875, 508.8486328125, 508.0649108886719),
    (1149.1407470703125, 1148.7989501953125, 519.8158569335938, 501.40264892578125),
    (1145.5113525390625, 1144.77392578125, 514.9682006835938, 501.88201904296875),
]


# ----------------------------------------------------------------------------- metrics / losses
def mpjpe_per_joint_sum(pred, target):
    """(B,J,3) x2 -> (J,): per-joint sum over the batch of L2 position errors (a metric, not a mean)."""
    return torch.linalg.vector_norm(pred - target, dim=-1).sum(dim=0)


def mpjpe_mean(pred, target):
    return torch.linalg.vector_norm(pred - target, dim=-1).mean()


def epoch_mpjpe_mm(metric_sum, n_samples, zero_centre=True, first=1):
    """train_1.py:100-104: mean over joints [first:17] of (sum/N), x 17/16 x 1000 when root is zeroed."""
    m = (metric_sum / n_samples)[first:17].mean()
    return m * (17.0 / 16.0) * 1000.0 if zero_centre else m


def mse_mean(pred, target):
    return ((pred - target) ** 2).mean()


def l1_mean(pred, target):
    return (pred - target).abs().mean()


def _centre_batch0(t):
    """The reference's ``t[1:] -= t[0]`` (losses.py:34,37,91-94): subtract sample 0 from samples 1..B-1."""
    out = t.clone()
    out[1:] = out[1:] - out[0]
    return out


def triangle_loss(pred2d, pred3d, lift_gt, lift_pred, gt2d, gt3d, proj_pred=None, project=False):
    """losses.py:24-53.  Returns (total, l2d, l3d, lift, proj)."""
    l2d = l1_mean(pred2d, gt2d)
    l3d = l1_mean(pred3d, gt3d)
    lift = l1_mean(lift_pred, pred3d)
    proj = torch.zeros(())
    if project:
        proj = l1_mean(_centre_batch0(proj_pred), _centre_batch0(pred2d))
        return l2d + l3d + lift + proj, l2d, l3d, lift, proj
    return l2d + l3d + lift, l2d, l3d, lift, proj


def triangle_loss_sep(pred2d, pred3d, lift_gt, lift_pred, gt2d, gt3d, proj_pred=None, proj_gt=None,
                      project=True):
    """losses.py:81-110.  ``project=False`` is a reference bug (UnboundLocalError); the
    oracle then returns proj = 0 and the 4-term sum the ``else`` branch computes."""
    l2d = l1_mean(pred2d, gt2d)
    l3d = l1_mean(pred3d, gt3d)
    gap = l1_mean(lift_pred, lift_gt)
    lift = l1_mean(lift_gt, gt3d)
    if project:
        pp, pg, g2 = _centre_batch0(proj_pred), _centre_batch0(proj_gt), _centre_batch0(gt2d)
        gap_proj = l1_mean(pp, pg)
        proj = l1_mean(pg, g2)
        return l2d + l3d + lift + gap + proj + gap_proj, l2d, l3d, lift, proj
    return l2d + l3d + lift + gap, l2d, l3d, lift, torch.zeros(())


# ----------------------------------------------------------------------------- flip
def flip_pose(data):
    """utils.py:372-396: mirror x (2D: 1-x, 3D: -x) and swap left/right joints."""
    out = data.clone()
    if data.shape[-1] == 2:
        out[..., 0] = 1 - out[..., 0]
    elif data.shape[-1] == 3:
        out[..., 0] = -out[..., 0]
    src = out.clone()
    out[..., LEFT_JOINTS + RIGHT_JOINTS, :] = src[..., RIGHT_JOINTS + LEFT_JOINTS, :]
    return out


def flip_average(lift_fn, x2d, n_joints=17):
    """(flip_pose(model(flip_pose(x))) + model(x)) / 2, train_1.py:128-134."""
    B = x2d.shape[0]
    a = lift_fn(x2d).reshape(B, n_joints, 3)
    b = flip_pose(lift_fn(flip_pose(x2d)).reshape(B, n_joints, 3))
    return (a + b) / 2


# ----------------------------------------------------------------------------- normalisation
def normalise(x, mean, std):
    return (x - mean) / std                  # main.py:334,337


def denormalise(y, mean, std):
    return y * std + mean                    # main.py:473-479, train_5.py:42-47


def normalise_minmax(x):
    return 2 * x - 1                         # main.py:331


def denormalise_minmax(y, vmax, vmin):
    return y * (vmax - vmin) + vmin          # train_5.py:35-40


# ----------------------------------------------------------------------------- PARITY UNPINNED
def pinhole_reprojection(joints_cam, intrinsics):
    """PARITY UNPINNED.  (B,J,3) camera-frame metres, (B,4)=(fx,fy,cx,cy) px -> (B,J,2) in pixel/1000."""
    fx, fy, cx, cy = [intrinsics[:, i].reshape(-1, 1) for i in range(4)]
    X, Y, Z = joints_cam[..., 0], joints_cam[..., 1], joints_cam[..., 2]
    u = (fx * X / Z + cx) / 1000.0
    v = (fy * Y / Z + cy) / 1000.0
    return torch.stack([u, v], dim=-1)


def reprojection_l1(joints_cam, intrinsics, target2d):
    """PARITY UNPINNED.  mean |project(joints) - target2d|."""
    return l1_mean(pinhole_reprojection(joints_cam, intrinsics), target2d)


def bone_lengths(joints):
    """PARITY UNPINNED.  (B,J,3) -> (B,16) lengths of the H36M-17 edges."""
    p = torch.tensor([e[0] for e in H36M_EDGES])
    c = torch.tensor([e[1] for e in H36M_EDGES])
    return torch.linalg.vector_norm(joints[:, c] - joints[:, p], dim=-1)


def bone_length_l1(pred, target):
    """PARITY UNPINNED.  mean |bone(pred) - bone(target)|."""
    return l1_mean(bone_lengths(pred), bone_lengths(target))


def coco2h36m(x):
    """COCO-17 -> H36M-17 joints, ``phase1_lifting/video2keypoints.py:14-57`` (x: (..., 17, d) array or tensor)."""
    a = [11, 12, 14, 16, 11, 13, 15, None, 5, 0, 1, 5, 7, 9, 6, 8, 10]
    b = [12, 12, 14, 16, 11, 13, 15, None, 6, 0, 2, 5, 7, 9, 6, 8, 10]
    x = np.asarray(x) if not torch.is_tensor(x) else x.numpy()       # input dtype kept: the sums round as in the reference
    y = np.zeros(x.shape)
    for j in range(17):
        if j != 7:
            y[..., j, :] = x[..., a[j], :] if a[j] == b[j] else (x[..., a[j], :] + x[..., b[j], :]) * 0.5
    y[..., 7, :] = (y[..., 0, :] + y[..., 8, :]) * 0.5
    return y


def ingest_keypoints(coco, mean=None, std=None, minmax=False, scale=1e-3):
    """coco2h36m, pixel / 1000, then ``(x - mean) / std`` (main.py:334) or ``2x - 1`` (main.py:331)."""
    y = coco2h36m(np.asarray(coco)[..., :2]) * scale
    if mean is not None:
        return (y - np.asarray(mean)) / np.asarray(std)
    return 2 * y - 1 if minmax else y
