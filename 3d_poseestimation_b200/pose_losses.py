"""Drop-in for ``phase5_loop/losses.py`` (``loss_MPJPE``, ``TriangleLoss``, ``TriangleLoss_sep``) plus the keypoint
math the training scripts wrap around the lifter -- ``flip_pose`` (``phase3_direct/my_HybrIK/utils.py:372-396``),
flip averaging (``train_1.py:128-134``, ``train_5.py:153-168``), mean/std (de)normalisation (``main.py:334,337,
473-479``; ``train_5.py:32-49``) -- and the fused MPJPE + reprojection + bone-length loss BASELINE.json's north_star
asks for.  Every function runs hand-written sm_100a kernels behind the C ABI (``include/lifter_b200.h``); CUDA
float32 tensors only, no CPU or PyTorch fallback.

Differences from the reference, all deliberate:
* the reference's ``TriangleLoss*`` centre ``proj[1:] -= proj[0]`` IN PLACE on the caller's tensors (and along the
  batch dimension); the arithmetic is reproduced bit for bit, the caller's tensors are left untouched;
* the per-call ``.cpu().item()`` of every term (four host syncs per step, losses.py:48-51) is replaced by detached
  device scalars in the same lists; ``report_losses`` reads them once;
* ``TriangleLoss_sep(Project=False)`` raises ``UnboundLocalError`` in the reference; here ``loss_proj`` is 0.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _cabi

LEFT_JOINTS = [4, 5, 6, 11, 12, 13]       # utils.py:382
RIGHT_JOINTS = [1, 2, 3, 14, 15, 16]      # utils.py:383
# utils.py:37 (sk_points, [5,6] listed twice): the 16 parent -> child edges of the H36M-17 skeleton
H36M_EDGES = [(0, 1), (1, 2), (2, 3), (0, 4), (4, 5), (5, 6), (0, 7), (7, 8), (8, 9), (9, 10),
              (8, 11), (11, 12), (12, 13), (8, 14), (14, 15), (15, 16)]
# utils.py:131-172: (fx, fy, cx, cy) in pixels of cameras 54138969, 55011271, 58860488, 60457274
H36M_INTRINSICS = [
    (1145.0494384765625, 1143.7811279296875, 512.54150390625, 515.4514770507812),
    (1149.6756591796875, 1147.5916748046875, 508.8486328125, 508.0649108886719),
    (1149.1407470703125, 1148.7989501953125, 519.8158569335938, 501.40264892578125),
    (1145.5113525390625, 1144.77392578125, 514.9682006835938, 501.88201904296875),
]


def _dev(t, what):
    if not torch.is_tensor(t):
        raise TypeError(f"{what}: expected a tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{what}: the sm_100a build has no CPU path; move the tensor to a B200 with .to('cuda')")
    if t.dtype != torch.float32:
        t = t.float()                      # the reference calls .float() on everything it feeds the losses
    return t.contiguous()


def _vec(v, n, dev, what):
    t = torch.as_tensor(np.asarray(v, dtype=np.float32) if not torch.is_tensor(v) else v, dtype=torch.float32,
                        device=dev).reshape(-1).contiguous()
    if t.numel() != n:
        raise ValueError(f"{what}: statistics vector has {t.numel()} entries, expected {n}")
    return t


# ------------------------------------------------------------------------------------------------ flip
class _FlipPose(torch.autograd.Function):
    @staticmethod
    def forward(ctx, data):
        x = _dev(data, "flip_pose")
        J, D = x.shape[-2], x.shape[-1]
        out = torch.empty_like(x)
        _cabi.check(_cabi.lib().lifter_flip_pose(_cabi.ptr(x), _cabi.ptr(out), x.numel() // (J * D), J, D, 0,
                                                 _cabi.stream_ptr()))
        return out

    @staticmethod
    def backward(ctx, g):
        g = _dev(g, "flip_pose")
        J, D = g.shape[-2], g.shape[-1]
        out = torch.empty_like(g)
        _cabi.check(_cabi.lib().lifter_flip_pose(_cabi.ptr(g), _cabi.ptr(out), g.numel() // (J * D), J, D, 1,
                                                 _cabi.stream_ptr()))
        return out


def flip_pose(data):
    """Horizontal flip of ``[..., 17, D]`` keypoints: ``x -> 1 - x`` (D == 2) or ``-x`` (D == 3), left/right joints
    swapped (utils.py:372-396).  Tensors go through the CUDA kernel (differentiable); the reference's 2-D numpy branch
    (``H36_dataset.__getitem__``) stays numpy."""
    if isinstance(data, np.ndarray):
        out = data.copy()
        if data.shape[-1] == 2:
            out[..., 0] = 1 - out[..., 0]
        elif data.shape[-1] == 3:
            out[..., 0] *= -1
        out[..., LEFT_JOINTS + RIGHT_JOINTS, :] = out[..., RIGHT_JOINTS + LEFT_JOINTS, :]
        return out
    if data.dim() < 2 or data.shape[-2] != 17 or data.shape[-1] not in (2, 3):
        raise ValueError(f"flip_pose expects [..., 17, 2|3] keypoints, got {tuple(data.shape)}")
    return _FlipPose.apply(data)


def flip_average(model, x2d, n_joints=17):
    """Flip test-time augmentation ``(flip_pose(model(flip_pose(x))) + model(x)) / 2`` (train_1.py:128-134)."""
    B = x2d.shape[0]
    a = model(x2d).reshape(B, n_joints, 3)
    b = flip_pose(model(flip_pose(x2d.reshape(B, n_joints, -1))).reshape(B, n_joints, 3))
    return (a + b) / 2


# ------------------------------------------------------------------------------------------------ keypoint ingest
_COCO_A = [11, 12, 14, 16, 11, 13, 15, -1, 5, 0, 1, 5, 7, 9, 6, 8, 10]
_COCO_B = [12, 12, 14, 16, 11, 13, 15, -1, 6, 0, 2, 5, 7, 9, 6, 8, 10]


def coco2h36m(x, scale=1.0, mean=None, std=None, minmax=False):
    """COCO-17 -> H36M-17 joint order (``phase1_lifting/video2keypoints.py:14-57``, ``phase2_opp_mb/run.py``).

    A numpy ``(17, d)`` array goes through the reference's dataset-side arithmetic in numpy, as there.  A CUDA tensor
    ``[..., 17, C]`` (C = 2, or 3 with a confidence column as in the AlphaPose JSON) goes through ONE kernel that also
    applies what the loader does next: ``scale`` (the ``/1000`` of pixel coordinates) and then either the mean/std
    standardisation (``main.py:334``) or ``2x - 1`` (``main.py:331``); the result ``[..., 17, 2]`` is the lifter input.
    """
    if isinstance(x, np.ndarray):
        y = np.zeros(x.shape)
        for j in range(17):
            if j != 7:
                y[j, :] = x[_COCO_A[j], :] if _COCO_A[j] == _COCO_B[j] else (x[_COCO_A[j], :] + x[_COCO_B[j], :]) * 0.5
        y[7, :] = (y[0, :] + y[8, :]) * 0.5
        return y
    x = _dev(x, "coco2h36m")
    if x.dim() < 2 or x.shape[-2] != 17 or x.shape[-1] not in (2, 3, 4):
        raise ValueError(f"coco2h36m expects [..., 17, 2|3] keypoints, got {tuple(x.shape)}")
    if (mean is None) != (std is None) or (minmax and mean is not None):
        raise ValueError("coco2h36m: pass mean AND std, or minmax=True, or neither")
    poses = x.numel() // (17 * x.shape[-1])
    out = torch.empty(x.shape[:-1] + (2,), dtype=torch.float32, device=x.device)
    m = _vec(mean, 34, x.device, "coco2h36m mean") if mean is not None else None
    sd = _vec(std, 34, x.device, "coco2h36m std") if std is not None else None
    mode = 0 if mean is not None else (2 if minmax else -1)
    _cabi.check(_cabi.lib().lifter_coco2h36m(_cabi.ptr(x), int(x.shape[-1]), _cabi.ptr(out), poses, float(scale),
                                              _cabi.ptr(m), _cabi.ptr(sd), mode, _cabi.stream_ptr()))
    return out


# ------------------------------------------------------------------------------------- dataset-side processing
def keypoint_stats(dataset, z_c=True):
    """Per (joint, coordinate) statistics of a key-point data set, ``phase1_lifting/main.py:275-289``.

    ``dataset``: CUDA fp32 ``[frames, joints, dims]``.  Returns a CUDA fp64 tensor ``[4, joints, dims]``: mean,
    POPULATION standard deviation (divide by the frame count, as the reference does), min and max
    (``phase5_loop/train_5.py:35-40``) of the values after the optional root-centring ``frame[1:] -= frame[0]``.
    One read of the data, fp64 accumulation, fixed summation order; nothing is copied to the host."""
    x = _dev(dataset, "keypoint_stats")
    if x.dim() != 3 or x.shape[0] < 1:
        raise ValueError(f"keypoint_stats expects a non-empty [frames, joints, dims] array, got {tuple(x.shape)}")
    n, J, D = (int(v) for v in x.shape)
    nbytes = _cabi.lib().lifter_keypoint_stats_scratch_bytes(n, J, D)
    if nbytes < 0:
        raise ValueError(f"keypoint_stats: {J} joints x {D} dims unsupported (at most 64 values per frame)")
    scratch = torch.empty(nbytes // 8, dtype=torch.float64, device=x.device)
    out = torch.empty((4, J, D), dtype=torch.float64, device=x.device)
    _cabi.check(_cabi.lib().lifter_keypoint_stats(_cabi.ptr(x), n, J, D, int(bool(z_c)), _cabi.ptr(scratch),
                                                  _cabi.ptr(out), _cabi.stream_ptr()))
    return out


def process_data(dataset, stats=None, is_train=True, standardize=False, z_c=True, minmax=False, num_of_joints=17,
                 save_dir=None):
    """``H36_dataset.process_data`` (``phase1_lifting/main.py:271-355``) on the device: root-centring, the training
    statistics, ``(x - mean) / std`` (``minmax=True`` on 2D data: ``2x - 1``, the reference's ``Normalize`` switch),
    the root row zeroed (or dropped when ``num_of_joints == 16``).

    ``dataset``: CUDA fp32 ``[frames, 17, dims]``.  ``is_train=True`` computes the statistics from this data (the
    reference then writes ``{mean,std}_train_{2,3}d.npy``; ``save_dir`` does the same, as float64 ``(17, dims)``);
    otherwise pass ``stats`` = what a training call returned (or a ``(mean, std)`` pair of arrays).  Returns
    ``(processed [frames, 17 | 16, dims] fp32 CUDA, stats [4, 17, dims] fp64 CUDA)``."""
    x = _dev(dataset, "process_data")
    if x.dim() != 3:
        raise ValueError(f"process_data expects [frames, joints, dims], got {tuple(x.shape)}")
    n, J, D = (int(v) for v in x.shape)
    if is_train:
        stats = keypoint_stats(x, z_c)
        if save_dir is not None:
            host = stats[:2].cpu().numpy()
            np.save(os.path.join(save_dir, f"mean_train_{D}d.npy"), host[0])
            np.save(os.path.join(save_dir, f"std_train_{D}d.npy"), host[1])
    elif stats is not None and not torch.is_tensor(stats):
        mean, std = (torch.as_tensor(np.asarray(v), dtype=torch.float64).reshape(J, D) for v in stats)
        stats = torch.stack([mean, std, mean, mean]).to(x.device)
    mode = -1
    if standardize:
        mode = 2 if (minmax and D == 2) else 0
        if mode == 0 and stats is None:
            raise ValueError("process_data: standardize=True needs is_train=True or stats")
    if stats is not None:
        stats = stats.to(device=x.device, dtype=torch.float64).contiguous()
        if stats.shape[0] < 2 or stats[0].numel() != J * D:
            raise ValueError(f"process_data: stats must be [>=2, {J}, {D}], got {tuple(stats.shape)}")
    root_mode = 2 if num_of_joints == 16 else (1 if z_c else 0)
    out = torch.empty((n, J - (root_mode == 2), D), dtype=torch.float32, device=x.device)
    _cabi.check(_cabi.lib().lifter_process_keypoints(_cabi.ptr(x), _cabi.ptr(out), n, J, D, int(bool(z_c)),
                                                     _cabi.ptr(stats) if mode == 0 else None, mode, root_mode,
                                                     _cabi.stream_ptr()))
    return out, stats


# ------------------------------------------------------------------------------------------------ (de)normalisation
def keypoints_to_heatmap_3D(keypoints, sigma=0.5, out=None, check=True):
    """Volumetric Gaussian heat-map targets, ``phase3_direct/my_HybrIK/H36_dataset.py:143-202`` (``xyz_to_uvw``,
    ``_keypoint_to_heatmap_3D``, ``keypoints_to_heatmap_3D``): key points ``[..., 3]`` in [-1, 1] (fp32 or fp64, CUDA)
    -> fp32 ``[..., 64, 64, 64]``, one volume per key point; the dataset calls it per frame with ``(17, 3)``.

    ONE streaming kernel writes every volume exactly once (1 MB per key point).  ``check``: mirror the reference's
    asserts on the coordinate range (one host read of a status word); ``check=False`` keeps the call asynchronous.
    A numpy array is uploaded, processed on the GPU and downloaded again (there is no CPU implementation)."""
    if isinstance(keypoints, np.ndarray):
        dev = torch.device("cuda", torch.cuda.current_device())
        return keypoints_to_heatmap_3D(torch.from_numpy(np.ascontiguousarray(keypoints)).to(dev), sigma, None, check).cpu().numpy()
    if not torch.is_tensor(keypoints) or not keypoints.is_cuda:
        raise RuntimeError("keypoints_to_heatmap_3D: the sm_100a build has no CPU path; pass a CUDA tensor or a numpy array")
    if keypoints.shape[-1] != 3:
        raise ValueError(f"keypoints_to_heatmap_3D expects [..., 3] key points, got {tuple(keypoints.shape)}")
    kp = keypoints if keypoints.dtype in (torch.float32, torch.float64) else keypoints.float()
    kp = kp.contiguous()
    n = kp.numel() // 3
    shape = tuple(kp.shape[:-1]) + (64, 64, 64)
    if out is None:
        out = torch.empty(shape, dtype=torch.float32, device=kp.device)
    elif tuple(out.shape) != shape or out.dtype != torch.float32 or not out.is_contiguous() or out.device != kp.device:
        raise ValueError(f"keypoints_to_heatmap_3D: out must be a contiguous fp32 tensor of shape {shape} on {kp.device}")
    status = torch.zeros(1, dtype=torch.int32, device=kp.device) if check else None
    _cabi.check(_cabi.lib().lifter_keypoint_heatmaps(_cabi.ptr(kp), int(kp.dtype == torch.float64), n, float(sigma),
                                                     _cabi.ptr(out), _cabi.ptr(status), _cabi.stream_ptr()))
    if check and int(status.item()) & 1:
        raise AssertionError("keypoints_to_heatmap_3D: key points must lie in [-1, 1] (H36_dataset.py:156-157)")
    return out


class _AffineCols(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, a, b, mode):
        x = _dev(x, "normalise")
        cols = a.numel() if a is not None else 1
        out = torch.empty_like(x)
        _cabi.check(_cabi.lib().lifter_affine_cols(_cabi.ptr(x), _cabi.ptr(out), x.numel() // cols, cols, _cabi.ptr(a),
                                                   _cabi.ptr(b), mode, _cabi.stream_ptr()))
        ctx.mode = mode
        ctx.save_for_backward(a, b)
        return out

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        g = _dev(g, "normalise")
        if ctx.mode == 2:
            return 2.0 * g, None, None, None
        zero = torch.zeros_like(a)
        out = torch.empty_like(g)
        if ctx.mode == 0:      # d/dx (x - a) / b = 1 / b
            args = (zero, b, 0)
        elif ctx.mode == 1:    # d/dx x * b + a = b
            args = (zero, b, 1)
        else:                  # d/dx x * (a - b) + b = a - b
            args = (zero, a - b, 1)
        _cabi.check(_cabi.lib().lifter_affine_cols(_cabi.ptr(g), _cabi.ptr(out), g.numel() // a.numel(), a.numel(),
                                                   _cabi.ptr(args[0]), _cabi.ptr(args[1]), args[2], _cabi.stream_ptr()))
        return out, None, None, None


def normalise(x, mean, std):
    """``(x - mean) / std`` per (joint, coordinate), broadcast over the batch (main.py:334,337)."""
    n = int(np.prod(np.shape(mean)))
    return _AffineCols.apply(x, _vec(mean, n, x.device, "normalise"), _vec(std, n, x.device, "normalise"), 0)


def denormalise(y, mean, std):
    """``y * std + mean`` (main.py:473-479, train_5.py:42-47).  The repo's ``std_train_*.npy`` are float64, which makes
    the reference's result float64; here statistics are cast to float32 explicitly."""
    n = int(np.prod(np.shape(mean)))
    return _AffineCols.apply(y, _vec(mean, n, y.device, "denormalise"), _vec(std, n, y.device, "denormalise"), 1)


def normalise_minmax(x):
    """``2x - 1`` (main.py:331)."""
    return _MinMaxNorm.apply(x)


class _MinMaxNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = _dev(x, "normalise_minmax")
        out = torch.empty_like(x)
        _cabi.check(_cabi.lib().lifter_affine_cols(_cabi.ptr(x), _cabi.ptr(out), x.numel(), 1, None, None, 2,
                                                   _cabi.stream_ptr()))
        return out

    @staticmethod
    def backward(ctx, g):
        return 2.0 * g


def denormalise_minmax(y, vmax, vmin):
    """``y * (max - min) + min`` (train_5.py:35-40)."""
    n = int(np.prod(np.shape(vmax)))
    return _AffineCols.apply(y, _vec(vmax, n, y.device, "denormalise_minmax"),
                             _vec(vmin, n, y.device, "denormalise_minmax"), 3)


# ------------------------------------------------------------------------------------------------ metric
def _pose_loss_call(pred, target, mean, std, root, intrinsics, target2d, w, want_grad, global_rows=0):
    lib = _cabi.lib()
    pred, target = _dev(pred, "pose loss"), _dev(target, "pose loss")
    if pred.shape != target.shape or pred.dim() != 3 or pred.shape[-1] != 3:
        raise ValueError(f"expected (B, J, 3) prediction and target, got {tuple(pred.shape)} / {tuple(target.shape)}")
    B, J, _ = pred.shape
    dev = pred.device
    d = _cabi.PoseLossDesc()
    keep = [pred, target]
    d.pred, d.target, d.rows, d.global_rows, d.joints = pred.data_ptr(), target.data_ptr(), B, global_rows, J
    if mean is not None:
        m, s = _vec(mean, 3 * J, dev, "pose loss"), _vec(std, 3 * J, dev, "pose loss")
        keep += [m, s]
        d.mean, d.stdv = m.data_ptr(), s.data_ptr()
    if root is not None:
        r = _dev(root, "pose loss").reshape(B, 3)
        keep.append(r)
        d.root = r.data_ptr()
    if intrinsics is not None:
        k, t2 = _dev(intrinsics, "pose loss").reshape(B, 4), _dev(target2d, "pose loss").reshape(B, J, 2)
        keep += [k, t2]
        d.intrinsics, d.target2d = k.data_ptr(), t2.data_ptr()
    d.w_mpjpe, d.w_reproj, d.w_bone = w
    out = torch.empty(4 + J, dtype=torch.float32, device=dev)
    scratch = torch.empty(max(1, _cabi.check(lib.lifter_pose_loss_scratch_bytes(B, J)) // 4), dtype=torch.float32,
                          device=dev)
    dpred = torch.empty_like(pred) if want_grad else None
    d.out, d.scratch, d.dpred = out.data_ptr(), scratch.data_ptr(), (dpred.data_ptr() if want_grad else None)
    _cabi.check(lib.lifter_pose_loss(C.byref(d), _cabi.stream_ptr()))
    return out, dpred


def loss_MPJPE(prediction, target):
    """``phase5_loop/losses.py:3-7`` (= ``train_1.py:19-23``): per-joint sum over the batch of L2 position errors,
    shape ``(J,)`` -- a metric (the scripts never back-propagate through it); returned without autograd history."""
    if prediction.shape[0] == 0:
        return torch.zeros(prediction.shape[1], dtype=torch.float32, device=prediction.device)
    out, _ = _pose_loss_call(prediction.detach(), target.detach(), None, None, None, None, None, (0.0, 0.0, 0.0), False)
    return out[4:]


def epoch_mpjpe_mm(metric_sum, n_samples, zero_centre=True, first=1):
    """Epoch reduction of the accumulated ``loss_MPJPE`` sums (train_1.py:100-104, 143-145): mean over joints
    ``[first:17]`` of ``sum / N``, times ``17/16 * 1000`` mm when the root joint is zeroed."""
    m = (metric_sum / n_samples)[first:17].mean()
    return m * (17.0 / 16.0) * 1000.0 if zero_centre else m


class _PoseLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target, mean, std, root, intrinsics, target2d, w, holder):
        out, dpred = _pose_loss_call(pred, target, mean, std, root, intrinsics, target2d, w, True)
        ctx.save_for_backward(dpred)
        ctx.shape = pred.shape
        holder.append(out)
        return out[0].clone()

    @staticmethod
    def backward(ctx, g):
        (dpred,) = ctx.saved_tensors
        return (dpred * g).reshape(ctx.shape), None, None, None, None, None, None, None, None


class PoseLoss(torch.nn.Module):
    """Fused ``w_mpjpe * mean MPJPE + w_reproj * reprojection L1 + w_bone * bone-length L1`` on (B, 17, 3) joints,
    one kernel pass (north_star (b)).  ``mean``/``std`` de-standardise prediction and target first.  The
    reprojection and bone terms have no reference implementation (PARITY UNPINNED); their math:
    ``u = (fx X / Z + cx) / 1000`` against 2D keypoints in pixel/1000, ``| ||J_c - J_p|| - target |`` over
    ``H36M_EDGES``.  ``self.terms`` holds ``[total, mpjpe, reproj, bone, per-joint MPJPE sums...]`` of the last call."""

    def __init__(self, w_mpjpe=1.0, w_reproj=0.0, w_bone=0.0, mean=None, std=None):
        super().__init__()
        self.w = (float(w_mpjpe), float(w_reproj), float(w_bone))
        self.mean, self.std = mean, std
        self.terms = None

    def forward(self, pred, target, intrinsics=None, target2d=None, root=None):
        if self.w[1] != 0.0 and intrinsics is None:
            raise ValueError("PoseLoss: w_reproj != 0 needs camera intrinsics and 2D targets")
        holder = []
        B = pred.shape[0]
        total = _PoseLoss.apply(pred.reshape(B, -1, 3), target.reshape(B, -1, 3), self.mean, self.std, root,
                                intrinsics if self.w[1] != 0.0 else None, target2d, self.w, holder)
        self.terms = holder[0]
        return total


# ------------------------------------------------------------------------------------------------ L1 / triangle
class _L1Mean(torch.autograd.Function):
    """``torch.nn.L1Loss()`` (losses.py:21) with the optional batch-dimension centring of both operands."""

    @staticmethod
    def forward(ctx, a, b, centre):
        a, b = _dev(a, "L1"), _dev(b, "L1")
        if a.shape != b.shape:
            raise RuntimeError(f"The size of tensor a {tuple(a.shape)} must match the size of tensor b {tuple(b.shape)}")
        n = a.numel()
        row_len = n // a.shape[0]
        out = torch.empty(1, dtype=torch.float32, device=a.device)
        scratch = torch.empty((n + 2047) // 2048 + 1, dtype=torch.float32, device=a.device)
        _cabi.check(_cabi.lib().lifter_l1_sum(_cabi.ptr(a), _cabi.ptr(b), n, row_len, int(centre), 1.0 / n,
                                              _cabi.ptr(scratch), _cabi.ptr(out), _cabi.stream_ptr()))
        ctx.save_for_backward(a, b)
        ctx.centre, ctx.shapes = int(centre), (a.shape, b.shape)
        return out.reshape(())

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        n = a.numel()
        g = g.reshape(1).float().contiguous()
        da, db = torch.empty_like(a), torch.empty_like(b)
        _cabi.check(_cabi.lib().lifter_l1_grad(_cabi.ptr(a), _cabi.ptr(b), n, n // a.shape[0], ctx.centre, _cabi.ptr(g),
                                               1.0 / n, _cabi.ptr(da), _cabi.ptr(db), _cabi.stream_ptr()))
        return (da if ctx.needs_input_grad[0] else None), (db if ctx.needs_input_grad[1] else None), None


def l1_loss(a, b, centre=False):
    return _L1Mean.apply(a, b, centre)


class TriangleLoss(torch.nn.Module):
    """``phase5_loop/losses.py:10-62``: ``L1(pred2d, gt2d) + L1(pred3d, gt3d) + L1(lift_pred, pred3d)
    [+ L1(centre(proj_pred), centre(pred2d))]`` where ``centre`` subtracts SAMPLE 0 from samples 1.. (the reference's
    ``t[1:] -= t[0]``).  Returns ``(total, loss_2d, loss_3d, loss_lift, loss_proj)``."""

    def __init__(self, Project=False):
        super().__init__()
        self.Project = Project
        self.loss_2d, self.loss_3d, self.loss_lift, self.loss_proj = [], [], [], []

    def forward(self, predicted_2d, predicted_3d, lift_2d_gt, lift_2d_pred, gt_2d, gt_3d, proj_3d_pred=None,
                proj_3d_gt=None):
        loss_2d_ = l1_loss(predicted_2d, gt_2d)
        loss_3d_ = l1_loss(predicted_3d, gt_3d)
        loss_lift = l1_loss(lift_2d_pred, predicted_3d)
        loss_proj = 0
        if self.Project:
            loss_proj = l1_loss(proj_3d_pred, predicted_2d, centre=True)
            returned_loss = loss_2d_ + loss_3d_ + loss_lift + loss_proj
        else:
            returned_loss = loss_2d_ + loss_3d_ + loss_lift
        self.loss_2d.append(loss_2d_.detach())
        self.loss_3d.append(loss_3d_.detach())
        self.loss_lift.append(loss_lift.detach())
        if self.Project:
            self.loss_proj.append(loss_proj.detach())
        return returned_loss, loss_2d_, loss_3d_, loss_lift, loss_proj

    def report_losses(self):
        print(float(sum(self.loss_2d) / len(self.loss_2d)), float(sum(self.loss_3d) / len(self.loss_3d)),
              float(sum(self.loss_lift) / len(self.loss_lift)))
        self.loss_2d, self.loss_3d, self.loss_lift = [], [], []
        self.loss_domain_gap = []


class TriangleLoss_sep(torch.nn.Module):
    """``phase5_loop/losses.py:65-119``: ``L1(pred2d, gt2d) + L1(pred3d, gt3d) + L1(lift_gt, gt3d) +
    L1(lift_pred, lift_gt) [+ L1(centre(proj_gt), centre(gt2d)) + L1(centre(proj_pred), centre(proj_gt))]``."""

    def __init__(self, Project=False):
        super().__init__()
        self.Project = Project
        self.loss_2d, self.loss_3d, self.loss_lift, self.loss_domain_gap = [], [], [], []
        self.loss_proj, self.loss_gap_proj = [], []

    def forward(self, predicted_2d, predicted_3d, lift_2d_gt, lift_2d_pred, gt_2d, gt_3d, proj_3d_pred=None,
                proj_3d_gt=None):
        loss_2d_ = l1_loss(predicted_2d, gt_2d)
        loss_3d_ = l1_loss(predicted_3d, gt_3d)
        domain_gap_loss = l1_loss(lift_2d_pred, lift_2d_gt)
        loss_lift = l1_loss(lift_2d_gt, gt_3d)
        loss_proj = 0
        if self.Project:
            loss_gap_proj = l1_loss(proj_3d_pred, proj_3d_gt, centre=True)
            loss_proj = l1_loss(proj_3d_gt, gt_2d, centre=True)
            returned_loss = loss_2d_ + loss_3d_ + loss_lift + domain_gap_loss + loss_proj + loss_gap_proj
        else:
            returned_loss = loss_2d_ + loss_3d_ + loss_lift + domain_gap_loss
        self.loss_2d.append(loss_2d_.detach())
        self.loss_3d.append(loss_3d_.detach())
        self.loss_lift.append(loss_lift.detach())
        self.loss_domain_gap.append(domain_gap_loss.detach())
        return returned_loss, loss_2d_, loss_3d_, loss_lift, loss_proj

    def report_losses(self):
        print(float(sum(self.loss_2d) / len(self.loss_2d)), float(sum(self.loss_3d) / len(self.loss_3d)),
              float(sum(self.loss_lift) / len(self.loss_lift)),
              float(sum(self.loss_domain_gap) / len(self.loss_domain_gap)))
        self.loss_2d, self.loss_3d, self.loss_lift, self.loss_domain_gap = [], [], [], []


# Every entry point launches on the device of its first CUDA tensor, whatever device is current (_cabi.device_guard).
for _name in ("coco2h36m", "keypoint_stats", "process_data", "keypoints_to_heatmap_3D", "normalise", "denormalise", "normalise_minmax",
              "denormalise_minmax", "loss_MPJPE", "_pose_loss_call", "l1_loss"):
    globals()[_name] = _cabi.device_guard(globals()[_name])
for _cls in (_FlipPose, _AffineCols, _MinMaxNorm, _PoseLoss, _L1Mean):
    _cls.forward = staticmethod(_cabi.device_guard(_cls.forward))
    _cls.backward = staticmethod(_cabi.device_guard(_cls.backward))
del _name, _cls
