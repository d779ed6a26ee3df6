"""ctypes binding of ``include/lifter_b200.h`` (the C ABI of the sm_100a kernels).

There is no CPU fallback and no PyTorch fallback: if the library cannot be loaded, or a call is made
without a B200, this module raises.
"""
from __future__ import annotations

import ctypes as C
import functools
import os

from . import build as _build

_LIB = None

FMT_F16, FMT_BF16 = 0, 1
OUT_16, OUT_32_SLABS, OUT_32_DIRECT = 0, 1, 2
ERR_INVALID, ERR_CUDA, ERR_BATCH1, ERR_NODEVICE = -1, -2, -3, -4


class LifterConfig(C.Structure):
    """``lifter_config_t``: the constructor arguments of LinearModel (baselineModel.py:51-56)."""
    _fields_ = [("i_dim", C.c_int32), ("o_dim", C.c_int32), ("linear_size", C.c_int32), ("num_stage", C.c_int32),
                ("use_bn", C.c_int32), ("p_dropout", C.c_float), ("bn_eps", C.c_float), ("bn_momentum", C.c_float)]


class GemmDesc(C.Structure):
    """``lifter_gemm_desc_t``"""
    _fields_ = [("A", C.c_void_p), ("lda", C.c_int64), ("B", C.c_void_p), ("ldb", C.c_int64),
                ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32), ("a_mn_major", C.c_int32),
                ("b_mn_major", C.c_int32), ("fmt", C.c_int32), ("out_mode", C.c_int32),
                ("D", C.c_void_p), ("ldd", C.c_int64), ("n_valid", C.c_int32), ("splits", C.c_int32),
                ("relu", C.c_int32), ("scale", C.c_void_p), ("shift", C.c_void_p), ("residual", C.c_void_p),
                ("ldr", C.c_int64), ("stat_sum", C.c_void_p), ("stat_sq", C.c_void_p), ("cluster", C.c_int32),
                ("timeline", C.c_void_p), ("max_stages", C.c_int32)]


class DpInfo(C.Structure):
    """``lifter_dp_t``: peer pointers of the symmetric SyncBN exchange buffer."""
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("sync", C.c_void_p * 8), ("grad", C.c_void_p * 8),
                ("params", C.c_void_p * 8), ("operands", C.c_void_p * 8), ("flags", C.c_void_p * 8),
                ("mc_grad", C.c_void_p), ("mc_params", C.c_void_p), ("mc_operands", C.c_void_p)]


class TrainCtx(C.Structure):
    """``lifter_train_ctx_t``: one forward/backward pair of the training path."""
    _fields_ = [("params", C.c_void_p), ("operands", C.c_void_p), ("running_mean", C.c_void_p),
                ("running_var", C.c_void_p), ("num_batches_tracked", C.c_void_p), ("x", C.c_void_p),
                ("out", C.c_void_p), ("saved", C.c_void_p), ("saved_bytes", C.c_int64), ("batch", C.c_int64),
                ("global_batch", C.c_int64), ("row_offset", C.c_int64), ("fmt", C.c_int32), ("bn_mode", C.c_int32),
                ("apply_dropout", C.c_int32), ("step", C.c_uint32), ("seed", C.c_uint64), ("grad_out", C.c_void_p),
                ("grad_params", C.c_void_p), ("grad_x", C.c_void_p), ("scratch", C.c_void_p),
                ("scratch_bytes", C.c_int64), ("step_state", C.c_void_p), ("dp", C.POINTER(DpInfo)),
                ("calls_per_step", C.c_int32), ("call_index", C.c_int32)]


class Optim(C.Structure):
    """``lifter_optim_t``: the Adam/AdamW state ``lifter_train_backward_step`` updates range by range."""
    _fields_ = [("params", C.c_void_p), ("operands", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p),
                ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("weight_decay", C.c_float),
                ("decoupled", C.c_int32), ("grad_scale", C.c_float)]


class PoseLossDesc(C.Structure):
    """``lifter_pose_loss_t``"""
    _fields_ = [("pred", C.c_void_p), ("target", C.c_void_p), ("mean", C.c_void_p), ("stdv", C.c_void_p),
                ("root", C.c_void_p), ("intrinsics", C.c_void_p), ("target2d", C.c_void_p), ("rows", C.c_int64),
                ("global_rows", C.c_int64), ("joints", C.c_int32), ("w_mpjpe", C.c_float), ("w_reproj", C.c_float),
                ("w_bone", C.c_float), ("dpred", C.c_void_p), ("scratch", C.c_void_p), ("out", C.c_void_p)]


class Phase5LossDesc(C.Structure):
    """``lifter_phase5_loss_t``"""
    _fields_ = [(k, C.c_void_p) for k in ("out_pred", "out_pred_flip", "out_gt", "out_gt_flip", "pred2d", "gt2d", "pred3d",
                                          "gt3d")] + [("mode", C.c_int32)] + \
               [(k, C.c_void_p) for k in ("mean3d", "std3d", "mean2d", "std2d", "root", "intrinsics")] + \
               [("w_reproj", C.c_float), ("w_bone", C.c_float), ("rows", C.c_int64), ("global_rows", C.c_int64),
                ("joints", C.c_int32)] + \
               [(k, C.c_void_p) for k in ("d_out_pred", "d_out_pred_flip", "d_out_gt", "d_out_gt_flip", "d_pred2d",
                                          "d_pred3d", "scratch", "out")]


class AdamTable(C.Structure):
    """``lifter_adam_table_t``"""
    _fields_ = [("n", C.c_int32), ("p_off", C.c_int64 * 16), ("numel", C.c_int64 * 16), ("op_off", C.c_int64 * 16)]


class VitDesc(C.Structure):
    """``lifter_vit_t``: shapes, fp32 parameters and 16-bit GEMM operands of the MyViT lifter."""
    _fields_ = [(k, C.c_int32) for k in ("tokens", "in_dim", "hidden", "heads", "blocks", "mlp_dim", "head_dim", "out_dim",
                                         "fmt")] + [("ln_eps", C.c_float)] + \
               [(k, C.c_void_p) for k in ("map_w", "map_b", "pos", "ln", "mlp_b1", "mlp_b2", "head_b1", "head_b2", "w_qkv",
                                          "w_out", "w_mlp1", "w_mlp2", "w_head1", "w_head2")]


class VitGrads(C.Structure):
    """``lifter_vit_grads_t``"""
    _fields_ = [(k, C.c_void_p) for k in ("map", "ln", "w_qkv", "w_out", "w_mlp1", "mlp_b1", "w_mlp2", "mlp_b2", "w_head1",
                                          "head_b1", "w_head2", "head_b2")]


LOSS_MSE, LOSS_L1, LOSS_MPJPE = 0, 1, 2
LOSS_STATIC_MAX_ROWS = 65536
DP_MAX_WORLD = 8
DP_FLAG_BYTES = 256

_P, _I64, _I32, _F = C.c_void_p, C.c_int64, C.c_int32, C.c_float
_CFG = C.POINTER(LifterConfig)
_CTX = C.POINTER(TrainCtx)

# name -> (restype, argtypes); must list every symbol include/lifter_b200.h declares (tests check this)
SIGNATURES = {
    "lifter_version": (_I32, []),
    "lifter_last_error": (C.c_char_p, []),
    "lifter_device_check": (_I32, []),
    "lifter_source_digest": (C.c_char_p, []),
    "lifter_launch_count": (C.c_longlong, []),
    "lifter_num_param_tensors": (_I32, [_CFG]),
    "lifter_param_offset": (_I64, [_CFG, _I32]),
    "lifter_param_numel": (_I64, [_CFG, _I32]),
    "lifter_param_total": (_I64, [_CFG]),
    "lifter_operand_bytes": (_I64, [_CFG]),
    "lifter_eval_workspace_bytes": (_I64, [_CFG, _I64]),
    "lifter_pack_weights": (_I32, [_CFG, _P, _P, _I32, _P]),
    "lifter_fold_bn": (_I32, [_CFG, _P, _P, _P, _P, _P]),
    "lifter_forward_eval": (_I32, [_CFG, _P, _P, _P, _P, _P, _I64, _P, _P, _P, _P, _P, _I64, _I32, _I64, _P]),
    "lifter_train_saved_bytes": (_I64, [_CFG, _I64]),
    "lifter_train_scratch_bytes": (_I64, [_CFG, _I64]),
    "lifter_train_fwd_sums_offset": (_I64, [_CFG, _I64, _I32]),
    "lifter_train_bwd_sums_offset": (_I64, [_CFG, _I64, _I32]),
    "lifter_train_fwd_layer_gemm": (_I32, [_CFG, _CTX, _I32, _P]),
    "lifter_train_fwd_layer_finish": (_I32, [_CFG, _CTX, _I32, _P]),
    "lifter_train_fwd_output": (_I32, [_CFG, _CTX, _P]),
    "lifter_train_forward": (_I32, [_CFG, _CTX, _P]),
    "lifter_train_bwd_begin": (_I32, [_CFG, _CTX, _P]),
    "lifter_train_bwd_layer_stats": (_I32, [_CFG, _CTX, _I32, _P]),
    "lifter_train_bwd_layer_finish": (_I32, [_CFG, _CTX, _I32, _P]),
    "lifter_train_backward": (_I32, [_CFG, _CTX, _P]),
    "lifter_train_backward_step": (_I32, [_CFG, _CTX, C.POINTER(Optim), _P]),
    "lifter_loss_scratch_bytes": (_I64, [_I64, _I32]),
    "lifter_loss": (_I32, [_P, _P, _I64, _I32, _I32, _I64, _P, _P, _P, _P, _I32, _P]),
    "lifter_flip_pose": (_I32, [_P, _P, _I64, _I32, _I32, _I32, _P]),
    "lifter_coco2h36m": (_I32, [_P, _I32, _P, _I64, _F, _P, _P, _I32, _P]),
    "lifter_keypoint_stats_scratch_bytes": (_I64, [_I64, _I32, _I32]),
    "lifter_keypoint_stats": (_I32, [_P, _I64, _I32, _I32, _I32, _P, _P, _P]),
    "lifter_process_keypoints": (_I32, [_P, _P, _I64, _I32, _I32, _I32, _P, _I32, _I32, _P]),
    "lifter_keypoint_heatmaps": (_I32, [_P, _I32, _I64, C.c_double, _P, _P, _P]),
    "lifter_affine_cols": (_I32, [_P, _P, _I64, _I32, _P, _P, _I32, _P]),
    "lifter_pose_loss_scratch_bytes": (_I64, [_I64, _I32]),
    "lifter_pose_loss": (_I32, [C.POINTER(PoseLossDesc), _P]),
    "lifter_loss_static": (_I32, [_P, _P, _I64, _I32, _I32, _I64, _P, _P, _P, _I32, _P]),
    "lifter_phase5_loss_scratch_bytes": (_I64, [_I64]),
    "lifter_phase5_loss": (_I32, [C.POINTER(Phase5LossDesc), _P]),
    "lifter_phase5_input_grad": (_I32, [_P, _P, _P, _I64, _I32, _P, _P]),
    "lifter_sum_buffers": (_I32, [_P, _P, _P, _P, _P, _I64, _P]),
    "lifter_l1_sum": (_I32, [_P, _P, _I64, _I32, _I32, _F, _P, _P, _P]),
    "lifter_l1_grad": (_I32, [_P, _P, _I64, _I32, _I32, _P, _F, _P, _P, _P]),
    "lifter_adam_step": (_I32, [_P, _P, _P, _P, _I64, _F, _F, _F, _F, _F, _I32, _I32, _F, _P]),
    "lifter_dp_sync_bytes": (_I64, [_CFG, _I32]),
    "lifter_step_advance": (_I32, [_P, _F, _F, _F, _P]),
    "lifter_adam_step_packed": (_I32, [_CFG, _P, _P, _P, _P, _F, _F, _F, _F, _F, _I32, _I32, _F, _P, _P, _I32, _P]),
    "lifter_adam_step_table": (_I32, [_P, _P, _P, _P, _I64, _F, _F, _F, _F, _F, _I32, _I32, _F, _P, _P, _I32, C.POINTER(AdamTable), _P]),
    "lifter_vit_map_grad": (_I32, [_P, _I32, _I32, _P, _P, _P]),
    "lifter_dp_adam_step_table": (_I32, [C.POINTER(DpInfo), _I64, _P, _P, _F, _F, _F, _F, _I32, _F, _P, _I32, C.POINTER(AdamTable), _P]),
    "lifter_dp_adam_step": (_I32, [_CFG, C.POINTER(DpInfo), _P, _P, _F, _F, _F, _F, _I32, _F, _P, _I32, _P]),
    "lifter_vit_workspace_bytes": (_I64, [C.POINTER(VitDesc), _I64]),
    "lifter_vit_forward": (_I32, [C.POINTER(VitDesc), _P, _I64, _P, _P, _P]),
    "lifter_vit_saved_bytes": (_I64, [C.POINTER(VitDesc), _I64]),
    "lifter_vit_bwd_scratch_bytes": (_I64, [C.POINTER(VitDesc), _I64]),
    "lifter_vit_train_forward": (_I32, [C.POINTER(VitDesc), _P, _I64, _P, _P, _P]),
    "lifter_vit_backward": (_I32, [C.POINTER(VitDesc), _P, _I64, _P, _P, _P, C.POINTER(VitGrads), _P, _P]),
    "lifter_gemm": (_I32, [C.POINTER(GemmDesc), _P]),
    "lifter_trace_begin": (_I32, [_P]),
    "lifter_trace_end": (_I32, [C.c_char_p, _I32]),
    "lifter_gemm_slab_rows": (_I32, [_I32]),
    "lifter_gemm_stat_rows": (_I32, [_I32]),
    "lifter_gemm_effective_splits": (_I32, [_I32, _I32]),
}


def lib():
    """The loaded C-ABI library; builds it in-tree with nvcc if the prebuilt .so is missing or stale."""
    global _LIB
    if _LIB is None:
        path = _build.lib_path()
        if not os.path.exists(path) or (_build.is_stale() and os.environ.get("LIFTER_NO_REBUILD") != "1"):
            path = _build.build()
        handle = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)        # AttributeError if the library does not export it
            fn.restype, fn.argtypes = res, args
        _LIB = handle
    return _LIB


def last_error():
    msg = lib().lifter_last_error()
    return msg.decode() if msg else ""


def check(code, what=""):
    """Map a negative return code to the exception the reference's PyTorch path would raise."""
    if code is None or code >= 0:
        return code
    msg = last_error() or what
    if code == ERR_BATCH1:
        raise ValueError(msg)
    if code == ERR_INVALID:
        raise ValueError(msg)
    raise RuntimeError(msg)


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    """The current stream of the CURRENT device (entry points run under ``device_guard``, which makes the device of
    their tensors current first)."""
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _cuda_device_of(obj, depth=0):
    import torch
    if torch.is_tensor(obj):
        return obj.device if obj.is_cuda else None
    flat = getattr(obj, "_flat", None)                       # LinearModel
    if not torch.is_tensor(flat):
        flat = getattr(obj, "flat", None)                    # VitTrainer
    if torch.is_tensor(flat):
        return flat.device if flat.is_cuda else None
    if depth == 0:
        model = getattr(obj, "model", None)                  # trainers
        if model is not None:
            return _cuda_device_of(model, 1)
        if isinstance(obj, torch.nn.Module):                 # MyViT and friends: device of the first parameter
            for p in obj.parameters():
                return p.device if p.is_cuda else None
        if isinstance(obj, (list, tuple)) and obj:
            return _cuda_device_of(obj[0], 1)
    return None


def device_guard(fn):
    """Every kernel launch, ``cudaGetDevice`` (SM count, side streams) and ``current_stream`` of the library refers to the
    CURRENT CUDA device.  This decorator makes the device of the call's first CUDA tensor / model current for the
    duration of the call, so a model on ``cuda:1`` works while ``cuda:0`` is current (it used to launch on GPU 0 with
    GPU-1 pointers)."""
    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        import torch
        dev = None
        for a in args:
            dev = _cuda_device_of(a)
            if dev is not None:
                break
        if dev is None:
            for a in kwargs.values():
                dev = _cuda_device_of(a)
                if dev is not None:
                    break
        if dev is None or dev.index is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(dev):
            return fn(*args, **kwargs)
    return wrapper
