"""autograd bridge of the training path: ``LinearModel.forward`` with gradients (train mode, or eval
mode with gradients flowing to the input as at the phase-5 call site ``train_5.py:145-168``).

Forward and backward each run as one sequence of sm_100a kernels behind the C ABI
(``lifter_train_forward`` / ``lifter_train_backward``); with a process group set on the model the
per-layer BatchNorm column sums are all-reduced over NCCL between the GEMM and the normalisation
(synchronised BatchNorm).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _cabi


def _dist_world(model):
    pg = getattr(model, "process_group", None)
    if pg is None:
        return None, 1, 0
    import torch.distributed as dist
    return pg, dist.get_world_size(pg), dist.get_rank(pg)


def _make_ctx(model, x, out, saved, fmt, training, step, global_batch, row_offset, step_state=None, calls=None):
    t = _cabi.TrainCtx()
    t.params = model._flat.data_ptr()
    t.operands = model._operands(fmt).data_ptr()
    t.running_mean, t.running_var = model._rm.data_ptr(), model._rv.data_ptr()
    t.num_batches_tracked = model._nbt.data_ptr() if training else None
    if not training and not model.BN:
        pass
    t.x = x.data_ptr() if x is not None else None
    t.out = out.data_ptr() if out is not None else None
    t.saved, t.saved_bytes = saved.data_ptr(), saved.numel()
    t.batch, t.global_batch, t.row_offset = x.shape[0] if x is not None else 0, global_batch, row_offset
    t.fmt, t.bn_mode, t.apply_dropout = fmt, (0 if training else 1), (1 if training else 0)
    t.step, t.seed = step & 0xFFFFFFFF, model.dropout_seed & 0xFFFFFFFFFFFFFFFF
    t.step_state = step_state.data_ptr() if step_state is not None else None
    if calls is not None:                    # (calls per optimiser step, index of this call): own masks, own exchange slots
        t.calls_per_step, t.call_index = int(calls[0]), int(calls[1])
    dp = getattr(model, "_dp", None)
    if dp is not None and step_state is not None and training:
        t.dp = C.pointer(dp)                 # in-kernel SyncBN exchange over the symmetric buffer
    return t


@_cabi.device_guard
def run_forward(model, x, training, saved=None, out=None, step=None, step_state=None, calls=None):
    """Forward with saved activations.  Returns (out, saved, ctx_struct)."""
    lib, cfg = _cabi.lib(), C.byref(model._cfg)
    fmt = model._fmt_code(model.train_format)
    B = x.shape[0]
    pg, world, rank = _dist_world(model)
    sync_bn = training and model.BN and world > 1
    global_batch = B * world if sync_bn else B
    row_offset = model.row_offset if model.row_offset else (rank * B if world > 1 else 0)
    if step is None:
        step = model.dropout_step
        if training:
            model.dropout_step += 1          # every training forward draws fresh masks, like torch.nn.Dropout
    nbytes = _cabi.check(lib.lifter_train_saved_bytes(cfg, B))
    if saved is None or saved.numel() < nbytes:
        saved = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    if out is None:
        out = torch.empty(B, model.output_size, dtype=torch.float32, device=x.device)
    t = _make_ctx(model, x, out, saved, fmt, training, step, global_batch, row_offset, step_state, calls)
    st = _cabi.stream_ptr()
    if not sync_bn or bool(t.dp):
        _cabi.check(lib.lifter_train_forward(cfg, C.byref(t), st))
    else:
        import torch.distributed as dist
        L, H = 2 * model.num_stage + 1, model.linear_size
        for l in range(L):
            _cabi.check(lib.lifter_train_fwd_layer_gemm(cfg, C.byref(t), l, st))
            off = lib.lifter_train_fwd_sums_offset(cfg, B, l)
            dist.all_reduce(saved[off:off + 16 * H].view(torch.float64), group=pg)
            _cabi.check(lib.lifter_train_fwd_layer_finish(cfg, C.byref(t), l, st))
        _cabi.check(lib.lifter_train_fwd_output(cfg, C.byref(t), st))
    if training:
        model._stats_epoch += 1              # running statistics changed: invalidate the eval-mode fold
    return out, saved, t


@_cabi.device_guard
def run_backward(model, t, grad_out, grad_flat, grad_x, scratch=None, optim=None):
    """Backward of the pair described by ``t``; writes ``grad_flat`` (parameter layout) and ``grad_x``.
    ``optim`` (``_cabi.Optim``): fold the optimiser into the backward pass (``lifter_train_backward_step``)."""
    lib, cfg = _cabi.lib(), C.byref(model._cfg)
    B = t.batch
    nbytes = _cabi.check(lib.lifter_train_scratch_bytes(cfg, B))
    if scratch is None or scratch.numel() < nbytes:
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=grad_out.device)
    t.grad_out, t.grad_params = grad_out.data_ptr(), grad_flat.data_ptr()
    t.grad_x = grad_x.data_ptr() if grad_x is not None else None
    t.scratch, t.scratch_bytes = scratch.data_ptr(), scratch.numel()
    st = _cabi.stream_ptr()
    pg, world, _ = _dist_world(model)
    sync_bn = t.bn_mode == 0 and model.BN and world > 1
    if optim is not None:
        if sync_bn and not bool(t.dp):
            raise RuntimeError("the fused backward + optimiser step needs the in-kernel data-parallel exchange")
        _cabi.check(lib.lifter_train_backward_step(cfg, C.byref(t), C.byref(optim), st))
    elif not sync_bn or bool(t.dp):
        _cabi.check(lib.lifter_train_backward(cfg, C.byref(t), st))
    else:
        import torch.distributed as dist
        L, H = 2 * model.num_stage + 1, model.linear_size
        _cabi.check(lib.lifter_train_bwd_begin(cfg, C.byref(t), st))
        for l in range(L - 1, -1, -1):
            _cabi.check(lib.lifter_train_bwd_layer_stats(cfg, C.byref(t), l, st))
            off = lib.lifter_train_bwd_sums_offset(cfg, B, l)
            dist.all_reduce(scratch[off:off + 16 * H].view(torch.float64), group=pg)
            _cabi.check(lib.lifter_train_bwd_layer_finish(cfg, C.byref(t), l, st))
    return scratch


class _LifterFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, training, x, *params):
        out, saved, t = run_forward(model, x, training)
        ctx.model, ctx.t, ctx.saved, ctx.x = model, t, saved, x
        ctx.versions = model._versions()
        return out

    @staticmethod
    def backward(ctx, grad_out):
        model, t = ctx.model, ctx.t
        if model._versions() != ctx.versions:
            raise RuntimeError("one of the variables needed for gradient computation has been modified by an "
                               "inplace operation (LinearModel parameters changed between forward and backward)")
        grad_out = grad_out.contiguous().float()
        need_x = ctx.needs_input_grad[2]
        grad_x = torch.empty_like(ctx.x) if need_x else None
        grad_flat = torch.zeros_like(model._flat)
        run_backward(model, t, grad_out, grad_flat, grad_x)
        grads = []
        for i, (p, (off, n)) in enumerate(zip(model._ordered_params(), model._offsets)):
            grads.append(grad_flat[off:off + n].view(p.shape) if ctx.needs_input_grad[3 + i] else None)
        ctx.saved = None
        return (None, None, grad_x, *grads)


def lifter_apply(model, x, needs_grad):
    if not needs_grad:                       # model.train() under no_grad: batch statistics + dropout, nothing kept
        out, _, _ = run_forward(model, x, model.training)
        return out
    return _LifterFunction.apply(model, model.training, x, *model._ordered_params())
