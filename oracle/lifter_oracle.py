"""CPU restatement of the reference lifter (``phase1_lifting/baselineModel.py``).

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Functional PyTorch fp32, no modules:
parameters and BatchNorm buffers travel in dicts keyed by the reference's
``state_dict()`` names, so a reference checkpoint is an oracle input.

Reference lines restated here:
* input stage  ``baselineModel.py:87-94``  flatten -> w1 -> batch_norm1 -> relu -> dropout
* residual block ``baselineModel.py:32-47`` (w1,bn1,relu,drop, w2,bn2,relu,drop, x+y), x num_stage
* output layer ``baselineModel.py:100``    w2, no activation
* BatchNorm1d train/eval semantics are torch's (momentum 0.1, eps 1e-5, biased batch
  variance for normalisation, unbiased for the running estimate; B == 1 in train mode
  raises ValueError) -- SURVEY.md §8 a5/a6
* ``weight_init`` ``baselineModel.py:10-12``: kaiming-normal (fan_in, gain sqrt 2) on Linear weights

Dropout masks are an *input* (``masks[site]`` boolean keep-masks, see ``oracle/philox.py``)
so that the oracle, the patched reference and the CUDA kernels share one mask stream.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch

BN_EPS = 1e-5
BN_MOMENTUM = 0.1


# ----------------------------------------------------------------------------- naming
def layer_names(num_stage):
    """(linear_prefix, bn_prefix) of every hidden layer in forward order; site index = position."""
    names = [("w1", "batch_norm1")]
    for i in range(num_stage):
        names.append((f"linear_stages.{i}.w1", f"linear_stages.{i}.batch_norm1"))
        names.append((f"linear_stages.{i}.w2", f"linear_stages.{i}.batch_norm2"))
    return names


def param_names(num_stage):
    """Reference ``model.parameters()`` order (registration order, baselineModel.py:68-85, 24-28)."""
    out = []
    for lin, bn in layer_names(num_stage):
        out += [lin + ".weight", lin + ".bias", bn + ".weight", bn + ".bias"]
    out += ["w2.weight", "w2.bias"]
    return out


def param_shapes(i_dim, o_dim, linear_size, num_stage):
    H = linear_size
    shapes = OrderedDict()
    for k, (lin, bn) in enumerate(layer_names(num_stage)):
        shapes[lin + ".weight"] = (H, i_dim if k == 0 else H)
        shapes[lin + ".bias"] = (H,)
        shapes[bn + ".weight"] = (H,)
        shapes[bn + ".bias"] = (H,)
    shapes["w2.weight"] = (o_dim, H)
    shapes["w2.bias"] = (o_dim,)
    return shapes


def numpy_init(i_dim, o_dim, linear_size, num_stage, seed, randomize_bn=True):
    """Deterministic, torch-RNG-independent weights (used where fixtures are too big to commit).

    Linear: U(+-1/sqrt(fan_in)) like torch's default; BN affine / running stats randomised so that
    folding and statistics code paths are exercised.
    """
    rs = np.random.RandomState(seed)
    params, buffers = OrderedDict(), OrderedDict()
    for name, shape in param_shapes(i_dim, o_dim, linear_size, num_stage).items():
        base = name.rsplit(".", 1)[0]
        if "batch_norm" in name:
            if name.endswith("weight"):
                v = rs.uniform(0.5, 1.5, shape) if randomize_bn else np.ones(shape)
            else:
                v = rs.normal(0, 0.2, shape) if randomize_bn else np.zeros(shape)
        else:
            fan_in = shape[1] if len(shape) == 2 else params[base + ".weight"].shape[1]
            bound = 1.0 / math.sqrt(fan_in)
            v = rs.uniform(-bound, bound, shape)
        params[name] = torch.from_numpy(np.asarray(v, dtype=np.float32))
    for _, bn in layer_names(num_stage):
        H = linear_size
        buffers[bn + ".running_mean"] = torch.from_numpy(
            (rs.normal(0, 0.3, H) if randomize_bn else np.zeros(H)).astype(np.float32))
        buffers[bn + ".running_var"] = torch.from_numpy(
            (rs.uniform(0.3, 1.8, H) if randomize_bn else np.ones(H)).astype(np.float32))
        buffers[bn + ".num_batches_tracked"] = torch.zeros((), dtype=torch.int64)
    return params, buffers


def kaiming_normal_(params, generator=None):
    """``weight_init`` (baselineModel.py:10-12): N(0, sqrt(2/fan_in)) on every Linear weight."""
    for name, p in params.items():
        if name.endswith(".weight") and p.dim() == 2:
            std = math.sqrt(2.0 / p.shape[1])
            p.copy_(torch.randn(p.shape, generator=generator) * std)
    return params


# ----------------------------------------------------------------------------- forward
def _batch_norm(z, gamma, beta, rm, rv, training, new_buffers, bn_name, nbt):
    if training:
        if z.shape[0] == 1:
            raise ValueError(
                f"Expected more than 1 value per channel when training, got input size {tuple(z.shape)}")
        mu = z.mean(dim=0)
        var = z.var(dim=0, unbiased=False)
        B = z.shape[0]
        new_buffers[bn_name + ".running_mean"] = (1 - BN_MOMENTUM) * rm + BN_MOMENTUM * mu.detach()
        new_buffers[bn_name + ".running_var"] = (1 - BN_MOMENTUM) * rv + BN_MOMENTUM * var.detach() * (B / (B - 1))
        new_buffers[bn_name + ".num_batches_tracked"] = nbt + 1
    else:
        mu, var = rm, rv
    return (z - mu) / torch.sqrt(var + BN_EPS) * gamma + beta


def lifter_forward(params, buffers, x, *, num_stage, training, use_bn=True, p_dropout=0.5, masks=None):
    """Returns (out (B, o_dim), new_buffers).  ``masks[site]`` = boolean keep-mask (B, H) per dropout site."""
    new_buffers = OrderedDict((k, v.clone()) for k, v in buffers.items())
    y = x.reshape(x.shape[0], -1)                                   # Flatten, baselineModel.py:89
    resid = None
    for site, (lin, bn) in enumerate(layer_names(num_stage)):
        if site >= 1 and site % 2 == 1:
            resid = y                                               # block input, baselineModel.py:32
        z = y @ params[lin + ".weight"].t() + params[lin + ".bias"]
        if use_bn:
            z = _batch_norm(z, params[bn + ".weight"], params[bn + ".bias"],
                            buffers[bn + ".running_mean"], buffers[bn + ".running_var"],
                            training, new_buffers, bn, buffers[bn + ".num_batches_tracked"])
        z = torch.relu(z)
        if training and p_dropout > 0:
            if p_dropout >= 1:
                z = z * 0
            else:
                keep = masks[site]
                keep = torch.as_tensor(np.asarray(keep)) if not torch.is_tensor(keep) else keep
                z = z * keep.to(z.dtype) / (1.0 - p_dropout)
        y = z
        if site >= 2 and site % 2 == 0:
            y = resid + y                                           # out = x + y, baselineModel.py:45
    out = y @ params["w2.weight"].t() + params["w2.bias"]           # baselineModel.py:100
    return out, new_buffers


def lifter_forward_backward(params, buffers, x, loss_fn, *, num_stage, training, use_bn=True,
                            p_dropout=0.5, masks=None, need_input_grad=False):
    """Runs forward, ``loss = loss_fn(out)``, autograd backward.  Returns dict with out, loss, grads, dx, buffers."""
    ps = OrderedDict((k, v.detach().clone().requires_grad_(True)) for k, v in params.items())
    xin = x.detach().clone().requires_grad_(need_input_grad)
    out, nb = lifter_forward(ps, buffers, xin, num_stage=num_stage, training=training, use_bn=use_bn,
                             p_dropout=p_dropout, masks=masks)
    loss = loss_fn(out)
    loss.backward()
    grads = OrderedDict((k, (v.grad if v.grad is not None else torch.zeros_like(v))) for k, v in ps.items())
    return dict(out=out.detach(), loss=loss.detach(), grads=grads,
                dx=xin.grad if need_input_grad else None, buffers=nb)
