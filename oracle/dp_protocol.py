"""TEST INFRASTRUCTURE (oracle side): CPU emulation of the data-parallel protocol the kernels implement
(csrc/train_kernels.cu exchange_column_sums, csrc/loss_adam.cu adam_dp_kernel), over any torch.distributed backend.

Per layer, forward : local column sums (sum z, sum z^2) -> all ranks' sums gathered, added in RANK ORDER -> batch
                     statistics of the GLOBAL batch.
Per layer, backward: local (sum dy, sum dy*zhat) -> dgamma/dbeta from the LOCAL sums; the gathered, rank-ordered sums
                     give c1, c2 of dz = gamma*rstd*(dy - c1 - zhat*c2).
Optimiser          : rank r adds slice r of every rank's flat gradient in rank order, runs Adam on that slice and
                     all-gathers the new parameters.
The reference has no distributed code; its oracle is the single-process run on the concatenated
batch, which tests/test_dp_cpu.py compares this against.
"""
from collections import OrderedDict

import torch
import torch.distributed as dist

from . import adam_oracle
from .lifter_oracle import BN_EPS, layer_names


def _gather_sum(t):
    """All ranks' tensors added in rank order (what the LL exchange computes on every GPU)."""
    parts = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, t.contiguous())
    s = torch.zeros_like(t)
    for p in parts:
        s = s + p
    return s


def dp_forward_backward(params, x, y, masks, *, num_stage, p_dropout, global_rows):
    """One rank's shard.  MSE(mean over the GLOBAL batch).  Returns (local loss share, local gradients)."""
    keep = 1.0 - p_dropout
    a, saved, resid = x, [], None
    for site, (lin, bn) in enumerate(layer_names(num_stage)):
        if site >= 1 and site % 2 == 1:
            resid = a
        z = a @ params[lin + ".weight"].t()                       # (the bias cancels in batch-statistics BN)
        sums = _gather_sum(torch.stack([z.sum(0), (z * z).sum(0)]))
        mean = sums[0] / global_rows
        var = (sums[1] / global_rows - mean * mean).clamp_min(0)
        rstd = 1.0 / torch.sqrt(var + BN_EPS)
        zhat = (z - mean) * rstd
        act = zhat * params[bn + ".weight"] + params[bn + ".bias"]
        on = (act > 0) & masks[site]
        out = torch.where(on, act / keep, torch.zeros_like(act))
        saved.append((a, zhat, rstd, on, lin, bn))
        a = out if not (site >= 2 and site % 2 == 0) else resid + out
    pred = a @ params["w2.weight"].t() + params["w2.bias"]
    diff = pred - y
    loss_share = (diff * diff).sum() / (global_rows * y.shape[1])
    g = 2.0 * diff / (global_rows * y.shape[1])
    grads = OrderedDict()
    grads["w2.weight"], grads["w2.bias"] = g.t() @ a, g.sum(0)
    g = g @ params["w2.weight"]                                    # gradient of the residual stream
    stream = None
    for site in range(len(saved) - 1, -1, -1):
        a_in, zhat, rstd, on, lin, bn = saved[site]
        if site >= 2 and site % 2 == 0:
            stream = g                                             # out = x + y: the block input gets g as well
        dy = torch.where(on, g / keep, torch.zeros_like(g))
        local = torch.stack([dy.sum(0), (dy * zhat).sum(0)])
        grads[bn + ".bias"], grads[bn + ".weight"] = local[0], local[1]      # LOCAL sums: the reduce adds shards
        glob = _gather_sum(local)
        dz = params[bn + ".weight"] * rstd * (dy - glob[0] / global_rows - zhat * glob[1] / global_rows)
        grads[lin + ".weight"] = dz.t() @ a_in
        grads[lin + ".bias"] = torch.zeros_like(params[lin + ".bias"])       # exactly zero in front of train-mode BN
        g = dz @ params[lin + ".weight"]
        if site >= 1 and site % 2 == 1:
            g = g + stream
    return loss_share, grads


def dp_adam(params, grads, m, v, step, lr):
    """Fused reduce-scatter + Adam + all-gather over the flat buffer, slice r on rank r."""
    rank, world = dist.get_rank(), dist.get_world_size()
    names = list(params.keys())
    flat_p = torch.cat([params[k].reshape(-1) for k in names])
    flat_g = torch.cat([grads[k].reshape(-1) for k in names])
    n = flat_p.numel()
    sl = (n + world - 1) // world
    lo, hi = min(rank * sl, n), min((rank + 1) * sl, n)
    all_g = [torch.zeros_like(flat_g) for _ in range(world)]
    dist.all_gather(all_g, flat_g)                                 # (the kernel reads only slice r of each peer)
    gsum = torch.zeros(hi - lo, dtype=flat_g.dtype)
    for q in range(world):
        gsum = gsum + all_g[q][lo:hi]
    p_new, m[lo:hi], v[lo:hi] = adam_oracle.adam_step(flat_p[lo:hi], gsum, m[lo:hi], v[lo:hi], step, lr)
    pad = torch.zeros(sl, dtype=flat_p.dtype)
    pad[:hi - lo] = p_new
    parts = [torch.zeros_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)
    flat_new = torch.cat(parts)[:n]
    out, off = OrderedDict(), 0
    for k in names:
        out[k] = flat_new[off:off + params[k].numel()].view_as(params[k]).clone()
        off += params[k].numel()
    return out
