"""CPU model of the CUDA training path WITH its 16-bit rounding points (test infrastructure).

The oracle (oracle/lifter_oracle.py) is the reference's fp32 arithmetic.  The kernels round to fp16/bf16
where they store GEMM operands; a ReLU whose pre-activation sits within that rounding of zero can switch,
which moves a single-sample gradient by a whole unit's worth -- legitimate mixed-precision behaviour, but
it makes "GPU vs fp32 oracle" a loose check of the backward arithmetic.  This model applies the SAME
roundings as csrc/lifter_train.cu (fp32 accumulation, 16-bit stores of a0, W, z, a, dOut, g, dz), so
"GPU vs model" is a tight check of every kernel, and "model vs oracle" is the documented precision cost.
"""
import torch

from oracle.lifter_oracle import BN_EPS, layer_names


def _r(t, dt):
    return t.to(dt).float()


def forward_backward(params, buffers, x, y, masks, *, num_stage, use_bn=True, p=0.5, dtype=torch.bfloat16,
                     training=True, loss="mse"):
    """Returns dict(out, loss, grads{name}, dx) following the kernels' dataflow (single process)."""
    names = layer_names(num_stage)
    L, B = len(names), x.shape[0]
    inv_keep = 1.0 / (1.0 - p) if (training and 0 < p < 1) else (0.0 if (training and p >= 1) else 1.0)
    r = lambda t: _r(t, dtype)
    W = [r(params[lin + ".weight"]) for lin, _ in names]
    W2 = r(params["w2.weight"])
    a_prev = r(x.reshape(B, -1))
    a_in, zs, acts, means, rstds = [], [], [], [], []
    a_list = []
    for l, (lin, bn) in enumerate(names):
        a_in.append(a_prev)
        z = r(a_prev @ W[l].t())                                   # GEMM epilogue stores z rounded
        bias, gamma, beta = params[lin + ".bias"], params[bn + ".weight"], params[bn + ".bias"]
        if use_bn:
            if training:
                zd = z.double()
                mean = zd.mean(0)
                var = ((zd * zd).mean(0) - mean * mean).clamp_min(0)
                mean_z, rstd = mean.float(), (1.0 / torch.sqrt(var + BN_EPS)).float()
            else:
                mean_z = buffers[bn + ".running_mean"] - bias
                rstd = 1.0 / torch.sqrt(buffers[bn + ".running_var"] + BN_EPS)
            scale = gamma * rstd
            shift = beta - mean_z * scale
        else:
            mean_z, rstd = torch.zeros_like(bias), torch.ones_like(bias)
            scale, shift = torch.ones_like(bias), bias
        yv = z * scale + shift
        keep = torch.as_tensor(masks[l]) if (training and p > 0) else torch.ones_like(yv, dtype=torch.bool)
        active = (yv > 0) & keep
        out = torch.where(active, yv * inv_keep, torch.zeros_like(yv))
        if l >= 2 and l % 2 == 0:
            out = out + a_list[l - 2]
        a_prev = r(out)
        a_list.append(a_prev)
        zs.append(z); acts.append(active); means.append(mean_z); rstds.append(rstd)
    out = a_prev @ W2.t() + params["w2.bias"]
    n = out.numel()
    if loss == "mse":
        lossv = ((out - y) ** 2).mean()
        dout = 2.0 * (out - y) / n
    else:
        raise ValueError(loss)

    grads = {}
    dO = r(dout)
    grads["w2.bias"] = dout.sum(0)
    grads["w2.weight"] = dO.t() @ a_list[L - 1]
    g_stream = r(dO @ W2)
    g_h = None
    dx = None
    for l in range(L - 1, -1, -1):
        lin, bn = names[l]
        g = g_h if l % 2 == 1 else g_stream
        dy = torch.where(acts[l], g * inv_keep, torch.zeros_like(g))
        zh = (zs[l] - means[l]) * rstds[l]
        if use_bn:
            sdy, sdyz = dy.double().sum(0), (dy * zh).double().sum(0)
            grads[bn + ".bias"], grads[bn + ".weight"] = sdy.float(), sdyz.float()
            k = params[bn + ".weight"] * rstds[l]
            c1, c2 = ((sdy / B).float(), (sdyz / B).float()) if training else (0.0, 0.0)
            dz = r(k * (dy - c1 - zh * c2))
        else:
            grads[bn + ".bias"] = torch.zeros_like(params[bn + ".bias"])
            grads[bn + ".weight"] = torch.zeros_like(params[bn + ".weight"])
            dz = r(dy)
        grads[lin + ".bias"] = dz.sum(0)
        grads[lin + ".weight"] = (dz.t() @ a_in[l])[:, :params[lin + ".weight"].shape[1]]
        if l == 0:
            dx = dz @ W[0]
        elif l % 2 == 0:
            g_h = r(dz @ W[l])
        else:
            g_stream = r(dz @ W[l] + g_stream)
    return dict(out=out, loss=lossv, grads=grads, dx=dx)
