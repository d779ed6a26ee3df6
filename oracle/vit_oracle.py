"""CPU restatement of the ``MyViT`` lifter, functional PyTorch fp32.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).

Follows ``phase1_lifting/baselineModel.py``: ``get_positional_embeddings`` ``:220-225``, ``Attention`` ``:228-254``
(LayerNorm, bias-free qkv, softmax(q k^T / sqrt(d_head)) v, bias-free output projection), ``MyViTBlock`` ``:286-305``
(``x + mhsa(norm1(x))`` -- so TWO LayerNorms in a row in front of the qkv projection -- then ``+ mlp(norm2(.))`` with an
exact-erf GELU), ``MyViT`` ``:307-362`` (per-joint Linear(2 -> hidden) + positional embedding, blocks, per-joint
Linear(hidden -> hidden/2) -> ReLU -> Linear(hidden/2 -> 3)).  Keyed by the reference's ``state_dict`` names.

Pinned against the imported reference module (``oracle/make_golden.py vit`` -> ``tests/golden/vit.npz``).
"""
import numpy as np
import torch
import torch.nn.functional as F


def positional_embeddings(sequence_length, d):
    """baselineModel.py:220-225 (the reference fills a float32 tensor from float64 numpy scalars)."""
    i = np.arange(sequence_length, dtype=np.float64)[:, None]
    j = np.arange(d)[None, :]
    even = np.sin(i / (1e4 ** (j / d)))
    odd = np.cos(i / (1e4 ** ((j - 1) / d)))
    return torch.from_numpy(np.where(j % 2 == 0, even, odd).astype(np.float32))


def param_shapes(n_joints=17, in_d=2, n_blocks=2, hidden_d=256, n_heads=4, out_d=3, mlp_ratio=4):
    H = hidden_d
    s = {"linear_mapper.weight": (H, in_d), "linear_mapper.bias": (H,), "pos_embed": (n_joints, H)}
    for b in range(n_blocks):
        p = f"blocks.{b}."
        s.update({p + "norm1.weight": (H,), p + "norm1.bias": (H,), p + "mhsa.norm.weight": (H,), p + "mhsa.norm.bias": (H,),
                  p + "mhsa.to_qkv.weight": (3 * H, H), p + "mhsa.to_out.weight": (H, H),
                  p + "norm2.weight": (H,), p + "norm2.bias": (H,),
                  p + "mlp.0.weight": (mlp_ratio * H, H), p + "mlp.0.bias": (mlp_ratio * H,),
                  p + "mlp.2.weight": (H, mlp_ratio * H), p + "mlp.2.bias": (H,)})
    s.update({"mlp.0.weight": (H // 2, H), "mlp.0.bias": (H // 2,), "mlp.2.weight": (out_d, H // 2), "mlp.2.bias": (out_d,)})
    return s


def numpy_init(seed, **cfg):
    """Deterministic weights independent of torch's RNG: N(0, 1/fan_in) matrices, small biases, LayerNorm gains near 1."""
    rs = np.random.RandomState(seed)
    out = {}
    for k, shp in param_shapes(**cfg).items():
        if k == "pos_embed":
            out[k] = positional_embeddings(*shp)
        elif "norm" in k and k.endswith("weight"):
            out[k] = torch.from_numpy((1.0 + 0.1 * rs.normal(size=shp)).astype(np.float32))
        elif k.endswith("bias"):
            out[k] = torch.from_numpy((0.1 * rs.normal(size=shp)).astype(np.float32))
        else:
            out[k] = torch.from_numpy((rs.normal(size=shp) / np.sqrt(shp[1])).astype(np.float32))
    return out


def vit_forward(params, x, *, n_blocks=2, n_heads=4):
    """x (B, joints, in_d) fp32 -> (B, joints, out_d).  baselineModel.py:344-362."""
    H = params["linear_mapper.weight"].shape[0]
    dh = H // n_heads
    out = F.linear(x, params["linear_mapper.weight"], params["linear_mapper.bias"]) + params["pos_embed"]
    B, N, _ = out.shape
    for b in range(n_blocks):
        p = f"blocks.{b}."
        u = F.layer_norm(out, (H,), params[p + "norm1.weight"], params[p + "norm1.bias"], 1e-5)
        u = F.layer_norm(u, (H,), params[p + "mhsa.norm.weight"], params[p + "mhsa.norm.bias"], 1e-5)
        q, k, v = F.linear(u, params[p + "mhsa.to_qkv.weight"]).chunk(3, dim=-1)
        q, k, v = (t.reshape(B, N, n_heads, dh).transpose(1, 2) for t in (q, k, v))
        attn = torch.softmax(q @ k.transpose(-1, -2) * dh ** -0.5, dim=-1)
        o = (attn @ v).transpose(1, 2).reshape(B, N, H)
        out = out + F.linear(o, params[p + "mhsa.to_out.weight"])
        u = F.layer_norm(out, (H,), params[p + "norm2.weight"], params[p + "norm2.bias"], 1e-5)
        m = F.gelu(F.linear(u, params[p + "mlp.0.weight"], params[p + "mlp.0.bias"]))
        out = out + F.linear(m, params[p + "mlp.2.weight"], params[p + "mlp.2.bias"])
    r = torch.relu(F.linear(out, params["mlp.0.weight"], params["mlp.0.bias"]))
    return F.linear(r, params["mlp.2.weight"], params["mlp.2.bias"])
