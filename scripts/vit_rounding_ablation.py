"""CPU ablation of the 16-bit rounding points of the MyViT eval path (fp32 oracle with fp16 roundings switched on one
group at a time): which stored tensor costs how much of the error budget after de-normalisation.  Test infrastructure
(imports oracle/); results quoted in DESIGN.md section 8."""
import sys, numpy as np, torch, torch.nn.functional as F
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from oracle import vit_oracle as V
st = dict(np.load(__import__("os").path.join(sys.path[0], "tests", "golden", "h36m_stats.npz")))
m3, s3 = torch.from_numpy(st["mean_train_3d"]).float(), torch.from_numpy(st["std_train_3d"]).float()
params = V.numpy_init(3)
B = 300
x = torch.randn(B, 17, 2, generator=torch.Generator().manual_seed(B))
ref = V.vit_forward(params, x) * s3 + m3
def r(t, on): return t.half().float() if on else t
def fwd(flags):
    H, nh, dh = 256, 4, 64
    W = lambda k: r(params[k], flags["w"])
    out = F.linear(x, params["linear_mapper.weight"], params["linear_mapper.bias"]) + params["pos_embed"]
    hf = out                    # fp32 value for first LN
    out = r(out, flags["h"])
    for b in range(2):
        p = f"blocks.{b}."
        src = hf if b == 0 else out
        u = F.layer_norm(src, (H,), params[p + "norm1.weight"], params[p + "norm1.bias"])
        u = r(F.layer_norm(u, (H,), params[p + "mhsa.norm.weight"], params[p + "mhsa.norm.bias"]), flags["u"])
        qkv = r(F.linear(u, W(p + "mhsa.to_qkv.weight")), flags["qkv"])
        q, k, v = (t.reshape(B, 17, nh, dh).transpose(1, 2) for t in qkv.chunk(3, -1))
        attn = torch.softmax(q @ k.transpose(-1, -2) * dh ** -0.5, -1)
        attn = r(attn, flags["p"])
        o = r((attn @ v).transpose(1, 2).reshape(B, 17, H), flags["o"])
        out = r(out + F.linear(o, W(p + "mhsa.to_out.weight")), flags["h"])
        u = r(F.layer_norm(out, (H,), params[p + "norm2.weight"], params[p + "norm2.bias"]), flags["u"])
        m = r(F.gelu(F.linear(u, W(p + "mlp.0.weight"), params[p + "mlp.0.bias"])), flags["m"])
        out = r(out + F.linear(m, W(p + "mlp.2.weight"), params[p + "mlp.2.bias"]), flags["h"])
    rr = r(torch.relu(F.linear(out, W("mlp.0.weight"), params["mlp.0.bias"])), flags["r"])
    return F.linear(rr, W("mlp.2.weight"), params["mlp.2.bias"]) * s3 + m3
keys = ["w", "h", "u", "qkv", "p", "o", "m", "r"]
def report(name, flags):
    out = fwd(flags); err = (out - ref).abs(); tol = 0.5e-3 + 1e-2 * ref.abs()
    print(f"{name:12s} rms {err.pow(2).mean().sqrt().item()*1e3:.3f} mm  max {err.max().item()*1e3:.3f} mm  outside {100*(err>tol).float().mean().item():.3f} %  worst err/tol {(err/tol).max().item():.2f}")
report("all", {k: True for k in keys})
for k in keys:
    report("only " + k, {kk: kk == k for kk in keys})
report("all but h", {k: k != "h" for k in keys})
report("all but h,w", {k: k not in ("h", "w") for k in keys})
