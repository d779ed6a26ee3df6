"""Development check of the MyViT training path against autograd on the fp32 oracle (run on a B200)."""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import vit_oracle as V  # noqa: E402
pkg = importlib.import_module("3d_poseestimation_b200")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
params = V.numpy_init(5)
m = pkg.MyViT()
m.load_state_dict(params)
m = m.cuda().train()
x = torch.randn(B, 17, 2, generator=torch.Generator().manual_seed(1))
y = torch.randn(B, 17, 3, generator=torch.Generator().manual_seed(2))
ref_p = {k: v.clone().requires_grad_(k != "pos_embed") for k, v in params.items()}
ref_out = V.vit_forward(ref_p, x)
ref_loss = torch.nn.functional.mse_loss(ref_out, y)
ref_loss.backward()
out = m(x.cuda())
loss = torch.nn.functional.mse_loss(out, y.cuda())
loss.backward()
print("loss", loss.item(), ref_loss.item(), "out rel", ((out.detach().cpu() - ref_out.detach()).norm() / ref_out.detach().norm()).item())
worst = 0.0
for n, p in m.named_parameters():
    if n == "pos_embed":
        assert p.grad is None
        continue
    g, r = p.grad.cpu().double().reshape(-1), ref_p[n].grad.double().reshape(-1)
    rel = ((g - r).norm() / (r.norm() + 1e-30)).item()
    cos = torch.nn.functional.cosine_similarity(g, r, dim=0).item()
    worst = max(worst, rel)
    print(f"{n:34s} rel {rel:.3e} cos {cos:.6f} |ref| {r.norm().item():.3e}")
print("worst rel", worst)
