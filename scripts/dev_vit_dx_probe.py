import importlib, sys, torch
sys.path.insert(0, "/root/repo")
from oracle import vit_oracle as V
pkg = importlib.import_module("3d_poseestimation_b200")
for (joints, in_d, hidden, heads, blocks, out_d, B) in [(21, 3, 512, 8, 2, 3, 9), (16, 2, 256, 8, 1, 2, 300), (17, 2, 256, 4, 2, 3, 512)]:
    for fmt in ("bf16", "fp16"):
        cfg = dict(n_joints=joints, in_d=in_d, n_blocks=blocks, hidden_d=hidden, n_heads=heads, out_d=out_d)
        params = V.numpy_init(13, **cfg)
        m = pkg.MyViT(chw=(1, joints, in_d), n_blocks=blocks, hidden_d=hidden, n_heads=heads, out_d=out_d)
        m.load_state_dict(params); m = m.cuda().train(); m.train_format = fmt
        x = torch.randn(B, joints, in_d, generator=torch.Generator().manual_seed(B)); y = torch.randn(B, joints, out_d, generator=torch.Generator().manual_seed(B + 1))
        xg = x.cuda().requires_grad_(True)
        torch.nn.functional.mse_loss(m(xg), y.cuda()).backward()
        xr = x.clone().requires_grad_(True)
        torch.nn.functional.mse_loss(V.vit_forward({k: v for k, v in params.items()}, xr, n_blocks=blocks, n_heads=heads), y).backward()
        g, r = xg.grad.cpu().reshape(-1).double(), xr.grad.reshape(-1).double()
        print(joints, hidden, B, fmt, "rel", float((g - r).norm() / r.norm()), "cos", float(torch.nn.functional.cosine_similarity(g, r, dim=0)))
