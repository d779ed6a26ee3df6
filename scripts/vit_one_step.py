"""Two eager VitTrainer steps at 8,192 poses (for ncu captures of the MyViT training kernels)."""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_poseestimation_b200")
VT = importlib.import_module("3d_poseestimation_b200.vit_trainer").VitTrainer
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
torch.manual_seed(0)
tr = VT(pkg.MyViT().cuda().train(), lr=1e-4, use_graph=False)
x, y = torch.randn(B, 17, 2, device="cuda"), torch.randn(B, 17, 3, device="cuda")
for _ in range(int(os.environ.get("STEPS", "2"))):
    tr.step(x, y)
torch.cuda.synchronize()
print("loss", tr.loss_value())
