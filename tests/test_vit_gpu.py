"""GPU parity of the MyViT eval-mode forward against the imported reference's golden
vectors and the fp32 CPU oracle.  fp16 tensor-core operands and 16-bit stored activations: agreement is to fp16
rounding through two transformer blocks, checked at north_star's eval tolerance after de-normalisation
(0.5 mm + 1e-2 relative with the repo's std_train_3d / mean_train_3d) and as a relative L2 error."""
import os

import numpy as np
import pytest
import torch

from oracle import vit_oracle as V

pytestmark = pytest.mark.gpu
T = torch.from_numpy


def _model(pkg, params, **cfg):
    m = pkg.MyViT(**cfg)
    m.load_state_dict(params)
    return m.cuda().eval()


def rel_l2(a, b):
    a, b = torch.as_tensor(a).double().reshape(-1), torch.as_tensor(b).double().reshape(-1)
    return float((a - b).norm() / (b.norm() + 1e-30))


@pytest.mark.parametrize("tag,cfg", [("default", {}), ("b3h8", dict(n_blocks=3, n_heads=8))])
def test_vit_eval_matches_reference_golden(pkg, golden_dir, tag, cfg):
    g = np.load(os.path.join(golden_dir, "vit.npz"))
    m = _model(pkg, V.numpy_init(11, **cfg), **cfg)
    with torch.no_grad():
        y = m(T(g[tag + "_x"]).cuda()).cpu().numpy()
    ref = g[tag + "_y"]
    assert y.shape == ref.shape
    assert rel_l2(y, ref) < 5e-3, rel_l2(y, ref)
    assert np.abs(y - ref).max() <= 5e-3 * np.abs(ref).max() + 2e-3


def test_vit_torch_seeded_construction_matches_reference_output(pkg, golden_dir):
    """Default torch initialisation under the same seed, then the reference's output on the same key points."""
    g = np.load(os.path.join(golden_dir, "vit.npz"))
    torch.manual_seed(5)
    m = pkg.MyViT().cuda().eval()
    with torch.no_grad():
        y = m(T(g["seeded_x"]).cuda()).cpu().numpy()
    assert rel_l2(y, g["seeded_y"]) < 5e-3


@pytest.mark.parametrize("B", [1, 7, 300, 2049])
def test_vit_eval_matches_oracle_after_denormalisation(pkg, golden_dir, B):
    st = dict(np.load(os.path.join(golden_dir, "h36m_stats.npz")))
    m3, s3 = T(st["mean_train_3d"]).float(), T(st["std_train_3d"]).float()
    params = V.numpy_init(3)
    m = _model(pkg, params)
    x = torch.randn(B, 17, 2, generator=torch.Generator().manual_seed(B))
    ref = V.vit_forward(params, x) * s3 + m3
    with torch.no_grad():
        out = m(x.cuda()).cpu() * s3 + m3
    err = (out - ref).abs()
    tol = 0.5e-3 + 1e-2 * ref.abs()                      # north_star: 0.5 mm absolute + 1e-2 relative (metres)
    frac_ok = float((err <= tol).float().mean())
    # 16-bit storage of every activation through two transformer blocks (~15 rounding points, ~1e-3 of the output
    # scale in total; these un-trained weights give outputs of several standard deviations): the north_star band holds
    # for all but a tail of small-valued coordinates, and everything is inside 2 mm + 1e-2 relative.  DESIGN.md section 8.
    assert frac_ok >= 0.995, f"{100 * (1 - frac_ok):.2f} % of the outputs outside 0.5 mm + 1e-2 relative"
    loose = 2e-3 + 1e-2 * ref.abs()
    assert bool((err <= loose).all()), f"max err {err.max().item() * 1e3:.3f} mm; worst err/tol {(err / loose).max().item():.2f}"
    assert rel_l2(out - m3, ref - m3) < 3e-3


@pytest.mark.parametrize("joints,in_d,hidden,heads,blocks,out_d", [
    (16, 2, 256, 4, 1, 3),      # one 16-row MMA tile, no padding rows
    (21, 3, 512, 8, 2, 3),      # 512 wide (two 256-column groups per lane), generic input width, 21 joints
    (32, 4, 256, 8, 1, 2),      # 32 tokens = the largest pose, head width 32
    (1, 2, 256, 4, 1, 5),       # a single token: softmax over one key
    (17, 2, 768, 12, 1, 3),     # 768 wide, twelve heads
    (17, 2, 256, 4, 0, 3),      # no transformer block at all: embedding + head
])
def test_vit_shapes_beyond_the_default(pkg, joints, in_d, hidden, heads, blocks, out_d):
    """Every template instantiation / shape branch of the transformer kernels against the fp32 oracle."""
    cfg = dict(n_joints=joints, in_d=in_d, n_blocks=blocks, hidden_d=hidden, n_heads=heads, out_d=out_d)
    params = V.numpy_init(21, **cfg)
    m = pkg.MyViT(chw=(1, joints, in_d), n_blocks=blocks, hidden_d=hidden, n_heads=heads, out_d=out_d)
    m.load_state_dict(params)
    m = m.cuda().eval()
    x = torch.randn(37, joints, in_d, generator=torch.Generator().manual_seed(joints))
    ref = V.vit_forward(params, x, n_blocks=blocks, n_heads=heads)
    with torch.no_grad():
        out = m(x.cuda()).cpu()
    assert tuple(out.shape) == (37, joints, out_d)
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)


def test_vit_chunking_weight_updates_and_formats(pkg):
    params = V.numpy_init(4)
    m = _model(pkg, params)
    x = torch.randn(777, 17, 2, device="cuda")
    a = m.infer(x)
    assert torch.equal(a, m.infer(x, chunk_poses=100)) and torch.equal(a, m.infer(x, chunk_poses=1))
    with torch.no_grad():
        m.mlp[2].bias.add_(1.0)                          # in-place update, like an optimiser step
        b = m(x)
    assert torch.allclose(b - a, torch.ones_like(a), atol=1e-5)
    m.infer_format = "bf16"
    c = m.infer(x)
    assert rel_l2(c.cpu(), b.cpu()) < 5e-2 and not torch.equal(c, b)
    assert tuple(m.infer(x[:0]).shape) == (0, 17, 3)


def test_vit_training_forward_is_not_a_silent_fallback(pkg):
    m = pkg.MyViT().cuda()
    x = torch.randn(4, 17, 2, device="cuda")
    with pytest.raises(NotImplementedError):
        m.train()(x)
    with pytest.raises(ValueError):
        m.eval()(torch.randn(4, 16, 2, device="cuda"))


@pytest.mark.parametrize("flip", [None, "input"])
def test_evaluate_drives_the_transformer_lifter(pkg, golden_dir, flip):
    """The validation loop train_1.py:112-145 with MyViT as `model_lift` (train_1.py:35): eval forward, flip TTA, fused
    loss + per-joint MPJPE with one host read, against the same loop on the CPU oracle."""
    import importlib
    from oracle import losses_oracle as L
    trainer = importlib.import_module("3d_poseestimation_b200.trainer")
    st = dict(np.load(os.path.join(golden_dir, "h36m_stats.npz")))
    s3 = T(st["std_train_3d"].astype(np.float32)).reshape(-1)
    params = V.numpy_init(8)
    m = _model(pkg, params)
    gen = torch.Generator().manual_seed(4)
    sizes = [300, 300, 77]
    batches = [(torch.rand(b, 17, 2, generator=gen), torch.randn(b, 17, 3, generator=gen)) for b in sizes]
    res = trainer.evaluate(m, [(x.pin_memory(), y.pin_memory()) for x, y in batches], loss="mse", flip=flip, metric_std=s3,
                           n_samples=sum(sizes))
    loss, metric = 0.0, torch.zeros(17)
    for x, y in batches:
        p = V.vit_forward(params, x)
        if flip == "input":
            p = (L.flip_pose(V.vit_forward(params, L.flip_pose(x))) + p) / 2
        loss += float(torch.nn.functional.mse_loss(p, y)) / len(batches)
        metric += L.mpjpe_per_joint_sum(p * s3.view(17, 3), y * s3.view(17, 3)) / sum(sizes)
    np.testing.assert_allclose(res["loss"], loss, rtol=3e-3)
    np.testing.assert_allclose(res["per_joint"].numpy(), metric.numpy(), rtol=3e-3)
    np.testing.assert_allclose(res["mpjpe_mm"], float(metric[1:].mean() * (17 / 16) * 1000), rtol=3e-3)
