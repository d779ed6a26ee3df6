"""GPU parity of eval-mode inference (LinearModel.forward under model.eval() + no_grad) against the CPU
oracle and the reference-generated golden vectors.  Tolerance (BASELINE.json north_star): 1e-2 relative /
0.5 mm absolute after de-normalisation; bit-level agreement is not expected (fp16 tensor-core operands)."""
import os

import numpy as np
import pytest
import torch

from oracle import lifter_oracle as lo
from oracle import losses_oracle as L

pytestmark = pytest.mark.gpu
T = torch.from_numpy


def _model_from(pkg, params, buffers, i_dim, o_dim, H, S, p, BN):
    m = pkg.LinearModel(i_dim, o_dim, linear_size=H, num_stage=S, p_dropout=p, BN=BN)
    m.load_state_dict({**params, **buffers})
    return m.cuda().eval()


@pytest.mark.parametrize("tag", ["small", "small_nobn", "h1024", "wide"])
def test_eval_matches_reference_golden(pkg, golden_dir, tag):
    g = dict(np.load(os.path.join(golden_dir, f"lifter_{tag}.npz")))
    i_dim, o_dim, H, S, B, BN, seed = [int(v) for v in g["cfg"]]
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed)
    m = _model_from(pkg, params, buffers, i_dim, o_dim, H, S, float(g["p"]), bool(BN))
    with torch.no_grad():
        out = m(T(g["x"]).cuda()).cpu().numpy()
    ref = g["eval_out"]
    err = np.abs(out - ref)
    assert err.max() <= 2e-3 + 2e-3 * np.abs(ref).max(), f"max err {err.max()} (ref absmax {np.abs(ref).max()})"


@pytest.mark.parametrize("B", [1, 64, 257, 20000])
def test_eval_matches_oracle_after_denormalisation(pkg, golden_dir, B):
    """Config C2 shape at oracle-sized batches: synthetic H36M keypoints from the repo statistics."""
    st = dict(np.load(os.path.join(golden_dir, "h36m_stats.npz")))
    m2, s2 = T(st["mean_train_2d"]).float(), T(st["std_train_2d"]).float()
    m3, s3 = T(st["mean_train_3d"]).float(), T(st["std_train_3d"]).float()
    params, buffers = lo.numpy_init(34, 51, 1024, 2, seed=9)
    m = _model_from(pkg, params, buffers, 34, 51, 1024, 2, 0.5, True)
    gen = torch.Generator().manual_seed(0)
    x = (m2 + s2 * torch.randn(B, 17, 2, generator=gen)).clamp(0, 1)
    xn = L.normalise(x, m2, s2)
    ref, _ = lo.lifter_forward(params, buffers, xn, num_stage=2, training=False)
    ref = L.denormalise(ref.reshape(B, 17, 3), m3, s3)
    with torch.no_grad():
        out = m(xn.cuda()).cpu().reshape(B, 17, 3)
    out = L.denormalise(out, m3, s3)
    err = (out - ref).abs()
    tol = 0.5e-3 + 1e-2 * ref.abs()           # 0.5 mm absolute + 1e-2 relative (metres)
    assert bool((err <= tol).all()), f"max err {err.max().item()*1e3:.3f} mm; worst err/tol {(err/tol).max().item():.2f}"
    # fused normalise / de-normalise path gives the same answer
    fused = m.infer(x.cuda(), in_stats=(m2, s2), out_stats=(m3, s3)).cpu().reshape(B, 17, 3)
    assert bool(((fused - ref).abs() <= tol).all())


def test_eval_chunking_is_invisible(pkg):
    params, buffers = lo.numpy_init(34, 51, 256, 1, seed=5)
    m = _model_from(pkg, params, buffers, 34, 51, 256, 1, 0.5, True)
    x = torch.randn(1000, 34, device="cuda")
    a = m.infer(x)
    b = m.infer(x, chunk_rows=128)
    c = m.infer(x, chunk_rows=300)
    assert torch.equal(a, b) and torch.equal(a, c)


def test_eval_sees_weight_updates(pkg):
    params, buffers = lo.numpy_init(34, 51, 128, 1, seed=6)
    m = _model_from(pkg, params, buffers, 34, 51, 128, 1, 0.5, True)
    x = torch.randn(32, 34, device="cuda")
    with torch.no_grad():
        a = m(x).clone()
        m.w2.bias.add_(1.0)                      # in-place update, like an optimiser step
        b = m(x)
    assert torch.allclose(b - a, torch.ones_like(a), atol=1e-5)


def test_cpu_input_raises(pkg):
    m = pkg.LinearModel(34, 51, linear_size=64, num_stage=1).cuda().eval()
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.zeros(4, 34))
