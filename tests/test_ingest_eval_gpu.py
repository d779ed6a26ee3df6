"""SURVEY §8(f) rows either side of the lifter: keypoint ingest (coco2h36m + scaling + normalisation in one kernel,
pinned by the reference's own function through tests/golden/ingest.npz) and the validation loop (flip test-time
augmentation, fused loss + MPJPE accumulation, one host read)."""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import lifter_oracle as lo, losses_oracle as L

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pl(pkg):
    return importlib.import_module("3d_poseestimation_b200.pose_losses")


def test_coco2h36m_matches_reference(pl, golden_dir):
    g = dict(np.load(os.path.join(golden_dir, "ingest.npz")))
    st = dict(np.load(os.path.join(golden_dir, "h36m_stats.npz")))
    kp = torch.from_numpy(g["coco"]).cuda()                                   # (N, 17, 3): x, y, confidence
    np.testing.assert_allclose(pl.coco2h36m(kp).cpu().numpy(), g["h36m"], rtol=1e-6, atol=1e-4)
    np.testing.assert_allclose(pl.coco2h36m(kp[..., :2].contiguous()).cpu().numpy(), g["h36m"], rtol=1e-6, atol=1e-4)
    got = pl.coco2h36m(kp, scale=1e-3, mean=st["mean_train_2d"], std=st["std_train_2d"])
    np.testing.assert_allclose(got.cpu().numpy(), g["standardised"], rtol=1e-5, atol=2e-5)
    np.testing.assert_allclose(pl.coco2h36m(kp, scale=1e-3, minmax=True).cpu().numpy(), g["minmax"], rtol=1e-5, atol=1e-6)
    # the dataset-side numpy branch is the reference's arithmetic
    np.testing.assert_allclose(pl.coco2h36m(g["coco"][0, :, :2]), g["h36m"][0], rtol=0, atol=0)
    assert pl.coco2h36m(kp[:0]).shape == (0, 17, 2)
    with pytest.raises(ValueError):
        pl.coco2h36m(torch.zeros(3, 16, 2, device="cuda"))
    with pytest.raises(RuntimeError):
        pl.coco2h36m(torch.zeros(3, 17, 2))


@pytest.mark.parametrize("flip", [None, "output", "input"])
@pytest.mark.parametrize("host", [False, True])
def test_evaluate_matches_the_reference_loop(pkg, pl, golden_dir, flip, host):
    trainer = importlib.import_module("3d_poseestimation_b200.trainer")
    st = dict(np.load(os.path.join(golden_dir, "h36m_stats.npz")))
    s3 = torch.from_numpy(st["std_train_3d"].astype(np.float32)).reshape(-1)
    params, buffers = lo.numpy_init(34, 51, 128, 2, seed=9, randomize_bn=True)
    m = pkg.LinearModel(34, 51, linear_size=128, num_stage=2, p_dropout=0.5, BN=True)
    m.load_state_dict({**params, **buffers})
    m = m.cuda().train()
    gen = torch.Generator().manual_seed(4)
    sizes = [300, 300, 300, 77]                                               # ragged last batch, as a DataLoader yields
    batches = [(torch.rand(b, 17, 2, generator=gen), torch.randn(b, 17, 3, generator=gen)) for b in sizes]
    feed = [(x.pin_memory(), y.pin_memory()) for x, y in batches] if host else [(x.cuda(), y.cuda()) for x, y in batches]
    res = trainer.evaluate(m, feed, loss="mse", flip=flip, metric_std=s3, n_samples=sum(sizes))
    assert m.training and res["n"] == sum(sizes)                              # mode restored

    # the loop of train_1.py:112-145 on the CPU oracle (eval-mode forward, flip, MSE, loss_MPJPE / len(test_set))
    def fwd(x):
        with torch.no_grad():
            return lo.lifter_forward(params, buffers, x.reshape(x.shape[0], -1), num_stage=2, training=False)[0]
    loss, metric = 0.0, torch.zeros(17)
    for x, y in batches:
        b = x.shape[0]
        p = fwd(x).reshape(b, 17, 3)
        if flip == "output":
            p = (L.flip_pose(p) + p) / 2
        elif flip == "input":
            p = (L.flip_pose(fwd(L.flip_pose(x)).reshape(b, 17, 3)) + p) / 2
        loss += float(torch.nn.functional.mse_loss(p, y)) / len(batches)
        metric += L.mpjpe_per_joint_sum(p * s3.view(17, 3), y * s3.view(17, 3)) / sum(sizes)
    want_mm = float(metric[1:].mean() * (17 / 16) * 1000)
    np.testing.assert_allclose(res["loss"], loss, rtol=2e-3)
    np.testing.assert_allclose(res["per_joint"].numpy(), metric.numpy(), rtol=2e-3)
    np.testing.assert_allclose(res["mpjpe_mm"], want_mm, rtol=2e-3)
