"""BASELINE.json configs[3] and configs[4] on the GPU at test sizes: the phase-5 loop body around the lifter
(backup_june8.py:120-175: four lifter forwards with flip averaging, TriangleLoss, gradient into the 2D input) and the
widened lifter (i_dim 16*34, 2048 wide, 4 residual blocks), each against the CPU oracle with the same dropout masks."""
import importlib

import numpy as np
import pytest
import torch

from oracle import adam_oracle, lifter_oracle as lo, losses_oracle as L
from oracle.philox import dropout_keep_mask

pytestmark = pytest.mark.gpu
T = torch.from_numpy


def rel_l2(a, b):
    a, b = torch.as_tensor(a).double().reshape(-1), torch.as_tensor(b).double().reshape(-1)
    return float((a - b).norm() / (b.norm() + 1e-30))


def test_config5_wide_lifter_training_curve(pkg):
    trainer_mod = importlib.import_module("3d_poseestimation_b200.trainer")
    i_dim, o_dim, H, S, B, seed, steps = 16 * 34, 51, 2048, 4, 128, 5, 6
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed, randomize_bn=False)
    m = pkg.LinearModel(i_dim, o_dim, linear_size=H, num_stage=S, p_dropout=0.5, BN=True)
    m.load_state_dict({**params, **buffers})
    m = m.cuda()
    assert sum(p.numel() for p in m.parameters()) == 34_828_339
    tr = trainer_mod.LifterTrainer(m, lr=1e-3, loss="mse", optimizer="adam", seed=seed + 7)
    rs = np.random.RandomState(seed + 1)
    xs = rs.normal(0, 1, (steps, B, i_dim)).astype(np.float32)
    ys = rs.normal(0, 1, (steps, B, o_dim)).astype(np.float32)
    mo = {k: torch.zeros_like(v) for k, v in params.items()}
    vo = {k: torch.zeros_like(v) for k, v in params.items()}
    got, exp = [], []
    for s in range(steps):
        got.append(float(tr.step(T(xs[s]).cuda(), T(ys[s]).cuda()).item()))
        masks = [dropout_keep_mask(B, H, site, s, seed + 7, 0.5) for site in range(2 * S + 1)]
        y = T(ys[s])
        r = lo.lifter_forward_backward(params, buffers, T(xs[s]), lambda o: ((o - y) ** 2).mean(), num_stage=S,
                                       training=True, p_dropout=0.5, masks=masks)
        buffers = r["buffers"]
        for k in params:
            params[k], mo[k], vo[k] = adam_oracle.adam_step(params[k], r["grads"][k], mo[k], vo[k], s + 1, 1e-3)
        exp.append(float(r["loss"]))
    np.testing.assert_allclose(got, exp, rtol=2e-2)
    assert rel_l2(m.state_dict()["w2.weight"].cpu(), params["w2.weight"]) < 3e-2


def test_config4_phase5_loop_body(pkg):
    """Lifter inside the phase-5 loop: lift(pred 2D) and lift(gt 2D), each averaged with its flipped twin, TriangleLoss
    against a (stand-in) direct 3D prediction; gradients reach the lifter parameters AND the predicted 2D input."""
    pl = importlib.import_module("3d_poseestimation_b200.pose_losses")
    i_dim, o_dim, H, S, B, seed = 34, 51, 256, 2, 64, 9
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed, randomize_bn=False)
    m = pkg.LinearModel(i_dim, o_dim, linear_size=H, num_stage=S, p_dropout=0.5, BN=True)
    m.load_state_dict({**params, **buffers})
    m = m.cuda().train()
    m.dropout_seed, m.dropout_step = 3, 0
    rs = np.random.RandomState(seed)
    y1_hat = T(rs.uniform(0, 1, (B, 17, 2)).astype(np.float32))           # model_2d(frame): pixel / 1000
    y1 = T(rs.uniform(0, 1, (B, 17, 2)).astype(np.float32))
    y2_hat = T(rs.normal(0, 0.3, (B, 17, 3)).astype(np.float32))           # model_3d(frame)
    y2 = T(rs.normal(0, 0.3, (B, 17, 3)).astype(np.float32))

    def run(lift, flip, tri, dev):
        a = y1_hat.clone().to(dev).requires_grad_(True)
        b3 = y2_hat.clone().to(dev).requires_grad_(True)
        g2, g3, gt2 = y1.to(dev), y2.to(dev), y1.to(dev)
        lift_pred = (flip(lift(flip(a)).reshape(B, 17, 3)) + lift(a).reshape(B, 17, 3)) / 2
        lift_gt = (flip(lift(flip(g2)).reshape(B, 17, 3)) + lift(g2).reshape(B, 17, 3)) / 2
        out = tri(a, b3, lift_gt, lift_pred, gt2, g3)
        out[0].backward()
        return [float(o) for o in out[:4]], a.grad.cpu(), b3.grad.cpu()
    # GPU: four forwards draw dropout steps 0..3 in call order
    got, da, db = run(m, pl.flip_pose, pl.TriangleLoss(Project=False), "cuda")
    # oracle: the same four forwards with the same mask streams; BN running stats advance call by call
    ps = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    state = {"buf": buffers, "step": 0}

    def lift_o(x):
        masks = [dropout_keep_mask(B, H, site, state["step"], 3, 0.5) for site in range(2 * S + 1)]
        state["step"] += 1
        out, state["buf"] = lo.lifter_forward(ps, state["buf"], x, num_stage=S, training=True, p_dropout=0.5, masks=masks)
        return out
    exp, da_o, db_o = run(lift_o, L.flip_pose, lambda *a: L.triangle_loss(*a), "cpu")
    np.testing.assert_allclose(got, exp, rtol=1e-2)
    # L1 gradients are signs: sign(lift_pred - pred3d) flips wherever the 16-bit lifter output crosses pred3d
    # (|difference| below the ~1e-2 output tolerance), so a fraction of a percent of the entries differ by 2/n
    assert rel_l2(db, db_o) < 0.15, rel_l2(db, db_o)
    assert rel_l2(da, da_o) < 0.25, rel_l2(da, da_o)                        # through the 16-bit lifter
    cos = torch.nn.functional.cosine_similarity(m.w2.weight.grad.cpu().reshape(-1), ps["w2.weight"].grad.reshape(-1), dim=0)
    assert float(cos) > 0.98
    assert int(m.state_dict()["batch_norm1.num_batches_tracked"]) == 4


def test_learned_projector_shape(pkg):
    """the phase-5 learned projector reuses the same module at a tiny width,
    ``LinearModel(i_dim=51, o_dim=34, linear_size=64)`` (backup_june8.py:72): eval output and one training step's
    gradients against the oracle."""
    i_dim, o_dim, H, S, B, seed = 51, 34, 64, 2, 96, 13
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed, randomize_bn=True)
    m = pkg.LinearModel(i_dim, o_dim, linear_size=H, num_stage=S, p_dropout=0.5, BN=True)
    m.load_state_dict({**params, **buffers})
    m = m.cuda()
    rs = np.random.RandomState(seed)
    x, y = T(rs.normal(0, 1, (B, 17, 3)).astype(np.float32)), T(rs.normal(0, 1, (B, o_dim)).astype(np.float32))
    m.eval()
    with torch.no_grad():
        out = m(x.cuda())
        ref = lo.lifter_forward(params, buffers, x, num_stage=S, training=False)[0]
    assert out.shape == (B, o_dim)
    assert rel_l2(out.cpu(), ref) < 5e-3
    m.train()
    m.dropout_seed, m.dropout_step = 21, 0
    loss = ((m(x.cuda()) - y.cuda()) ** 2).mean()
    loss.backward()
    masks = [dropout_keep_mask(B, H, site, 0, 21, 0.5) for site in range(2 * S + 1)]
    r = lo.lifter_forward_backward(params, buffers, x, lambda o: ((o - y) ** 2).mean(), num_stage=S, training=True,
                                   p_dropout=0.5, masks=masks)
    np.testing.assert_allclose(float(loss), float(r["loss"]), rtol=2e-2)
    g = {k: p.grad.cpu() for k, p in m.named_parameters()}
    assert rel_l2(g["w2.weight"], r["grads"]["w2.weight"]) < 5e-2
    assert rel_l2(g["w1.weight"], r["grads"]["w1.weight"]) < 0.15
