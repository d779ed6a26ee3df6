"""Pins the CPU oracle (oracle/) against vectors produced by the imported reference
(oracle/make_golden.py ran /root/reference; the reference holds no tests of its own)."""
import os

import numpy as np
import pytest
import torch

from oracle import adam_oracle, lifter_oracle as lo, losses_oracle as L
from oracle.philox import dropout_keep_mask, philox4x32_10

T = torch.from_numpy


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name)))


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, exp in kat:
        got = tuple(int(np.asarray(v).ravel()[0]) for v in philox4x32_10(*ctr, *key))
        assert got == exp


def test_dropout_mask_stream_properties():
    m = dropout_keep_mask(512, 1024, site=2, step=5, seed=1234, p=0.5)
    assert abs(m.mean() - 0.5) < 0.01
    # shards of a data-parallel run reproduce the global mask
    top = dropout_keep_mask(256, 1024, 2, 5, 1234, 0.5)
    bot = dropout_keep_mask(256, 1024, 2, 5, 1234, 0.5, row_offset=256)
    assert np.array_equal(np.concatenate([top, bot]), m)
    assert dropout_keep_mask(8, 64, 0, 0, 1, 0.0).all()
    assert not dropout_keep_mask(8, 64, 0, 0, 1, 1.0).any()
    assert not np.array_equal(m, dropout_keep_mask(512, 1024, 3, 5, 1234, 0.5))


def _case(golden_dir, tag):
    g = _load(golden_dir, f"lifter_{tag}.npz")
    i_dim, o_dim, H, S, B, BN, seed = [int(v) for v in g["cfg"]]
    p = float(g["p"])
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed)
    masks = [dropout_keep_mask(B, H, s, 0, seed + 7, p) for s in range(2 * S + 1)]
    return g, dict(num_stage=S, use_bn=bool(BN), p_dropout=p), params, buffers, masks


@pytest.mark.parametrize("tag", ["small", "small_nobn", "h1024", "wide"])
def test_lifter_oracle_matches_reference(golden_dir, tag):
    g, kw, params, buffers, masks = _case(golden_dir, tag)
    x, y = T(g["x"]), T(g["y"])
    if "param/w1.weight" in g:                      # numpy_init is what the generator used
        for k in params:
            assert np.array_equal(g["param/" + k], params[k].numpy())
    out, _ = lo.lifter_forward(params, buffers, x, training=False, **kw)
    np.testing.assert_allclose(out.numpy(), g["eval_out"], rtol=1e-4, atol=2e-5)

    mse = lambda o: ((o - y) ** 2).mean()
    r = lo.lifter_forward_backward(params, buffers, x, mse, training=False, need_input_grad=True, **kw)
    np.testing.assert_allclose(r["dx"].numpy(), g["eval_dx"], rtol=2e-3, atol=1e-6)

    r = lo.lifter_forward_backward(params, buffers, x, mse, training=True, masks=masks, need_input_grad=True, **kw)
    np.testing.assert_allclose(r["out"].numpy(), g["train_out"], rtol=1e-4, atol=5e-5)
    np.testing.assert_allclose(float(r["loss"]), float(g["train_loss"]), rtol=1e-5)
    np.testing.assert_allclose(r["dx"].numpy(), g["train_dx"], rtol=5e-3, atol=2e-6)
    for k in buffers:
        np.testing.assert_allclose(r["buffers"][k].numpy(), g["buf_after/" + k], rtol=1e-5, atol=1e-6)
    for k, gr in r["grads"].items():
        if "grad/" + k in g:
            np.testing.assert_allclose(gr.numpy(), g["grad/" + k], rtol=5e-3, atol=2e-6, err_msg=k)
        else:
            np.testing.assert_allclose(gr.numpy().reshape(-1)[:64], g["gradhead/" + k], rtol=5e-3, atol=2e-6, err_msg=k)
            s = np.array([gr.numpy().sum(dtype=np.float64), np.abs(gr.numpy()).sum(dtype=np.float64)])
            if kw["use_bn"] and k.endswith(".bias") and "batch_norm" not in k and k != "w2.bias":
                # a bias in front of train-mode BatchNorm has an analytically zero gradient: float noise only
                assert s[1] < 1e-4 and g["gradsum/" + k][1] < 1e-4
            else:
                np.testing.assert_allclose(s[1], g["gradsum/" + k][1], rtol=1e-3, err_msg=k)


def test_train_mode_batch_of_one_raises():
    params, buffers = lo.numpy_init(34, 51, 64, 1, 0)
    with pytest.raises(ValueError, match="Expected more than 1 value per channel"):
        lo.lifter_forward(params, buffers, torch.zeros(1, 34), num_stage=1, training=True,
                          masks=[np.ones((1, 64), bool)] * 3)


def test_weight_init_statistics(golden_dir):
    g = _load(golden_dir, "weight_init.npz")
    params, _ = lo.numpy_init(34, 51, 256, 1, 0)
    lo.kaiming_normal_(params, generator=torch.Generator().manual_seed(0))
    for name, std, fan_in in zip(g["names"], g["stds"], g["fan_in"]):
        assert abs(float(params[str(name)].std()) - np.sqrt(2.0 / fan_in)) < 0.05 * np.sqrt(2.0 / fan_in)
        assert abs(std - np.sqrt(2.0 / fan_in)) < 0.05 * np.sqrt(2.0 / fan_in)     # the reference agrees


def test_losses_match_reference(golden_dir):
    g = _load(golden_dir, "losses.npz")
    t = {k: T(v) for k, v in g.items()}
    np.testing.assert_allclose(L.mpjpe_per_joint_sum(t["pred3d"], t["gt3d"]).numpy(), g["mpjpe_sum"], rtol=1e-6)
    args = (t["pred2d"], t["pred3d"], t["lift_gt"], t["lift_pred"], t["gt2d"], t["gt3d"])
    for proj in (False, True):
        r = L.triangle_loss(*args, proj_pred=t["proj_pred"], project=proj)
        np.testing.assert_allclose([float(v) for v in r], g[f"triangle_{int(proj)}"], rtol=1e-6)
    r = L.triangle_loss_sep(*args, proj_pred=t["proj_pred"], proj_gt=t["proj_gt"], project=True)
    np.testing.assert_allclose([float(v) for v in r], g["triangle_sep_1"], rtol=1e-6)
    # gradients of TriangleLoss(Project=True)
    ins = [t[k].clone().requires_grad_(True) for k in ("pred2d", "pred3d", "lift_pred", "proj_pred")]
    L.triangle_loss(ins[0], ins[1], t["lift_gt"], ins[2], t["gt2d"], t["gt3d"], proj_pred=ins[3], project=True)[0].backward()
    for v, k in zip(ins, ("pred2d", "pred3d", "lift_pred", "proj_pred")):
        np.testing.assert_allclose(v.grad.numpy(), g["triangle_1_grad_" + k], rtol=1e-5, atol=1e-8)
    assert np.array_equal(L.flip_pose(t["pred2d"]).numpy(), g["flip2d"])
    assert np.array_equal(L.flip_pose(t["pred3d"]).numpy(), g["flip3d"])
    assert torch.equal(L.flip_pose(L.flip_pose(t["pred3d"])), t["pred3d"])          # involution
    for name, fn in (("mse", L.mse_mean), ("l1", L.l1_mean), ("mpjpe_mean", L.mpjpe_mean)):
        p = t["pred3d"].clone().requires_grad_(True)
        l = fn(p, t["gt3d"])
        l.backward()
        np.testing.assert_allclose(float(l.detach()), float(g[name]), rtol=1e-6)
        np.testing.assert_allclose(p.grad.numpy(), g[name + "_grad"], rtol=1e-5, atol=1e-9)
    acc = L.mpjpe_per_joint_sum(t["pred3d"], t["gt3d"])
    np.testing.assert_allclose(float(L.epoch_mpjpe_mm(acc, 6)), float(g["epoch_mm"]), rtol=1e-6)
    st = _load(golden_dir, "h36m_stats.npz")
    y = L.denormalise(t["pred3d"].double(), T(st["mean_train_3d"]).double(), T(st["std_train_3d"]))
    np.testing.assert_allclose(y.numpy(), g["destd3d"], rtol=1e-12)
    xn = L.normalise(t["gt2d"], T(st["mean_train_2d"]), T(st["std_train_2d"]))     # fp32 subtract, fp64 divide
    np.testing.assert_allclose(xn.numpy(), g["std2d"], rtol=1e-12)


def test_unpinned_terms_are_self_consistent():
    """PARITY UNPINNED terms (no reference implementation): basic geometric sanity only."""
    j = torch.randn(4, 17, 3) * 0.3 + torch.tensor([0.0, 0.0, 5.0])
    K = torch.tensor(L.H36M_INTRINSICS)[torch.tensor([0, 1, 2, 3])]
    uv = L.pinhole_reprojection(j, K)
    assert uv.shape == (4, 17, 2) and float(L.reprojection_l1(j, K, uv)) == 0.0
    assert L.bone_lengths(j).shape == (4, 16)
    assert float(L.bone_length_l1(j, j + 1.0)) < 1e-5      # translation invariant


def test_adam_matches_torch_optim(golden_dir):
    g = _load(golden_dir, "adam.npz")
    for tag, kw in (("adam", {}), ("adam_l2", dict(weight_decay=1e-2)),
                    ("adamw", dict(weight_decay=1e-2, decoupled=True))):
        p = T(g["p0"].copy())
        m = torch.zeros_like(p)
        v = torch.zeros_like(p)
        for s in range(g["grads"].shape[0]):
            p, m, v = adam_oracle.adam_step(p, T(g["grads"][s]), m, v, s + 1, 1e-3, **kw)
            np.testing.assert_allclose(p.numpy(), g[tag][s], rtol=2e-6, atol=1e-7, err_msg=f"{tag} step {s}")


def test_plateau_scheduler_matches_torch():
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1e-4)
    sch = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, factor=0.7, patience=3, cooldown=2, min_lr=5e-6)
    mine = adam_oracle.ReduceLROnPlateauOracle(1e-4)
    rs = np.random.RandomState(0)
    for i in range(80):
        metric = 1.0 / (1 + i // 9) + 0.01 * rs.rand()
        sch.step(metric)
        assert abs(mine.step(metric) - opt.param_groups[0]["lr"]) < 1e-12


def test_training_curve_matches_reference(golden_dir):
    """Loop body train_1.py:73-100 (MSE + Adam + dropout 0.5) re-enacted by the oracle: same curve."""
    g = _load(golden_dir, "train_curve.npz")
    i_dim, o_dim, H, S, B, seed, steps = [int(v) for v in g["cfg"]]
    p = float(g["p"])
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed, randomize_bn=False)
    rs = np.random.RandomState(seed + 1)
    xs = rs.normal(0, 1, (steps, B, i_dim)).astype(np.float32)
    ys = np.tanh(xs @ g["wtrue"]).astype(np.float32)
    m = {k: torch.zeros_like(v) for k, v in params.items()}
    v = {k: torch.zeros_like(w) for k, w in params.items()}
    losses = []
    for s in range(steps):
        masks = [dropout_keep_mask(B, H, site, s, seed + 7, p) for site in range(2 * S + 1)]
        y = T(ys[s])
        r = lo.lifter_forward_backward(params, buffers, T(xs[s]), lambda o: ((o - y) ** 2).mean(),
                                       num_stage=S, training=True, p_dropout=p, masks=masks)
        buffers = r["buffers"]
        for k in params:
            params[k], m[k], v[k] = adam_oracle.adam_step(params[k], r["grads"][k], m[k], v[k], s + 1, 1e-3)
        losses.append(float(r["loss"]))
    np.testing.assert_allclose(losses, g["losses"], rtol=2e-3)
    np.testing.assert_allclose(params["w2.bias"].numpy(), g["final_w2_bias"], rtol=1e-2, atol=2e-4)


def test_keypoint_ingest_matches_reference(golden_dir):
    """coco2h36m (video2keypoints.py:14-57) + /1000 + the normalisations of main.py:331-337."""
    g = dict(np.load(os.path.join(golden_dir, "ingest.npz")))
    st = dict(np.load(os.path.join(golden_dir, "h36m_stats.npz")))
    np.testing.assert_array_equal(L.coco2h36m(g["coco"][..., :2]), g["h36m"])
    np.testing.assert_allclose(L.ingest_keypoints(g["coco"], st["mean_train_2d"], st["std_train_2d"]), g["standardised"],
                               rtol=1e-12)
    np.testing.assert_allclose(L.ingest_keypoints(g["coco"], minmax=True), g["minmax"], rtol=1e-12)


@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("zc", [0, 1])
def test_dataset_processing_matches_reference_process_data(golden_dir, dim, zc):
    """oracle/dataset_oracle.py against the reference's own process_data (main.py:271-355)."""
    from oracle import dataset_oracle as D
    g = _load(golden_dir, "process_data.npz")
    raw, tag = g[f"raw{dim}"], f"d{dim}_zc{zc}"
    out, mean, std = D.process_data(raw, standardize=True, z_c=bool(zc))
    np.testing.assert_allclose(mean, g[tag + "_mean"], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(std, g[tag + "_sd"], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(out, g[tag + "_std"], rtol=1e-10, atol=1e-12)
    plain, _, _ = D.process_data(raw, mean, std, standardize=False, z_c=bool(zc))
    np.testing.assert_allclose(plain, g[tag + "_plain"], rtol=1e-12, atol=1e-14)
    if zc:
        j16, _, _ = D.process_data(raw, standardize=True, z_c=True, num_of_joints=16)
        np.testing.assert_allclose(j16, g[f"d{dim}_j16"], rtol=1e-10, atol=1e-12)
        if dim == 2:
            mm, _, _ = D.process_data(raw, standardize=True, z_c=True, minmax=True)
            np.testing.assert_allclose(mm, g["d2_minmax"], rtol=1e-12, atol=1e-14)
    m2, s2, lo_, hi_ = D.keypoint_stats(raw, bool(zc))
    assert np.array_equal(m2, mean) and np.array_equal(s2, std)
    assert np.all(lo_ <= m2) and np.all(m2 <= hi_)


@pytest.mark.parametrize("tag,cfg", [("default", {}), ("b3h8", dict(n_blocks=3, n_heads=8))])
def test_vit_oracle_matches_reference_module(golden_dir, tag, cfg):
    """oracle/vit_oracle.py against the imported reference MyViT (baselineModel.py:307-362)."""
    from oracle import vit_oracle as V
    g = np.load(os.path.join(golden_dir, "vit.npz"), allow_pickle=False)
    params = V.numpy_init(11, **cfg)
    y = V.vit_forward(params, T(g[tag + "_x"]), **cfg).numpy()
    np.testing.assert_allclose(y, g[tag + "_y"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(V.positional_embeddings(17, 256).numpy(), g["pos_embed"], rtol=0, atol=1e-7)


@pytest.mark.parametrize("tag,cfg", [("default", {}), ("b3h8", dict(n_blocks=3, n_heads=8))])
def test_vit_oracle_gradients_match_reference_module(golden_dir, tag, cfg):
    """autograd through oracle/vit_oracle.py == loss.backward() on the imported reference MyViT (train_1.py:94-95)."""
    from oracle import vit_oracle as V
    g = np.load(os.path.join(golden_dir, "vit.npz"), allow_pickle=False)
    params = {k: v.clone().requires_grad_(k != "pos_embed") for k, v in V.numpy_init(11, **cfg).items()}
    loss = torch.nn.functional.mse_loss(V.vit_forward(params, T(g[tag + "_x"]), **cfg), T(g[tag + "_target"]))
    loss.backward()
    np.testing.assert_allclose(loss.item(), float(g[tag + "_loss"]), rtol=1e-5)
    names = [str(n) for n in g[tag + "_grad_names"]]
    assert names == [k for k in params if k != "pos_embed"]
    np.testing.assert_allclose([float(params[n].grad.double().abs().sum()) for n in names], g[tag + "_grad_abs"], rtol=1e-4)
    np.testing.assert_allclose([float(params[n].grad.double().sum()) for n in names], g[tag + "_grad_sums"], rtol=1e-3, atol=1e-5)
    np.testing.assert_allclose(params["linear_mapper.weight"].grad.numpy(), g[tag + "_grad_map_w"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(params["mlp.2.weight"].grad.numpy(), g[tag + "_grad_head_w"], rtol=1e-4, atol=1e-6)


def test_dropout_stream_lane_widths_and_shards():
    """oracle/philox.py (twin of csrc/philox.cuh): the lane width follows p, the keep rate is 1 - p, a data-parallel
    shard reproduces its rows of the global stream at any offset, and sixteen-bit lanes are the round-1 definition."""
    from oracle.philox import dropout_lanes
    assert dropout_lanes(0.5) == (1, 16, 1) and dropout_lanes(0.25) == (8, 2, 64) and dropout_lanes(0.3)[:2] == (16, 1)
    for p in (0.5, 0.25, 0.3, 0.0, 1.0):
        g = dropout_keep_mask(4096, 64, 3, 11, 0x1234_5678_9ABC, p)
        assert abs(g.mean() - (1 - p)) < 0.01
        for off, n in ((0, 100), (16, 50), (23, 77), (4000, 96)):
            assert (dropout_keep_mask(n, 64, 3, 11, 0x1234_5678_9ABC, p, row_offset=off) == g[off:off + n]).all()
    # k = 16: counter (row, col // 8, site, step), lane e = (r[e // 2] >> 16 (e % 2)) & 0xFFFF, keep = lane >= round(p 65536)
    rows, cols, site, step, seed, p = 9, 24, 2, 7, 99, 0.3
    r = philox4x32_10(np.arange(rows, dtype=np.uint64)[:, None], np.arange(cols // 8, dtype=np.uint64)[None, :], site, step,
                      seed & 0xFFFFFFFF, seed >> 32)
    want = np.empty((rows, cols // 8, 8), dtype=bool)
    for e in range(8):
        want[:, :, e] = ((r[e // 2] >> np.uint32(16 * (e % 2))) & np.uint32(0xFFFF)) >= np.uint32(19661)
    assert (dropout_keep_mask(rows, cols, site, step, seed, p) == want.reshape(rows, cols)).all()


def test_heatmap_oracle_matches_reference_golden(golden_dir):
    """oracle/heatmap_oracle.py against volumes produced by the reference's own keypoints_to_heatmap_3D
    (H36_dataset.py:143-202, exec of the source slice; stored as their non-zero voxels): bit for bit, for float64 and
    float32 key points, including windows clipped at one, two and three faces and half-integer voxel coordinates."""
    from oracle import heatmap_oracle as HO
    g = dict(np.load(os.path.join(golden_dir, "heatmaps.npz")))
    for tag, key in (("f64", "poses64"), ("f32", "poses32")):
        vol = HO.keypoints_to_heatmap_3d(g[key])
        assert vol.shape == (6, 17, 64, 64, 64) and vol.dtype == np.float32
        kp_i, vox_i, val = HO.to_sparse(vol.reshape(-1, 64, 64, 64))
        assert np.array_equal(kp_i, g[tag + "_kp"]) and np.array_equal(vox_i, g[tag + "_vox"])
        assert np.array_equal(val, g[tag + "_val"])
    assert HO.window(0.5) == 3 and HO.window(1.75) == 11
    with pytest.raises(AssertionError):
        HO.keypoints_to_heatmap_3d(np.array([[0.0, 1.5, 0.0]]))
