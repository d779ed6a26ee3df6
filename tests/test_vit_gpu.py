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
    # The residual stream is fp32 in the eval path (round 2); GEMM operands and weights are 16-bit.  Round 1 kept the stream
    # in 16 bits too and left ~0.5 % of the coordinates outside the band (all inside 2 mm): the CPU ablation
    # (scripts/vit_rounding_ablation.py) put the worst error at 0.64 of the tolerance with an fp32 stream.
    worst = float((err / tol).max())
    print(f"[B={B}] worst err/tol {worst:.2f}, max err {err.max().item() * 1e3:.3f} mm, in band {100 * frac_ok:.3f} %")
    assert frac_ok == 1.0, f"{100 * (1 - frac_ok):.3f} % of the outputs outside 0.5 mm + 1e-2 relative (worst err/tol {worst:.2f})"
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


def test_vit_unsupported_calls_raise(pkg):
    m = pkg.MyViT().cuda()
    with pytest.raises(RuntimeError):
        m(torch.randn(4, 17, 2))                                             # CPU tensor: no fallback
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


# ---------------------------------------------------------------------------------------------------- training path
def _oracle_grads(params, x, y, **cfg):
    p = {k: v.clone().requires_grad_(k != "pos_embed") for k, v in params.items()}
    xr = x.clone().requires_grad_(True)
    loss = torch.nn.functional.mse_loss(V.vit_forward(p, xr, **cfg), y)
    loss.backward()
    g = {k: v.grad for k, v in p.items() if k != "pos_embed"}
    g["__x__"] = xr.grad
    return float(loss.detach()), g


def _check_grads(m, ref, rel_tol=3e-2):
    for n, p in m.named_parameters():
        if n == "pos_embed":
            assert p.grad is None                                     # requires_grad = False (baselineModel.py:330)
            continue
        g, r = p.grad.cpu().double().reshape(-1), ref[n].double().reshape(-1)
        assert tuple(p.grad.shape) == tuple(ref[n].shape), n
        cos = torch.nn.functional.cosine_similarity(g, r, dim=0).item()
        assert rel_l2(g, r) < rel_tol and cos > 0.999, f"{n}: rel {rel_l2(g, r):.3e} cos {cos:.6f}"


@pytest.mark.parametrize("switches", [{}, {"LIFTER_VIT_FUSE_GELU_BWD": "1"}, {"LIFTER_VIT_WGRAD_STREAM": "0"}],
                         ids=["default", "gelu-bwd-in-dgrad", "one-stream"])
@pytest.mark.parametrize("fmt", ["bf16", "fp16"])
@pytest.mark.parametrize("tag,cfg", [("default", {}), ("b3h8", dict(n_blocks=3, n_heads=8))])
def test_vit_backward_matches_reference_golden(pkg, golden_dir, tag, cfg, fmt, switches, monkeypatch):
    """loss.backward() (train_1.py:94-95) on the reference's golden inputs: the loss against the imported reference's,
    every parameter gradient against autograd through the oracle (itself pinned to the reference's gradients), and two
    gradient tensors directly against the reference's.  16-bit operands and stored gradients: agreement to ~1e-2.
    ``switches``: the library's A/B paths (GELU backward in the data-gradient GEMM's epilogue; weight gradients on the
    chain's own stream) must give the same gradients."""
    for k, v in switches.items():
        monkeypatch.setenv(k, v)
    g = np.load(os.path.join(golden_dir, "vit.npz"))
    params = V.numpy_init(11, **cfg)
    m = pkg.MyViT(**cfg)
    m.load_state_dict(params)
    m = m.cuda().train()
    m.train_format = fmt
    x, y = T(g[tag + "_x"]), T(g[tag + "_target"])
    loss = torch.nn.functional.mse_loss(m(x.cuda()), y.cuda())
    loss.backward()
    np.testing.assert_allclose(loss.item(), float(g[tag + "_loss"]), rtol=1e-2)
    _, ref = _oracle_grads(params, x, y, **cfg)
    _check_grads(m, ref, 3e-2)       # fp16 has more mantissa but small gradients reach its subnormals (no loss scaling): same band
    assert rel_l2(m.linear_mapper.weight.grad.cpu(), g[tag + "_grad_map_w"]) < 3e-2
    assert rel_l2(m.mlp[2].weight.grad.cpu(), g[tag + "_grad_head_w"]) < 3e-2


@pytest.mark.parametrize("joints,in_d,hidden,heads,blocks,out_d,B", [(21, 3, 512, 8, 2, 3, 9), (16, 2, 256, 8, 1, 2, 300), (1, 2, 256, 4, 1, 3, 5)])
def test_vit_backward_other_shapes(pkg, joints, in_d, hidden, heads, blocks, out_d, B):
    cfg = dict(n_joints=joints, in_d=in_d, n_blocks=blocks, hidden_d=hidden, n_heads=heads, out_d=out_d)
    params = V.numpy_init(13, **cfg)
    m = pkg.MyViT(chw=(1, joints, in_d), n_blocks=blocks, hidden_d=hidden, n_heads=heads, out_d=out_d)
    m.load_state_dict(params)
    m = m.cuda().train()
    x = torch.randn(B, joints, in_d, generator=torch.Generator().manual_seed(B))
    y = torch.randn(B, joints, out_d, generator=torch.Generator().manual_seed(B + 1))
    xg = x.cuda().requires_grad_(True)          # the phase-5 loop feeds the lifter from a 2D model: gradient w.r.t. the key points
    torch.nn.functional.mse_loss(m(xg), y.cuda()).backward()
    _, ref = _oracle_grads(params, x, y, n_blocks=blocks, n_heads=heads)
    _check_grads(m, ref)
    assert tuple(xg.grad.shape) == tuple(x.shape)
    cos = torch.nn.functional.cosine_similarity(xg.grad.cpu().reshape(-1), ref["__x__"].reshape(-1), dim=0).item()
    # dx = dh0 W_map projects 256-512 bf16-stored gradient columns onto 2-4 inputs with heavy cancellation: the 4e-3 storage
    # rounding of dh0 shows as several percent here (measured 6e-2), against ~1e-2 on the parameter gradients
    assert rel_l2(xg.grad.cpu(), ref["__x__"]) < 0.3 and cos > 0.95, (rel_l2(xg.grad.cpu(), ref["__x__"]), cos)


def test_vit_training_loop_follows_the_oracle_curve(pkg):
    """The loop body train_1.py:73-100 with model_lift = MyViT and torch.optim.Adam over model.parameters(): 150 steps
    against the same loop on the fp32 oracle (same data, same initial weights).  30-step window means within 3 % (measured:
    0.2-2.4 %; bf16 activations put ~0.7 % on the evaluated loss itself once it is small), single steps within 8 % (measured worst 5.1 %); afterwards
    the no-grad forward sees the trained weights."""
    params = V.numpy_init(17)
    m = pkg.MyViT()
    m.load_state_dict(params)
    m = m.cuda().train()
    lr = 3e-4        # at 1e-3 this un-warmed-up transformer is chaotic: 1 % gradient noise moves the fp32 curve itself by 15 %
    opt = torch.optim.Adam(m.parameters(), lr=lr)
    ref_p = {k: v.clone().requires_grad_(k != "pos_embed") for k, v in params.items()}
    ref_opt = torch.optim.Adam([v for k, v in ref_p.items() if k != "pos_embed"], lr=lr)
    rs = np.random.RandomState(3)
    wtrue = rs.normal(0, 0.5, (2, 3)).astype(np.float32)
    steps, B, got, exp = 150, 64, [], []
    for s in range(steps):
        x = rs.normal(0, 1, (B, 17, 2)).astype(np.float32)
        y = np.tanh(x @ wtrue + 0.3 * x.sum(1, keepdims=True) @ wtrue).astype(np.float32)
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(m(T(x).cuda()), T(y).cuda())
        loss.backward()
        opt.step()
        got.append(loss.detach())
        ref_opt.zero_grad()
        rl = torch.nn.functional.mse_loss(V.vit_forward(ref_p, T(x)), T(y))
        rl.backward()
        ref_opt.step()
        exp.append(float(rl.detach()))
    got = torch.stack(got).cpu().double().numpy()
    exp = np.asarray(exp)
    assert exp[-20:].mean() < 0.5 * exp[:20].mean()                  # the curve actually descends
    np.testing.assert_allclose(got.reshape(-1, 30).mean(1), exp.reshape(-1, 30).mean(1), rtol=3e-2)
    np.testing.assert_allclose(got, exp, rtol=8e-2)
    xe = torch.randn(33, 17, 2)
    with torch.no_grad():
        out = m.eval()(xe.cuda()).cpu()
        ref = V.vit_forward({k: v.detach() for k, v in ref_p.items()}, xe)
    assert rel_l2(out, ref) < 3e-2


@pytest.mark.parametrize("optimizer,use_graph", [("adam", True), ("adamw", True), ("adam", False)])
def test_vit_fused_trainer_matches_the_autograd_loop(pkg, optimizer, use_graph):
    """VitTrainer (flat parameters, fused MSE, one table-driven Adam launch that also refreshes the 16-bit operands, CUDA-graph
    replay) against the loop train_1.py:91-96 written with autograd + torch.optim over the SAME kernels: the forward /
    backward launches are identical, so the curves differ only by fp32 summation order in the loss and the update
    (measured: identical loss on step 1, 7e-6 on step 2, 6e-4 on step 3 -- Adam's first updates are ~lr * sign(g), so
    near-zero gradient entries amplify the summation-order noise -- and inside 2 % over 40 steps at lr 3e-4).  Afterwards the re-homed module parameters
    hold the trained weights (eval forward of the trainer's model == eval forward of the autograd-trained twin)."""
    import importlib
    VT = importlib.import_module("3d_poseestimation_b200.vit_trainer").VitTrainer
    params = V.numpy_init(17)
    a, b = pkg.MyViT(), pkg.MyViT()
    a.load_state_dict(params)
    b.load_state_dict(params)
    a, b = a.cuda().train(), b.cuda().train()
    lr = 3e-4
    opt = (torch.optim.Adam if optimizer == "adam" else torch.optim.AdamW)(a.parameters(), lr=lr)
    tr = VT(b, lr=lr, optimizer=optimizer, use_graph=use_graph)
    rs = np.random.RandomState(5)
    wtrue = rs.normal(0, 0.5, (2, 3)).astype(np.float32)
    got, exp = [], []
    for s in range(40):
        B = 64 if s != 20 else 48                                # a second batch size mid-run: eager step, stale device state
        x = T(rs.normal(0, 1, (B, 17, 2)).astype(np.float32)).cuda()
        y = torch.tanh(x @ T(wtrue).cuda())
        if s == 30:                                              # ReduceLROnPlateau-style change between replays
            tr.set_lr(lr * 0.5)
            for g in opt.param_groups:
                g["lr"] = lr * 0.5
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(a(x), y)
        loss.backward()
        opt.step()
        exp.append(loss.detach())
        got.append(tr.step(x, y).clone())
    got, exp = torch.stack(got).cpu().double().numpy(), torch.stack(exp).cpu().double().numpy()
    assert exp[-5:].mean() < 0.7 * exp[:5].mean()
    np.testing.assert_allclose(got[:2], exp[:2], rtol=1e-4)
    np.testing.assert_allclose(got[:3], exp[:3], rtol=3e-3)
    np.testing.assert_allclose(got, exp, rtol=2e-2)
    assert tr.step_count == 40 and tr.graph_launches() > 0
    xe = torch.randn(33, 17, 2).cuda()
    with torch.no_grad():
        oa, ob = a.eval()(xe), b.eval()(xe)
    assert rel_l2(ob.cpu(), oa.cpu()) < 2e-2
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        assert rel_l2(q.detach().cpu(), p.detach().cpu()) < 2e-2, n
    # the module's parameters ARE the flat buffer (state_dict / checkpoints see the trained values)
    assert b.linear_mapper.weight.data_ptr() == tr.flat.data_ptr()
    with pytest.raises(ValueError):
        tr.step(xe[:, :5], xe)


def _vt():
    import importlib
    return importlib.import_module("3d_poseestimation_b200.vit_trainer").VitTrainer


def test_vit_fused_trainer_graph_replay_equals_eager_across_batch_sizes(pkg):
    """Graph replay (device step state, one captured graph per batch size, at most three resident sizes) against the same
    kernel sequence launched eagerly with host scalars: identical arithmetic, so the losses agree to fp32 round-off while
    the batch size walks through five values (evicting and re-capturing graphs) and the learning rate changes."""
    VT = _vt()
    params = V.numpy_init(3)
    a, b = pkg.MyViT(), pkg.MyViT()
    a.load_state_dict(params)
    b.load_state_dict(params)
    ta, tb = VT(a.cuda().train(), lr=3e-4, optimizer="adamw"), VT(b.cuda().train(), lr=3e-4, optimizer="adamw", use_graph=False)
    g = torch.Generator().manual_seed(9)
    la, lb = [], []
    for i, B in enumerate([32, 48, 64, 80, 96, 32, 32, 48, 48, 96, 96, 96]):
        x, y = torch.randn(B, 17, 2, generator=g).cuda(), torch.randn(B, 17, 3, generator=g).cuda()
        if i == 6:
            ta.set_lr(1e-4)
            tb.set_lr(1e-4)
        la.append(ta.step(x, y).clone())
        lb.append(tb.step(x, y).clone())
    la, lb = torch.stack(la).cpu().numpy(), torch.stack(lb).cpu().numpy()
    np.testing.assert_allclose(la, lb, rtol=2e-5)
    assert len(ta._graphs) <= VT.MAX_RESIDENT_BATCH_SIZES and len(ta._bufs) <= VT.MAX_RESIDENT_BATCH_SIZES
    assert rel_l2(ta.flat.cpu(), tb.flat.cpu()) < 1e-5
    # optimiser state round trip into a fresh trainer + host-side parameter writes (load_state_dict copies into the re-homed views)
    c = pkg.MyViT()
    tc = VT(c.cuda().train(), lr=1e-4, optimizer="adamw")
    c.load_state_dict(a.state_dict())
    tc.refresh_operands()
    tc.load_state_dict(ta.state_dict())
    x, y = torch.randn(64, 17, 2, generator=g).cuda(), torch.randn(64, 17, 3, generator=g).cuda()
    np.testing.assert_allclose(float(tc.step(x, y)), float(ta.step(x, y)), rtol=2e-5)
    assert rel_l2(tc.flat.cpu(), ta.flat.cpu()) < 1e-5


def test_vit_fused_trainer_at_the_bench_shape(pkg):
    """The exact path ``bench.py --workload vit`` times (8,192 poses, bf16, graph replay): three steps against the autograd
    loop over the same kernels, and run_epoch (host batches in, host losses out) against step()."""
    VT = _vt()
    torch.manual_seed(1)
    a = pkg.MyViT().cuda().train()
    b = pkg.MyViT().cuda().train()
    b.load_state_dict(a.state_dict())
    opt = torch.optim.Adam(a.parameters(), lr=1e-4)
    tr = VT(b, lr=1e-4)
    g = torch.Generator().manual_seed(2)
    hosts = [(torch.randn(8192, 17, 2, generator=g).pin_memory(), torch.randn(8192, 17, 3, generator=g).pin_memory()) for _ in range(3)]
    exp = []
    for x, y in hosts:
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(a(x.cuda()), y.cuda())
        loss.backward()
        opt.step()
        exp.append(float(loss))
    got = tr.run_epoch(hosts).numpy()
    np.testing.assert_allclose(got[:2], exp[:2], rtol=1e-4)
    np.testing.assert_allclose(got, exp, rtol=3e-3)
    assert tr.graph_launches() >= 60 and tr.step_count == 3
    with pytest.raises(RuntimeError):
        tr.run_epoch([(hosts[0][0].cuda(), hosts[0][1].cuda())])
