"""GPU parity of the training path (forward with batch statistics + Philox dropout, backward, fused loss,
fused Adam) against the CPU oracle and the reference-generated golden vectors.  The kernels use bf16
tensor-core operands with fp32 accumulation, so agreement is to bf16 rounding, not bit-exact: outputs and
gradients are compared by relative L2 error and cosine; the loss curve within 1 % (north_star)."""
import importlib
import os

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


def _case(golden_dir, tag):
    g = dict(np.load(os.path.join(golden_dir, f"lifter_{tag}.npz")))
    i_dim, o_dim, H, S, B, BN, seed = [int(v) for v in g["cfg"]]
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed)
    return g, (i_dim, o_dim, H, S, B, bool(BN), seed, float(g["p"])), params, buffers


def _model(pkg, params, buffers, i_dim, o_dim, H, S, p, BN):
    m = pkg.LinearModel(i_dim, o_dim, linear_size=H, num_stage=S, p_dropout=p, BN=BN)
    m.load_state_dict({**params, **buffers})
    return m.cuda()


@pytest.mark.parametrize("fmt", ["fp16", "bf16"])
@pytest.mark.parametrize("tag", ["small", "small_nobn", "h1024", "wide"])
def test_train_step_matches_reference_golden(pkg, golden_dir, tag, fmt):
    """One training step (forward, MSE, backward) on the reference's golden inputs.

    Tight check: against tests/rounding_model.py (the oracle's arithmetic with the kernels' 16-bit rounding
    points) -- every output, gradient and statistic to ~1e-3.  Loose check: against the fp32 golden vectors of
    the reference itself; the gap is the precision cost of 16-bit operands (ReLU decisions within rounding of
    zero can switch, which moves a single-sample gradient by one unit's worth), see DESIGN.md.
    """
    import rounding_model as rm
    g, (i_dim, o_dim, H, S, B, BN, seed, p), params, buffers = _case(golden_dir, tag)
    m = _model(pkg, params, buffers, i_dim, o_dim, H, S, p, BN).train()
    m.train_format = fmt
    m.dropout_seed, m.dropout_step = seed + 7, 0                 # the mask stream the golden run used
    x = T(g["x"]).cuda().requires_grad_(True)
    y = T(g["y"]).cuda()
    out = m(x)
    loss = torch.nn.MSELoss(reduction="mean")(out, y)            # train_1.py:37,94
    loss.backward()
    masks = [dropout_keep_mask(B, H, s, 0, seed + 7, p) for s in range(2 * S + 1)]
    dt = torch.float16 if fmt == "fp16" else torch.bfloat16
    mod = rm.forward_backward(params, buffers, T(g["x"]), T(g["y"]), masks, num_stage=S, use_bn=BN, p=p, dtype=dt)
    # --- tight: kernels vs the rounding model
    assert rel_l2(out.detach().cpu(), mod["out"]) < 2e-3
    assert abs(loss.item() - float(mod["loss"])) < 1e-3 * float(mod["loss"])
    # (fp32 summation order differs between the GPU and the model, so a handful of 16-bit roundings -- and with
    #  them the odd ReLU decision -- still differ: 1e-2-level agreement, against 1e-1-level for the fp32 oracle)
    # (measured 0.4e-2 .. 3.1e-2 over the cases and mask streams: at 16..32 rows one switched ReLU is visible)
    assert rel_l2(x.grad.cpu(), mod["dx"]) < 4e-2, f"dx vs model {rel_l2(x.grad.cpu(), mod['dx'])}"
    scale = max(float(v.norm()) for v in mod["grads"].values())
    for name, prm in m.named_parameters():
        gm = mod["grads"][name]
        if not BN and "batch_norm" in name:
            assert prm.grad is None or float(prm.grad.abs().max()) == 0.0
            continue
        lin_bias = BN and name.endswith(".bias") and "batch_norm" not in name and not name.startswith("w2.")
        if lin_bias or float(gm.norm()) < 1e-4 * scale:
            # a Linear bias in front of train-mode BN has an analytically zero gradient (the reference returns
            # fp32 rounding noise, the rounding model 16-bit noise); the kernels write exact zeros
            assert float(prm.grad.norm()) < 1e-3 * scale, name
            continue
        assert rel_l2(prm.grad.cpu(), gm) < 5e-2, f"{name}: rel l2 vs model {rel_l2(prm.grad.cpu(), gm)}"
    allg = torch.cat([p_.grad.cpu().reshape(-1) for n_, p_ in m.named_parameters() if p_.grad is not None and (BN or "batch_norm" not in n_)])
    allm = torch.cat([mod["grads"][n_].reshape(-1) for n_, p_ in m.named_parameters() if p_.grad is not None and (BN or "batch_norm" not in n_)])
    # (measured 0.3e-2 .. 1.9e-2 over the cases, formats and the round-2 mask stream; bf16 at 32 rows is the largest)
    assert rel_l2(allg, allm) < 3e-2, f"all gradients vs model {rel_l2(allg, allm)}"
    # --- loose: vs the reference's own fp32 vectors
    eps = 2 ** -10 if fmt == "fp16" else 2 ** -7
    assert rel_l2(out.detach().cpu(), g["train_out"]) < 2 * eps
    assert abs(loss.item() - float(g["train_loss"])) < eps * float(g["train_loss"])
    assert rel_l2(x.grad.cpu(), g["train_dx"]) < 0.2
    sd = m.state_dict()
    for k in buffers:
        np.testing.assert_allclose(sd[k].cpu().numpy(), g["buf_after/" + k], rtol=4 * eps, atol=eps, err_msg=k)


@pytest.mark.parametrize("tag", ["small", "h1024"])
def test_eval_mode_backward_to_input(pkg, golden_dir, tag):
    """Phase-5 call site: frozen lifter in eval mode, gradients flow through it to its input."""
    g, (i_dim, o_dim, H, S, B, BN, seed, p), params, buffers = _case(golden_dir, tag)
    m = _model(pkg, params, buffers, i_dim, o_dim, H, S, p, BN).eval()
    x = T(g["x"]).cuda().requires_grad_(True)
    out = m(x)
    torch.nn.functional.mse_loss(out, T(g["y"]).cuda()).backward()
    print(f"[{tag}] eval-grad out rel {rel_l2(out.detach().cpu(), g['eval_out']):.2e} dx rel {rel_l2(x.grad.cpu(), g['eval_dx']):.2e}")
    assert rel_l2(out.detach().cpu(), g["eval_out"]) < 2e-2
    assert rel_l2(x.grad.cpu(), g["eval_dx"]) < 6e-2
    sd = m.state_dict()
    for k in buffers:                                            # eval mode leaves the running statistics alone
        assert torch.equal(sd[k].cpu(), buffers[k]), k


def test_train_mode_batch_of_one_raises(pkg):
    m = pkg.LinearModel(34, 51, linear_size=64, num_stage=1).cuda().train()
    with pytest.raises(ValueError, match="Expected more than 1 value per channel"):
        m(torch.zeros(1, 34, device="cuda"))


def test_dropout_stream_and_no_grad_train_forward(pkg):
    params, buffers = lo.numpy_init(34, 51, 256, 1, seed=3)
    m = _model(pkg, params, buffers, 34, 51, 256, 1, 0.5, True).train()
    x = torch.randn(64, 34, device="cuda")
    m.dropout_seed, m.dropout_step = 5, 0
    with torch.no_grad():
        a = m(x)
        b = m(x)                                 # step advanced: fresh masks
        m.dropout_step = 0
        c = m(x)
    assert not torch.allclose(a, b) and torch.allclose(a, c)
    assert int(m.batch_norm1.num_batches_tracked) == 3


@pytest.mark.parametrize("p", [0.5, 0.25, 0.3])
@pytest.mark.parametrize("row_offset", [0, 16, 24, 7])
def test_dropout_lane_widths_and_row_offsets(pkg, p, row_offset):
    """The mask stream of csrc/philox.cuh for every lane width (p = 0.5: 1-bit lanes, 16 rows per call; 0.25: 8-bit, 2
    rows; 0.3: 16-bit, 1 row), with a shard offset that is / is not a multiple of the call's row footprint and a row
    count that is not a multiple of anything: a single wrong keep decision moves an output by O(1)."""
    H, S, B, seed = 128, 1, 203, 12
    params, buffers = lo.numpy_init(34, 51, H, S, seed)
    m = _model(pkg, params, buffers, 34, 51, H, S, p, True).train()
    m.dropout_seed, m.dropout_step, m.row_offset = 77, 5, row_offset
    x = T(np.random.RandomState(3).normal(0, 1, (B, 34)).astype(np.float32))
    with torch.no_grad():
        out = m(x.cuda()).cpu()
    masks = [dropout_keep_mask(B, H, s, 5, 77, p, row_offset=row_offset) for s in range(2 * S + 1)]
    ref, _ = lo.lifter_forward(params, buffers, x, num_stage=S, training=True, p_dropout=p, masks=masks)
    assert rel_l2(out, ref) < 2e-2, rel_l2(out, ref)


@pytest.mark.parametrize("kind", ["mse", "l1", "mpjpe"])
def test_fused_loss_matches_oracle(pkg, golden_dir, kind):
    cabi = importlib.import_module("3d_poseestimation_b200._cabi")
    st = dict(np.load(os.path.join(golden_dir, "h36m_stats.npz")))
    std3 = T(st["std_train_3d"]).float()
    B, J = 1000, 17
    g = torch.Generator().manual_seed(4)
    pred, tgt = torch.randn(B, J, 3, generator=g), torch.randn(B, J, 3, generator=g)
    fn = {"mse": L.mse_mean, "l1": L.l1_mean, "mpjpe": L.mpjpe_mean}[kind]
    p = pred.clone().requires_grad_(True)
    ref = fn(p, tgt)
    ref.backward()
    lib = cabi.lib()
    out = torch.zeros(1 + J, device="cuda")
    dpred = torch.empty(B, J * 3, device="cuda")
    scratch = torch.empty(lib.lifter_loss_scratch_bytes(B, J), dtype=torch.uint8, device="cuda")
    pc, tc, sc = pred.reshape(B, -1).cuda(), tgt.reshape(B, -1).cuda(), std3.reshape(-1).cuda()
    code = {"mse": 0, "l1": 1, "mpjpe": 2}[kind]
    cabi.check(lib.lifter_loss(cabi.ptr(pc), cabi.ptr(tc), B, J, code, B, cabi.ptr(sc), cabi.ptr(dpred),
                               cabi.ptr(scratch), cabi.ptr(out), 0, cabi.stream_ptr()))
    np.testing.assert_allclose(out[0].item(), float(ref), rtol=1e-5)
    np.testing.assert_allclose(dpred.cpu().numpy(), p.grad.reshape(B, -1).numpy(), rtol=1e-4, atol=1e-9)
    metric = L.mpjpe_per_joint_sum(pred * std3, tgt * std3)      # de-standardised MPJPE sums, losses.py:3-7
    np.testing.assert_allclose(out[1:].cpu().numpy(), metric.numpy(), rtol=1e-4)
    cabi.check(lib.lifter_loss(cabi.ptr(pc), cabi.ptr(tc), B, J, code, B, None, None, cabi.ptr(scratch),
                               cabi.ptr(out), 1, cabi.stream_ptr()))          # accumulate, no std, no gradient
    plain = L.mpjpe_per_joint_sum(pred, tgt)
    np.testing.assert_allclose(out[1:].cpu().numpy(), (metric + plain).numpy(), rtol=1e-4)


def test_fused_adam_matches_torch_optim_golden(pkg, golden_dir):
    cabi = importlib.import_module("3d_poseestimation_b200._cabi")
    g = dict(np.load(os.path.join(golden_dir, "adam.npz")))
    n = g["p0"].shape[0]
    npad = (n + 3) // 4 * 4
    for tag, wd, dec in (("adam", 0.0, 0), ("adam_l2", 1e-2, 0), ("adamw", 1e-2, 1)):
        p = torch.zeros(npad, device="cuda")
        p[:n] = T(g["p0"]).cuda()
        m, v = torch.zeros_like(p), torch.zeros_like(p)
        for s in range(g["grads"].shape[0]):
            gr = torch.zeros(npad, device="cuda")
            gr[:n] = T(g["grads"][s]).cuda()
            cabi.check(cabi.lib().lifter_adam_step(cabi.ptr(p), cabi.ptr(gr), cabi.ptr(m), cabi.ptr(v), npad, 1e-3, 0.9,
                                                   0.999, 1e-8, wd, dec, s + 1, 1.0, cabi.stream_ptr()))
            np.testing.assert_allclose(p[:n].cpu().numpy(), g[tag][s], rtol=2e-6, atol=2e-7, err_msg=f"{tag} step {s}")


def test_external_torch_optimizer_drives_the_module(pkg, golden_dir):
    """train_1.py:39,75,94-96 unchanged: torch.optim over model.parameters(); loss curve of the golden run."""
    g = dict(np.load(os.path.join(golden_dir, "train_curve.npz")))
    i_dim, o_dim, H, S, B, seed, steps = [int(v) for v in g["cfg"]]
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed, randomize_bn=False)
    m = _model(pkg, params, buffers, i_dim, o_dim, H, S, float(g["p"]), True).train()
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    rs = np.random.RandomState(seed + 1)
    xs = rs.normal(0, 1, (steps, B, i_dim)).astype(np.float32)
    ys = np.tanh(xs @ g["wtrue"]).astype(np.float32)
    m.dropout_seed = seed + 7
    losses = []
    for s in range(steps):
        m.dropout_step = s
        opt.zero_grad()
        loss = torch.nn.MSELoss(reduction="mean")(m(T(xs[s]).cuda()), T(ys[s]).cuda())
        loss.backward()
        opt.step()
        losses.append(loss.item())
    np.testing.assert_allclose(losses, g["losses"], rtol=2e-2)
    assert abs(np.mean(losses[-10:]) - np.mean(g["losses"][-10:])) < 1e-2 * np.mean(g["losses"][-10:])


@pytest.mark.parametrize("optimizer", ["adam", "adamw"])
def test_fused_trainer_follows_the_oracle_curve(pkg, optimizer):
    """Loop body train_1.py:73-100 as the fused step (forward, MSE, backward, Adam) vs the oracle, same masks."""
    trainer_mod = importlib.import_module("3d_poseestimation_b200.trainer")
    i_dim, o_dim, H, S, B, seed, steps = 34, 51, 256, 2, 64, 31, 60
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed, randomize_bn=False)
    m = _model(pkg, params, buffers, i_dim, o_dim, H, S, 0.5, True)
    tr = trainer_mod.LifterTrainer(m, lr=1e-3, loss="mse", optimizer=optimizer, seed=seed + 7)
    rs = np.random.RandomState(seed + 1)
    xs = rs.normal(0, 1, (steps, B, i_dim)).astype(np.float32)
    wtrue = rs.normal(0, 0.3, (i_dim, o_dim)).astype(np.float32)
    ys = np.tanh(xs @ wtrue).astype(np.float32)
    mo = {k: torch.zeros_like(v) for k, v in params.items()}
    vo = {k: torch.zeros_like(v) for k, v in params.items()}
    kw = dict(weight_decay=0.01, decoupled=True) if optimizer == "adamw" else {}
    got, exp = [], []
    for s in range(steps):
        got.append(float(tr.step(T(xs[s]).cuda(), T(ys[s]).cuda()).item()))
        masks = [dropout_keep_mask(B, H, site, s, seed + 7, 0.5) for site in range(2 * S + 1)]
        y = T(ys[s])
        r = lo.lifter_forward_backward(params, buffers, T(xs[s]), lambda o: ((o - y) ** 2).mean(), num_stage=S,
                                       training=True, p_dropout=0.5, masks=masks)
        buffers = r["buffers"]
        for k in params:
            params[k], mo[k], vo[k] = adam_oracle.adam_step(params[k], r["grads"][k], mo[k], vo[k], s + 1, 1e-3, **kw)
        exp.append(float(r["loss"]))
    np.testing.assert_allclose(got, exp, rtol=2e-2)
    assert abs(np.mean(got[-10:]) - np.mean(exp[-10:])) < 1e-2 * np.mean(exp[-10:])
    sd = m.state_dict()
    assert rel_l2(sd["w2.weight"].cpu(), params["w2.weight"]) < 2e-2
    assert rel_l2(sd["batch_norm1.running_var"].cpu(), buffers["batch_norm1.running_var"]) < 2e-2
    assert int(sd["batch_norm1.num_batches_tracked"]) == steps
    # the eval path sees the trained weights and running statistics
    m.eval()
    xe = T(xs[0])
    ref, _ = lo.lifter_forward(params, buffers, xe, num_stage=S, training=False)
    with torch.no_grad():
        out = m(xe.cuda()).cpu()
    assert rel_l2(out, ref) < 1e-1


def test_fused_trainer_loss_curve_over_1k_steps(pkg):
    """north_star: "training must agree on loss curves within 1 % over 1k steps with a fixed seed".

    1,000 fused steps (train_1.py:73-100: forward, MSE, backward, Adam; dropout 0.5, batch statistics) against
    the fp32 oracle on the same data, weights and dropout masks.  Tolerances: every 50-step window mean within
    1 %; single steps within 3 % (a CPU simulation of the kernels' bf16 rounding points over the same 1,000
    steps shows 0.17 % on the windows and 1.8 % on the worst single step)."""
    trainer_mod = importlib.import_module("3d_poseestimation_b200.trainer")
    i_dim, o_dim, H, S, B, seed, steps, win = 34, 51, 256, 2, 64, 31, 1000, 50
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed, randomize_bn=False)
    m = _model(pkg, params, buffers, i_dim, o_dim, H, S, 0.5, True)
    tr = trainer_mod.LifterTrainer(m, lr=1e-3, loss="mse", optimizer="adam", seed=seed + 7)
    rs = np.random.RandomState(seed + 1)
    wtrue = rs.normal(0, 0.3, (i_dim, o_dim)).astype(np.float32)
    mo = {k: torch.zeros_like(v) for k, v in params.items()}
    vo = {k: torch.zeros_like(v) for k, v in params.items()}
    got, exp = [], []
    for s in range(steps):
        x = rs.normal(0, 1, (B, i_dim)).astype(np.float32)
        y = T(np.tanh(x @ wtrue).astype(np.float32))
        got.append(tr.step(T(x).cuda(), y.cuda()).clone())          # no host read inside the loop
        masks = [dropout_keep_mask(B, H, site, s, seed + 7, 0.5) for site in range(2 * S + 1)]
        r = lo.lifter_forward_backward(params, buffers, T(x), lambda o: ((o - y) ** 2).mean(), num_stage=S,
                                       training=True, p_dropout=0.5, masks=masks)
        buffers = r["buffers"]
        for k in params:
            params[k], mo[k], vo[k] = adam_oracle.adam_step(params[k], r["grads"][k], mo[k], vo[k], s + 1, 1e-3)
        exp.append(float(r["loss"]))
    got = torch.stack([g.reshape(()) for g in got]).cpu().double().numpy()
    exp = np.asarray(exp)
    assert exp[-win:].mean() < 0.3 * exp[:win].mean()                 # the curve actually descends
    gw, ew = got.reshape(-1, win).mean(1), exp.reshape(-1, win).mean(1)
    np.testing.assert_allclose(gw, ew, rtol=1e-2, err_msg="50-step window means")
    np.testing.assert_allclose(got, exp, rtol=3e-2, err_msg="single steps")
    assert int(m.state_dict()["batch_norm1.num_batches_tracked"]) == steps
