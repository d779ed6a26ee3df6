"""Generate ``tests/golden/*.npz`` by RUNNING THE REFERENCE ITSELF (read-only import).

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Run once in the build container:

    python oracle/make_golden.py            # needs /root/reference; writes tests/golden/

The reference has no golden vectors of its own, so parity is pinned on outputs of
the imported reference modules: ``phase1_lifting/baselineModel.py`` (LinearModel, weight_init),
``phase5_loop/losses.py`` (loss_MPJPE, TriangleLoss, TriangleLoss_sep) and, by source slice because
its file imports matplotlib/seaborn, ``phase3_direct/my_HybrIK/utils.py:371-396`` (flip_pose).
torch.optim.Adam/AdamW (the reference's optimiser, ``train_1.py:39``) pin the fused-Adam math.
Dropout is patched to consume the Philox keep-masks of ``oracle/philox.py`` so the reference, the
oracle and the CUDA kernels see identical masks.
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.environ.get("LIFTER_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(REF, "phase1_lifting"))
sys.path.insert(0, os.path.join(REF, "phase5_loop"))

from oracle import lifter_oracle as lo          # noqa: E402
from oracle.philox import dropout_keep_mask     # noqa: E402

import baselineModel as ref_model               # noqa: E402  (the reference)
import losses as ref_losses                     # noqa: E402  (the reference)

GOLD = os.path.join(ROOT, "tests", "golden")
warnings.filterwarnings("ignore")


def ref_flip_pose():
    src = open(os.path.join(REF, "phase3_direct/my_HybrIK/utils.py")).read().splitlines()
    ns = {}
    exec("\n".join(src[370:396]), ns)      # utils.py:371-396 (import copy + def flip_pose)
    return ns["flip_pose"]


def ref_coco2h36m():
    src = open(os.path.join(REF, "phase1_lifting/video2keypoints.py")).read().splitlines()
    ns = {"np": np}
    exec("\n".join(src[13:57]), ns)       # video2keypoints.py:14-57 (def coco2h36m; the file imports cv2/matplotlib)
    return ns["coco2h36m"]


def ingest_case():
    """Keypoint ingest: the reference's coco2h36m on AlphaPose-shaped (17, 3) rows, then /1000 and the two
    normalisations of main.py:331-337."""
    f = ref_coco2h36m()
    rs = np.random.RandomState(23)
    N = 9
    kp = np.concatenate([rs.uniform(0, 1000, (N, 17, 2)), rs.uniform(0, 1, (N, 17, 1))], axis=2).astype(np.float32)
    h36m = np.stack([f(kp[i, :, :2]) for i in range(N)])                    # float64, reference arithmetic
    P = os.path.join(REF, "phase1_lifting")
    m2, s2 = np.load(os.path.join(P, "mean_train_2d.npy")), np.load(os.path.join(P, "std_train_2d.npy"))
    scaled = h36m / 1000
    np.savez(os.path.join(GOLD, "ingest.npz"), coco=kp, h36m=h36m, standardised=(scaled - m2) / s2,
             minmax=2 * scaled - 1)
    print("wrote ingest", h36m.shape)


def ref_process_data(num_of_joints=17, normalize=False):
    """main.py:271-355 (the method ``process_data``), taken by exec of the source slice: main.py itself imports cv2,
    wandb and the dataset paths.  The method reads module globals (sample, zero_centre, Normalize, num_of_joints,
    Samples) and writes / reads ``{mean,std}_train_{2,3}d.npy`` in the working directory."""
    import textwrap
    src = open(os.path.join(REF, "phase1_lifting", "main.py")).read().splitlines()
    ns = {"np": np, "sample": False, "zero_centre": True, "Normalize": normalize, "num_of_joints": num_of_joints,
          "Samples": None}
    exec(textwrap.dedent("\n".join(src[270:355])), ns)
    return ns["process_data"]


def process_data_case():
    """Dataset-side statistics and standardisation through the reference's own process_data."""
    import tempfile
    rs = np.random.RandomState(41)
    N = 257
    out = {}
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            for dim in (2, 3):
                scale = np.array([0.25, 0.2] if dim == 2 else [0.6, 0.5, 0.4])
                offs = np.array([0.5, 0.45] if dim == 2 else [1.8, -0.95, 4.7])
                raw = (offs + scale * rs.normal(0, 1, (N, 17, dim))).astype(np.float32)
                out[f"raw{dim}"] = raw
                for zc in (0, 1):
                    f = ref_process_data()
                    got = f(None, raw.astype(np.float64).copy(), is_train=True, standardize=True, z_c=bool(zc))
                    tag = f"d{dim}_zc{zc}"
                    out[tag + "_std"] = got
                    out[tag + "_mean"] = np.load(f"mean_train_{dim}d.npy")
                    out[tag + "_sd"] = np.load(f"std_train_{dim}d.npy")
                    out[tag + "_plain"] = ref_process_data()(None, raw.astype(np.float64).copy(), is_train=False,
                                                             standardize=False, z_c=bool(zc))
                # 16-joint variant (root dropped) and the 2x - 1 mode of main.py:331
                out[f"d{dim}_j16"] = ref_process_data(num_of_joints=16)(None, raw.astype(np.float64).copy(), is_train=True,
                                                                        standardize=True, z_c=True)
            out["d2_minmax"] = ref_process_data(normalize=True)(None, out["raw2"].astype(np.float64).copy(), is_train=True,
                                                               standardize=True, z_c=True)
        finally:
            os.chdir(cwd)
    np.savez(os.path.join(GOLD, "process_data.npz"), **out)
    print("wrote process_data", {k: v.shape for k, v in out.items()})


def vit_case():
    """MyViT (baselineModel.py:307-362) through the imported reference module: default configuration with the oracle's
    numpy-seeded weights, a torch-seeded default construction (pins the constructor's RNG consumption and the positional
    embedding), and a 3-block / 8-head variant."""
    from oracle import vit_oracle as V
    out = {}
    rs = np.random.RandomState(77)
    for tag, cfg, B in (("default", dict(), 5), ("b3h8", dict(n_blocks=3, n_heads=8), 3)):
        params = V.numpy_init(11, **cfg)
        m = ref_model.MyViT(**cfg)
        m.load_state_dict({k: v.clone() for k, v in params.items()}, strict=True)
        m.eval()
        x = rs.normal(0, 1, (B, 17, 2)).astype(np.float32)
        with torch.no_grad():
            y = m(torch.from_numpy(x))
        out[tag + "_x"], out[tag + "_y"] = x, y.numpy()
        # training: MSE against a random target, loss.backward() on the reference module (train_1.py:94-95)
        tgt = rs.normal(0, 1, y.shape).astype(np.float32)
        m.train()
        m.zero_grad()
        loss = torch.nn.functional.mse_loss(m(torch.from_numpy(x)), torch.from_numpy(tgt))
        loss.backward()
        names = [n for n, p_ in m.named_parameters() if p_.grad is not None]
        out[tag + "_target"], out[tag + "_loss"] = tgt, np.float32(loss.item())
        out[tag + "_grad_names"] = np.array(names)
        out[tag + "_grad_sums"] = np.array([float(dict(m.named_parameters())[n].grad.double().sum()) for n in names])
        out[tag + "_grad_abs"] = np.array([float(dict(m.named_parameters())[n].grad.double().abs().sum()) for n in names])
        out[tag + "_grad_map_w"] = m.linear_mapper.weight.grad.numpy().copy()
        out[tag + "_grad_head_w"] = m.mlp[2].weight.grad.numpy().copy()
        m.eval()
    torch.manual_seed(5)
    m = ref_model.MyViT()
    sd = m.state_dict()
    out["seeded_names"] = np.array(list(sd.keys()))
    out["seeded_sums"] = np.array([float(v.double().sum()) for v in sd.values()])
    out["seeded_abs"] = np.array([float(v.double().abs().sum()) for v in sd.values()])
    out["pos_embed"] = sd["pos_embed"].numpy()
    x = rs.normal(0, 1, (4, 17, 2)).astype(np.float32)
    with torch.no_grad():
        out["seeded_x"], out["seeded_y"] = x, m.eval()(torch.from_numpy(x)).numpy()
    np.savez(os.path.join(GOLD, "vit.npz"), **out)
    print("wrote vit", {k: v.shape for k, v in out.items()})


def ref_heatmap_methods():
    """H36_dataset.py:143-202 (xyz_to_uvw, _keypoint_to_heatmap_3D, keypoints_to_heatmap_3D) by exec of the source slice
    inside a stand-in class: the file itself imports cv2 and the dataset paths."""
    import math
    import textwrap
    src = open(os.path.join(REF, "phase3_direct/my_HybrIK/H36_dataset.py")).read().splitlines()
    body = textwrap.dedent("\n".join(src[142:202]))
    ns = {"np": np, "math": math}
    exec("class _Ref:\n" + textwrap.indent(body, "    "), ns)
    return ns["_Ref"]()


def heatmaps_case():
    """Heat-map targets through the reference's own methods; volumes stored sparsely (the 27
    non-zero voxels of each 64^3 volume)."""
    from oracle import heatmap_oracle as HO
    ref = ref_heatmap_methods()
    rs = np.random.RandomState(57)
    poses = rs.uniform(-1, 1, (6, 17, 3))
    # border cases: exactly +-1 (window clipped on one, two and three faces), half-integer voxel coordinates (np.rint
    # rounds half to even), the centre
    poses[0, 0] = [1.0, 1.0, 1.0]
    poses[0, 1] = [-1.0, -1.0, -1.0]
    poses[0, 2] = [1.0, -1.0, 0.3]
    poses[0, 3] = [0.0, 0.0, 0.0]
    poses[0, 4] = [(16.5 / 31.5) - 1, -((20.5 / 31.5) - 1), -((2.5 / 31.5) - 1)]
    poses[0, 5] = [0.999, -0.999, 0.97]
    poses32 = poses.astype(np.float32)
    out = {"poses64": poses, "poses32": poses32}
    for tag, arr in (("f64", poses), ("f32", poses32)):
        vols = np.stack([ref.keypoints_to_heatmap_3D(arr[i]) for i in range(arr.shape[0])])      # (6, 17, 64, 64, 64)
        assert vols.dtype == np.float32
        kp_i, vox_i, val = HO.to_sparse(vols.reshape(-1, 64, 64, 64))
        out[tag + "_kp"], out[tag + "_vox"], out[tag + "_val"] = kp_i, vox_i, val
    np.savez_compressed(os.path.join(GOLD, "heatmaps.npz"), **out)
    print("wrote heatmaps", {k: v.shape for k, v in out.items()})


class MaskedDropout(torch.nn.Module):
    """Stands in for ``torch.nn.Dropout(p)`` in the reference module; same arithmetic, external mask."""

    def __init__(self, p, masks, site):
        super().__init__()
        self.p, self.masks, self.site = p, masks, site

    def forward(self, x):
        if not self.training or self.p == 0:
            return x
        return x * torch.from_numpy(self.masks[self.site]).to(x.dtype) / (1.0 - self.p)


def build_reference(i_dim, o_dim, H, S, p, BN, params, buffers, masks):
    m = ref_model.LinearModel(i_dim, o_dim, linear_size=H, num_stage=S, p_dropout=p, BN=BN)
    sd = {**{k: v.clone() for k, v in params.items()}, **{k: v.clone() for k, v in buffers.items()}}
    m.load_state_dict(sd, strict=True)
    # one Dropout module object is shared by call sites; give every call site its own masked stand-in
    # input stage uses model.dropout (site 0); block i uses .dropout (site 2i+1) and .dropout2 (site 2i+2)
    m.dropout = MaskedDropout(p, masks, 0)
    for i, blk in enumerate(m.linear_stages):
        blk.dropout = MaskedDropout(p, masks, 2 * i + 1)
        blk.dropout2 = MaskedDropout(p, masks, 2 * i + 2)
    return m


def lifter_case(tag, i_dim, o_dim, H, S, B, p, BN, seed, store_weights):
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed)
    rs = np.random.RandomState(seed + 100)
    x = torch.from_numpy(rs.normal(0, 1, (B, i_dim)).astype(np.float32))
    y = torch.from_numpy(rs.normal(0, 1, (B, o_dim)).astype(np.float32))
    masks = [dropout_keep_mask(B, H, site, step=0, seed=seed + 7, p=p) for site in range(2 * S + 1)]
    out = {"x": x.numpy(), "y": y.numpy(), "cfg": np.array([i_dim, o_dim, H, S, B, int(BN), seed]), "p": np.array(p)}

    m = build_reference(i_dim, o_dim, H, S, p, BN, params, buffers, masks)
    m.eval()
    with torch.no_grad():
        out["eval_out"] = m(x).numpy()
    # eval-mode backward w.r.t. the input (phase-5 call site, backup_june8.py:137 / train_5.py:145-168)
    xe = x.clone().requires_grad_(True)
    m.zero_grad()
    torch.nn.functional.mse_loss(m(xe), y).backward()
    out["eval_dx"] = xe.grad.numpy()

    m.train()
    m.zero_grad()
    xt = x.clone().requires_grad_(True)
    yt = m(xt)
    loss = torch.nn.MSELoss(reduction="mean")(yt, y)          # train_1.py:37,94
    loss.backward()
    out["train_out"] = yt.detach().numpy()
    out["train_loss"] = loss.detach().numpy()
    out["train_dx"] = xt.grad.numpy()
    sd = m.state_dict()
    for k in buffers:
        out["buf_after/" + k] = sd[k].numpy()
    grads = {k: (v.grad.numpy() if v.grad is not None else np.zeros(v.shape, np.float32))
             for k, v in m.named_parameters()}     # BN=False leaves the BN affine without a gradient
    if store_weights:
        for k, v in params.items():
            out["param/" + k] = v.numpy()
        for k, v in buffers.items():
            out["buf/" + k] = v.numpy()
        for k, g in grads.items():
            out["grad/" + k] = g
    else:                                   # big model: digest the gradients
        for k, g in grads.items():
            out["gradsum/" + k] = np.array([g.sum(dtype=np.float64), np.abs(g).sum(dtype=np.float64)])
            out["gradhead/" + k] = g.reshape(-1)[:64].copy()
    np.savez_compressed(os.path.join(GOLD, f"lifter_{tag}.npz"), **out)
    print("wrote", tag, {k: v.shape for k, v in out.items() if not k.startswith(("param", "grad", "buf"))})


def weight_init_case():
    torch.manual_seed(5)
    m = ref_model.LinearModel(34, 51, linear_size=256, num_stage=1)
    m.apply(ref_model.weight_init)              # main.py:396
    stds = {k: float(v.std()) for k, v in m.named_parameters() if v.dim() == 2}
    np.savez(os.path.join(GOLD, "weight_init.npz"), names=np.array(list(stds)), stds=np.array(list(stds.values())),
             fan_in=np.array([dict(m.named_parameters())[k].shape[1] for k in stds]))
    print("wrote weight_init", stds)


def losses_case():
    flip = ref_flip_pose()
    rs = np.random.RandomState(11)
    B = 6
    t = lambda *s: torch.from_numpy(rs.normal(0, 1, s).astype(np.float32))
    pred2d, gt2d, proj_pred, proj_gt = t(B, 17, 2), t(B, 17, 2), t(B, 17, 2), t(B, 17, 2)
    pred3d, gt3d, lift_gt, lift_pred = t(B, 17, 3), t(B, 17, 3), t(B, 17, 3), t(B, 17, 3)
    out = dict(pred2d=pred2d, gt2d=gt2d, proj_pred=proj_pred, proj_gt=proj_gt, pred3d=pred3d, gt3d=gt3d,
               lift_gt=lift_gt, lift_pred=lift_pred)
    out = {k: v.numpy().copy() for k, v in out.items()}
    out["mpjpe_sum"] = ref_losses.loss_MPJPE(pred3d, gt3d).numpy()
    c = lambda x: x.clone()
    for proj in (False, True):
        r = ref_losses.TriangleLoss(Project=proj)(c(pred2d), c(pred3d), c(lift_gt), c(lift_pred), c(gt2d), c(gt3d),
                                                  c(proj_pred), c(proj_gt))
        out[f"triangle_{int(proj)}"] = np.array([float(v) for v in r])
    r = ref_losses.TriangleLoss_sep(Project=True)(c(pred2d), c(pred3d), c(lift_gt), c(lift_pred), c(gt2d), c(gt3d),
                                                  c(proj_pred), c(proj_gt))
    out["triangle_sep_1"] = np.array([float(v) for v in r])
    # gradients of the TriangleLoss(Project=True) total w.r.t. its differentiable inputs
    ins = [c(v).requires_grad_(True) for v in (pred2d, pred3d, lift_pred)]
    pp = c(proj_pred).requires_grad_(True)
    tot = ref_losses.TriangleLoss(Project=True)(ins[0], ins[1], c(lift_gt), ins[2], c(gt2d), c(gt3d), pp * 1.0, None)[0]
    tot.backward()
    out["triangle_1_grad_pred2d"], out["triangle_1_grad_pred3d"] = ins[0].grad.numpy(), ins[1].grad.numpy()
    out["triangle_1_grad_lift_pred"], out["triangle_1_grad_proj_pred"] = ins[2].grad.numpy(), pp.grad.numpy()
    out["flip2d"] = flip(pred2d).numpy()
    out["flip3d"] = flip(pred3d).numpy()
    # losses + grads used for training (torch modules the reference instantiates)
    for name, fn in (("mse", torch.nn.MSELoss(reduction="mean")), ("l1", torch.nn.L1Loss())):
        p = c(pred3d).requires_grad_(True)
        l = fn(p, gt3d)
        l.backward()
        out[name], out[name + "_grad"] = l.detach().numpy(), p.grad.numpy()
    # two2three.py:387-390 mean-MPJPE (cannot import that file: undefined names at import) -- same formula
    p = c(pred3d).requires_grad_(True)
    l = torch.mean(torch.norm(p - gt3d, dim=len(gt3d.shape) - 1))
    l.backward()
    out["mpjpe_mean"], out["mpjpe_mean_grad"] = l.detach().numpy(), p.grad.numpy()
    # epoch metric reduction, train_1.py:100-104
    acc = ref_losses.loss_MPJPE(pred3d, gt3d) / B
    out["epoch_mm"] = (torch.mean(acc[1:17]) * (17 / 16) * 1000).numpy()
    # de-standardisation, main.py:473-479 with the committed statistics (std is float64 in the reference)
    P = os.path.join(REF, "phase1_lifting")
    stats = {n: np.load(os.path.join(P, n + ".npy")) for n in
             ("mean_train_2d", "std_train_2d", "mean_train_3d", "std_train_3d")}
    y = torch.mul(pred3d, torch.from_numpy(stats["std_train_3d"])) + torch.from_numpy(stats["mean_train_3d"])
    out["destd3d"] = y.numpy()
    xn = np.divide(gt2d.numpy() - stats["mean_train_2d"], stats["std_train_2d"])       # main.py:334
    out["std2d"] = xn
    np.savez_compressed(os.path.join(GOLD, "losses.npz"), **out)
    np.savez(os.path.join(GOLD, "h36m_stats.npz"), **stats)
    print("wrote losses, h36m_stats")


def adam_case():
    rs = np.random.RandomState(3)
    n, steps = 257, 6
    p0 = rs.normal(0, 1, n).astype(np.float32)
    gs = rs.normal(0, 1, (steps, n)).astype(np.float32)
    out = {"p0": p0, "grads": gs}
    for tag, ctor in (("adam", lambda p: torch.optim.Adam([p], lr=1e-3)),
                      ("adam_l2", lambda p: torch.optim.Adam([p], lr=1e-3, weight_decay=1e-2)),
                      ("adamw", lambda p: torch.optim.AdamW([p], lr=1e-3))):     # train_1.py:39 defaults
        p = torch.nn.Parameter(torch.from_numpy(p0.copy()))
        opt = ctor(p)
        traj = []
        for s in range(steps):
            p.grad = torch.from_numpy(gs[s].copy())
            opt.step()
            traj.append(p.detach().numpy().copy())
        out[tag] = np.stack(traj)
    np.savez_compressed(os.path.join(GOLD, "adam.npz"), **out)
    print("wrote adam")


def train_curve_case():
    """Re-enacts the loop body train_1.py:73-100 (MSE + Adam, dropout 0.5) on the reference module."""
    i_dim, o_dim, H, S, B, p, seed, steps = 34, 51, 128, 2, 32, 0.5, 21, 40
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed, randomize_bn=False)
    masks = [None] * (2 * S + 1)
    m = build_reference(i_dim, o_dim, H, S, p, True, params, buffers, masks)
    m.train()
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    rs = np.random.RandomState(seed + 1)
    xs = rs.normal(0, 1, (steps, B, i_dim)).astype(np.float32)
    wtrue = rs.normal(0, 0.3, (i_dim, o_dim)).astype(np.float32)
    ys = np.tanh(xs @ wtrue).astype(np.float32)
    losses = []
    for s in range(steps):
        for site in range(2 * S + 1):
            masks[site] = dropout_keep_mask(B, H, site, step=s, seed=seed + 7, p=p)
        opt.zero_grad()
        loss = torch.nn.MSELoss(reduction="mean")(m(torch.from_numpy(xs[s])), torch.from_numpy(ys[s]))
        loss.backward()
        opt.step()
        losses.append(float(loss))
    np.savez_compressed(os.path.join(GOLD, "train_curve.npz"), cfg=np.array([i_dim, o_dim, H, S, B, seed, steps]),
                        p=np.array(p), wtrue=wtrue, losses=np.array(losses),
                        final_w2_bias=m.w2.bias.detach().numpy())
    print("wrote train_curve", losses[:3], losses[-3:])


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(8)
    if len(sys.argv) > 1 and sys.argv[1] == "ingest":      # only the keypoint-ingest vectors
        ingest_case()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "vit":
        vit_case()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "heatmaps":
        heatmaps_case()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "process_data":
        process_data_case()
        sys.exit(0)
    lifter_case("small", 34, 51, 64, 2, 16, 0.5, True, seed=1, store_weights=True)
    lifter_case("small_nobn", 34, 51, 64, 1, 8, 0.25, False, seed=2, store_weights=True)
    lifter_case("h1024", 34, 51, 1024, 2, 32, 0.5, True, seed=3, store_weights=False)
    lifter_case("wide", 544, 51, 256, 4, 24, 0.5, True, seed=4, store_weights=False)
    weight_init_case()
    losses_case()
    adam_case()
    train_curve_case()
    ingest_case()
    process_data_case()
    vit_case()
    heatmaps_case()
