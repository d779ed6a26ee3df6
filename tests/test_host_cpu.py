"""Host-side logic that needs no GPU: the C-ABI library loads and exports every declared symbol, the
module mirrors the reference's constructor / state_dict contract, and errors are the reference's."""
import ctypes as C
import importlib
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def cabi(pkg):
    return importlib.import_module("3d_poseestimation_b200._cabi")


def test_library_exports_every_declared_symbol(cabi):
    header = open(os.path.join(ROOT, "include", "lifter_b200.h")).read()
    declared = set(re.findall(r"\b(lifter_[a-z0-9_]+)\s*\(", header))
    declared -= {"lifter_config", "lifter_gemm_desc"}
    lib = cabi.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/lifter_b200.h but not exported"
        assert name in cabi.SIGNATURES, f"{name} has no ctypes signature"
    assert lib.lifter_version() >= 1


def test_layout_queries_match_reference_parameter_count(cabi):
    cfg = cabi.LifterConfig(34, 51, 1024, 2, 1, 0.5, 1e-5, 0.1)
    lib = cabi.lib()
    n = lib.lifter_num_param_tensors(C.byref(cfg))
    assert n == 22                                              # 22 parameter tensors
    numel = sum(lib.lifter_param_numel(C.byref(cfg), t) for t in range(n))
    assert numel == 4296755                                     # 4,296,755 parameters
    offs = [lib.lifter_param_offset(C.byref(cfg), t) for t in range(n)]
    assert all(o % 32 == 0 for o in offs) and offs == sorted(offs)
    wide = cabi.LifterConfig(16 * 34, 51, 2048, 4, 1, 0.5, 1e-5, 0.1)
    assert sum(lib.lifter_param_numel(C.byref(wide), t) for t in range(lib.lifter_num_param_tensors(C.byref(wide)))) == 34828339


def test_unsupported_shapes_raise_value_error(pkg):
    with pytest.raises(ValueError, match="linear_size"):
        pkg.LinearModel(34, 51, linear_size=1000)
    with pytest.raises(ValueError, match="o_dim"):
        pkg.LinearModel(34, 200)
    with pytest.raises(ValueError, match="dropout probability"):
        pkg.LinearModel(34, 51, p_dropout=1.5)


def test_state_dict_contract(pkg):
    m = pkg.LinearModel(34, 51, linear_size=64, num_stage=2, p_dropout=0.5, BN=True)
    ref = torch.nn.ModuleDict()         # expected keys, SURVEY §8b
    keys = ["w1.weight", "w1.bias"] + [f"batch_norm1.{k}" for k in ("weight", "bias", "running_mean", "running_var", "num_batches_tracked")]
    for i in range(2):
        for lin, bn in (("w1", "batch_norm1"), ("w2", "batch_norm2")):
            keys += [f"linear_stages.{i}.{lin}.weight", f"linear_stages.{i}.{lin}.bias"]
            keys += [f"linear_stages.{i}.{bn}.{k}" for k in ("weight", "bias", "running_mean", "running_var", "num_batches_tracked")]
    keys += ["w2.weight", "w2.bias"]
    sd = m.state_dict()
    assert sorted(sd.keys()) == sorted(keys)
    assert sd["w1.weight"].shape == (64, 34) and sd["w2.weight"].shape == (51, 64)
    assert sd["batch_norm1.num_batches_tracked"].dtype == torch.int64 and sd["batch_norm1.num_batches_tracked"].shape == ()
    assert all(v.dtype == torch.float32 for k, v in sd.items() if "num_batches" not in k)
    # parameters are leaf fp32 nn.Parameters in the reference's order, views of one flat buffer
    names = [n for n, _ in m.named_parameters()]
    assert names[:4] == ["w1.weight", "w1.bias", "batch_norm1.weight", "batch_norm1.bias"] and names[-2:] == ["w2.weight", "w2.bias"]
    assert all(p.is_leaf and p.requires_grad for p in m.parameters())
    assert m._flat_is_intact()
    # load_state_dict round trip keeps the flat views
    sd2 = {k: torch.randn_like(v) if v.dtype == torch.float32 else v + 3 for k, v in sd.items()}
    m.load_state_dict(sd2)
    assert m._flat_is_intact()
    for k, v in m.state_dict().items():
        assert torch.equal(v, sd2[k])
    # checkpoint envelope of train_1.py:186 round-trips through torch.save
    import io
    buf = io.BytesIO()
    torch.save({"epoch": 3, "batch_size": 64, "model": m.state_dict(), "optimizer": {}}, buf)
    buf.seek(0)
    m2 = pkg.LinearModel(34, 51, linear_size=64, num_stage=2)
    m2.load_state_dict(torch.load(buf)["model"])
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), m2.state_dict().values()))


def test_same_seed_same_init_as_torch_modules(pkg):
    """Constructor consumes the torch RNG exactly like the reference (Linear, BN, blocks, Linear)."""
    torch.manual_seed(1)
    m = pkg.LinearModel(34, 51, linear_size=64, num_stage=1)
    torch.manual_seed(1)
    w1 = torch.nn.Linear(34, 64)
    blk = [torch.nn.Linear(64, 64), torch.nn.Linear(64, 64)]
    w2 = torch.nn.Linear(64, 51)
    assert torch.equal(m.w1.weight, w1.weight) and torch.equal(m.linear_stages[0].w2.bias, blk[1].bias)
    assert torch.equal(m.w2.weight, w2.weight)


def test_weight_init_apply(pkg):
    torch.manual_seed(0)
    m = pkg.LinearModel(34, 51, linear_size=256, num_stage=1)
    b0 = m.w1.bias.clone()
    m.apply(pkg.weight_init)                       # main.py:396
    assert m._flat_is_intact()
    assert abs(float(m.linear_stages[0].w1.weight.std()) - np.sqrt(2 / 256)) < 0.05 * np.sqrt(2 / 256)
    assert torch.equal(m.w1.bias, b0)              # biases untouched


def test_top_level_import_name(pkg):
    """``sys.path.append(<dir>); from baselineModel import LinearModel`` (train_1.py:15-17)."""
    d = os.path.join(ROOT, "3d_poseestimation_b200")
    sys.path.insert(0, d)
    try:
        sys.modules.pop("baselineModel", None)
        bm = importlib.import_module("baselineModel")
        assert bm.LinearModel is pkg.LinearModel and bm.weight_init is pkg.weight_init
    finally:
        sys.path.remove(d)
        sys.modules.pop("baselineModel", None)


def test_cpu_forward_raises_no_fallback(pkg):
    m = pkg.LinearModel(34, 51, linear_size=64, num_stage=1).eval()
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.zeros(2, 17, 2))


def test_fused_transformer_trainer_has_no_cpu_path(pkg):
    """VitTrainer (the fused MyViT step) refuses a model that is not on a GPU, and argument errors come before any CUDA call."""
    import importlib
    VT = importlib.import_module("3d_poseestimation_b200.vit_trainer").VitTrainer
    with pytest.raises(RuntimeError, match="no CPU path"):
        VT(pkg.MyViT())
    with pytest.raises(TypeError):
        VT(pkg.LinearModel(34, 51, linear_size=64, num_stage=1))
    with pytest.raises(ValueError, match="out_d"):
        VT(pkg.MyViT(chw=(1, 17, 3), out_d=2))
    with pytest.raises(ValueError, match="optimizer"):
        VT(pkg.MyViT(), optimizer="sgd")


def test_product_does_not_import_oracle():
    pkgdir = os.path.join(ROOT, "3d_poseestimation_b200")
    for dirpath, _, files in os.walk(pkgdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f"{f} imports the oracle"


def test_myvit_constructor_mirrors_the_reference(pkg, golden_dir):
    """baselineModel.py:307-342: same state_dict keys / shapes, same torch-RNG consumption (seed 5 => the reference's
    initial weights, pinned by per-tensor checksums from the imported reference), same positional embedding; no CPU
    path and no silent PyTorch fallback for training."""
    import numpy as np
    g = np.load(os.path.join(golden_dir, "vit.npz"))
    torch.manual_seed(5)
    m = pkg.MyViT()
    sd = m.state_dict()
    assert list(sd.keys()) == [str(k) for k in g["seeded_names"]]
    np.testing.assert_allclose([float(v.double().sum()) for v in sd.values()], g["seeded_sums"], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose([float(v.double().abs().sum()) for v in sd.values()], g["seeded_abs"], rtol=1e-9)
    np.testing.assert_allclose(sd["pos_embed"].numpy(), g["pos_embed"], atol=1e-7)
    assert not m.pos_embed.requires_grad and sum(p.numel() for p in m.parameters()) == sum(v.numel() for v in sd.values())
    with pytest.raises(RuntimeError):
        m(torch.zeros(2, 17, 2))                                   # CPU tensor
    with pytest.raises(ValueError):
        pkg.MyViT(hidden_d=100)
    from importlib import import_module
    bm = import_module("3d_poseestimation_b200.baselineModel")
    assert bm.MyViT is pkg.MyViT and callable(bm.get_positional_embeddings)
