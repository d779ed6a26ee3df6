"""The rounding model used by the GPU training tests reduces to the oracle when nothing is rounded."""
import os

import numpy as np
import pytest
import torch

import rounding_model as rm
from oracle import lifter_oracle as lo
from oracle.philox import dropout_keep_mask


@pytest.mark.parametrize("tag", ["small", "small_nobn", "wide"])
def test_fp32_rounding_model_equals_oracle(golden_dir, tag):
    g = dict(np.load(os.path.join(golden_dir, f"lifter_{tag}.npz")))
    i_dim, o_dim, H, S, B, BN, seed = [int(v) for v in g["cfg"]]
    p = float(g["p"])
    params, buffers = lo.numpy_init(i_dim, o_dim, H, S, seed)
    masks = [dropout_keep_mask(B, H, s, 0, seed + 7, p) for s in range(2 * S + 1)]
    X, Y = torch.from_numpy(g["x"]), torch.from_numpy(g["y"])
    ref = lo.lifter_forward_backward(params, buffers, X, lambda o: ((o - Y) ** 2).mean(), num_stage=S, training=True,
                                     use_bn=bool(BN), p_dropout=p, masks=masks, need_input_grad=True)
    mod = rm.forward_backward(params, buffers, X, Y, masks, num_stage=S, use_bn=bool(BN), p=p, dtype=torch.float32)
    rel = lambda a, b: float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
    assert rel(mod["out"], ref["out"]) < 1e-5 and rel(mod["dx"], ref["dx"]) < 1e-5
    for k, v in ref["grads"].items():
        if float(v.norm()) > 1e-5:
            assert rel(mod["grads"][k], v) < 1e-5, k
