"""GPU parity of the volumetric heat-map targets (phase3_direct/my_HybrIK/H36_dataset.py:143-202)
against the oracle, which is pinned bit for bit to volumes produced by the reference's own methods."""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import heatmap_oracle as HO

pytestmark = pytest.mark.gpu


def _pl():
    return importlib.import_module("3d_poseestimation_b200.pose_losses")


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_heatmaps_match_reference_golden(golden_dir, dtype):
    g = dict(np.load(os.path.join(golden_dir, "heatmaps.npz")))
    tag, key = ("f64", "poses64") if dtype == np.float64 else ("f32", "poses32")
    vol = _pl().keypoints_to_heatmap_3D(torch.from_numpy(g[key]).cuda())
    assert vol.shape == (6, 17, 64, 64, 64) and vol.dtype == torch.float32
    kp_i, vox_i, val = HO.to_sparse(vol.cpu().numpy().reshape(-1, 64, 64, 64))
    # the same voxels are non-zero (clipping at the borders, np.rint's half-to-even) ...
    assert np.array_equal(kp_i, g[tag + "_kp"]) and np.array_equal(vox_i, g[tag + "_vox"])
    # ... and hold the reference's values to the last float32 bit but one (device exp vs numpy's, both fp64)
    np.testing.assert_allclose(val, g[tag + "_val"], rtol=1.5e-7, atol=0)


@pytest.mark.parametrize("n,sigma", [(1, 0.5), (17, 0.5), (403, 0.5), (51, 1.75), (5, 1.0)])
def test_heatmaps_match_oracle(n, sigma):
    rs = np.random.RandomState(n)
    kp = rs.uniform(-1, 1, (n, 3))
    kp[0] = [1.0, -1.0, 1.0]
    got = _pl().keypoints_to_heatmap_3D(torch.from_numpy(kp).cuda(), sigma=sigma).cpu().numpy()
    ref = HO.keypoints_to_heatmap_3d(kp, sigma=sigma)
    assert np.array_equal(got != 0, ref != 0)
    np.testing.assert_allclose(got, ref, rtol=1.5e-7, atol=0)
    assert int((got != 0).reshape(n, -1).sum(1).max()) <= HO.window(sigma) ** 3


def test_heatmaps_shapes_range_check_and_numpy_round_trip():
    pl = _pl()
    kp = torch.zeros(2, 17, 3, device="cuda")
    out = torch.full((2, 17, 64, 64, 64), 7.0, device="cuda")          # every voxel is overwritten, not accumulated
    assert pl.keypoints_to_heatmap_3D(kp, out=out) is out
    assert float(out.sum()) > 0 and float(out.max()) <= 1.0 and int((out != 0).sum()) == 2 * 17 * 27
    assert pl.keypoints_to_heatmap_3D(torch.zeros(0, 3, device="cuda")).shape == (0, 64, 64, 64)
    with pytest.raises(AssertionError):
        pl.keypoints_to_heatmap_3D(torch.tensor([[0.0, 1.01, 0.0]], device="cuda"))
    bad = pl.keypoints_to_heatmap_3D(torch.tensor([[0.0, 3.0, 0.0]], device="cuda"), check=False)   # window outside: zeros
    assert float(bad.abs().sum()) == 0.0
    a = np.random.RandomState(0).uniform(-1, 1, (17, 3))
    h = pl.keypoints_to_heatmap_3D(a)                                   # the dataset's per-frame call (numpy in / out)
    assert isinstance(h, np.ndarray) and h.shape == (17, 64, 64, 64)
    np.testing.assert_allclose(h, HO.keypoints_to_heatmap_3d(a), rtol=1.5e-7)
    with pytest.raises(RuntimeError):
        pl.keypoints_to_heatmap_3D(torch.zeros(3, 3))
