"""GPU parity of the dataset-side keypoint processing against the reference's own
``process_data`` (golden vectors, ``phase1_lifting/main.py:271-355``) and the numpy oracle at larger sizes.
Tolerances: the kernels stream in fp32 (non-tensor fp64 is a few operations per clock per SM on B200) and fold groups
of eight shifted values into fp64 accumulators, so the mean agrees to ~1e-6 of the column's standard deviation and the
standard deviation to 1e-5 relative -- both below the quantisation of the fp32 key points themselves.  Processed key
points are fp32 against the reference's float64 result: 2e-6 (absolute, in units of the standard deviation)."""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import dataset_oracle as D

pytestmark = pytest.mark.gpu
T = torch.from_numpy


@pytest.fixture(scope="module")
def P():
    return importlib.import_module("3d_poseestimation_b200.pose_losses")


@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("zc", [0, 1])
def test_process_data_matches_reference_golden(P, golden_dir, dim, zc, tmp_path):
    g = dict(np.load(os.path.join(golden_dir, "process_data.npz")))
    raw, tag = T(g[f"raw{dim}"]).cuda(), f"d{dim}_zc{zc}"
    out, stats = P.process_data(raw, is_train=True, standardize=True, z_c=bool(zc), save_dir=str(tmp_path))
    st = stats.cpu().numpy()
    np.testing.assert_allclose(st[0], g[tag + "_mean"], rtol=1e-7, atol=1e-6 * g[tag + "_sd"].max())
    np.testing.assert_allclose(st[1], g[tag + "_sd"], rtol=1e-5, atol=1e-12)
    np.testing.assert_allclose(out.cpu().numpy(), g[tag + "_std"], rtol=2e-6, atol=2e-6)
    # the files the reference writes (main.py:291-299), same names / dtype / shape
    saved = np.load(os.path.join(tmp_path, f"mean_train_{dim}d.npy"))
    assert saved.dtype == np.float64 and saved.shape == (17, dim)
    np.testing.assert_allclose(saved, g[tag + "_mean"], rtol=1e-7, atol=1e-6 * g[tag + "_sd"].max())
    # is_train=False, standardize=False: centring + root zeroing only; statistics passed as a (mean, std) pair
    plain, _ = P.process_data(raw, stats=(g[tag + "_mean"], g[tag + "_sd"]), is_train=False, standardize=False, z_c=bool(zc))
    np.testing.assert_allclose(plain.cpu().numpy(), g[tag + "_plain"], rtol=1e-6, atol=1e-7)
    again, _ = P.process_data(raw, stats=stats, is_train=False, standardize=True, z_c=bool(zc))
    assert torch.equal(again, out)
    if zc:
        j16, _ = P.process_data(raw, is_train=True, standardize=True, z_c=True, num_of_joints=16)
        assert tuple(j16.shape) == (raw.shape[0], 16, dim)
        np.testing.assert_allclose(j16.cpu().numpy(), g[f"d{dim}_j16"], rtol=2e-6, atol=2e-6)
        assert float(out[:, 0].abs().max()) == 0.0
        if dim == 2:
            mm, _ = P.process_data(raw, is_train=True, standardize=True, z_c=True, minmax=True)
            np.testing.assert_allclose(mm.cpu().numpy(), g["d2_minmax"], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("frames,J,dim", [(1, 17, 3), (2, 17, 2), (6, 17, 3), (40, 17, 3), (4099, 17, 2), (390_001, 17, 3), (1000, 21, 3), (333, 1, 2)])
def test_keypoint_stats_vs_oracle_ragged_sizes(P, frames, J, dim):
    """One frame, fewer frames than row groups, sizes that are no multiple of any block shape, the H36M scale
    (~390 k frames), the largest frame width."""
    rs = np.random.RandomState(frames % 1000 + dim)
    raw = (rs.normal(0, 1, (frames, J, dim)) * rs.uniform(0.1, 2, (J, dim)) + rs.uniform(-5, 5, (J, dim))).astype(np.float32)
    for zc in (False, True):
        st = P.keypoint_stats(T(raw).cuda(), z_c=zc).cpu().numpy()
        mean, std, lo, hi = D.keypoint_stats(raw, zc)
        np.testing.assert_allclose(st[0], mean, rtol=1e-7, atol=1e-6 * max(std.max(), 1e-3))
        np.testing.assert_allclose(st[1], std, rtol=1e-5, atol=1e-10)
        # min / max: exact, up to the fp32 rounding of the centred difference (rounding is monotone)
        assert np.array_equal(st[2], lo.astype(np.float32)) and np.array_equal(st[3], hi.astype(np.float32))
        out, _ = P.process_data(T(raw).cuda(), is_train=True, standardize=frames > 1, z_c=zc)
        exp, _, _ = D.process_data(raw, standardize=frames > 1, z_c=zc)
        np.testing.assert_allclose(out.cpu().numpy(), exp, rtol=2e-5, atol=2e-5)


def test_keypoint_stats_is_deterministic_and_shift_invariant(P):
    """Same bits run after run (fixed-order partial sums); a large common offset does not cost precision (the sums
    are taken around the first frame)."""
    rs = np.random.RandomState(5)
    raw = (1000.0 + 1e-2 * rs.normal(0, 1, (50_000, 17, 3))).astype(np.float32)
    x = T(raw).cuda()
    a, b = P.keypoint_stats(x, z_c=False), P.keypoint_stats(x, z_c=False)
    assert torch.equal(a, b)
    _, std, _, _ = D.keypoint_stats(raw, False)
    np.testing.assert_allclose(a[1].cpu().numpy(), std, rtol=1e-5)


def test_keypoint_stats_unaligned_input_takes_the_plain_load_kernel(P):
    """The statistics pass stages frames with bulk copies, which need a 16-byte aligned source; a view that starts 4
    bytes into an allocation goes through the plain-load kernel and must give the same numbers."""
    rs = np.random.RandomState(9)
    raw = rs.normal(0.3, 0.7, (5000, 17, 3)).astype(np.float32)
    base = torch.empty(raw.size + 1, device="cuda")
    x = base[1:].view(5000, 17, 3)
    x.copy_(T(raw))
    assert x.data_ptr() % 16 != 0 and x.is_contiguous()
    a = P.keypoint_stats(x).cpu().numpy()
    b = P.keypoint_stats(T(raw).cuda()).cpu().numpy()
    pa, _ = P.process_data(x, is_train=True, standardize=True)
    pb, _ = P.process_data(T(raw).cuda(), is_train=True, standardize=True)
    np.testing.assert_allclose(pa.cpu().numpy(), pb.cpu().numpy(), rtol=0, atol=2e-6)
    np.testing.assert_allclose(pb.cpu().numpy(), D.process_data(raw, standardize=True)[0], rtol=2e-5, atol=2e-5)
    mean, std, lo, hi = D.keypoint_stats(raw)
    for st in (a, b):
        np.testing.assert_allclose(st[0], mean, rtol=1e-7, atol=1e-6 * std.max())
        np.testing.assert_allclose(st[1], std, rtol=1e-5)
        assert np.array_equal(st[2], lo.astype(np.float32)) and np.array_equal(st[3], hi.astype(np.float32))


def test_dataset_processing_errors(P):
    with pytest.raises(RuntimeError):
        P.keypoint_stats(torch.zeros(4, 17, 3))                                  # CPU tensor: no fallback
    with pytest.raises(ValueError):
        P.keypoint_stats(torch.zeros(0, 17, 3, device="cuda"))
    with pytest.raises(ValueError):
        P.keypoint_stats(torch.zeros(4, 33, 3, device="cuda"))                  # 99 values per frame
    with pytest.raises(ValueError):
        P.process_data(torch.zeros(4, 17, 3, device="cuda"), is_train=False, standardize=True)
    out, _ = P.process_data(torch.zeros(0, 17, 3, device="cuda"), is_train=False, standardize=False)
    assert tuple(out.shape) == (0, 17, 3)
