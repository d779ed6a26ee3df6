"""GPU parity of the tcgen05 GEMM (csrc/gemm_sm100.cu) through the C ABI (lifter_gemm), against a
float64 matmul of the same 16-bit-rounded operands.  Covers every operand major, output mode and
epilogue option the forward / dgrad / wgrad calls use."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cabi(pkg):
    import importlib
    c = importlib.import_module("3d_poseestimation_b200._cabi")
    assert torch.cuda.is_available(), "GPU tests need a B200"
    c.check(c.lib().lifter_device_check())
    return c


def _dt(fmt):
    return torch.float16 if fmt == 0 else torch.bfloat16


def run_gemm(c, A, B, M, N, K, a_mn, b_mn, fmt, out_mode, D, ldd, n_valid=0, splits=1, relu=0, scale=None,
             shift=None, residual=None, stat=None, cluster=0):
    d = c.GemmDesc()
    d.A, d.lda, d.B, d.ldb = A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0)
    d.M, d.N, d.K, d.a_mn_major, d.b_mn_major, d.fmt = M, N, K, a_mn, b_mn, fmt
    d.out_mode, d.D, d.ldd, d.n_valid, d.splits, d.relu = out_mode, D.data_ptr(), ldd, n_valid, splits, relu
    d.cluster = cluster
    d.scale = scale.data_ptr() if scale is not None else None
    d.shift = shift.data_ptr() if shift is not None else None
    if residual is not None:
        d.residual, d.ldr = residual.data_ptr(), residual.stride(0)
    if stat is not None:
        d.stat_sum, d.stat_sq = stat[0].data_ptr(), stat[1].data_ptr()
    c.check(c.lib().lifter_gemm(C.byref(d), c.stream_ptr()))
    torch.cuda.synchronize()


def _operands(M, N, K, a_mn, b_mn, fmt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = (torch.randn(M, K, generator=g, device="cuda") * 0.5).to(_dt(fmt))
    b = (torch.randn(N, K, generator=g, device="cuda") * 0.5).to(_dt(fmt))
    A = a.t().contiguous() if a_mn else a          # MN-major: stored [K][M]
    B = b.t().contiguous() if b_mn else b
    ref = a.double() @ b.double().t()
    return A, B, ref


@pytest.mark.parametrize("cluster", [1, 2])
@pytest.mark.parametrize("fmt", [0, 1])
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (128, 256, 64), (256, 256, 256), (384, 1024, 1024), (640, 1024, 256),
                                   (1000, 128, 192), (77, 512, 128), (9472, 1024, 1024)])
def test_kmajor_16bit_output(cabi, M, N, K, fmt, cluster):
    A, B, ref = _operands(M, N, K, 0, 0, fmt, 1)
    D = torch.full((M, N), float("nan"), device="cuda").to(_dt(fmt))
    run_gemm(cabi, A, B, M, N, K, 0, 0, fmt, cabi.OUT_16, D, N, cluster=cluster)
    err = (D.double() - ref).abs().max().item()
    tol = (2 ** -10 if fmt == 0 else 2 ** -7) * ref.abs().max().item() + 1e-3
    assert torch.isfinite(D.float()).all(), "unwritten / non-finite outputs"
    assert err <= tol, f"max err {err} > {tol}"


@pytest.mark.parametrize("cluster", [1, 2])
@pytest.mark.parametrize("a_mn,b_mn", [(0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 256, 128), (1024, 1024, 512), (200, 320, 100), (200, 320, 104)])
def test_mn_major_operands(cabi, M, N, K, a_mn, b_mn, cluster):
    fmt = 1
    if (a_mn and M % 8) or (b_mn and N % 8) or ((not a_mn or not b_mn) and K % 8):
        pytest.skip("row stride must be a multiple of 16 bytes")
    A, B, ref = _operands(M, N, K, a_mn, b_mn, fmt, 2)
    D = torch.full((M, N), float("nan"), device="cuda").to(_dt(fmt))
    run_gemm(cabi, A, B, M, N, K, a_mn, b_mn, fmt, cabi.OUT_16, D, N, cluster=cluster)
    err = (D.double() - ref).abs().max().item()
    tol = 2 ** -7 * ref.abs().max().item() + 1e-3
    assert torch.isfinite(D.float()).all()
    assert err <= tol, f"max err {err} > {tol}"


@pytest.mark.parametrize("splits", [1, 3, 4, 16])
def test_split_k_fp32_slabs_wgrad_shape(cabi, splits):
    """dW[N_out][K_in] = dZ^T X: both operands MN-major, contraction over the batch, split-K slabs."""
    M, N, K, fmt = 1024, 1024, 2048 + 40, 1
    A, B, ref = _operands(M, N, K, 1, 1, fmt, 3)
    eff = cabi.lib().lifter_gemm_effective_splits(K, splits)
    slab = cabi.lib().lifter_gemm_slab_rows(M)
    D = torch.full((eff * slab, N), float("nan"), device="cuda")
    run_gemm(cabi, A, B, M, N, K, 1, 1, fmt, cabi.OUT_32_SLABS, D, N, splits=splits)
    got = D.view(eff, slab, N)[:, :M].double().sum(0)
    err = (got - ref).abs().max().item()
    assert err <= 1e-4 * ref.abs().max().item() + 1e-3, f"max err {err}"


def test_small_m_slab_padding(cabi):
    """wgrad of the output layer: M = 64 valid rows inside a 128-row slab."""
    M, N, K, fmt = 64, 1024, 700, 1
    A, B, ref = _operands(M, N, K, 1, 1, fmt, 4)
    eff = cabi.lib().lifter_gemm_effective_splits(K, 4)
    D = torch.zeros(eff * 128, N, device="cuda")
    run_gemm(cabi, A, B, M, N, K, 1, 1, fmt, cabi.OUT_32_SLABS, D, N, splits=4)
    got = D.view(eff, 128, N)[:, :M].double().sum(0)
    assert (got - ref).abs().max().item() <= 1e-4 * ref.abs().max().item() + 1e-3


@pytest.mark.parametrize("M,n_valid,K", [(128, 51, 1024), (1000, 51, 1024), (333, 34, 1024), (64, 64, 64)])
def test_direct_fp32_output_with_affine(cabi, M, n_valid, K):
    fmt = 0
    N = 64
    A, B, ref = _operands(M, N, K, 0, 0, fmt, 5)
    scale = torch.rand(N, device="cuda") + 0.5
    shift = torch.randn(N, device="cuda")
    D = torch.full((M, n_valid), float("nan"), device="cuda")
    guard = torch.full((M + 4, n_valid), 7.0, device="cuda")      # detect writes past the last row
    run_gemm(cabi, A, B, M, N, K, 0, 0, fmt, cabi.OUT_32_DIRECT, guard, n_valid, n_valid=n_valid, scale=scale, shift=shift)
    exp = (ref * scale.double() + shift.double())[:, :n_valid]
    assert (guard[:M].double() - exp).abs().max().item() <= 1e-3 * exp.abs().max().item() + 1e-3
    assert (guard[M:] == 7.0).all()


def test_epilogue_affine_relu_residual(cabi):
    M, N, K, fmt = 300, 1024, 1024, 0
    A, B, ref = _operands(M, N, K, 0, 0, fmt, 6)
    scale = torch.rand(N, device="cuda") + 0.5
    shift = torch.randn(N, device="cuda")
    res = torch.randn(M, N, device="cuda").half()
    D = torch.zeros(M, N, device="cuda").half()
    run_gemm(cabi, A, B, M, N, K, 0, 0, fmt, cabi.OUT_16, D, N, relu=1, scale=scale, shift=shift, residual=res)
    exp = torch.relu(ref * scale.double() + shift.double()) + res.double()
    err = (D.double() - exp).abs().max().item()
    assert err <= 2 ** -10 * exp.abs().max().item() + 2e-3, f"max err {err}"


@pytest.mark.parametrize("M", [128, 8192, 1000])
def test_column_statistics(cabi, M):
    N, K, fmt = 1024, 1024, 1
    A, B, ref = _operands(M, N, K, 0, 0, fmt, 7)
    P = cabi.lib().lifter_gemm_stat_rows(M)
    ssum = torch.zeros(P, N, device="cuda")
    ssq = torch.zeros(P, N, device="cuda")
    D = torch.zeros(M, N, device="cuda").bfloat16()
    run_gemm(cabi, A, B, M, N, K, 0, 0, fmt, cabi.OUT_16, D, N, stat=(ssum, ssq))
    z = D.double()                      # statistics are those of the stored (rounded) values
    np.testing.assert_allclose(ssum.double().sum(0).cpu().numpy(), z.sum(0).cpu().numpy(), rtol=1e-4, atol=1e-2)
    np.testing.assert_allclose(ssq.double().sum(0).cpu().numpy(), (z * z).sum(0).cpu().numpy(), rtol=1e-4, atol=1e-2)
    # and per 32-row slab
    rows = min(M, 32)
    np.testing.assert_allclose(ssum[0].double().cpu().numpy(), z[:rows].sum(0).cpu().numpy(), rtol=1e-4, atol=1e-3)


def test_bad_arguments_raise(cabi):
    A = torch.zeros(128, 64, device="cuda").half()
    D = torch.zeros(128, 64, device="cuda").half()
    with pytest.raises(ValueError):
        run_gemm(cabi, A, A, 0, 64, 64, 0, 0, 0, cabi.OUT_16, D, 64)
    with pytest.raises(ValueError):
        run_gemm(cabi, A, A, 128, 64, 64, 0, 0, 0, cabi.OUT_16, D, 64, splits=2)
