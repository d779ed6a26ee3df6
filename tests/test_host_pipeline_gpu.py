"""The host-facing pipelines (``LifterTrainer.run_epoch`` and ``LinearModel.infer_host``) must give exactly what the
device-tensor calls give: same kernels, only the copies move to side streams."""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu


def _model(pkg, seed=3):
    torch.manual_seed(seed)
    return pkg.LinearModel(34, 51, linear_size=256, num_stage=2, p_dropout=0.5, BN=True).cuda()


@pytest.mark.parametrize("use_graph", [True, False])
def test_run_epoch_equals_step_loop(pkg, use_graph):
    trainer_mod = importlib.import_module("3d_poseestimation_b200.trainer")
    g = torch.Generator().manual_seed(0)
    batches = [(torch.randn(192, 17, 2, generator=g).pin_memory(), torch.randn(192, 17, 3, generator=g).pin_memory())
               for _ in range(7)]
    batches.append((torch.randn(64, 17, 2, generator=g), torch.randn(64, 17, 3, generator=g)))   # ragged, unpinned tail
    a, b = _model(pkg), _model(pkg)
    ta = trainer_mod.LifterTrainer(a, lr=1e-3, loss="mse", optimizer="adam", seed=5, use_graph=use_graph)
    tb = trainer_mod.LifterTrainer(b, lr=1e-3, loss="mse", optimizer="adam", seed=5, use_graph=use_graph)
    want = []
    for x, y in batches:
        ta.step(x.cuda(), y.cuda())
        want.append(ta.loss_value())
    got = tb.run_epoch(iter(batches))
    assert got.shape == (len(batches),) and not got.is_cuda
    assert torch.equal(got, torch.tensor(want, dtype=torch.float32))
    assert torch.equal(a._flat, b._flat)
    assert tb.run_epoch([]).numel() == 0
    with pytest.raises(RuntimeError):
        tb.run_epoch([(batches[0][0].cuda(), batches[0][1].cuda())])


@pytest.mark.parametrize("B,chunk", [(1, 16384), (1000, 256), (5000, 1024), (40000, 16384)])
def test_infer_host_equals_infer(pkg, B, chunk):
    m = _model(pkg).eval()
    x = torch.randn(B, 17, 2, generator=torch.Generator().manual_seed(B)).pin_memory()
    want = m.infer(x.cuda()).cpu()
    got = m.infer_host(x, chunk_rows=chunk)
    assert got.shape == (B, 51) and not got.is_cuda
    assert torch.equal(got, want)
    out = torch.empty(B, 51)
    assert m.infer_host(x, out=out, chunk_rows=chunk) is out and torch.equal(out, want)
    with pytest.raises(RuntimeError):
        m.infer_host(x.cuda())


def test_infer_host_batches_is_one_pipeline(pkg):
    m = _model(pkg).eval()
    gen = torch.Generator().manual_seed(8)
    xs = [torch.randn(b, 17, 2, generator=gen).pin_memory() for b in (700, 1, 3000, 1024)]
    want = [m.infer(x.cuda()).cpu() for x in xs]
    got = m.infer_host_batches(xs, chunk_rows=512)
    assert len(got) == 4 and all(torch.equal(g, w) for g, w in zip(got, want))
    outs = [torch.empty(x.shape[0], 51).pin_memory() for x in xs]
    again = m.infer_host_batches(iter(xs), outs=outs, chunk_rows=4096)
    assert all(a is o for a, o in zip(again, outs)) and all(torch.equal(g, w) for g, w in zip(outs, want))
    assert m.infer_host_batches([]) == []
