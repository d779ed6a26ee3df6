"""Data-parallel protocol on CPU (gloo, world_size 2): the exchange / reduce / slice logic the kernels implement,
emulated by oracle/dp_protocol.py over torch.distributed, must equal the single-process oracle on the concatenated
batch ("DP + SyncBN + gradient averaging over N equal shards == one run on the global batch")."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import adam_oracle, dp_protocol, lifter_oracle as lo
        from oracle.philox import dropout_keep_mask
        torch.set_num_threads(1)
        I, O, H, S, B, p = 34, 51, 32, 1, 24, 0.5                  # B rows per rank
        params, buffers = lo.numpy_init(I, O, H, S, seed=4)
        params = {k: v.double() for k, v in params.items()}
        rs = np.random.RandomState(1)
        xg = torch.from_numpy(rs.normal(0, 1, (world * B, I)))
        yg = torch.from_numpy(rs.normal(0, 1, (world * B, O)))
        L = 2 * S + 1
        # single process on the global batch (masks addressed by GLOBAL rows)
        gmasks = [torch.from_numpy(np.asarray(dropout_keep_mask(world * B, H, s, 0, 9, p))) for s in range(L)]
        ref = lo.lifter_forward_backward(params, {k: v.double() if v.is_floating_point() else v for k, v in buffers.items()},
                                         xg, lambda o: ((o - yg) ** 2).mean(), num_stage=S, training=True,
                                         p_dropout=p, masks=gmasks)
        # this rank's shard: the same mask stream through row_offset (csrc/philox.cuh: `row` is the global row)
        lmasks = [torch.from_numpy(np.asarray(dropout_keep_mask(B, H, s, 0, 9, p, row_offset=rank * B))) for s in range(L)]
        for s in range(L):
            assert torch.equal(lmasks[s], gmasks[s][rank * B:(rank + 1) * B]), "shard masks differ from the global stream"
        share, grads = dp_protocol.dp_forward_backward(params, xg[rank * B:(rank + 1) * B], yg[rank * B:(rank + 1) * B],
                                                       lmasks, num_stage=S, p_dropout=p, global_rows=world * B)
        loss = share.clone()
        dist.all_reduce(loss)
        assert abs(float(loss) - float(ref["loss"])) < 1e-12 * abs(float(ref["loss"])) + 1e-14
        scale = max(float(g.norm()) for g in ref["grads"].values())
        for k in params:                                           # sum of the shard gradients == global gradient
            tot = grads[k].clone()
            dist.all_reduce(tot)
            assert float((tot - ref["grads"][k]).norm()) < 1e-9 * scale, k
        # optimiser: slice-wise reduce + Adam + gather == Adam on the summed gradient
        n = sum(v.numel() for v in params.values())
        m, v = torch.zeros(n, dtype=torch.float64), torch.zeros(n, dtype=torch.float64)
        new = dp_protocol.dp_adam(params, grads, m, v, 1, 1e-3)
        for k in params:
            want, _, _ = adam_oracle.adam_step(params[k], ref["grads"][k], torch.zeros_like(params[k]),
                                               torch.zeros_like(params[k]), 1, 1e-3)
            # (a bias in front of train-mode BN: the reference's ~1e-17 rounding noise makes Adam take a full step;
            #  the protocol's exact zero does not -- both leave the loss unchanged)
            if k.endswith(".bias") and "batch_norm" not in k and not k.startswith("w2."):
                continue
            assert float((new[k] - want).abs().max()) < 1e-9, k
        ret[rank] = "ok"
    except Exception as e:                                          # noqa: BLE001
        ret[rank] = f"{type(e).__name__}: {e}"
    finally:
        dist.destroy_process_group()


def test_dp_protocol_matches_single_process_oracle():
    world = 2
    ret = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(world, 29541, ret), nprocs=world, join=True)
    assert [ret.get(r) for r in range(world)] == ["ok"] * world, dict(ret)
