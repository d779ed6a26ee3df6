"""Training step of the MyViT lifter as ONE replayable kernel sequence.

The reference drives MyViT exactly as it drives LinearModel: ``optimizer.zero_grad(); loss = MSELoss(model(x), y);
loss.backward(); optimizer.step()`` (``phase1_lifting/train_1.py:39,91-96`` with ``baselineModel.py:307-362`` as the model).
``MyViT.forward`` under autograd already runs the fused forward / backward kernels, but leaves the loss, the 34 small
per-parameter optimiser updates and the 16-bit operand re-packing to framework kernels.  ``VitTrainer`` removes those:

  * every trainable parameter of the module is RE-HOMED as a view into one flat fp32 buffer laid out like the kernel
    descriptor (``lifter_vit_t``): small tables first, then the six GEMM weight regions;
  * ``lifter_vit_backward`` writes its gradients into a second flat buffer of the same layout;
  * one ``lifter_adam_step_table`` launch updates all parameters AND refreshes the 16-bit operand copies of the six weight
    regions in the same pass (no ``at::`` kernel, no packing on the next forward);
  * ``lifter_loss`` gives MSE + d loss / d prediction in one launch; the step's scalars live in the device step state, so
    the whole sequence is captured once per batch size as a CUDA graph and replayed.

No CPU path: the constructor raises without the sm_100a library or off a CUDA device.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _cabi
from .vit import MyViT


def _round(n, k=64):
    return (n + k - 1) // k * k


class VitTrainer:
    """``VitTrainer(model, lr=1e-4, optimizer="adam")`` then ``loss = trainer.step(x, y)`` per batch.

    ``x``: (B, joints, in_d) key points, ``y``: (B, joints, 3) targets, both fp32 on the model's device.  ``step`` returns
    the device scalar of the MSE loss (``float(loss)`` synchronises).  ``world_size > 1`` (one process per GPU of one NVSwitch
    box, ``torch.distributed`` initialised with NCCL for the plumbing): every rank steps on its own B rows, the loss is the
    mean over all ``B * world_size`` rows (``step`` returns this rank's share, ``loss_value()`` the sum), and the optimiser
    launch becomes the fused NVLink reduce-scatter / Adam / all-gather kernel of the lifter (``lifter_dp_adam_step_table``):
    the model has no BatchNorm, so that kernel is the step's only exchange.  ``optimizer``: "adam" (L2 ``weight_decay``, default 0,
    ``torch.optim.Adam``) or "adamw" (decoupled, default 0.01, ``torch.optim.AdamW``); ``set_lr`` is what
    ``ReduceLROnPlateau`` would drive (train_1.py:41,106)."""

    MAX_RESIDENT_BATCH_SIZES = 3

    def __init__(self, model: MyViT, lr=1e-4, optimizer="adam", betas=(0.9, 0.999), eps=1e-8, weight_decay=None,
                 use_graph=True, world_size=1, rank=0, process_group=None):
        if not isinstance(model, MyViT):
            raise TypeError("VitTrainer drives the MyViT lifter")
        if model.out_d != 3:
            raise ValueError("VitTrainer: the fused loss kernel works on 3-D joints (out_d == 3)")
        if optimizer not in ("adam", "adamw"):
            raise ValueError(f"VitTrainer: optimizer must be 'adam' or 'adamw', got {optimizer!r}")
        dev = model.linear_mapper.weight.device
        if dev.type != "cuda":
            raise RuntimeError("VitTrainer: the sm_100a build has no CPU path; move the model to a B200 first")
        _cabi.lib()                                   # raises when the CUDA library is missing
        self.model, self.device = model, dev
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.decoupled = optimizer == "adamw"
        self.weight_decay = float((0.01 if self.decoupled else 0.0) if weight_decay is None else weight_decay)
        self.use_graph = bool(use_graph)
        self.step_count = 0
        self.joints = model.chw[1]
        self.world, self.rank, self.pg = int(world_size), int(rank), process_group
        if self.world > 1:
            import torch.distributed as dist
            if not dist.is_initialized():
                raise RuntimeError("VitTrainer(world_size > 1) needs torch.distributed (backend 'nccl') to be initialised")
            if self.world > _cabi.DP_MAX_WORLD:
                raise RuntimeError(f"VitTrainer: the NVLink exchange serves at most {_cabi.DP_MAX_WORLD} ranks (one NVSwitch box)")
            if self.pg is None:
                self.pg = dist.group.WORLD
        self._dp = None
        with torch.cuda.device(dev):
            self._flatten()
            self.state = torch.zeros(8, dtype=torch.int32, device=dev)
            self.state.view(torch.float32)[2] = self.lr
            self.loss_out = torch.zeros(1 + self.joints, dtype=torch.float32, device=dev)
        self._bufs, self._graphs, self._order, self._warm = {}, {}, [], set()
        self._launches, self._state_stale = 0, True

    # ------------------------------------------------------------------------------------------------ flat layout
    def _regions(self):
        m = self.model
        H, nb, ind = m.hidden_d, m.n_block, m.input_d
        small = [("map_w", H * ind), ("map_b", H), ("ln", nb * 6 * H), ("mlp_b1", nb * 4 * H), ("mlp_b2", nb * H),
                 ("head_b1", H // 2), ("head_b2", 64)]
        weights = [("w_qkv", nb * 3 * H * H), ("w_out", nb * H * H), ("w_mlp1", nb * 4 * H * H), ("w_mlp2", nb * 4 * H * H),
                   ("w_head1", (H // 2) * H), ("w_head2", 64 * (H // 2))]
        return small, weights

    def _flatten(self):
        m, dev = self.model, self.device
        H, nb, od = m.hidden_d, m.n_block, m.out_d
        small, weights = self._regions()
        off, o = {}, 0
        for k, n in small + weights:
            off[k] = (o, n)
            o += _round(n)
        self.n = o
        self.off = off
        self.w0 = off["w_qkv"][0]                                   # the weight regions are the tail [w0, n)
        n_op = self.n - self.w0
        if self.world > 1:
            flat, self.grad, sym_ops = self._symmetric_buffers(self.n, n_op)
        else:
            flat, self.grad, sym_ops = torch.zeros(self.n, dtype=torch.float32, device=dev), torch.zeros(self.n, dtype=torch.float32, device=dev), None
        self.flat = flat
        self.m_buf, self.v_buf = torch.zeros(self.n, dtype=torch.float32, device=dev), torch.zeros(self.n, dtype=torch.float32, device=dev)
        reg = lambda k, *shape: flat[off[k][0]:off[k][0] + off[k][1]].view(*shape)

        def rehome(p, view):
            with torch.no_grad():
                view.copy_(p.detach().to(device=dev, dtype=torch.float32))
            p.data = view

        rehome(m.linear_mapper.weight, reg("map_w", H, m.input_d))
        rehome(m.linear_mapper.bias, reg("map_b", H))
        ln, b1, b2 = reg("ln", nb, 6, H) if nb else None, reg("mlp_b1", nb, 4 * H) if nb else None, reg("mlp_b2", nb, H) if nb else None
        wq = reg("w_qkv", nb, 3 * H, H) if nb else None
        wo = reg("w_out", nb, H, H) if nb else None
        w1 = reg("w_mlp1", nb, 4 * H, H) if nb else None
        w2 = reg("w_mlp2", nb, H, 4 * H) if nb else None
        for b, blk in enumerate(m.blocks):
            for j, p in enumerate((blk.norm1.weight, blk.norm1.bias, blk.mhsa.norm.weight, blk.mhsa.norm.bias,
                                   blk.norm2.weight, blk.norm2.bias)):
                rehome(p, ln[b, j])
            rehome(blk.mlp[0].bias, b1[b])
            rehome(blk.mlp[2].bias, b2[b])
            rehome(blk.mhsa.to_qkv.weight, wq[b])
            rehome(blk.mhsa.to_out.weight, wo[b])
            rehome(blk.mlp[0].weight, w1[b])
            rehome(blk.mlp[2].weight, w2[b])
        rehome(m.mlp[0].bias, reg("head_b1", H // 2))
        rehome(m.mlp[2].bias, reg("head_b2", 64)[:od])
        rehome(m.mlp[0].weight, reg("w_head1", H // 2, H))
        rehome(m.mlp[2].weight, reg("w_head2", 64, H // 2)[:od])
        self.pos = m.pos_embed.detach().to(device=dev, dtype=torch.float32).contiguous()
        # 16-bit GEMM operands of the weight tail, refreshed by the optimiser pass from here on
        self.fmt = 0 if m.train_format == "fp16" else 1
        dt = torch.float16 if self.fmt == 0 else torch.bfloat16
        if self.world > 1:
            # every replica starts from rank 0's parameters (what DistributedDataParallel does at construction)
            import torch.distributed as dist
            dist.broadcast(flat, src=dist.get_global_rank(self.pg, 0), group=self.pg)
            self.operands = sym_ops.view(dt)
            self.operands.copy_(flat[self.w0:])
            torch.cuda.synchronize(dev)
            dist.barrier(group=self.pg)
        else:
            self.operands = flat[self.w0:].to(dt)
        self.gmap = torch.zeros((1 + m.input_d) * H, dtype=torch.float32, device=dev)
        d = _cabi.VitDesc()
        d.tokens, d.in_dim, d.hidden, d.heads, d.blocks = m.chw[1], m.input_d, H, m.n_heads, nb
        d.mlp_dim, d.head_dim, d.out_dim = 4 * H, H // 2, od
        d.fmt, d.ln_eps = self.fmt, 1e-5
        for k, _ in small:
            setattr(d, k, flat.data_ptr() + 4 * off[k][0] if off[k][1] else None)
        d.pos = self.pos.data_ptr()
        for k, _ in weights:
            setattr(d, k, self.operands.data_ptr() + 2 * (off[k][0] - self.w0) if off[k][1] else None)
        self.desc = d
        G = _cabi.VitGrads()
        G.map = self.gmap.data_ptr()
        for k, _ in small[2:] + weights:
            setattr(G, k, self.grad.data_ptr() + 4 * off[k][0] if off[k][1] else None)
        self.G = G
        t = _cabi.AdamTable()
        live = [(k, n) for k, n in weights if n]
        t.n = len(live)
        for i, (k, n) in enumerate(live):
            t.p_off[i], t.numel[i], t.op_off[i] = off[k][0], n, off[k][0] - self.w0
        self.table = t
        m._packed.clear()

    def _symmetric_buffers(self, n, n_op):
        """Gradient, parameter and operand buffers in ONE NVLink symmetric allocation (peer-mapped on every rank), plus the
        flag block of the fused reduce-scatter / Adam / all-gather kernel (``lifter_dp_adam_step_table``)."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        dev = self.device

        def up(v):
            return (v + 255) // 256 * 256
        off_grad = _cabi.DP_FLAG_BYTES
        off_par = off_grad + up(4 * n)
        off_ops = off_par + up(4 * n)
        total = off_ops + up(2 * n_op)
        try:
            buf = symm.empty(total, dtype=torch.uint8, device=dev)
            buf.zero_()
            hdl = symm.rendezvous(buf, self.pg)
        except Exception as e:                                   # noqa: BLE001
            raise RuntimeError(f"VitTrainer(world_size={self.world}): NVLink symmetric memory is unavailable ({e}); there is "
                               "no NCCL fallback for this trainer") from e
        torch.cuda.synchronize(dev)
        dist.barrier(group=self.pg)                              # every rank's buffer is zeroed before first use
        dp = _cabi.DpInfo()
        dp.world, dp.rank = self.world, self.rank
        for p in range(self.world):
            base = int(hdl.buffer_ptrs[p])
            dp.flags[p], dp.grad[p], dp.params[p], dp.operands[p] = base, base + off_grad, base + off_par, base + off_ops
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)          # NVSwitch multicast view of the same buffer (0: none)
        if mc:
            dp.mc_grad, dp.mc_params, dp.mc_operands = mc + off_grad, mc + off_par, mc + off_ops
        self._sym, self._sym_hdl, self._dp = buf, hdl, dp
        return (buf[off_par:off_par + 4 * n].view(torch.float32), buf[off_grad:off_grad + 4 * n].view(torch.float32),
                buf[off_ops:off_ops + 2 * n_op])

    def refresh_operands(self):
        """Call after writing parameters from the host side (``load_state_dict``, manual edits)."""
        self.operands.copy_(self.flat[self.w0:])
        self.model._packed.clear()

    # ------------------------------------------------------------------------------------------------ step
    def _buffers(self, B):
        b = self._bufs.get(B)
        if b is not None:
            self._order.remove(B)
            self._order.append(B)
            return b
        while len(self._order) >= self.MAX_RESIDENT_BATCH_SIZES:
            old = self._order.pop(0)
            self._bufs.pop(old, None)
            self._graphs.pop(old, None)
            self._warm.discard(old)
        m, lib, dev = self.model, _cabi.lib(), self.device
        J = self.joints
        u8 = lambda n: torch.empty(max(int(n), 1), dtype=torch.uint8, device=dev)
        saved = lib.lifter_vit_saved_bytes(C.byref(self.desc), B)
        scratch = lib.lifter_vit_bwd_scratch_bytes(C.byref(self.desc), B)
        if saved < 0 or scratch < 0:
            raise ValueError("VitTrainer: " + _cabi.last_error())
        b = dict(x=torch.empty((B, J, m.input_d), dtype=torch.float32, device=dev),
                 y=torch.empty((B, J, 3), dtype=torch.float32, device=dev),
                 out=torch.empty((B, J, 3), dtype=torch.float32, device=dev),
                 dpred=torch.empty((B, J, 3), dtype=torch.float32, device=dev),
                 saved=u8(saved), scratch=u8(scratch),
                 lscratch=u8(_cabi.check(lib.lifter_loss_scratch_bytes(B, J))))
        self._bufs[B] = b
        self._order.append(B)
        return b

    def _enqueue(self, b, B, state):
        lib, st, d = _cabi.lib(), _cabi.stream_ptr(), C.byref(self.desc)
        m = self.model
        wd = self.weight_decay
        if state is not None:
            _cabi.check(lib.lifter_step_advance(_cabi.ptr(state), self.betas[0], self.betas[1], wd if self.decoupled else 0.0, st))
        _cabi.check(lib.lifter_vit_train_forward(d, _cabi.ptr(b["x"]), B, _cabi.ptr(b["saved"]), _cabi.ptr(b["out"]), st))
        _cabi.check(lib.lifter_loss(_cabi.ptr(b["out"]), _cabi.ptr(b["y"]), B, self.joints, _cabi.LOSS_MSE, B * self.world, None,
                                    _cabi.ptr(b["dpred"]), _cabi.ptr(b["lscratch"]), _cabi.ptr(self.loss_out), 0, st))
        _cabi.check(lib.lifter_vit_backward(d, _cabi.ptr(b["x"]), B, _cabi.ptr(b["saved"]), _cabi.ptr(b["dpred"]),
                                            _cabi.ptr(b["scratch"]), C.byref(self.G), None, st))
        _cabi.check(lib.lifter_vit_map_grad(_cabi.ptr(self.gmap), m.input_d, m.hidden_d,
                                            C.c_void_p(self.grad.data_ptr() + 4 * self.off["map_w"][0]),
                                            C.c_void_p(self.grad.data_ptr() + 4 * self.off["map_b"][0]), st))
        if self._dp is not None:
            # data parallel: rank r sums slice r of every rank's gradient over NVLink, runs Adam on it and stores the new
            # parameters and operand copies into every rank's buffers -- one kernel, no collective call
            _cabi.check(lib.lifter_dp_adam_step_table(C.byref(self._dp), self.n, _cabi.ptr(self.m_buf), _cabi.ptr(self.v_buf),
                                                      self.betas[0], self.betas[1], self.eps, wd, int(self.decoupled), 1.0,
                                                      _cabi.ptr(state), self.fmt, C.byref(self.table), st))
            return
        _cabi.check(lib.lifter_adam_step_table(_cabi.ptr(self.flat), _cabi.ptr(self.grad), _cabi.ptr(self.m_buf),
                                               _cabi.ptr(self.v_buf), self.n, self.lr, self.betas[0], self.betas[1], self.eps, wd,
                                               int(self.decoupled), self.step_count + 1, 1.0,
                                               _cabi.ptr(state) if state is not None else None, _cabi.ptr(self.operands),
                                               self.fmt, C.byref(self.table), st))

    def _check(self, x, y):
        m = self.model
        for t, last, name in ((x, m.input_d, "x"), (y, 3, "y")):
            if not torch.is_tensor(t) or t.device != self.device or t.dtype != torch.float32:
                raise ValueError(f"VitTrainer.step: {name} must be an fp32 tensor on {self.device}")
            if t.dim() != 3 or t.shape[1] != self.joints or t.shape[2] != last:
                raise ValueError(f"VitTrainer.step: {name} must be (B, {self.joints}, {last}), got {tuple(t.shape)}")
        if x.shape[0] != y.shape[0]:
            raise ValueError("VitTrainer.step: x and y disagree on the batch size")

    @_cabi.device_guard
    def step(self, x, y):
        """One optimiser step on the batch; returns the device scalar of the loss."""
        self._check(x, y)
        B = int(x.shape[0])
        if B == 0:
            return self.loss_out[0]
        if self.model.linear_mapper.weight.data_ptr() != self.flat.data_ptr() + 4 * self.off["map_w"][0]:
            raise RuntimeError("VitTrainer: the model's parameters were moved (model.to / load with assign) after the trainer "
                               "re-homed them; build a new VitTrainer")
        b = self._buffers(B)
        b["x"].copy_(x, non_blocking=True)
        b["y"].copy_(y, non_blocking=True)
        if not self.use_graph or B not in self._warm:
            # the first step at a batch size runs eagerly (plans, kernel attributes); it is a real step.  Data parallel: the
            # exchange kernel takes its epoch from the device step state, so the eager step advances that state too
            lib = _cabi.lib()
            n0 = lib.lifter_launch_count()
            if self._dp is not None:
                if self._state_stale:
                    self._sync_state()
                self._enqueue(b, B, self.state)
            else:
                self._enqueue(b, B, None)
                self._state_stale = True
            self._launches = int(lib.lifter_launch_count() - n0)
            self._warm.add(B)
        else:
            g = self._graphs.get(B)
            if g is None:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._enqueue(b, B, self.state)
                self._graphs[B] = g
            if self._state_stale:
                self._sync_state()
            g.replay()
        self.step_count += 1
        self.model._param_epoch = getattr(self.model, "_param_epoch", 0) + 1
        return self.loss_out[0]

    def run_epoch(self, batches):
        """The epoch loop of ``train_1.py:73-100`` over HOST batches ``(x, y)`` (fp32 CPU tensors, pinned for the upload to
        be asynchronous): per step an H2D copy of the batch, the step, and the loss into pinned host memory; ONE wait at the
        end (the reference blocks on ``loss.cpu().item()`` every step, train_1.py:98).  Returns the losses (CPU tensor)."""
        batches = list(batches)
        if not batches:
            return torch.empty(0)
        out = torch.empty(len(batches), dtype=torch.float32).pin_memory()
        with torch.cuda.device(self.device):
            for i, (x, y) in enumerate(batches):
                if x.is_cuda or y.is_cuda:
                    raise RuntimeError("run_epoch takes host batches; call step() with device tensors")
                xd = x.to(self.device, non_blocking=True)
                yd = y.to(self.device, non_blocking=True)
                out[i:i + 1].copy_(self.step(xd, yd).reshape(1), non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            if self.world > 1:                                   # each rank holds its share of the global mean
                import torch.distributed as dist
                t = out.to(self.device)
                dist.all_reduce(t, group=self.pg)
                return t.cpu()
        return out.clone()

    def _sync_state(self):
        self.state[1] = self.step_count
        self.state.view(torch.float32)[2] = self.lr
        self._state_stale = False

    def set_lr(self, lr):
        self.lr = float(lr)
        self.state.view(torch.float32)[2] = self.lr

    def loss_value(self):
        """Host read of the last step's loss (synchronises); data parallel: the sum of the ranks' shares = the global mean."""
        if self.world > 1:
            import torch.distributed as dist
            t = self.loss_out[:1].clone()
            dist.all_reduce(t, group=self.pg)
            return float(t.item())
        return float(self.loss_out[0].item())

    def graph_launches(self):
        """Kernels of one step (the library's launch counter over the eager first step; the graph replays the same list)."""
        return self._launches

    def state_dict(self):
        """Optimiser state.  Data parallel: rank r owns (and updates) slice r of ``m`` / ``v`` only -- save and load it per rank,
        with the same world size."""
        return {"step": self.step_count, "lr": self.lr, "m": self.m_buf, "v": self.v_buf, "world": self.world, "rank": self.rank}

    def load_state_dict(self, sd):
        if int(sd.get("world", self.world)) != self.world or int(sd.get("rank", self.rank)) != self.rank:
            raise ValueError("VitTrainer.load_state_dict: the optimiser state was saved by another rank / world size")
        self.step_count, self.lr = int(sd["step"]), float(sd["lr"])
        self.m_buf.copy_(sd["m"])
        self.v_buf.copy_(sd["v"])
        self._sync_state()
