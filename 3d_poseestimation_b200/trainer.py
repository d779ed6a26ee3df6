"""Fused training step of the lifter: the loop body of ``phase1_lifting/train_1.py:73-100``
(``zero_grad -> forward -> loss -> backward -> optimizer.step -> metric``) as one sequence of sm_100a
kernels with no host synchronisation: forward with saved activations, fused loss + gradient + MPJPE
metric, backward, (data parallel: NCCL all-reduce of the flat gradient and of the BatchNorm column sums),
fused flat-buffer Adam/AdamW, refresh of the 16-bit operand copies.

The reference reads ``loss.cpu().item()`` every step (train_1.py:98); here the loss stays on the device and
``loss_value()`` reads it when the caller wants it.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _autograd, _cabi

_LOSS = {"mse": _cabi.LOSS_MSE, "l1": _cabi.LOSS_L1, "mpjpe": _cabi.LOSS_MPJPE}


class LifterTrainer:
    """``LifterTrainer(model, lr=1e-4, loss="mse", optimizer="adamw")``.

    optimizer: "adam" (torch.optim.Adam: weight_decay is L2) or "adamw" (decoupled, default 0.01 as at
    train_1.py:39).  ``process_group`` enables data parallelism: every rank calls ``step`` with its shard of
    the global batch; gradients are summed over ranks (the loss is a mean over the GLOBAL batch) and
    BatchNorm uses global batch statistics.
    """

    def __init__(self, model, lr=1e-4, loss="mse", optimizer="adamw", betas=(0.9, 0.999), eps=1e-8,
                 weight_decay=None, seed=0, process_group=None, world_size=None, rank=None, metric_std=None,
                 use_graph=True, require_symmetric=False):
        if loss not in _LOSS:
            raise ValueError(f"unknown loss {loss!r}: use 'mse', 'l1' or 'mpjpe'")
        if optimizer not in ("adam", "adamw"):
            raise ValueError(f"unknown optimizer {optimizer!r}: use 'adam' or 'adamw'")
        if model.output_size % 3 != 0:
            raise ValueError("the fused loss works on (J, 3) joint outputs: o_dim must be a multiple of 3")
        if not model._flat.is_cuda:
            raise RuntimeError("LifterTrainer needs the model on a B200 (no CPU path)")
        self.model = model
        self.lr, self.betas, self.eps = float(lr), betas, float(eps)
        self.decoupled = optimizer == "adamw"
        self.weight_decay = float(weight_decay if weight_decay is not None else (0.01 if self.decoupled else 0.0))
        self.loss_kind = _LOSS[loss]
        self.joints = model.output_size // 3
        self.step_count = 0
        model.dropout_seed = int(seed)
        self.pg = process_group
        if process_group is None and world_size is not None and world_size > 1:
            import torch.distributed as dist
            self.pg = dist.group.WORLD
        model.process_group = self.pg
        if self.pg is not None:
            import torch.distributed as dist
            self.world, self.rank = dist.get_world_size(self.pg), dist.get_rank(self.pg)
        else:
            self.world, self.rank = 1, 0
        dev = model._flat.device
        self.m = torch.zeros_like(model._flat)
        self.v = torch.zeros_like(model._flat)
        self.grad = torch.zeros_like(model._flat)
        self.loss_out = torch.zeros(1 + self.joints, dtype=torch.float32, device=dev)     # [loss, metric per joint]
        self.metric_std = None if metric_std is None else torch.as_tensor(
            metric_std, dtype=torch.float32, device=dev).reshape(-1).contiguous()
        self._bufs = {}
        # device step state (lifter_step_state_t): lets ONE captured CUDA graph serve every step
        self.state = torch.zeros(8, dtype=torch.int32, device=dev)
        self.state.view(torch.float32)[2] = self.lr
        self.use_graph = bool(use_graph)
        # data parallel: capture the NCCL gradient all-reduce into the step graph too (LIFTER_DP_GRAPH=0 turns it off)
        self.graph_collectives = os.environ.get("LIFTER_DP_GRAPH", "1") != "0"
        # per-layer optimiser ranges inside backward (LIFTER_FUSE_OPTIM=0: one Adam launch after backward)
        # Data parallel, measured (profiles/r02_dp_breakdown_dp{2,8}.jsonl): with ONE peer a single closing reduce-scatter /
        # Adam / all-gather launch beats the ranges (0.535 vs 0.547 ms per step), with seven peers the ranges win (0.572 vs
        # 0.610 ms): the default is the closing launch for a world of exactly two.
        self.fuse_optimizer = os.environ.get("LIFTER_FUSE_OPTIM", "0" if self.world == 2 else "1") != "0"
        self._state_stale = False
        self._sym = None
        self._dp_adam = False
        self._pipe = None
        self._launches = None                    # kernel launches of one step (counted by the library at capture)
        self.require_symmetric = bool(require_symmetric)
        if self.world > 1:
            with torch.cuda.device(dev):
                self._setup_symmetric(dev)
        self._graphs = {}

    # ------------------------------------------------------------------ data parallel: NVLink symmetric memory
    def _setup_symmetric(self, dev):
        """Symmetric (peer-mapped) buffer for the in-kernel SyncBN exchange; falls back to NCCL all-reduces of the
        BatchNorm sums if symmetric memory is unavailable (more than 8 ranks, no P2P)."""
        import torch.distributed as dist
        model = self.model
        model._dp = None
        # every replica starts from rank 0's parameters and BatchNorm buffers (what DistributedDataParallel does at
        # construction): ranks built from different seeds would otherwise run the first step on diverging weights
        with torch.no_grad():
            for t in (model._flat, model._rm, model._rv, model._nbt):
                dist.broadcast(t, src=dist.get_global_rank(self.pg, 0), group=self.pg)
        model._param_epoch += 1
        model._stats_epoch += 1
        if self.world > _cabi.DP_MAX_WORLD:
            if self.require_symmetric:
                raise RuntimeError(f"data parallel over {self.world} ranks: the in-kernel NVLink exchange serves at most "
                                   f"{_cabi.DP_MAX_WORLD} (one NVSwitch box)")
            return
        lib, cfg = _cabi.lib(), C.byref(model._cfg)
        fmt = model._fmt_code(model.train_format)
        n = model._flat.numel()

        def up(v):
            return (v + 255) // 256 * 256
        sync_b = up(_cabi.check(lib.lifter_dp_sync_bytes(cfg, self.world)) * self._bn_calls_per_step())
        op_b = up(lib.lifter_operand_bytes(cfg))
        off_flags = sync_b
        off_grad = off_flags + _cabi.DP_FLAG_BYTES
        off_par = off_grad + up(4 * n)
        off_ops = off_par + up(4 * n)
        total = off_ops + op_b
        try:
            import torch.distributed._symmetric_memory as symm
            buf = symm.empty(total, dtype=torch.uint8, device=dev)
            buf.zero_()
            hdl = symm.rendezvous(buf, self.pg)
            torch.cuda.synchronize()
            dist.barrier(group=self.pg)                          # every rank's buffer is zeroed before first use
        except Exception as e:                                   # noqa: BLE001 -- plumbing fallback, kernels unchanged
            if self.require_symmetric:
                raise RuntimeError(f"NVLink symmetric memory is unavailable ({e}) and require_symmetric=True: refusing "
                                   "to fall back to NCCL all-reduces under the same configuration name") from e
            import warnings
            warnings.warn(f"symmetric memory unavailable ({e}); BatchNorm sums and gradients go through NCCL")
            return
        dp = _cabi.DpInfo()
        dp.world, dp.rank = self.world, self.rank
        for p in range(self.world):
            base = int(hdl.buffer_ptrs[p])
            dp.sync[p] = base
            dp.flags[p] = base + off_flags
            dp.grad[p] = base + off_grad
            dp.params[p] = base + off_par
            dp.operands[p] = base + off_ops
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)             # NVSwitch multicast view of the same buffer (0: none)
        if mc:
            dp.mc_grad, dp.mc_params, dp.mc_operands = mc + off_grad, mc + off_par, mc + off_ops
        self._sym, self._sym_hdl = buf, hdl
        # gradients, parameters and the training operand copy live in the symmetric buffer from now on
        self.grad = buf[off_grad:off_grad + 4 * n].view(torch.float32)
        model._adopt_storage(buf[off_par:off_par + 4 * n].view(torch.float32), buf[off_ops:off_ops + op_b], fmt)
        torch.cuda.synchronize()
        dist.barrier(group=self.pg)
        model._dp = dp
        self._dp_adam = os.environ.get("LIFTER_DP_NCCL_GRADS") != "1"      # (env: fall back to NCCL all-reduce + Adam)

    def _bn_calls_per_step(self):
        """Lifter forwards per optimiser step: each needs its own slots in the SyncBN exchange buffer."""
        return 1

    # ------------------------------------------------------------------ buffers sized by the batch
    MAX_RESIDENT_BATCH_SIZES = 3

    def _alloc_buffers(self, B, dev):
        lib, cfg = _cabi.lib(), C.byref(self.model._cfg)
        return dict(
            saved=torch.empty(_cabi.check(lib.lifter_train_saved_bytes(cfg, B)), dtype=torch.uint8, device=dev),
            scratch=torch.empty(_cabi.check(lib.lifter_train_scratch_bytes(cfg, B)), dtype=torch.uint8, device=dev),
            out=torch.empty(B, self.model.output_size, dtype=torch.float32, device=dev),
            dpred=torch.empty(B, self.model.output_size, dtype=torch.float32, device=dev),
            lscratch=torch.empty(_cabi.check(lib.lifter_loss_scratch_bytes(B, self.joints)), dtype=torch.uint8,
                                 device=dev))

    def _buffers(self, B, dev):
        """Workspaces of batch size B.  The last few batch sizes stay resident (a trailing partial batch does not evict
        the full-size buffers); a captured graph holds the ADDRESSES of its buffers, so graphs are dropped together with
        the buffers they were captured on."""
        b = self._bufs.pop(B, None)
        if b is None:
            b = self._alloc_buffers(B, dev)
            while len(self._bufs) >= self.MAX_RESIDENT_BATCH_SIZES:
                old = next(iter(self._bufs))                      # least recently used
                torch.cuda.synchronize(dev)                       # nothing in flight may still use it
                del self._bufs[old]
                for key in [k for k in self._graphs if k[0] == old]:
                    del self._graphs[key]
        self._bufs[B] = b                                         # most recently used last
        return b

    def graph_launches(self):
        """Kernel launches of one training step as the library counted them while the step was enqueued (captured)."""
        return self._launches

    # ------------------------------------------------------------------ one optimiser step
    def _enqueue(self, inputs, b, accumulate_metric, state):
        """The kernel sequence of one step on the current stream.  ``state`` (device step state) set: every
        step-dependent scalar is read from device memory, so the sequence can be captured once and replayed."""
        x, y = inputs
        model = self.model
        lib, cfg = _cabi.lib(), C.byref(model._cfg)
        B = x.shape[0]
        st = _cabi.stream_ptr()
        fmt = model._fmt_code(model.train_format)
        if state is not None:
            _cabi.check(lib.lifter_step_advance(_cabi.ptr(state), self.betas[0], self.betas[1],
                                                self.weight_decay if self.decoupled else 0.0, st))
        # forward (saved activations), dropout stream addressed by the optimiser step
        out, _, t = _autograd.run_forward(model, x, True, saved=b["saved"], out=b["out"], step=self.step_count,
                                          step_state=state)
        # fused loss + dL/dpred + per-joint MPJPE sums; the mean is over the GLOBAL batch
        _cabi.check(lib.lifter_loss(_cabi.ptr(out), _cabi.ptr(y), B, self.joints, self.loss_kind, B * self.world,
                                    _cabi.ptr(self.metric_std), _cabi.ptr(b["dpred"]), _cabi.ptr(b["lscratch"]),
                                    _cabi.ptr(self.loss_out), int(accumulate_metric), st))
        self._optimise(t, b["dpred"], b["scratch"], state, fused_backward=True)

    def _optimise(self, t, dpred, scratch, state, fused_backward):
        """Backward of the forward described by ``t`` + the optimiser step.  ``fused_backward``: the backward pass has not
        run yet and may carry the per-layer optimiser ranges; otherwise ``self.grad`` already holds the gradient."""
        model = self.model
        lib, cfg = _cabi.lib(), C.byref(model._cfg)
        st = _cabi.stream_ptr()
        fmt = model._fmt_code(model.train_format)
        ops = model._operands(fmt)
        if fused_backward:
            if state is not None and self.fuse_optimizer and (self.world == 1 or self._dp_adam):
                # backward with the optimiser folded in: each layer's Adam range (data parallel: its NVLink reduce-scatter
                # + Adam + all-gather) runs on a side stream as soon as that layer's gradients exist
                opt = _cabi.Optim(model._flat.data_ptr(), ops.data_ptr(), self.m.data_ptr(), self.v.data_ptr(),
                                  self.betas[0], self.betas[1], self.eps, self.weight_decay, int(self.decoupled), 1.0)
                _autograd.run_backward(model, t, dpred, self.grad, None, scratch=scratch, optim=opt)
                return
            _autograd.run_backward(model, t, dpred, self.grad, None, scratch=scratch)
        if self._dp_adam and state is not None:
            # data parallel: reduce-scatter of the gradients + Adam + all-gather of the parameters in ONE kernel over
            # NVLink peer memory (sum of shard gradients = gradient of the global batch)
            _cabi.check(lib.lifter_dp_adam_step(cfg, C.byref(model._dp), _cabi.ptr(self.m), _cabi.ptr(self.v),
                                                self.betas[0], self.betas[1], self.eps, self.weight_decay,
                                                int(self.decoupled), 1.0, _cabi.ptr(state), fmt, st))
            return
        if self.world > 1:
            import torch.distributed as dist                      # (use_graph=False, or no symmetric memory)
            dist.all_reduce(self.grad, group=self.pg)             # sum of shard gradients = global-batch gradient
        # Adam / AdamW over the flat buffer; the same pass rewrites the 16-bit operand copies of the weights
        _cabi.check(lib.lifter_adam_step_packed(cfg, _cabi.ptr(model._flat), _cabi.ptr(self.grad), _cabi.ptr(self.m),
                                                _cabi.ptr(self.v), self.lr, self.betas[0], self.betas[1], self.eps,
                                                self.weight_decay, int(self.decoupled), self.step_count + 1, 1.0,
                                                _cabi.ptr(state), _cabi.ptr(ops), fmt, st))

    def _after_step(self):
        model = self.model
        self.step_count += 1
        model._param_epoch += 1                                   # the flat parameters changed under torch's feet
        model._operands_refreshed(model._fmt_code(model.train_format))

    def _check_inputs(self, x, y):
        model = self.model
        x = model._check_input(x)
        y = y.reshape(y.shape[0], -1)
        if y.shape != (x.shape[0], model.output_size) or y.dtype != torch.float32 or not y.is_cuda:
            raise RuntimeError(f"target must be a float32 CUDA tensor of {x.shape[0]} x {model.output_size} values")
        return x, y.contiguous()

    @_cabi.device_guard
    def step(self, *tensors, accumulate_metric=False, _loss_copy=None):
        """x: (B, 17, 2) or (B, i_dim) fp32 CUDA, y: (B, 17, 3) or (B, o_dim).  Returns the device loss scalar.
        (``_loss_copy``: a one-element device tensor that also receives the loss -- ``run_epoch``'s per-slot mailbox.)"""
        model = self.model
        inputs = self._check_inputs(*tensors)
        B = inputs[0].shape[0]
        model.train()
        if self.use_graph and B >= 2 and (self.world == 1 or self.graph_collectives):
            return self._step_graph(inputs, bool(accumulate_metric), _loss_copy)
        loss = self._step_eager(inputs, B, accumulate_metric)
        if _loss_copy is not None:
            _loss_copy.copy_(loss.reshape(1))
        return loss

    def _counted(self, fn):
        before = _cabi.lib().lifter_launch_count()
        fn()
        self._launches = int(_cabi.lib().lifter_launch_count() - before)

    def _step_eager(self, inputs, B, accumulate_metric):
        b = self._buffers(B, inputs[0].device)
        if self.world > 1 and self.use_graph and B >= 2:
            # data parallel: same kernel sequence driven by the device step state (it is also the epoch of the
            # in-kernel BatchNorm exchange); launched eagerly because of the NCCL gradient all-reduce in the middle
            if self._state_stale:
                self._sync_state()
            self._counted(lambda: self._enqueue(inputs, b, accumulate_metric, self.state))
            self._after_step()
            return self.loss_out[0]
        self._counted(lambda: self._enqueue(inputs, b, accumulate_metric, None))
        self._after_step()
        self._state_stale = True                                  # the device step state did not see this step
        return self.loss_out[0]

    def _capture(self, inputs, B, accumulate_metric, loss_copy=None):
        model = self.model
        b = self._buffers(B, inputs[0].device)
        model._operands(model._fmt_code(model.train_format))              # packed before the capture starts
        self._sync_state()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())

        def body():
            self._enqueue(inputs, b, accumulate_metric, self.state)
            if loss_copy is not None:
                loss_copy.copy_(self.loss_out[:1])
        with torch.cuda.graph(graph, stream=side):
            self._counted(body)
        return graph

    def _step_graph(self, inputs, accumulate_metric, loss_copy=None):
        """Replay of the captured step: no per-kernel host work.  One graph per batch size reads its inputs from static
        buffers the batch is copied into; input tensors that come back at the same addresses (a prefetch slot of
        ``run_epoch``, a resident benchmark batch) get a graph of their own that reads them in place."""
        model = self.model
        B = inputs[0].shape[0]
        key = (B, accumulate_metric, model._flat.data_ptr())
        g = self._graphs.get(key)
        if g is None:
            static = tuple(torch.empty_like(t) for t in inputs)
            g = {"generic": (self._capture(static, B, accumulate_metric), static), "in_place": {}, "seen": {}}
            self._graphs[key] = g
        else:
            self._buffers(B, inputs[0].device)                            # LRU touch
        if self._state_stale:
            self._sync_state()
        # the graph reads the 16-bit operand copy: re-validate it against the fp32 parameters (load_state_dict,
        # .apply(weight_init) or an in-place edit between steps re-packs into the same buffer; otherwise a version check)
        model._operands(model._fmt_code(model.train_format))
        ptrs = tuple(t.data_ptr() for t in inputs) + (0 if loss_copy is None else loss_copy.data_ptr(),)
        graph = g["in_place"].get(ptrs)
        if graph is None:
            seen = g["seen"]
            seen[ptrs] = seen.get(ptrs, 0) + 1
            if seen[ptrs] >= 3 and len(g["in_place"]) < 4:                # the same buffers for the third time
                graph = g["in_place"][ptrs] = self._capture(inputs, B, accumulate_metric, loss_copy)
            elif len(seen) > 256:
                seen.clear()
        if graph is None:
            graph, static = g["generic"]
            for dst, src in zip(static, inputs):
                dst.copy_(src, non_blocking=True)
            graph.replay()
            if loss_copy is not None:
                loss_copy.copy_(self.loss_out[:1])
        else:
            graph.replay()
        model._stats_epoch += 1
        self._after_step()
        return self.loss_out[0]

    @_cabi.device_guard
    def run_epoch(self, batches, accumulate_metric=False):
        """The epoch loop of ``train_1.py:73-100`` over an iterable of HOST batches ``(x, y)`` (fp32 CPU tensors;
        pinned, as a ``DataLoader(pin_memory=True)`` yields them, for the upload to be asynchronous).

        The upload of batch i+1 runs on a copy stream while step i computes; each step's loss is read back into
        pinned host memory without stalling the queue (the reference blocks on ``loss.cpu().item()`` every step,
        train_1.py:98).  Returns the per-step losses as a float32 CPU tensor (data parallel: summed over ranks, like
        ``loss_value``).  Data parallel: the ranks meet at a host barrier first, so that an epoch never starts against
        a peer that is still checkpointing or validating (the in-kernel exchange waits, but not for ever).
        """
        model = self.model
        dev = model._flat.device
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.pg)
        cur = torch.cuda.current_stream(dev)
        pipe = self._pipe
        if pipe is None:
            pipe = {"copy": torch.cuda.Stream(dev), "down": torch.cuda.Stream(dev), "slots": {},
                    "ready": [torch.cuda.Event() for _ in range(2)], "free": [torch.cuda.Event() for _ in range(2)],
                    "read": [torch.cuda.Event() for _ in range(2)]}
            self._pipe = pipe
        copy, down = pipe["copy"], pipe["down"]
        copy.wait_stream(cur)
        down.wait_stream(cur)
        losses, n = None, 0
        for i, host in enumerate(batches):
            if any(t.is_cuda for t in host):
                raise RuntimeError("run_epoch takes host batches; call step() with device tensors")
            host = [t.reshape(t.shape[0], -1) for t in host]
            s = i & 1
            key = (s,) + tuple((tuple(t.shape), t.dtype) for t in host)
            slot = pipe["slots"].get(s)
            if slot is None or slot[0] != key:
                if slot is not None:
                    torch.cuda.synchronize(dev)                    # a batch-size change: let the old slot drain
                slot = (key, [torch.empty(t.shape, dtype=t.dtype, device=dev) for t in host],
                        torch.zeros(1, dtype=torch.float32, device=dev))
                pipe["slots"][s] = slot
            _, devt, lossd = slot
            with torch.cuda.stream(copy):
                if i >= 2:
                    copy.wait_event(pipe["free"][s])               # step i-2 has consumed this slot ...
                    copy.wait_event(pipe["read"][s])               # ... and its loss has left the slot's mailbox
                for d, h in zip(devt, host):
                    d.copy_(h, non_blocking=True)
                pipe["ready"][s].record(copy)
            cur.wait_event(pipe["ready"][s])
            self.step(*devt, accumulate_metric=accumulate_metric, _loss_copy=lossd)   # the step leaves its loss in lossd
            pipe["free"][s].record(cur)
            if losses is None or n == losses.numel():
                grown = torch.empty(max(256, 2 * n), dtype=torch.float32).pin_memory()
                if n:
                    torch.cuda.synchronize(dev)
                    grown[:n] = losses[:n]
                losses = grown
            with torch.cuda.stream(down):
                # D2H of this step's loss on a third stream: neither the compute stream nor the uploads wait for it
                down.wait_event(pipe["free"][s])
                losses[n:n + 1].copy_(lossd, non_blocking=True)
                pipe["read"][s].record(down)
            n += 1
        cur.wait_stream(down)
        cur.synchronize()
        out = losses[:n].clone() if n else torch.empty(0)
        if self.world > 1 and n:
            import torch.distributed as dist
            t = out.to(dev)
            dist.all_reduce(t, group=self.pg)
            out = t.cpu()
        return out

    def _sync_state(self):
        """Host -> device: optimiser step count and learning rate."""
        self.state[1] = self.step_count
        self.state.view(torch.float32)[2] = self.lr
        self._state_stale = False

    def loss_value(self):
        """Host read of the last step's loss (synchronises); with data parallelism, the sum over ranks."""
        if self.world > 1:
            import torch.distributed as dist
            t = self.loss_out[:1].clone()
            dist.all_reduce(t, group=self.pg)
            return float(t.item())
        return float(self.loss_out[0].item())

    def metric_sums(self):
        """Per-joint sums of L2 errors (``loss_MPJPE``, losses.py:3-7) accumulated since the last reset."""
        return self.loss_out[1:]

    def reset_metric(self):
        self.loss_out[1:].zero_()

    def set_lr(self, lr):
        """The per-step host scalar ReduceLROnPlateau drives (train_1.py:41,106)."""
        self.lr = float(lr)
        self.state.view(torch.float32)[2] = self.lr

    def state_dict(self):
        return {"step": self.step_count, "m": self.m, "v": self.v, "lr": self.lr}

    def load_state_dict(self, sd):
        self.step_count, self.lr = int(sd["step"]), float(sd["lr"])
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        self._sync_state()


class Phase5Trainer(LifterTrainer):
    """The loop body of the phase-5 script around the lifter as ONE replayable kernel sequence
    (``phase5_loop/backup_june8.py:120-175``; same shape in ``train_5.py:123-205``):

        lift_pred = (flip_pose(lift(flip_pose(y1_hat))) + lift(y1_hat)) / 2        four train-mode lifter forwards,
        lift_gt   = (flip_pose(lift(flip_pose(y1))) + lift(y1)) / 2                each with its own BatchNorm batch
        loss = TriangleLoss(y1_hat, y2_hat, lift_gt, lift_pred, y1, y2)            statistics and dropout masks
        loss.backward(); optimizer_lift.step()

    ``y1_hat`` / ``y2_hat`` are what the image models predict (out of scope here: they arrive as tensors); the step
    leaves ``d loss / d y1_hat`` and ``d loss / d y2_hat`` in ``grad_pred2d`` / ``grad_pred3d`` for them, and updates the
    lifter with AdamW (``backup_june8.py:83``).  ``loss="triangle"`` is ``TriangleLoss`` (the lifts of the ground-truth
    key points run -- they move the BatchNorm running statistics -- but get no gradient), ``"triangle_sep"``
    ``TriangleLoss_sep`` (``train_5.py:73``).  ``w_reproj`` / ``w_bone`` add the pinhole-reprojection and bone-length
    terms on ``lift_pred`` -- PARITY UNPINNED, there is no reference implementation -- which need
    ``stats3d=(mean, std)``, and for the reprojection ``intrinsics`` per pose, ``stats2d`` and optionally ``root``.
    ``step(y1_hat, y2_hat, y1, y2[, intrinsics])``; ``losses`` holds ``[total, loss_2d, loss_3d, loss_lift, gap,
    reprojection, bone, 0]`` on the device (the reference reads four of them with ``.cpu().item()`` every step).
    """

    def __init__(self, model, lr=1e-4, loss="triangle", optimizer="adamw", flip=True, w_reproj=0.0, w_bone=0.0,
                 stats3d=None, stats2d=None, root=None, **kw):
        if loss not in ("triangle", "triangle_sep"):
            raise ValueError("Phase5Trainer: loss must be 'triangle' (TriangleLoss) or 'triangle_sep' (TriangleLoss_sep)")
        if model.input_size % 2 != 0 or model.output_size // 3 * 2 != model.input_size:
            raise ValueError("Phase5Trainer lifts (J, 2) key points to (J, 3) joints")
        self.mode = 1 if loss == "triangle_sep" else 0
        self.flip = bool(flip)
        super().__init__(model, lr=lr, loss="l1", optimizer=optimizer, **kw)
        dev = model._flat.device
        J = self.joints
        self.w_reproj, self.w_bone = float(w_reproj), float(w_bone)

        def vec(v, n):
            return None if v is None else torch.as_tensor(v, dtype=torch.float32, device=dev).reshape(-1).contiguous()[:n]
        self.mean3d, self.std3d = (vec(stats3d[0], 3 * J), vec(stats3d[1], 3 * J)) if stats3d is not None else (None, None)
        self.mean2d, self.std2d = (vec(stats2d[0], 2 * J), vec(stats2d[1], 2 * J)) if stats2d is not None else (None, None)
        self.root = vec(root, 3)
        if (self.w_reproj or self.w_bone) and self.mean3d is None:
            raise ValueError("Phase5Trainer: the reprojection / bone terms work on de-standardised joints: pass stats3d")
        self.losses = torch.zeros(8, dtype=torch.float32, device=dev)
        self.loss_out = self.losses                                        # [0] is the total, like the base class
        self.grad_pred2d = self.grad_pred3d = None

    def _bn_calls_per_step(self):
        return 4 if getattr(self, "flip", True) else 2

    def _alloc_buffers(self, B, dev):
        lib, cfg = _cabi.lib(), C.byref(self.model._cfg)
        O, I, n = self.model.output_size, self.model.input_size, self._bn_calls_per_step()
        nb = _cabi.check(lib.lifter_train_saved_bytes(cfg, B))
        f32 = dict(dtype=torch.float32, device=dev)
        return dict(
            saved=[torch.empty(nb, dtype=torch.uint8, device=dev) for _ in range(n)],
            scratch=torch.empty(_cabi.check(lib.lifter_train_scratch_bytes(cfg, B)), dtype=torch.uint8, device=dev),
            out=[torch.empty(B, O, **f32) for _ in range(n)], dout=[torch.empty(B, O, **f32) for _ in range(n)],
            xflip=[torch.empty(B, I, **f32) for _ in range(2)], dx=[torch.empty(B, I, **f32) for _ in range(2)],
            d2_own=torch.empty(B, I, **f32), d2=torch.empty(B, I, **f32), d3=torch.empty(B, O, **f32),
            grads=[torch.empty_like(self.model._flat) for _ in range(n)],
            root=None if self.root is None else self.root.reshape(1, 3).expand(B, 3).contiguous(),
            lscratch=torch.empty(_cabi.check(lib.lifter_phase5_loss_scratch_bytes(B)), dtype=torch.uint8, device=dev))

    def _check_inputs(self, y1_hat, y2_hat, y1, y2, intrinsics=None):
        model = self.model
        a = model._check_input(y1_hat)
        g2 = model._check_input(y1)
        B = a.shape[0]
        out = [a]
        for t, width, what in ((y2_hat, model.output_size, "predicted 3D pose"), (g2, model.input_size, "2D key points"),
                               (y2, model.output_size, "3D pose")):
            t = t.reshape(t.shape[0], -1)
            if t.shape != (B, width) or t.dtype != torch.float32 or not t.is_cuda:
                raise RuntimeError(f"{what}: expected a float32 CUDA tensor of {B} x {width} values")
            out.append(t.contiguous())
        if intrinsics is not None:
            k = intrinsics.reshape(intrinsics.shape[0], -1)
            if k.shape != (B, 4) or k.dtype != torch.float32 or not k.is_cuda:
                raise RuntimeError(f"intrinsics: expected a float32 CUDA tensor of {B} x 4 values (fx, fy, cx, cy)")
            out.append(k.contiguous())
        elif self.w_reproj:
            raise RuntimeError("Phase5Trainer(w_reproj != 0): step() needs the camera intrinsics of every pose")
        return tuple(out)

    def _enqueue(self, inputs, b, accumulate_metric, state):
        y1_hat, y2_hat, y1, y2 = inputs[:4]
        intr = inputs[4] if len(inputs) > 4 else None
        model = self.model
        lib, cfg = _cabi.lib(), C.byref(model._cfg)
        B, J, st = y1_hat.shape[0], self.joints, _cabi.stream_ptr()
        n = self._bn_calls_per_step()
        if state is not None:
            _cabi.check(lib.lifter_step_advance(_cabi.ptr(state), self.betas[0], self.betas[1],
                                                self.weight_decay if self.decoupled else 0.0, st))
        xs = [y1_hat, y1]                                  # backup_june8.py:137-138, then the flipped twins :152-153
        if self.flip:
            for k, src in enumerate((y1_hat, y1)):
                _cabi.check(lib.lifter_flip_pose(_cabi.ptr(src), _cabi.ptr(b["xflip"][k]), B, J, 2, 0, st))
            xs += b["xflip"]
        ctxs = []
        for k, x in enumerate(xs):                         # every call: own batch statistics, masks, running-stat update
            _, _, t = _autograd.run_forward(model, x, True, saved=b["saved"][k], out=b["out"][k], step=self.step_count,
                                            step_state=state, calls=(n, k))
            ctxs.append(t)
        d = _cabi.Phase5LossDesc()
        d.out_pred, d.out_gt = b["out"][0].data_ptr(), b["out"][1].data_ptr()
        d.d_out_pred, d.d_out_gt = b["dout"][0].data_ptr(), b["dout"][1].data_ptr()
        if self.flip:
            d.out_pred_flip, d.out_gt_flip = b["out"][2].data_ptr(), b["out"][3].data_ptr()
            d.d_out_pred_flip, d.d_out_gt_flip = b["dout"][2].data_ptr(), b["dout"][3].data_ptr()
        d.pred2d, d.gt2d, d.pred3d, d.gt3d = y1_hat.data_ptr(), y1.data_ptr(), y2_hat.data_ptr(), y2.data_ptr()
        d.mode = self.mode
        for name in ("mean3d", "std3d", "mean2d", "std2d"):
            v = getattr(self, name)
            setattr(d, name, None if v is None else v.data_ptr())
        d.root = None if b["root"] is None else b["root"].data_ptr()
        d.intrinsics = None if (intr is None or not self.w_reproj) else intr.data_ptr()
        d.w_reproj, d.w_bone = self.w_reproj, self.w_bone
        d.rows, d.global_rows, d.joints = B, B * self.world, J
        d.d_pred2d, d.d_pred3d = b["d2_own"].data_ptr(), b["d3"].data_ptr()
        d.scratch, d.out = b["lscratch"].data_ptr(), self.losses.data_ptr()
        _cabi.check(lib.lifter_phase5_loss(C.byref(d), st))
        # backward of every forward that reached the loss: the lifts of the predicted key points also return the
        # gradient w.r.t. their input (it flows on into the 2D model)
        live = [k for k in range(n) if self.mode == 1 or k % 2 == 0]
        for k in live:
            _autograd.run_backward(model, ctxs[k], b["dout"][k], b["grads"][k], b["dx"][k // 2] if k % 2 == 0 else None,
                                   scratch=b["scratch"])
        g = [b["grads"][k] for k in live]
        if len(g) == 1:
            self.grad.copy_(g[0])
        else:
            _cabi.check(lib.lifter_sum_buffers(_cabi.ptr(self.grad), _cabi.ptr(g[0]), _cabi.ptr(g[1]),
                                               _cabi.ptr(g[2]) if len(g) > 2 else None,
                                               _cabi.ptr(g[3]) if len(g) > 3 else None, self.grad.numel(), st))
        _cabi.check(lib.lifter_phase5_input_grad(_cabi.ptr(b["d2_own"]), _cabi.ptr(b["dx"][0]),
                                                 _cabi.ptr(b["dx"][1]) if self.flip else None, B, J, _cabi.ptr(b["d2"]), st))
        self.grad_pred2d, self.grad_pred3d = b["d2"], b["d3"]
        self._optimise(None, None, None, state, fused_backward=False)


@_cabi.device_guard
def evaluate(model, batches, loss="mse", flip=None, metric_std=None, n_samples=None, zero_centre=True):
    """The validation loop of ``train_1.py:112-145`` without its per-batch host syncs.

    ``batches`` yields ``(x, y)`` -- CUDA tensors, or CPU tensors (pinned for overlap), which are uploaded on a copy
    stream one batch ahead.  Per batch: eval-mode forward, optional flip test-time augmentation, then ONE fused kernel
    for the loss and the per-joint MPJPE sums, accumulated on the device; the host reads the totals once at the end.

    flip: ``None``; ``"output"`` = ``(flip_pose(model(x)) + model(x)) / 2``, literally what ``train_1.py:128-134`` computes
    (it flips the input only AFTER the second forward); ``"input"`` = ``(flip_pose(model(flip_pose(x))) + model(x)) / 2``
    (``train_5.py:153-168``).  Returns ``{"loss": mean of the batch losses, "per_joint": sums / n_samples (J,),
    "mpjpe_mm": mean(per_joint[1:]) * J/(J-1) * 1000 when zero_centre else mean(per_joint) * 1000, "n": samples}``.
    """
    from . import pose_losses as pl
    if loss not in _LOSS:
        raise ValueError(f"unknown loss {loss!r}: use 'mse', 'l1' or 'mpjpe'")
    if flip not in (None, "output", "input"):
        raise ValueError("flip must be None, 'output' or 'input'")
    if model.output_size % 3 != 0:
        raise ValueError("the fused loss works on (J, 3) joint outputs: o_dim must be a multiple of 3")
    dev = next(model.parameters()).device                   # LinearModel or MyViT (train_1.py:35 validates either)
    if dev.type != "cuda":
        raise RuntimeError("evaluate needs the model on a B200 (no CPU path)")
    lib = _cabi.lib()
    J = model.output_size // 3
    std = None if metric_std is None else torch.as_tensor(metric_std, dtype=torch.float32, device=dev).reshape(-1).contiguous()
    was_training = model.training
    model.eval()
    cur = torch.cuda.current_stream(dev)
    copy = torch.cuda.Stream(dev)
    copy.wait_stream(cur)
    ready, free = [torch.cuda.Event() for _ in range(2)], [torch.cuda.Event() for _ in range(2)]
    slots = {}
    acc = torch.zeros(1 + J, dtype=torch.float32, device=dev)          # [sum of batch losses, per-joint sums]
    out = torch.zeros(1 + J, dtype=torch.float32, device=dev)
    scratch, n_batches, n = None, 0, 0
    try:
        for i, (x, y) in enumerate(batches):
            B = x.shape[0]
            if not x.is_cuda:                                          # host batch: upload one batch ahead
                s = i & 1
                xs, ys = x.reshape(B, -1), y.reshape(B, -1)
                slot = slots.get(s)
                if slot is None or slot[0].shape != xs.shape or slot[1].shape != ys.shape:
                    if slot is not None:
                        torch.cuda.synchronize(dev)
                    slot = (torch.empty(xs.shape, dtype=torch.float32, device=dev),
                            torch.empty(ys.shape, dtype=torch.float32, device=dev))
                    slots[s] = slot
                with torch.cuda.stream(copy):
                    if i >= 2:
                        copy.wait_event(free[s])
                    slot[0].copy_(xs, non_blocking=True)
                    slot[1].copy_(ys, non_blocking=True)
                    ready[s].record(copy)
                cur.wait_event(ready[s])
                x, y = slot
            x = x.reshape(B, -1).float()
            y = y.reshape(B, -1).float().contiguous()
            pred = model.infer(x)
            if flip == "output":
                pred = (pl.flip_pose(pred.reshape(B, J, 3)) + pred.reshape(B, J, 3)) * 0.5
            elif flip == "input":
                other = model.infer(pl.flip_pose(x.reshape(B, J, -1)).reshape(B, -1))
                pred = (pl.flip_pose(other.reshape(B, J, 3)) + pred.reshape(B, J, 3)) * 0.5
            pred = pred.reshape(B, -1).contiguous()
            need = _cabi.check(lib.lifter_loss_scratch_bytes(B, J))
            if scratch is None or scratch.numel() < need:
                scratch = torch.empty(need, dtype=torch.uint8, device=dev)
            _cabi.check(lib.lifter_loss(_cabi.ptr(pred), _cabi.ptr(y), B, J, _LOSS[loss], B, _cabi.ptr(std), None,
                                        _cabi.ptr(scratch), _cabi.ptr(out), 1, _cabi.stream_ptr()))
            acc[0] += out[0]                                           # out[1:] accumulates inside the kernel
            if slots:
                free[i & 1].record(cur)
            n_batches += 1
            n += B
    finally:
        if was_training:
            model.train()
    if n_batches == 0:
        raise ValueError("evaluate: no batches")
    acc[1:] = out[1:]
    host = acc.cpu()                                                   # the one host read
    n_samples = n if n_samples is None else int(n_samples)
    per_joint = host[1:] / n_samples
    mm = (per_joint[1:].mean() * (J / (J - 1)) if zero_centre else per_joint.mean()) * 1000.0
    return {"loss": float(host[0]) / n_batches, "per_joint": per_joint, "mpjpe_mm": float(mm), "n": n}
