"""train.train_explainn -- drop-in for explainn/train/train.py:13-142 of the reference.

Two execution paths, same results:
  * generic: any criterion / optimizer; forward+backward run on the fused kernels through autograd
    (autograd_fn.py), the optimizer is the caller's.
  * fused (engaged automatically when the criterion is BCEWithLogitsLoss / MSELoss with mean reduction and
    the optimizer is a plain torch.optim.Adam over exactly the model's parameters): parameters, gradients
    and Adam moments live in flat buffers, the loss and its gradient come from one kernel, the whole
    Adam update is one kernel, and under torch.distributed the step does ONE all-reduce of the flat
    gradient buffer (data parallel, one process per GPU).
The per-step `loss.item()` sync of the reference (train.py:73) is replaced by one device->host read of
all step losses per epoch; the reported epoch means are the same numbers.
"""
import copy
import os

import torch
import torch.nn as nn

from . import _cabi, autograd_fn, parallel
from ._cabi import call, ptr, stream


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


class _Captured:
    """One captured training step (CUDA graph, its static input / output tensors, launch counts)."""
    __slots__ = ("graph", "graph2", "gx", "gy", "gloss", "launches", "launches2", "keep")

    def __init__(self):
        self.graph = self.graph2 = self.gx = self.gy = self.gloss = self.keep = None
        self.launches = self.launches2 = 0


class FusedTrainer:
    """Flat-buffer training step: forward, fused loss, backward, (all-reduce), fused Adam."""

    LOSSES = {nn.BCEWithLogitsLoss: _cabi.LOSS_BCE, nn.MSELoss: _cabi.LOSS_MSE}

    @staticmethod
    def supports(model, criterion, optimizer):
        from .networks import ExplaiNN
        if not isinstance(model, ExplaiNN) or type(criterion) not in FusedTrainer.LOSSES:
            return False
        if getattr(criterion, "reduction", "mean") != "mean":
            return False
        if getattr(criterion, "weight", None) is not None or getattr(criterion, "pos_weight", None) is not None:
            return False
        if type(optimizer) is not torch.optim.Adam or len(optimizer.param_groups) != 1:
            return False
        g = optimizer.param_groups[0]
        if g.get("amsgrad") or g.get("weight_decay", 0) != 0 or g.get("maximize") or g.get("capturable") \
                or g.get("differentiable"):
            return False
        mp = list(model.parameters())
        if len(g["params"]) != len(mp) or any(a is not b for a, b in zip(g["params"], mp)):
            return False
        if not all(p.requires_grad and p.is_cuda and p.dtype == torch.float32 for p in mp):
            return False
        return True

    def __init__(self, model, criterion, optimizer, dropout_seed=None, use_graph=True):
        self.model, self.criterion, self.optimizer = model, criterion, optimizer
        self.kind = self.LOSSES[type(criterion)]
        self.params = list(autograd_fn._params(model))
        dev = self.params[0].device
        sizes = [p.numel() for p in self.params]
        # every slice starts on a 16-byte boundary
        self.offsets, off = [], 0
        for n in sizes:
            self.offsets.append(off)
            off += (n + 3) // 4 * 4
        self.n = off
        d0 = _dist()
        if d0 is not None and d0.get_world_size() > 1:
            parallel.broadcast_model_state(model)       # all replicas start from rank 0's weights and BatchNorm buffers
        # world 2..8 on one node: all-reduce + Adam run as ONE kernel that reads the peers' gradients over NVLink
        # (csrc/p2p.cu; 2.37 M vs 2.22 M seq/s with NCCL at 2 GPUs); the flat gradient buffer is then a CUDA-IPC
        # exportable allocation of parallel.PeerGradients.  XB_P2P_ADAM=0, or any rank that cannot map its peers,
        # selects the NCCL all-reduce + Adam pair on all ranks.
        self.p2p = None
        if d0 is not None and 2 <= d0.get_world_size() <= 8 and d0.get_backend() == "nccl" \
                and os.environ.get("XB_P2P_ADAM", "1") != "0":
            pg = parallel.PeerGradients(off, dev)
            if pg.ok:
                self.p2p = pg
            else:
                import warnings
                warnings.warn("explainn_b200: peer-memory optimizer step unavailable (%s); using the NCCL all-reduce" % pg.error)
        if self.p2p is not None:
            self.flat_g = self.p2p.flat_grad
            self.flat_p = self.p2p.flat_param          # peers store their updated slices straight into it
        else:
            self.flat_g = torch.zeros(off, dtype=torch.float32, device=dev)
            self.flat_p = torch.zeros(off, dtype=torch.float32, device=dev)
        self.p2p_timeout_s = float(os.environ.get("XB_P2P_TIMEOUT_S", "1800"))
        # the FC-stack part of the peer-memory step overlaps the conv backward (XB_P2P_OVERLAP=0: one call after it)
        self.p2p_overlap = os.environ.get("XB_P2P_OVERLAP", "1") != "0"
        self._p2p_fc_started = False
        if self.p2p is not None:
            self._p2p_stream = torch.cuda.Stream(dev)
            self._p2p_ev, self._p2p_ev2 = torch.cuda.Event(), torch.cuda.Event()
        self.flat_m = torch.zeros(off, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(off, dtype=torch.float32, device=dev)
        self.step_t = torch.zeros(1, dtype=torch.float32, device=dev)
        self.views_p, self.views_g = [], []
        state = optimizer.state
        for p, o, n in zip(self.params, self.offsets, sizes):
            v = self.flat_p[o:o + n].view_as(p)
            v.copy_(p.data)
            p.data = v                      # parameters now alias the flat buffer (same attribute paths)
            self.views_p.append(v)
            self.views_g.append(self.flat_g[o:o + n].view_as(p))
            st = state.get(p)
            if st:                          # continue from existing Adam state
                self.flat_m[o:o + n].view_as(p).copy_(st["exp_avg"])
                self.flat_v[o:o + n].view_as(p).copy_(st["exp_avg_sq"])
                self.step_t.fill_(float(st["step"]))
            # expose the moments through the optimizer so that optimizer.state_dict() stays meaningful
            state[p] = {"step": torch.tensor(float(self.step_t.item())),
                        "exp_avg": self.flat_m[o:o + n].view_as(p),
                        "exp_avg_sq": self.flat_v[o:o + n].view_as(p)}
        self.steps_done = int(self.step_t.item())
        self.dropout_seed = dropout_seed
        # CUDA-graph replay of forward+loss+backward (the step is launch-bound at batch 512, SURVEY 7.3 g)
        self.use_graph = use_graph
        if getattr(model, "sync_bn", False) and d0 is not None and d0.get_world_size() > 1:
            self.use_graph = False      # the synchronised-BatchNorm step has NCCL all-reduces between its kernels: eager
        self._graph = None
        self._graphs = {}
        self._graph_failed = False
        self._seed_ctr = torch.zeros(1, dtype=torch.int64, device=dev)
        self._loss_scratch = torch.zeros(_cabi.lib.explainn_loss_scratch_bytes(), dtype=torch.uint8, device=dev)
        self.world = 1
        d = _dist()
        if d is not None:
            self.world = d.get_world_size()
        # data parallel: the gradients of the FC stack (everything after the four conv / BatchNorm-1 tensors in the
        # flat buffer) are final before the conv backward starts, so their all-reduce runs beside it
        self.n_conv = self.offsets[4]
        # opt-in (XB_DP_OVERLAP=1): measured on 2 B200s the NCCL kernel competing with the conv backward for SMs costs more
        # (2.15 M seq/s) than the exposed 4.6 MB all-reduce of the default single collective (2.21 M seq/s)
        self.dp_overlap = os.environ.get("XB_DP_OVERLAP", "0") == "1"
        self._pending = None

    def _still_aliased(self):
        return all(p.data_ptr() == v.data_ptr() for p, v in zip(self.params, self.views_p))

    def _fwd_bwd(self, x, y, seed, seed_dev=None):
        """forward + fused loss + backward into the flat gradient buffer; returns the loss tensor.  With `seed_dev` (a
        captured step) the forward first increments that counter, so every replay draws a new dropout mask."""
        ps = self.views_p
        ws = autograd_fn.train_forward(self.model, x, ps, seed=seed, seed_dev=seed_dev, bump_seed=seed_dev is not None)
        B, C = ws.logits.shape
        if tuple(y.shape) != (B, C):
            raise ValueError("labels of shape %s do not match logits %s" % (tuple(y.shape), (B, C)))
        loss = torch.empty(1, dtype=torch.float32, device=ws.logits.device)
        dlogits = torch.empty_like(ws.logits)
        call("explainn_loss_fwd_bwd", ptr(ws.logits), ptr(y), B, C, self.kind, 1.0, ptr(loss), ptr(dlogits),
             ptr(self._loss_scratch), stream())
        autograd_fn.train_backward(self.model, ws, dlogits, ps, self.views_g, between=self._between)
        return loss

    # ---- data-parallel overlap -------------------------------------------------------------------------
    def _between(self, ws, gpool, params, grads, wrt_unit):
        """Runs after the FC-stack backward.  Eager: start the all-reduce of the FC-stack gradients.  During graph
        capture (world > 1): end the first graph here; the conv backward is captured as a second graph."""
        if self.world == 1:
            return True
        if self.p2p is not None:
            if not self.p2p_overlap:
                return True
            if torch.cuda.is_current_stream_capturing():
                self._split = (ws, gpool)
                return False
            self._p2p_fc_part()             # eager: the FC-stack update starts here, beside the conv backward
            return True
        if not self.dp_overlap:
            return True
        if torch.cuda.is_current_stream_capturing():
            self._split = (ws, gpool)
            return False
        self._start_fc_allreduce()
        return True

    # ---- peer-memory optimizer step in two parts ------------------------------------------------------------
    def _p2p_call(self, offset, count, tick, seq):
        import ctypes
        g = self.optimizer.param_groups[0]
        b1, b2 = g["betas"]
        call("explainn_p2p_adam_step", ctypes.cast(self.p2p.p_array, ctypes.c_void_p), ptr(self.flat_m), ptr(self.flat_v),
             offset, count, ptr(self.step_t), tick, ctypes.cast(self.p2p.g_array, ctypes.c_void_p),
             ctypes.cast(self.p2p.f_array, ctypes.c_void_p), self.world, self.p2p.rank, seq, float(g["lr"]), float(b1), float(b2),
             float(g["eps"]), self.p2p_timeout_s, stream())

    def _p2p_fc_part(self):
        """Reduce-scatter + Adam + all-gather of everything behind the conv / BatchNorm-1 tensors (99 % of the
        parameters; their gradients are final once the FC-stack backward is done) on a side stream, so that it runs
        beside the conv backward of this rank and of its peers."""
        main = torch.cuda.current_stream()
        self._p2p_ev.record(main)
        self._p2p_stream.wait_event(self._p2p_ev)
        with torch.cuda.stream(self._p2p_stream):
            self._p2p_call(self.n_conv, self.n - self.n_conv, 1, 2 * (self.steps_done + 1) - 1)
            self._p2p_ev2.record(self._p2p_stream)
        self._p2p_fc_started = True

    def _p2p_finish(self):
        """The conv / BatchNorm-1 part after the conv backward (and after the FC part: calls must not overlap)."""
        if self._p2p_fc_started:
            torch.cuda.current_stream().wait_event(self._p2p_ev2)
            self._p2p_call(0, self.n_conv, 0, 2 * (self.steps_done + 1))
        else:
            self._p2p_call(0, self.n, 1, 2 * (self.steps_done + 1))
        self._p2p_fc_started = False

    def _start_fc_allreduce(self):
        d = _dist()
        self._pending = d.all_reduce(self.flat_g[self.n_conv:], op=d.ReduceOp.SUM, async_op=True)

    def _finish_allreduce(self):
        """All-reduce of the (small) conv gradients, then wait for the FC-stack one.  Returns the Adam scale."""
        if self.world == 1:
            return 1.0
        d = _dist()
        if not self.dp_overlap:             # one collective over the whole flat buffer after the backward
            return parallel.allreduce_mean_gradients(self.flat_g)
        if self._pending is None:           # defensive: the first half was not started
            self._start_fc_allreduce()
        d.all_reduce(self.flat_g[:self.n_conv], op=d.ReduceOp.SUM)
        self._pending.wait()
        self._pending = None
        return 1.0 / self.world

    def _capture(self, x, y, seed):
        """Capture forward+loss+backward for this batch shape / dtype.  Capturing executes nothing, so the BatchNorm
        running statistics are untouched until the first replay.  One graph per input signature is kept (fp32 one-hot
        and ASCII text batches, or a short last batch, each get their own), at most four."""
        c = _Captured()
        c.gx = torch.empty_like(x)
        c.gy = torch.empty_like(y)
        if x.dtype == torch.uint8:
            self.model._pack(x)          # creates the cached offset vector of this shape outside the capture
        g = torch.cuda.CUDAGraph()
        l0 = _cabi.LAUNCHES
        try:
            self._split = None
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                c.gloss = self._fwd_bwd(c.gx, c.gy, seed, self._seed_ctr)
            c.launches = _cabi.LAUNCHES - l0
            if self._split is not None:      # world > 1: the conv backward is a second graph (see _between)
                ws, gpool = self._split
                c.graph2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(c.graph2, pool=g.pool(), capture_error_mode="thread_local"):
                    autograd_fn.train_backward_conv(self.model, ws, gpool, self.views_p, self.views_g)
                c.launches2 = _cabi.LAUNCHES - l0 - c.launches
                c.keep = (ws, gpool)         # keep the workspace of the captured graphs alive
            _cabi.LAUNCHES = l0              # captured, not executed
        except Exception as e:      # keep training eagerly rather than fail
            import warnings
            warnings.warn("explainn_b200: CUDA-graph capture failed (%s); continuing without graphs" % (e,))
            self._graph_failed = True
            self._graphs.clear()
            self._graph = None
            torch.cuda.synchronize()
            return None
        c.graph = g
        if len(self._graphs) >= 4:
            self._graphs.pop(next(iter(self._graphs)))
        return c

    @_cabi.on_device_of(1)
    def step(self, x, y):
        """One optimisation step (train.py:59-67).  Returns the loss as a 1-element device tensor.  `x`: fp32 / fp16
        one-hot (B,4,L) or uint8 ASCII text (B,L) on the device."""
        if not self._still_aliased():       # somebody re-assigned a parameter: re-adopt it
            for p, v in zip(self.params, self.views_p):
                if p.data_ptr() != v.data_ptr():
                    v.copy_(p.data)
                    p.data = v
            self._graphs.clear()
            self._graph = None
        base = self.dropout_seed if self.dropout_seed is not None else (torch.initial_seed() & 0xFFFFFFFFFFFF)
        seed = (base + 0x9E3779B97F4A7C15 * (self.steps_done + 1)) & 0xFFFFFFFFFFFFFFFF
        y = y.to(x.device, dtype=torch.float32).contiguous()
        x = x.contiguous()
        # the first step runs eagerly (it is the warm-up); a graph is captured the first time an input signature is
        # seen after that
        key = (tuple(x.shape), x.dtype, tuple(y.shape))
        c = self._graphs.get(key)
        if c is None and self.use_graph and not self._graph_failed and self.steps_done >= 1:
            c = self._capture(x, y, base)
            if c is not None:
                self._graphs[key] = c
        if c is not None and self.use_graph:
            self._graph = c.graph            # (kept for callers that test "is a graph in use")
            c.gx.copy_(x, non_blocking=True)
            c.gy.copy_(y, non_blocking=True)
            c.graph.replay()
            _cabi.LAUNCHES += c.launches
            if c.graph2 is not None:
                # the FC-stack gradients are final: their reduction (peer-memory kernel or NCCL) runs beside the conv backward
                if self.p2p is not None:
                    self._p2p_fc_part()
                else:
                    self._start_fc_allreduce()
                c.graph2.replay()
                _cabi.LAUNCHES += c.launches2
            loss = c.gloss.clone()
        else:
            loss = self._fwd_bwd(x, y, seed)
        g = self.optimizer.param_groups[0]
        b1, b2 = g["betas"]
        if self.p2p is not None:
            self._p2p_finish()
            self.steps_done += 1
            return loss
        scale = self._finish_allreduce()     # SUM over ranks (NCCL over NVLink); Adam divides by the world size
        call("explainn_adam_step", ptr(self.flat_p), ptr(self.flat_g), ptr(self.flat_m), ptr(self.flat_v), self.n,
             ptr(self.step_t), float(g["lr"]), float(b1), float(b2), float(g["eps"]), scale, stream())
        self.steps_done += 1
        return loss

    def sync_buffers(self):
        """Data parallel: every rank normalises with its own batch statistics, so the BatchNorm running statistics drift
        apart between replicas.  Average them over the ranks (call before validation / interpretation so that every
        rank evaluates the same model; DistributedDataParallel broadcasts rank 0's buffers instead)."""
        d = _dist()
        if d is None or self.world == 1:
            return
        for name, buf in self.model.named_buffers():
            if buf.dtype.is_floating_point:
                d.all_reduce(buf, op=d.ReduceOp.SUM)
                buf.div_(self.world)

    def check_health(self):
        """Raise if the peer-memory optimizer step ever gave up waiting for a rank (one small synchronous read)."""
        if self.p2p is not None:
            self.p2p.check()

    def sync_optimizer_state(self):
        if self.p2p is not None:
            # the Adam moments of a slice live on its owner (csrc/p2p.cu): gather them so that optimizer.state_dict()
            # is complete on every rank
            self.check_health()
            torch.cuda.current_stream().wait_stream(self._p2p_stream)
            ranges = ([self.p2p.slice_range(self.n_conv), self.p2p.slice_range(self.n - self.n_conv, self.n_conv)]
                      if self.p2p_overlap else [self.p2p.slice_range(self.n)])
            own = torch.zeros(self.n, dtype=torch.bool, device=self.flat_m.device)
            for lo, hi in ranges:
                own[lo:hi] = True
            d = _dist()
            for buf in (self.flat_m, self.flat_v):
                buf.mul_(own)
                d.all_reduce(buf, op=d.ReduceOp.SUM)
        for p in self.params:
            self.optimizer.state[p]["step"] = torch.tensor(float(self.steps_done))
        for p, gview in zip(self.params, self.views_g):
            p.grad = gview


class DevicePrefetcher:
    """Iterate a loader of (sequences, labels) host batches one batch ahead: the host-to-device copies of batch i+1
    are issued on a copy stream while batch i trains (asynchronous when the loader yields pinned memory).  Sequences
    may be fp32 / fp16 one-hot (B,4,L) or uint8 ASCII text (B,L); the latter is 16x fewer bytes on PCIe."""

    def __init__(self, loader, device):
        self.loader, self.device = loader, torch.device(device)
        self.stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None

    def _fetch(self, it):
        try:
            seqs, labels = next(it)
        except StopIteration:
            return None
        if self.stream is None:
            return seqs.to(self.device), labels.to(self.device), None
        with torch.cuda.stream(self.stream):
            x = seqs.to(self.device, non_blocking=True)
            y = labels.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return x, y, ev

    def __iter__(self):
        it = iter(self.loader)
        nxt = self._fetch(it)
        while nxt is not None:
            x, y, ev = nxt
            nxt = self._fetch(it)
            if ev is not None:
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(ev)
                x.record_stream(cur)
                y.record_stream(cur)
            yield x, y

    def __len__(self):
        return len(self.loader)


def _epoch_mean(losses):
    """Mean of per-step losses exactly as train.py:73-76 accumulates them (python floats)."""
    if not losses:
        return float("nan")
    vals = torch.cat([l.detach().reshape(1).float() for l in losses]).cpu().tolist()   # one sync per epoch
    total = 0.0
    for v in vals:
        total += v
    return total / len(vals)


def train_explainn(train_loader, test_loader, model, device, criterion,
                   optimizer, num_epochs=100, weights_folder="./",
                   name_ind=None, verbose=False, trim_weights=False,
                   checkpoint=1, patience=0):
    """Epoch loop of the reference (explainn/train/train.py:13-142) on the fused kernels; same arguments, files and
    return value, so callers of the reference need no change.

    train_loader / test_loader yield (sequences (B,4,L) one-hot, labels (B,C)) batches; `criterion` and `optimizer`
    are the caller's objects.  BCEWithLogitsLoss / MSELoss (mean) with a plain torch.optim.Adam over all model
    parameters take the flat-buffer fast path (FusedTrainer); anything else runs forward and backward on the same
    kernels through autograd and steps the caller's optimizer.  Per epoch: mean of the step losses, validation in
    eval() mode, best-so-far weights by validation loss.  `checkpoint` = k > 0 writes the best-so-far weights as
    model_epoch_{e}[_{name_ind}].pth every k epochs (and, like the reference, reloads them into the model at that
    point); `patience` = n > 0 stops n epochs after the best one; `trim_weights` clamps final.weight at zero after
    every step.  At the end the best weights are loaded, written as model_epoch_best_{e}[_{name_ind}].pth and
    (model, train losses, validation losses) is returned; the model is left in eval() mode.

    Under an initialised torch.distributed group (one process per GPU, each loader yielding that rank's shard) the
    gradients are averaged over ranks every step, all ranks start from rank 0's weights, the validation loss is
    averaged over ranks before the best-epoch / early-stopping decisions (so every rank takes the same ones) and
    only rank 0 writes files."""
    train_error, test_error = [], []
    d = _dist()
    world = d.get_world_size() if d is not None else 1
    rank = d.get_rank() if d is not None else 0
    fused = FusedTrainer(model, criterion, optimizer) if FusedTrainer.supports(model, criterion, optimizer) else None
    if world > 1 and fused is None:
        import warnings
        warnings.warn("explainn_b200: criterion / optimizer not recognised by the fused trainer; data-parallel training "
                      "continues on the autograd path with one gradient all-reduce per parameter per step")
        parallel.broadcast_model_state(model)
    best_model_wts = copy.deepcopy(model.state_dict())
    best_loss_valid = float('inf')
    best_epoch = 1

    for epoch in range(1, num_epochs + 1):
        step_losses = []
        model.train()
        for x, labels in DevicePrefetcher(train_loader, device):
            if fused is not None:
                loss = fused.step(x, labels)
            else:
                optimizer.zero_grad()
                outputs = model(x)
                loss = criterion(outputs, labels)
                loss.backward()
                if world > 1:
                    for p in model.parameters():
                        if p.grad is not None:
                            d.all_reduce(p.grad, op=d.ReduceOp.SUM)
                            p.grad.div_(world)
                optimizer.step()
            if trim_weights:        # non-negative output layer weights (train.py:70-71)
                model.final.weight.data.clamp_(0)
            step_losses.append(loss)
        epoch_loss = _epoch_mean(step_losses)
        train_error.append(epoch_loss)

        if fused is not None:
            fused.sync_buffers()
        val_losses = []
        with torch.no_grad():
            model.eval()
            for seqs, labels in test_loader:
                x = seqs.to(device, non_blocking=True)
                y = labels.to(device, non_blocking=True)
                val_losses.append(criterion(model(x), y))
        test_loss = _epoch_mean(val_losses)
        test_error.append(test_loss)
        if hasattr(model, "check_inputs"):
            model.check_inputs()
        if fused is not None:
            fused.check_health()

        if verbose:
            print('Epoch [{}], Current Train Loss: {:.5f}, Current Val Loss: {:.5f}'
                  .format(epoch, epoch_loss, test_loss))

        decide_loss = parallel.allreduce_mean_scalar(test_loss, device) if world > 1 else test_loss
        if decide_loss < best_loss_valid:
            best_loss_valid = decide_loss
            best_epoch = epoch
            best_model_wts = copy.deepcopy(model.state_dict())

        if checkpoint:
            if epoch % checkpoint == 0:
                model.load_state_dict(best_model_wts)
                f = f"model_epoch_{epoch}_{name_ind}.pth" if name_ind else f"model_epoch_{epoch}.pth"
                if rank == 0:
                    torch.save(best_model_wts, os.path.join(weights_folder, f))

        if patience:
            if epoch >= best_epoch + patience:
                if verbose:
                    print('Early stopping, Current Epoch {}, Best Epoch: {}, Patience: {}'
                          .format(epoch, best_epoch, patience))
                break

    if fused is not None:
        fused.sync_optimizer_state()
    model.load_state_dict(best_model_wts)
    f = f"model_epoch_best_{best_epoch}_{name_ind}.pth" if name_ind else f"model_epoch_best_{best_epoch}.pth"
    if rank == 0:
        torch.save(best_model_wts, os.path.join(weights_folder, f))
    return model, train_error, test_error
