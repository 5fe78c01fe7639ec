"""Training-mode forward/backward of ExplaiNN on the fused kernels, exposed to autograd.

Reference semantics (explainn/models/networks.py:96-116,147,152-162 under `model.train()`;
explainn/train/train.py:61-65): BatchNorm uses batch statistics (biased variance) and updates its
running statistics with momentum 0.1 and the unbiased variance; dropout p=0.3 is active unless the
Dropout module itself is in eval mode.  Parameters stay ordinary nn.Parameters: any criterion and any
optimizer work on top of this function, gradient hooks on `model.linears[0].weight` fire as usual
(explainn/utils/tools.py:248-264), `requires_grad=False` is honoured (scripts/finetune.py:216-221).
"""
import torch

from . import functional as Fn
from ._cabi import call, ptr, stream

H = Fn.HIDDEN

_SEED_STATE = {"counter": 0}


def _next_dropout_seed():
    """A fresh 64-bit seed per forward, derived from torch's CPU generator state so that
    torch.manual_seed makes runs repeatable."""
    _SEED_STATE["counter"] += 1
    base = torch.initial_seed() & 0xFFFFFFFFFFFF
    return (base * 0x9E3779B1 + _SEED_STATE["counter"] * 0x85EBCA77C2B2AE63) & 0xFFFFFFFFFFFFFFFF


def _num_fc(model):
    return model._options.get("num_fc", 2)


def _is_variant(model):
    """True for the units served by variants.py: num_fc = 0 or > 2, or a pool other than MaxPool1d(7, 7)."""
    o = model._options
    return o.get("num_fc", 2) not in (1, 2) or o.get("pool_size", 7) != Fn.POOL or o.get("pool_stride", 7) != Fn.POOL


def _params(model):
    """Parameters in state_dict order; the conv / BatchNorm-1 tensors always come first."""
    if _is_variant(model):
        from . import variants
        return variants.params(model)
    L = model.linears
    if _num_fc(model) == 0:        # networks.py:73-80 (eval only here: _check_modes refuses to train it)
        return (L[0].weight, L[0].bias, L[1].weight, L[1].bias, model.final.weight, model.final.bias)
    if _num_fc(model) == 1:        # networks.py:82-95
        return (L[0].weight, L[0].bias, L[1].weight, L[1].bias, L[6].weight, L[6].bias, L[7].weight, L[7].bias,
                model.final.weight, model.final.bias)
    return (L[0].weight, L[0].bias, L[1].weight, L[1].bias, L[6].weight, L[6].bias, L[7].weight, L[7].bias,
            L[10].weight, L[10].bias, L[11].weight, L[11].bias, model.final.weight, model.final.bias)


def _check_modes(model):
    L = model.linears
    nfc = _num_fc(model)
    if nfc not in (1, 2):
        raise NotImplementedError("explainn_b200: training is built for num_fc = 1 and 2; num_fc = %d runs in eval() mode only" % nfc)
    bns = (L[1], L[7]) if nfc == 1 else (L[1], L[7], L[11])
    if not all(bn.training for bn in bns):
        raise NotImplementedError("explainn_b200: in train() mode all BatchNorm layers must be in training mode")
    for bn in bns:
        if bn.momentum is None or not bn.track_running_stats or not bn.affine:
            raise NotImplementedError("explainn_b200: BatchNorm layers must keep the reference configuration "
                                      "(momentum=0.1, affine, track_running_stats)")


def _sync_group(model):
    """(dist, world) when the model asks for synchronised BatchNorm statistics (model.sync_bn = True) inside an
    initialised process group of more than one rank, else (None, 1)."""
    if not getattr(model, "sync_bn", False):
        return None, 1
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return None, 1
    return dist, dist.get_world_size()


def _allreduce_avg(d, t, world):
    d.all_reduce(t, op=d.ReduceOp.SUM)
    t.mul_(1.0 / world)


class TrainWorkspace:
    """Everything the backward needs, produced by one training forward."""
    __slots__ = ("pk", "B", "U", "P", "C", "F", "M", "mean1", "var1", "Mw", "pooled", "xhat", "idx", "h1", "mean2",
                 "invstd2", "h2", "mean3", "invstd3", "unit", "logits", "pdrop", "seed", "seed_dev", "nchunk", "dropmask",
                 "sync_world")


def _bump_counters(bn2, bn3, seed_dev, bump_seed):
    """num_batches_tracked of BatchNorm-2 / -3 (BatchNorm-1's is incremented by its statistics kernel) and, for a captured
    step, the per-replay dropout counter: one launch."""
    call("explainn_counters_inc", ptr(bn2.num_batches_tracked), ptr(bn3.num_batches_tracked) if bn3 is not None else None,
         ptr(seed_dev) if bump_seed else None, None, stream())


def train_forward(model, x, params=None, seed=None, seed_dev=None, bump_seed=False):
    """Runs the training-mode forward kernels.  Returns a TrainWorkspace (logits in ws.logits)."""
    if _is_variant(model):
        from . import variants
        if _sync_group(model)[0] is not None:
            raise NotImplementedError("explainn_b200: sync_bn is built for the default unit (num_fc = 2, MaxPool1d(7, 7))")
        if params is None:
            params = [p.detach() for p in _params(model)]
        if bump_seed:
            seed_dev.add_(1)
        return variants.train_forward(model, x, params, _next_dropout_seed() if seed is None else seed, seed_dev)
    _check_modes(model)
    L = model.linears
    if params is None:
        params = [p.detach() for p in _params(model)]
    if _num_fc(model) == 1:
        if bump_seed:
            seed_dev.add_(1)
        return _train_forward_fc1(model, x, params)
    (cw, cb, g1, b1n, W1, b1, g2, be2, W2, b2, g3, be3, Wf, bf) = params
    o = model._options
    U, F, C = o["num_cnns"], o["filter_size"], o["num_classes"]
    pk = model._pack(x)
    B = pk.B
    if B < 2:
        # same failure BatchNorm raises in train mode (worked around at scripts/utils/__init__.py:121-132)
        raise ValueError("Expected more than 1 value per channel when training, got input size %s" % (tuple(x.shape),))
    dev = pk.codes.device
    bn1, bn2, bn3, drop = L[1], L[7], L[11], L[9]
    _bump_counters(bn2, bn3, seed_dev, bump_seed)          # before the first kernel that reads the dropout counter
    ws = TrainWorkspace()
    ws.pk, ws.B, ws.U, ws.C, ws.F = pk, B, U, C, F
    S = pk.L - F + 1
    P = ws.P = model.pooled_length
    # synchronised BatchNorm (model.sync_bn under a process group): every additive statistic is all-reduced between the
    # kernel that produces it and the one that consumes it, so the step equals the reference's on the concatenated batch
    dsync, world = _sync_group(model)
    ws.sync_world = world
    ws.M = Fn.input_cooc_stats(pk, F)
    if dsync is not None:
        dsync.all_reduce(ws.M, op=dsync.ReduceOp.SUM)
    ws.mean1, ws.var1, ws.Mw = Fn.bn1_batch_stats(ws.M, cw, cb, B * S * world, bn1.momentum, bn1.running_mean, bn1.running_var,
                                                  bn1.num_batches_tracked)
    ones, zeros = _const(dev, U)
    # the conv is prepared with gamma=1, beta=0 and the batch statistics, so the scan yields x_hat = (y - mu) / sigma
    if getattr(model, "conv_backend", "tc") != "lut":      # training always uses the two-piece (fp32-accurate) weights
        tcp = Fn.conv_tc_prepare(cw, cb, ones, zeros, ws.mean1, ws.var1, bn1.eps, sign_of=g1)
        ws.pooled, ws.xhat, ws.idx = Fn.conv_pool_fwd_train_tc(pk, tcp, g1, b1n)
    else:
        prep = Fn.conv_prepare(cw, cb, ones, zeros, ws.mean1, ws.var1, bn1.eps)
        ws.pooled, ws.xhat, ws.idx = Fn.conv_pool_fwd_train(pk, prep, g1, b1n)
    nchunk = ws.nchunk = Fn.fc_nchunk(B)
    f32 = dict(dtype=torch.float32, device=dev)
    ws.h1 = torch.empty((U, B, H), **f32)
    part2 = torch.empty((U, nchunk, 2, H), **f32)
    ws.mean2 = torch.empty(U * H, **f32)
    ws.invstd2 = torch.empty(U * H, **f32)
    ws.h2 = torch.empty((U, B), **f32)
    part3 = torch.empty((U, nchunk, 2), **f32)
    ws.mean3 = torch.empty(U, **f32)
    ws.invstd3 = torch.empty(U, **f32)
    ws.unit = torch.empty((B, U), **f32)
    ws.logits = torch.empty((B, C), **f32)
    ws.pdrop = float(drop.p) if drop.training else 0.0
    ws.seed = _next_dropout_seed() if seed is None else seed
    ws.seed_dev = seed_dev
    ws.dropmask = torch.empty((U, B, 4), dtype=torch.int32, device=dev) if ws.pdrop > 0 else None
    fargs = (ptr(ws.pooled), B, U, P, C, ptr(W1), ptr(b1), ptr(g2), ptr(be2),
             ptr(bn2.running_mean), ptr(bn2.running_var), ptr(W2), ptr(b2), ptr(g3), ptr(be3), ptr(bn3.running_mean),
             ptr(bn3.running_var), float(bn2.momentum), float(bn2.eps), ws.pdrop, ws.seed, ptr(seed_dev), ptr(ws.dropmask), ptr(Wf), ptr(bf), ptr(ws.h1),
             ptr(part2), ptr(ws.mean2), ptr(ws.invstd2), ptr(ws.h2), ptr(part3), ptr(ws.mean3), ptr(ws.invstd3), ptr(ws.unit),
             ptr(ws.logits), stream())
    if dsync is None:
        call("explainn_fc_stack_fwd_train", *fargs)
    else:
        call("explainn_fc_stack_fwd_train_phase", 1, B * world, *fargs)
        dsync.all_reduce(part2, op=dsync.ReduceOp.SUM)
        call("explainn_fc_stack_fwd_train_phase", 2, B * world, *fargs)
        dsync.all_reduce(part3, op=dsync.ReduceOp.SUM)
        call("explainn_fc_stack_fwd_train_phase", 4, B * world, *fargs)
    return ws


def _train_forward_conv(model, x, cw, cb, g1, b1n):
    """Packing, BatchNorm-1 batch statistics and the fused conv -> BN -> exp -> pool forward (shared by the variants)."""
    o = model._options
    U, F, C = o["num_cnns"], o["filter_size"], o["num_classes"]
    pk = model._pack(x)
    B = pk.B
    if B < 2:
        raise ValueError("Expected more than 1 value per channel when training, got input size %s" % (tuple(x.shape),))
    dev = pk.codes.device
    bn1 = model.linears[1]
    ws = TrainWorkspace()
    ws.pk, ws.B, ws.U, ws.C, ws.F = pk, B, U, C, F
    S = pk.L - F + 1
    ws.P = model.pooled_length
    ws.M = Fn.input_cooc_stats(pk, F)
    ws.mean1, ws.var1, ws.Mw = Fn.bn1_batch_stats(ws.M, cw, cb, B * S, bn1.momentum, bn1.running_mean, bn1.running_var,
                                                  bn1.num_batches_tracked)
    ones, zeros = _const(dev, U)
    if getattr(model, "conv_backend", "tc") != "lut":      # training always uses the two-piece (fp32-accurate) weights
        tcp = Fn.conv_tc_prepare(cw, cb, ones, zeros, ws.mean1, ws.var1, bn1.eps, sign_of=g1)
        ws.pooled, ws.xhat, ws.idx = Fn.conv_pool_fwd_train_tc(pk, tcp, g1, b1n)
    else:
        prep = Fn.conv_prepare(cw, cb, ones, zeros, ws.mean1, ws.var1, bn1.eps)
        ws.pooled, ws.xhat, ws.idx = Fn.conv_pool_fwd_train(pk, prep, g1, b1n)
    ws.nchunk = Fn.fc_nchunk(B)
    ws.pdrop, ws.seed, ws.seed_dev, ws.dropmask = 0.0, 0, None, None
    return ws


def _train_forward_fc1(model, x, params):
    """num_fc = 1 (networks.py:82-95): pooled -> FC (P -> 1) -> BatchNorm -> ReLU -> final Linear."""
    if _sync_group(model)[0] is not None:
        raise NotImplementedError("explainn_b200: sync_bn is built for the default unit (num_fc = 2, MaxPool1d(7, 7))")
    (cw, cb, g1, b1n, W1, b1, g3, be3, Wf, bf) = params
    ws = _train_forward_conv(model, x, cw, cb, g1, b1n)
    bn3 = model.linears[7]
    B, U, P, C = ws.B, ws.U, ws.P, ws.C
    f32 = dict(dtype=torch.float32, device=ws.pooled.device)
    ws.h2 = torch.empty((U, B), **f32)
    part3 = torch.empty((U, ws.nchunk, 2), **f32)
    ws.mean3 = torch.empty(U, **f32)
    ws.invstd3 = torch.empty(U, **f32)
    ws.unit = torch.empty((B, U), **f32)
    ws.logits = torch.empty((B, C), **f32)
    call("explainn_fc1_stack_fwd_train", ptr(ws.pooled), B, U, P, C, ptr(W1), ptr(b1), ptr(g3), ptr(be3), ptr(bn3.running_mean),
         ptr(bn3.running_var), float(bn3.momentum), float(bn3.eps), ptr(Wf), ptr(bf), ptr(ws.h2), ptr(part3), ptr(ws.mean3),
         ptr(ws.invstd3), ptr(ws.unit), ptr(ws.logits), stream())
    bn3.num_batches_tracked.add_(1)
    return ws


def _train_backward_fc1(model, ws, gout, params, grads, wrt_unit, between):
    (cw, cb, g1, b1n, W1, b1, g3, be3, Wf, bf) = params
    if grads is None:
        grads = [torch.empty_like(p) for p in params]
    (dcw, dcb, dg1, db1n, dW1, db1, dg3, dbe3, dWf, dbf) = grads
    B, U, P, C = ws.B, ws.U, ws.P, ws.C
    f32 = dict(dtype=torch.float32, device=gout.device)
    gout = gout.contiguous()
    gh2 = torch.empty((U, B), **f32)
    dWp = torch.empty((ws.nchunk, U, P), **f32)
    gpool = torch.empty((B, U, P), **f32)
    call("explainn_fc1_stack_bwd", ptr(gout), B, U, P, C, ptr(ws.pooled), ptr(ws.h2), ptr(ws.unit), ptr(W1), ptr(g3), ptr(ws.mean3),
         ptr(ws.invstd3), None if wrt_unit else ptr(Wf), ptr(gh2), ptr(dWp), ptr(gpool), ptr(dW1), ptr(db1), ptr(dg3), ptr(dbe3),
         None if wrt_unit else ptr(dWf), None if wrt_unit else ptr(dbf), stream())
    if between is not None and between(ws, gpool, params, grads, wrt_unit) is False:
        return grads
    return train_backward_conv(model, ws, gpool, params, grads, wrt_unit)


_CONST = {}


def _const(dev, U):
    """(ones, zeros) of length U, cached per device.  Under CUDA-graph capture nothing is cached: tensors made there live
    in the graph's private pool and are filled by captured kernels, so they must not outlive the graph."""
    key = (str(dev), U)
    if key in _CONST:
        return _CONST[key]
    pair = (torch.ones(U, dtype=torch.float32, device=dev), torch.zeros(U, dtype=torch.float32, device=dev))
    if not torch.cuda.is_current_stream_capturing():
        _CONST[key] = pair
    return pair


def train_backward(model, ws, gout, params=None, grads=None, wrt_unit=False, between=None):
    """Backward kernels.  gout: (B,C) gradient w.r.t. the logits, or (B,U) w.r.t. the unit outputs when
    `wrt_unit`.  `grads`: optional list of 14 preallocated gradient tensors (e.g. views into a flat buffer).
    `between`: optional callable run after the FC-stack backward (grads[4:] final) and before the conv backward
    (grads[:4]); the data-parallel trainer uses it to start the all-reduce of the FC-stack gradients early, or to
    split the step into two CUDA graphs (it may return False to stop after the first half)."""
    L = model.linears
    if params is None:
        params = [p.detach() for p in _params(model)]
    if _is_variant(model):
        from . import variants
        return variants.train_backward(model, ws, gout, params, grads, wrt_unit, between)
    if _num_fc(model) == 1:
        return _train_backward_fc1(model, ws, gout, params, grads, wrt_unit, between)
    (cw, cb, g1, b1n, W1, b1, g2, be2, W2, b2, g3, be3, Wf, bf) = params
    B, U, P, C, F = ws.B, ws.U, ws.P, ws.C, ws.F
    dev = gout.device
    f32 = dict(dtype=torch.float32, device=dev)
    if grads is None:
        grads = [torch.empty_like(p) for p in params]
    (dcw, dcb, dg1, db1n, dW1, db1, dg2, dbe2, dW2, db2, dg3, dbe3, dWf, dbf) = grads
    gout = gout.contiguous()
    gh2 = torch.empty((U, B), **f32)
    pb2 = torch.empty((U, ws.nchunk, 3, H), **f32)
    dW1p = torch.empty((ws.nchunk, U, H, P), **f32)
    gpool = torch.empty((B, U, P), **f32)
    bargs = (ptr(gout), B, U, P, C, ptr(ws.pooled), ptr(ws.h1), ptr(ws.h2), ptr(ws.unit), ptr(W1),
             ptr(g2), ptr(be2), ptr(ws.mean2), ptr(ws.invstd2), ptr(W2), ptr(g3), ptr(ws.mean3), ptr(ws.invstd3),
             None if wrt_unit else ptr(Wf), ws.pdrop, ws.seed, ptr(ws.seed_dev), ptr(ws.dropmask), ptr(gh2), ptr(pb2), ptr(dW1p), ptr(gpool), ptr(dW1), ptr(db1),
             ptr(dg2), ptr(dbe2), ptr(dW2), ptr(db2), ptr(dg3), ptr(dbe3), None if wrt_unit else ptr(dWf),
             None if wrt_unit else ptr(dbf), stream())
    dsync, world = _sync_group(model) if getattr(ws, "sync_world", 1) > 1 else (None, 1)
    if dsync is None:
        call("explainn_fc_stack_bwd", *bargs)
    else:
        # sums of the BatchNorm backward passes, averaged over the ranks between producer and consumer (the kernels divide
        # by the local batch: mean of local sums / local batch = global sum / global batch)
        call("explainn_fc_stack_bwd_phase", 1, *bargs)
        _allreduce_avg(dsync, pb2.view(-1)[:2 * U], world)
        call("explainn_fc_stack_bwd_phase", 2, *bargs)
        _allreduce_avg(dsync, pb2, world)
        call("explainn_fc_stack_bwd_phase", 4, *bargs)
    if between is not None and between(ws, gpool, params, grads, wrt_unit) is False:
        return grads
    return train_backward_conv(model, ws, gpool, params, grads, wrt_unit)


def train_backward_conv(model, ws, gpool, params, grads, wrt_unit=False):
    """Second half of the backward: sparse conv / BatchNorm-1 gradients from d(pooled)."""
    if _is_variant(model):
        from . import variants
        return variants.train_backward_conv(model, ws, gpool, params, grads, wrt_unit)
    L = model.linears
    (cw, cb, g1, b1n) = params[:4]
    (dcw, dcb, dg1, db1n) = grads[:4]
    dWf, dbf = grads[-2], grads[-1]
    B, U, F = ws.B, ws.U, ws.F
    f32 = dict(dtype=torch.float32, device=gpool.device)
    dsync, world = _sync_group(model) if getattr(ws, "sync_world", 1) > 1 else (None, 1)
    nsplit = Fn._cabi.lib.explainn_conv_bwd_nsplit(U)          # one slice per SM: the tensor-core scatter
    if dsync is not None:
        # `partial` is all-reduced between the two phases: few slices (the shared-memory scatter kernel)
        sms = torch.cuda.get_device_properties(gpool.device).multi_processor_count
        nsplit = min(nsplit, max(1, -(-5 * sms // U)))
    partial = torch.empty((nsplit, U, 4 * F + 2), **f32)
    cargs = (ptr(ws.pk.codes), ptr(ws.pk.nmask), B, ws.pk.L, U, F, ptr(gpool), ptr(ws.pooled),
             ptr(ws.xhat), ptr(ws.idx), ptr(ws.M), ptr(ws.Mw), ptr(cb), ptr(ws.mean1), ptr(ws.var1), ptr(g1),
             float(L[1].eps), ptr(partial), nsplit, ptr(dcw), ptr(dcb), ptr(dg1), ptr(db1n), stream())
    if dsync is None:
        call("explainn_conv_pool_bwd", *cargs)
    else:
        n_pos = B * world * (ws.pk.L - F + 1)
        call("explainn_conv_pool_bwd_phase", 1, n_pos, *cargs)
        _allreduce_avg(dsync, partial, world)
        call("explainn_conv_pool_bwd_phase", 2, n_pos, *cargs)
    if wrt_unit:
        dWf.zero_()
        dbf.zero_()
    return grads


class _ExplaiNNTrainFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, x, wrt_unit, *params):
        ps = [p.detach().contiguous() for p in params]
        ws = train_forward(model, x, ps)
        ctx.model, ctx.ws, ctx.ps, ctx.wrt_unit = model, ws, ps, wrt_unit
        out = ws.unit if wrt_unit else ws.logits
        return out

    @staticmethod
    def backward(ctx, gout):
        grads = train_backward(ctx.model, ctx.ws, gout, ctx.ps, wrt_unit=ctx.wrt_unit)
        need = ctx.needs_input_grad[3:]
        out = [g if n else None for g, n in zip(grads, need)]
        if ctx.wrt_unit:
            out[-2] = out[-1] = None
        ctx.ws = None
        return (None, None, None) + tuple(out)


def forward_train(model, x):
    """model(x) in train() mode."""
    if torch.is_grad_enabled() and any(p.requires_grad for p in _params(model)):
        return _ExplaiNNTrainFn.apply(model, x, False, *_params(model))
    return train_forward(model, x).logits


def unit_outputs_train(model, x):
    """model.linears(x) in train() mode."""
    if torch.is_grad_enabled() and any(p.requires_grad for p in _params(model)):
        return _ExplaiNNTrainFn.apply(model, x, True, *_params(model))
    return train_forward(model, x).unit
