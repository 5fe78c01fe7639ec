"""Training-mode forward / backward of the model variants beside the default unit: any MaxPool1d(pool_size,
pool_stride), num_fc = 0 (global max-pool, explainn/models/networks.py:73-80) and num_fc > 2 (extra hidden layers,
networks.py:117-145).  Same contract as autograd_fn.train_forward / train_backward (which keep the default pool with
num_fc = 1, 2): a workspace object carries what the backward needs; parameters and gradients are flat lists in
state_dict order.

The conv stage of these variants materialises the (B,U,S) map of x_hat (table kernel) and pools it with
`explainn_pool_general`; the hidden layers are `explainn_hidden_*` blocks (csrc/generic.cu); the last hidden layer plus
`last_linear` reuse the default FC-stack kernels with P = 100.
"""
import torch

from . import functional as Fn
from ._cabi import call, ptr, stream

H = Fn.HIDDEN


class Workspace:
    pass


def layout(model):
    """Module handles of a variant, in state_dict order: (conv, bn1, fcs, bns, drops, last_fc, last_bn).
    fcs[i] / bns[i]: hidden FC layers (P -> 100, then 100 -> 100); drops[i]: the Dropout BEHIND hidden layer i (the one
    the reference puts at the head of the next layer); last_fc / last_bn: the 100 -> 1 layer (None for num_fc < 2)."""
    L = model.linears
    nfc = model._options["num_fc"]
    if nfc == 0:
        return L[0], L[1], [], [], [], None, None
    if nfc == 1:
        return L[0], L[1], [], [], [], L[6], L[7]
    if nfc == 2:
        return L[0], L[1], [L[6]], [L[7]], [L[9]], L[10], L[11]
    fcs, bns, drops = [L[6]], [L[7]], []
    for blk in model.linears_bg:
        drops.append(blk[0])
        fcs.append(blk[1])
        bns.append(blk[2])
    drops.append(model.last_linear[0])
    return L[0], L[1], fcs, bns, drops, model.last_linear[1], model.last_linear[2]


def params(model):
    conv, bn1, fcs, bns, drops, last_fc, last_bn = layout(model)
    out = [conv.weight, conv.bias, bn1.weight, bn1.bias]
    for fc, bn in zip(fcs, bns):
        out += [fc.weight, fc.bias, bn.weight, bn.bias]
    if last_fc is not None:
        out += [last_fc.weight, last_fc.bias, last_bn.weight, last_bn.bias]
    out += [model.final.weight, model.final.bias]
    return tuple(out)


def check_modes(model):
    conv, bn1, fcs, bns, drops, last_fc, last_bn = layout(model)
    all_bn = [bn1] + bns + ([last_bn] if last_bn is not None else [])
    if not all(bn.training for bn in all_bn):
        raise NotImplementedError("explainn_b200: in train() mode all BatchNorm layers must be in training mode")
    for bn in all_bn:
        if bn.momentum is None or not bn.track_running_stats or not bn.affine:
            raise NotImplementedError("explainn_b200: BatchNorm layers must keep the reference configuration "
                                      "(momentum=0.1, affine, track_running_stats)")


def _pdrop(mod):
    return float(mod.p) if mod.training else 0.0


def train_forward(model, x, ps, seed, seed_dev=None):
    from . import autograd_fn
    check_modes(model)
    conv, bn1, fcs, bns, drops, last_fc, last_bn = layout(model)
    o = model._options
    U, F, C, nfc = o["num_cnns"], o["filter_size"], o["num_classes"], o["num_fc"]
    pk = model._pack(x)
    B = pk.B
    if B < 2:
        raise ValueError("Expected more than 1 value per channel when training, got input size %s" % (tuple(x.shape),))
    dev = pk.codes.device
    f32 = dict(dtype=torch.float32, device=dev)
    S = pk.L - F + 1
    ws = Workspace()
    ws.pk, ws.B, ws.U, ws.C, ws.F, ws.nfc = pk, B, U, C, F, nfc
    ws.psize, ws.pstride = (S, S) if nfc == 0 else (o["pool_size"], o["pool_stride"])
    cw, cb, g1, b1n = ps[:4]
    ws.M = Fn.input_cooc_stats(pk, F)
    ws.mean1, ws.var1, ws.Mw = Fn.bn1_batch_stats(ws.M, cw, cb, B * S, bn1.momentum, bn1.running_mean, bn1.running_var,
                                                  bn1.num_batches_tracked)
    ones, zeros = autograd_fn._const(dev, U)
    # tables with gamma = 1, beta = 0 and the batch statistics: the scan yields x_hat = (y - mu) / sigma
    prep = Fn.conv_prepare(cw, cb, ones, zeros, ws.mean1, ws.var1, bn1.eps)
    z = Fn.unit_activations(pk, prep, preact=True)
    ws.pooled, ws.xhat, ws.idx16 = Fn.pool_general(z, ws.psize, ws.pstride, g1, b1n)
    del z
    P = ws.P = ws.pooled.shape[2]
    Wf, bf = ps[-2], ps[-1]
    ws.seed, ws.seed_dev = seed, seed_dev
    ws.nchunk = Fn.fc_nchunk(B)
    ws.blocks = []
    if nfc == 0:
        ws.unit = ws.pooled.reshape(B, U)
        ws.logits = Fn.linear_fwd(ws.unit, Wf, bf)
        return ws
    if nfc == 1:
        W1, b1, g3, be3 = ps[4:8]
        ws.h2 = torch.empty((U, B), **f32)
        part3 = torch.empty((U, ws.nchunk, 2), **f32)
        ws.mean3, ws.invstd3 = torch.empty(U, **f32), torch.empty(U, **f32)
        ws.unit, ws.logits = torch.empty((B, U), **f32), torch.empty((B, C), **f32)
        call("explainn_fc1_stack_fwd_train", ptr(ws.pooled), B, U, P, C, ptr(W1), ptr(b1), ptr(g3), ptr(be3), ptr(last_bn.running_mean),
             ptr(last_bn.running_var), float(last_bn.momentum), float(last_bn.eps), ptr(Wf), ptr(bf), ptr(ws.h2), ptr(part3),
             ptr(ws.mean3), ptr(ws.invstd3), ptr(ws.unit), ptr(ws.logits), stream())
        last_bn.num_batches_tracked.add_(1)
        return ws
    # hidden blocks 0 .. n-2 are materialised; the last hidden layer runs inside the FC-stack kernels
    X, Pin = ws.pooled, P
    n = len(fcs)
    for i in range(n - 1):
        W, b, g, be = ps[4 + 4 * i: 8 + 4 * i]
        blk = Workspace()
        blk.X, blk.Pin, blk.pdrop = X, Pin, _pdrop(drops[i])
        blk.seed = (seed + 0x632BE59BD9B4E019 * (i + 1)) & 0xFFFFFFFFFFFFFFFF
        blk.Y = torch.empty((U, B, H), **f32)
        blk.mean, blk.invstd = torch.empty(U * H, **f32), torch.empty(U * H, **f32)
        out = torch.empty((B, U, H), **f32)
        bn = bns[i]
        call("explainn_hidden_fwd_train", ptr(X), B, U, Pin, ptr(W), ptr(b), ptr(g), ptr(be), ptr(bn.running_mean), ptr(bn.running_var),
             float(bn.momentum), float(bn.eps), blk.pdrop, blk.seed, ptr(seed_dev), ptr(blk.Y), ptr(blk.mean), ptr(blk.invstd), ptr(out),
             stream())
        bn.num_batches_tracked.add_(1)
        ws.blocks.append(blk)
        X, Pin = out, H
    W1, b1, g2, be2, W2, b2, g3, be3 = ps[4 + 4 * (n - 1): 12 + 4 * (n - 1)]
    bn2 = bns[n - 1]
    ws.X, ws.Pin = X, Pin
    ws.pdrop = _pdrop(drops[n - 1])
    ws.h1 = torch.empty((U, B, H), **f32)
    part2 = torch.empty((U, ws.nchunk, 2, H), **f32)
    ws.mean2, ws.invstd2 = torch.empty(U * H, **f32), torch.empty(U * H, **f32)
    ws.h2 = torch.empty((U, B), **f32)
    part3 = torch.empty((U, ws.nchunk, 2), **f32)
    ws.mean3, ws.invstd3 = torch.empty(U, **f32), torch.empty(U, **f32)
    ws.unit, ws.logits = torch.empty((B, U), **f32), torch.empty((B, C), **f32)
    ws.dropmask = torch.empty((U, B, 4), dtype=torch.int32, device=dev) if ws.pdrop > 0 else None
    call("explainn_fc_stack_fwd_train", ptr(X), B, U, Pin, C, ptr(W1), ptr(b1), ptr(g2), ptr(be2), ptr(bn2.running_mean),
         ptr(bn2.running_var), ptr(W2), ptr(b2), ptr(g3), ptr(be3), ptr(last_bn.running_mean), ptr(last_bn.running_var),
         float(bn2.momentum), float(bn2.eps), ws.pdrop, seed, ptr(seed_dev), ptr(ws.dropmask), ptr(Wf), ptr(bf), ptr(ws.h1), ptr(part2),
         ptr(ws.mean2), ptr(ws.invstd2), ptr(ws.h2), ptr(part3), ptr(ws.mean3), ptr(ws.invstd3), ptr(ws.unit), ptr(ws.logits), stream())
    bn2.num_batches_tracked.add_(1)
    last_bn.num_batches_tracked.add_(1)
    return ws


def train_backward(model, ws, gout, ps, grads=None, wrt_unit=False, between=None):
    """gout: (B,C) gradient w.r.t. the logits, or (B,U) w.r.t. the unit outputs when `wrt_unit`."""
    conv, bn1, fcs, bns, drops, last_fc, last_bn = layout(model)
    B, U, P, C, F, nfc = ws.B, ws.U, ws.P, ws.C, ws.F, ws.nfc
    dev = gout.device
    f32 = dict(dtype=torch.float32, device=dev)
    if grads is None:
        grads = [torch.empty_like(p) for p in ps]
    gout = gout.contiguous()
    Wf = ps[-2]
    dWf, dbf = grads[-2], grads[-1]
    if nfc == 0:
        if wrt_unit:
            gpool = gout.reshape(B, U, 1).clone()
        else:
            gunit = torch.empty((B, U), **f32)
            call("explainn_linear_bwd", ptr(gout), ptr(ws.unit), ptr(Wf), B, U, C, ptr(gunit), ptr(dWf), ptr(dbf), stream())
            gpool = gunit.reshape(B, U, 1)
    elif nfc == 1:
        W1, g3 = ps[4], ps[6]
        dW1, db1, dg3, dbe3 = grads[4:8]
        gh2 = torch.empty((U, B), **f32)
        dWp = torch.empty((ws.nchunk, U, P), **f32)
        gpool = torch.empty((B, U, P), **f32)
        call("explainn_fc1_stack_bwd", ptr(gout), B, U, P, C, ptr(ws.pooled), ptr(ws.h2), ptr(ws.unit), ptr(W1), ptr(g3), ptr(ws.mean3),
             ptr(ws.invstd3), None if wrt_unit else ptr(Wf), ptr(gh2), ptr(dWp), ptr(gpool), ptr(dW1), ptr(db1), ptr(dg3), ptr(dbe3),
             None if wrt_unit else ptr(dWf), None if wrt_unit else ptr(dbf), stream())
    else:
        n = len(fcs)
        o0 = 4 + 4 * (n - 1)
        W1, b1, g2, be2, W2, b2, g3, be3 = ps[o0:o0 + 8]
        dW1, db1, dg2, dbe2, dW2, db2, dg3, dbe3 = grads[o0:o0 + 8]
        Pin = ws.Pin
        gh2 = torch.empty((U, B), **f32)
        pb2 = torch.empty((U, ws.nchunk, 3, H), **f32)
        dW1p = torch.empty((ws.nchunk, U, H, Pin), **f32)
        gX = torch.empty((B, U, Pin), **f32)
        call("explainn_fc_stack_bwd", ptr(gout), B, U, Pin, C, ptr(ws.X), ptr(ws.h1), ptr(ws.h2), ptr(ws.unit), ptr(W1), ptr(g2), ptr(be2),
             ptr(ws.mean2), ptr(ws.invstd2), ptr(W2), ptr(g3), ptr(ws.mean3), ptr(ws.invstd3), None if wrt_unit else ptr(Wf), ws.pdrop,
             ws.seed, ptr(ws.seed_dev), ptr(ws.dropmask), ptr(gh2), ptr(pb2), ptr(dW1p), ptr(gX), ptr(dW1), ptr(db1), ptr(dg2), ptr(dbe2),
             ptr(dW2), ptr(db2), ptr(dg3), ptr(dbe3), None if wrt_unit else ptr(dWf), None if wrt_unit else ptr(dbf), stream())
        for i in range(n - 2, -1, -1):
            blk = ws.blocks[i]
            W, b, g, be = ps[4 + 4 * i: 8 + 4 * i]
            dW, db, dg, dbe = grads[4 + 4 * i: 8 + 4 * i]
            sums = torch.empty(2 * U * H, **f32)
            G = torch.empty((U, B, H), **f32)
            gin = torch.empty((B, U, blk.Pin), **f32)
            call("explainn_hidden_bwd", ptr(gX), B, U, blk.Pin, ptr(blk.X), ptr(blk.Y), ptr(W), ptr(g), ptr(be), ptr(blk.mean),
                 ptr(blk.invstd), blk.pdrop, blk.seed, ptr(ws.seed_dev), ptr(sums), ptr(G), ptr(gin), ptr(dW), ptr(db), ptr(dg), ptr(dbe),
                 stream())
            gX = gin
        gpool = gX
    if wrt_unit:
        dWf.zero_()
        dbf.zero_()
    if between is not None and between(ws, gpool, ps, grads, wrt_unit) is False:
        return grads
    return train_backward_conv(model, ws, gpool, ps, grads)


def train_backward_conv(model, ws, gpool, ps, grads, wrt_unit=False):
    cw, cb, g1, b1n = ps[:4]
    dcw, dcb, dg1, db1n = grads[:4]
    B, U, F = ws.B, ws.U, ws.F
    nsplit = Fn._cabi.lib.explainn_conv_bwd_nsplit(U)
    partial = torch.empty((nsplit, U, 4 * F + 2), dtype=torch.float32, device=gpool.device)
    call("explainn_conv_pool_bwd_general", ptr(ws.pk.codes), ptr(ws.pk.nmask), B, ws.pk.L, U, F, int(ws.psize), int(ws.pstride),
         ptr(gpool.contiguous()), ptr(ws.pooled), ptr(ws.xhat), ptr(ws.idx16), ptr(ws.M), ptr(ws.Mw), ptr(cb), ptr(ws.mean1),
         ptr(ws.var1), ptr(g1), float(model.linears[1].eps), ptr(partial), nsplit, ptr(dcw), ptr(dcb), ptr(dg1), ptr(db1n), stream())
    return grads
