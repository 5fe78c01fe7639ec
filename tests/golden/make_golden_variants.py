"""Golden vectors for the remaining constructor variants of the unit (explainn/models/networks.py:48-49,73-80,102,117-145):
num_fc = 0 in training mode, num_fc > 2, and MaxPool1d sizes / strides other than 7 (overlapping and skipping windows),
produced by the UNMODIFIED reference.  Dropout probabilities are set to 0 in the training cases (Philox stream).

    python tests/golden/make_golden_variants.py          # rewrites tests/golden/var_*.npz
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import HERE, make_codes, networks, onehot, randomise_bn, sd_to_np  # noqa: E402  (sets up the reference import)


def build(U, L, C, F, num_fc, ps, st, seed):
    torch.manual_seed(seed)
    model = networks.ExplaiNN(U, L, C, F, num_fc=num_fc, pool_size=ps, pool_stride=st)
    randomise_bn(model, seed + 1)
    return model


def golden_eval(name, num_fc, ps, st, U, L, C, F, n, seed, n_frac=0.02):
    model = build(U, L, C, F, num_fc, ps, st, seed).eval()
    codes = make_codes(n, L, seed + 2, n_frac)
    x = onehot(codes)
    with torch.no_grad():
        logits = model(x)
        xr = x.repeat(1, U, 1)
        unit = model.linears(xr)
        pooled = model.linears[:4](xr)
    out = dict(num_fc=num_fc, pool_size=ps, pool_stride=st, U=U, L=L, C=C, F=F, codes=codes, logits=logits.numpy(),
               unit=unit.numpy(), pooled=pooled.numpy())
    out.update(sd_to_np(model.state_dict()))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "logits", tuple(logits.shape), "linears(x)", tuple(unit.shape), "pooled", tuple(pooled.shape))


def golden_train(name, num_fc, ps, st, U, L, C, F, B, seed, loss_kind, n_frac=0.02):
    model = build(U, L, C, F, num_fc, ps, st, seed).train()
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    codes = make_codes(B, L, seed + 2, n_frac)
    x = onehot(codes)
    g = torch.Generator().manual_seed(seed + 3)
    if loss_kind == "bce":
        y = (torch.rand(B, C, generator=g) < 0.5).float()
        crit = torch.nn.BCEWithLogitsLoss()
    else:
        y = torch.randn(B, C, generator=g)
        crit = torch.nn.MSELoss()
    out = dict(num_fc=num_fc, pool_size=ps, pool_stride=st, U=U, L=L, C=C, F=F, codes=codes, labels=y.numpy(), loss_kind=loss_kind)
    out.update(sd_to_np(model.state_dict(), "sd0."))
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    losses = []
    for step in range(2):
        opt.zero_grad()
        logits = model(x)
        loss = crit(logits, y)
        loss.backward()
        if step == 0:
            out["logits"] = logits.detach().numpy()
            for k, p in model.named_parameters():
                out["grad." + k] = p.grad.detach().numpy().copy()
        opt.step()
        losses.append(float(loss))
        sd = sd_to_np(model.state_dict(), "sd%d." % (step + 1))
        if step == 0:      # after the first step only the BatchNorm buffers are compared
            sd = {k: v for k, v in sd.items() if "running" in k or "num_batches" in k}
        out.update(sd)
    out["losses"] = np.array(losses, np.float64)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "losses", losses)


if __name__ == "__main__":
    golden_eval("var_eval_fc3", 3, 7, 7, 6, 90, 4, 19, n=40, seed=101)
    golden_eval("var_eval_fc5_pool44", 5, 4, 4, 3, 70, 3, 19, n=32, seed=102)
    golden_eval("var_eval_fc2_pool53", 2, 5, 3, 33, 80, 5, 20, n=40, seed=103)
    golden_eval("var_eval_fc1_pool29", 1, 2, 9, 10, 75, 2, 8, n=32, seed=104, n_frac=0.05)
    golden_train("var_train_fc0", 0, 7, 7, 20, 90, 4, 19, B=32, seed=111, loss_kind="bce")
    golden_train("var_train_fc3", 3, 7, 7, 5, 90, 4, 19, B=32, seed=112, loss_kind="bce")
    golden_train("var_train_fc4_pool44", 4, 4, 4, 3, 70, 3, 19, B=24, seed=113, loss_kind="mse")
    golden_train("var_train_fc2_pool53", 2, 5, 3, 10, 80, 5, 20, B=32, seed=114, loss_kind="mse")
    golden_train("var_train_fc1_pool92", 1, 9, 2, 10, 75, 2, 8, B=24, seed=115, loss_kind="bce", n_frac=0.05)
