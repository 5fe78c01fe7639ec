"""Golden vectors for the num_fc = 0 and num_fc = 1 variants of the unit (explainn/models/networks.py:73-95), produced
by the UNMODIFIED reference.

    python tests/golden/make_golden_numfc.py          # rewrites tests/golden/numfc*.npz
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import HERE, make_codes, networks, onehot, randomise_bn, sd_to_np  # noqa: E402  (sets up the reference import)


def golden_eval(name, num_fc, U, L, C, F, n, seed, n_frac):
    torch.manual_seed(seed)
    model = networks.ExplaiNN(U, L, C, F, num_fc=num_fc)
    randomise_bn(model, seed + 1)
    model.eval()
    codes = make_codes(n, L, seed + 2, n_frac)
    x = onehot(codes)
    with torch.no_grad():
        logits = model(x)
        xr = x.repeat(1, U, 1)
        unit = model.linears(xr)
        pooled = model.linears[:4](xr)
    out = dict(num_fc=num_fc, U=U, L=L, C=C, F=F, codes=codes, logits=logits.numpy(), unit=unit.numpy(), pooled=pooled.numpy())
    out.update(sd_to_np(model.state_dict()))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "logits", logits.shape, "pooled", tuple(pooled.shape))


def golden_train(name, num_fc, U, L, C, F, B, seed, loss_kind, n_frac=0.02):
    torch.manual_seed(seed)
    model = networks.ExplaiNN(U, L, C, F, num_fc=num_fc)
    randomise_bn(model, seed + 1)
    model.train()
    codes = make_codes(B, L, seed + 2, n_frac)
    x = onehot(codes)
    g = torch.Generator().manual_seed(seed + 3)
    if loss_kind == "bce":
        y = (torch.rand(B, C, generator=g) < 0.5).float()
        crit = torch.nn.BCEWithLogitsLoss()
    else:
        y = torch.randn(B, C, generator=g)
        crit = torch.nn.MSELoss()
    out = dict(num_fc=num_fc, U=U, L=L, C=C, F=F, codes=codes, labels=y.numpy(), loss_kind=loss_kind)
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
        out.update(sd_to_np(model.state_dict(), "sd%d." % (step + 1)))
    out["losses"] = np.array(losses, np.float64)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "losses", losses)


if __name__ == "__main__":
    golden_eval("numfc0_eval_u40_f19", 0, 40, 100, 5, 19, n=64, seed=81, n_frac=0.02)
    golden_eval("numfc0_eval_u7_f8", 0, 7, 61, 2, 8, n=32, seed=82, n_frac=0.05)
    golden_eval("numfc1_eval_u40_f19", 1, 40, 100, 5, 19, n=64, seed=83, n_frac=0.02)
    golden_eval("numfc1_eval_u33_f24", 1, 33, 90, 4, 24, n=40, seed=84, n_frac=0.0)
    golden_train("numfc1_train_u6_bce", 1, 6, 60, 3, 19, B=16, seed=85, loss_kind="bce")
    golden_train("numfc1_train_u40_mse", 1, 40, 100, 5, 20, B=48, seed=86, loss_kind="mse")
