"""Generate golden vectors by running the UNMODIFIED reference (/root/reference) in the build
container.  The reference is imported with empty stand-ins for h5py / matplotlib (its
explainn/utils/tools.py imports them at module level; nothing on the hot path uses them).

    python tests/golden/make_golden.py          # rewrites tests/golden/*.npz

The GPU box has no /root/reference, so tests only ever read the committed .npz files.
Sequences are stored as uint8 base codes (0..3 = A,C,G,T; 4 = N / all-zero column).
"""
import os
import sys
import types

import numpy as np
import torch

for _m in ("h5py", "matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(_m, types.ModuleType(_m))
sys.path.insert(0, "/root/reference")
from explainn import networks, interpretation  # noqa: E402
from explainn.train import train as ref_train  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def make_codes(n, L, seed, n_frac=0.0, pad=0):
    g = np.random.default_rng(seed)
    c = g.integers(0, 4, size=(n, L)).astype(np.uint8)
    if n_frac > 0:
        c[g.random((n, L)) < n_frac] = 4
    if pad:  # centre-padding with N as scripts/utils/__init__.py:100-108 does for short sequences
        for i in range(0, n, 5):
            k = int(g.integers(1, pad))
            c[i, :k] = 4
            c[i, L - k:] = 4
    return c


def onehot(codes):
    n, L = codes.shape
    x = np.zeros((n, 4, L), np.float32)
    for c in range(4):
        x[:, c, :] = (codes == c)
    return torch.from_numpy(x)


def sd_to_np(sd, prefix="sd."):
    return {prefix + k: v.detach().cpu().numpy().copy() for k, v in sd.items()}


def randomise_bn(model, seed):
    """Give BatchNorm non-trivial affine parameters and running statistics."""
    g = torch.Generator().manual_seed(seed)
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm1d):
            m.weight.data = 0.5 + torch.rand(m.weight.shape, generator=g)
            m.weight.data[::7] *= -1.0          # a few negative gammas
            m.bias.data = 0.2 * torch.randn(m.bias.shape, generator=g)
            m.running_mean.data = 0.1 * torch.randn(m.running_mean.shape, generator=g)
            m.running_var.data = 0.5 + torch.rand(m.running_var.shape, generator=g)


def golden_eval(name, U, L, C, F, n, seed, sd=None, n_frac=0.01, pad=0, n_act=8):
    torch.manual_seed(seed)
    model = networks.ExplaiNN(U, L, C, F)
    if sd is not None:
        model.load_state_dict(sd)
    else:
        randomise_bn(model, seed + 1)
    model.eval()
    codes = make_codes(n, L, seed + 2, n_frac, pad)
    x = onehot(codes)
    with torch.no_grad():
        logits = model(x)
        xr = x.repeat(1, U, 1)
        unit = model.linears(xr)
        act = model.linears[:3](xr)
        pooled = model.linears[:4](xr)
    act16 = act.numpy().astype(np.half)
    pwm = interpretation.get_pwms_explainn(act16, x, F)
    S = L - F + 1
    P = pooled.shape[2]
    win = act[:, :, :P * 7].reshape(n, U, P, 7)
    out = dict(U=U, L=L, C=C, F=F, codes=codes, logits=logits.numpy(), unit=unit.numpy(),
               pooled=pooled[:n_act].numpy(), argmax=win.argmax(-1).numpy().astype(np.uint8),
               act16=act16[:n_act], amax16=np.amax(act16, axis=(0, 2)), pwm=pwm)
    out.update(sd_to_np(model.state_dict()))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "logits", logits.shape, "hits>thr", int((act16 > 0.5 * np.amax(act16, axis=(0, 2))[None, :, None]).sum()))


def golden_train(name, U, L, C, F, B, seed, loss_kind, n_frac=0.02):
    torch.manual_seed(seed)
    model = networks.ExplaiNN(U, L, C, F)
    randomise_bn(model, seed + 1)
    model.train()
    model.linears[9].eval()     # dropout off: its Philox stream cannot be matched (SURVEY 7.3 e)
    codes = make_codes(B, L, seed + 2, n_frac)
    x = onehot(codes)
    g = torch.Generator().manual_seed(seed + 3)
    if loss_kind == "bce":
        y = (torch.rand(B, C, generator=g) < 0.5).float()
        crit = torch.nn.BCEWithLogitsLoss()
    else:
        y = torch.randn(B, C, generator=g)
        crit = torch.nn.MSELoss()
    out = dict(U=U, L=L, C=C, F=F, codes=codes, labels=y.numpy(), loss_kind=loss_kind)
    out.update(sd_to_np(model.state_dict(), "sd0."))
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    losses = []
    for step in range(2):       # two optimisation steps, train.py:59-67
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


def main():
    ck = torch.load("/root/reference/data/test/weights/model_epoch_best_4.pth", map_location="cpu")
    # README TF-binding model (cfg 1) with the shipped trained weights
    golden_eval("eval_cfg1_ckpt", 100, 200, 3, 19, n=384, seed=11, sd=ck, n_frac=0.01, pad=40, n_act=4)
    # random-init models: partial unit tile, F=20 (paper notebooks), F not multiple of 4, short F
    golden_eval("eval_u40_f20", 40, 100, 5, 20, n=96, seed=21, n_frac=0.02, pad=20, n_act=8)
    golden_eval("eval_u7_f8", 7, 64, 2, 8, n=48, seed=31, n_frac=0.05, n_act=8)
    golden_eval("eval_u33_f24", 33, 90, 4, 24, n=40, seed=41, n_frac=0.0, n_act=8)
    golden_train("train_u6_bce", 6, 60, 3, 19, B=16, seed=51, loss_kind="bce")
    golden_train("train_u40_mse", 40, 100, 5, 20, B=48, seed=61, loss_kind="mse")
    golden_train("train_u33_bce", 33, 75, 81, 11, B=32, seed=71, loss_kind="bce", n_frac=0.0)


if __name__ == "__main__":
    main()
