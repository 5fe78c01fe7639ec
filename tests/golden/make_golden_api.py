"""Golden vectors for the driver-level drop-ins, made by running the UNMODIFIED reference (/root/reference):

  api_importance.npz   get_median_unit_importance / get_specific_unit_importance
                       (explainn/interpretation/interpretation.py:261-340) on a 100-unit, 50-class model -- the only
                       shape the reference's hard-coded loops accept
  api_train_loop.npz   train_explainn (explainn/train/train.py:13-142): three epochs, Adam, BCE, dropout probability
                       set to 0 (its Philox stream cannot be matched), epoch losses and the returned best weights

    python tests/golden/make_golden_api.py
"""
import os
import sys
import tempfile
import types

import numpy as np
import torch

for _m in ("h5py", "matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(_m, types.ModuleType(_m))
sys.path.insert(0, "/root/reference")
from explainn import networks, interpretation  # noqa: E402
from explainn.train import train as ref_train  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import make_codes, onehot, randomise_bn, sd_to_np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def golden_importance():
    U, L, C, F, n = 100, 80, 50, 19, 64
    torch.manual_seed(81)
    model = networks.ExplaiNN(U, L, C, F)
    randomise_bn(model, 82)
    model.eval()
    codes = make_codes(n, L, 83, 0.01)
    x = onehot(codes)
    with torch.no_grad():
        xr = x.repeat(1, U, 1)
        act16 = model.linears[:3](xr).numpy().astype(np.half)
        unit = model.linears(xr).numpy()
    labels = ["c%d" % i for i in range(C)]
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        med = interpretation.get_median_unit_importance(act16, model, unit, labels)
        spec = interpretation.get_specific_unit_importance(act16, model, unit, 3, labels)
    out = dict(U=U, L=L, C=C, F=F, codes=codes, median=med, specific_filt=3,
               specific=np.stack([np.asarray(spec[k]) for k in labels]))
    out.update(sd_to_np(model.state_dict()))
    np.savez_compressed(os.path.join(HERE, "api_importance.npz"), **out)
    print("api_importance", med.shape, "nan rows", int(np.isnan(med).any(axis=1).sum()), "specific", out["specific"].shape)


def golden_train_loop():
    U, L, C, F, n = 12, 70, 3, 19, 96
    torch.manual_seed(91)
    model = networks.ExplaiNN(U, L, C, F)
    model.linears[9].p = 0.0          # dropout off for the whole run (train_explainn calls model.train() itself)
    codes = make_codes(n, L, 92, 0.01)
    x = onehot(codes)
    y = (torch.rand(n, C, generator=torch.Generator().manual_seed(93)) < 0.5).float()
    ds = torch.utils.data.TensorDataset(x, y)
    tl = torch.utils.data.DataLoader(ds, batch_size=32, shuffle=False)
    vl = torch.utils.data.DataLoader(ds, batch_size=48, shuffle=False)
    out = dict(U=U, L=L, C=C, F=F, codes=codes, labels=y.numpy())
    out.update(sd_to_np(model.state_dict(), "sd0."))
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    with tempfile.TemporaryDirectory() as d:
        m, tr, te = ref_train.train_explainn(tl, vl, model, "cpu", torch.nn.BCEWithLogitsLoss(), opt, num_epochs=3,
                                             weights_folder=d, name_ind="g", verbose=False, checkpoint=1, patience=0)
        files = sorted(os.listdir(d))
    out["train_error"] = np.array(tr, np.float64)
    out["test_error"] = np.array(te, np.float64)
    out["files"] = np.array(files)
    out.update(sd_to_np(m.state_dict(), "sd_final."))
    np.savez_compressed(os.path.join(HERE, "api_train_loop.npz"), **out)
    print("api_train_loop", tr, te, files)


if __name__ == "__main__":
    golden_importance()
    golden_train_loop()
