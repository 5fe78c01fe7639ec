"""CPU: the port's num_fc = 0 / 1 variants (networks.py:73-95) against golden vectors of the unmodified reference, and
the state_dict layout of the drop-in module for those variants."""
import numpy as np
import pytest
import torch

from oracle import torch_port as tp
from util import load_golden, onehot_from_codes, state_dict_from

EVAL = ["numfc0_eval_u40_f19", "numfc0_eval_u7_f8", "numfc1_eval_u40_f19", "numfc1_eval_u33_f24"]
TRAIN = ["numfc1_train_u6_bce", "numfc1_train_u40_mse"]


@pytest.mark.parametrize("name", EVAL)
def test_port_eval_matches_reference(name):
    g = load_golden(name)
    m = tp.PortExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), num_fc=int(g["num_fc"]))
    m.load_state_dict(state_dict_from(g))
    m.eval()
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    with torch.no_grad():
        assert np.array_equal(m(x).numpy(), g["logits"])
        assert np.array_equal(m.unit_outputs(x).numpy(), g["unit"])
        assert np.array_equal(m.pooled(x).numpy(), g["pooled"])


@pytest.mark.parametrize("name", TRAIN)
def test_port_train_step_matches_reference(name):
    g = load_golden(name)
    m = tp.PortExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), num_fc=int(g["num_fc"]))
    m.load_state_dict(state_dict_from(g, "sd0."))
    m.train()
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    y = torch.from_numpy(g["labels"])
    crit = torch.nn.BCEWithLogitsLoss() if str(g["loss_kind"]) == "bce" else torch.nn.MSELoss()
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    loss = tp.train_step(m, x, y, crit, opt)
    assert abs(float(loss.detach()) - g["losses"][0]) <= 1e-6 * abs(g["losses"][0])
    for k, v in m.state_dict().items():
        assert np.allclose(v.numpy(), g["sd1." + k], rtol=1e-5, atol=1e-7), k


@pytest.mark.parametrize("name", ["numfc0_eval_u7_f8", "numfc1_eval_u33_f24"])
def test_dropin_state_dict_layout(name):
    from explainn_b200 import networks
    g = load_golden(name)
    m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), num_fc=int(g["num_fc"]))
    sd = state_dict_from(g)
    assert list(m.state_dict().keys()) == list(sd.keys())
    assert all(tuple(m.state_dict()[k].shape) == tuple(v.shape) for k, v in sd.items())
    m.load_state_dict(sd)
    assert m._options["num_fc"] == int(g["num_fc"])
    with pytest.raises(NotImplementedError):
        networks.ExplaiNN(4, 100, 2, 25)                 # filter_size > 24: limit of the k-mer tables


VAR_EVAL = ["var_eval_fc3", "var_eval_fc5_pool44", "var_eval_fc2_pool53", "var_eval_fc1_pool29"]
VAR_TRAIN = ["var_train_fc0", "var_train_fc3", "var_train_fc4_pool44", "var_train_fc2_pool53", "var_train_fc1_pool92"]


def _var_port(g, prefix):
    m = tp.PortExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), pool_size=int(g["pool_size"]),
                        pool_stride=int(g["pool_stride"]), num_fc=int(g["num_fc"]))
    m.load_state_dict(state_dict_from(g, prefix))
    return m


@pytest.mark.parametrize("name", VAR_EVAL)
def test_port_variants_eval_match_reference(name):
    """num_fc > 2 and other pool sizes (networks.py:102,117-145): the port reproduces the reference bit for bit, and the
    drop-in module has the reference's state_dict layout (keys, order, shapes) for the variant."""
    from explainn_b200 import networks
    g = load_golden(name)
    m = _var_port(g, "sd.").eval()
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    with torch.no_grad():
        assert np.array_equal(m(x).numpy(), g["logits"])
        assert np.array_equal(m.unit_outputs(x).numpy(), g["unit"])
        assert np.array_equal(m.pooled(x).numpy(), g["pooled"])
    d = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), num_fc=int(g["num_fc"]),
                          pool_size=int(g["pool_size"]), pool_stride=int(g["pool_stride"]))
    sd = state_dict_from(g)
    assert list(d.state_dict().keys()) == list(sd.keys())
    assert all(tuple(d.state_dict()[k].shape) == tuple(v.shape) for k, v in sd.items())
    d.load_state_dict(sd)


@pytest.mark.parametrize("name", VAR_TRAIN)
def test_port_variants_train_step_matches_reference(name):
    g = load_golden(name)
    m = _var_port(g, "sd0.").train()
    for mod in m.modules():
        if isinstance(mod, torch.nn.Dropout):
            mod.p = 0.0
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    y = torch.from_numpy(g["labels"])
    crit = torch.nn.BCEWithLogitsLoss() if str(g["loss_kind"]) == "bce" else torch.nn.MSELoss()
    out = m(x)
    loss = crit(out, y)
    loss.backward()
    assert np.array_equal(out.detach().numpy(), g["logits"])
    for k, p in m.named_parameters():
        assert np.array_equal(p.grad.numpy(), g["grad." + k]), k
