"""GPU parity of the num_fc = 0 and num_fc = 1 variants of the unit (explainn/models/networks.py:73-95)
against golden vectors of the unmodified reference and the CPU port."""
import numpy as np
import pytest
import torch

from util import load_golden, onehot_from_codes, state_dict_from

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
EVAL = ["numfc0_eval_u40_f19", "numfc0_eval_u7_f8", "numfc1_eval_u40_f19", "numfc1_eval_u33_f24"]
TRAIN = ["numfc1_train_u6_bce", "numfc1_train_u40_mse"]
ZERO_GRAD = ("linears.0.bias", "linears.6.bias")      # biases in front of a train-mode BatchNorm
# BatchNorm-1 beta scales every pooled value of a unit by exp(beta); the train-mode BatchNorm behind the linear FC removes
# that scale up to its eps, so the gradient is ~1e-4 of its siblings: compared on the scale of d gamma (cf. test_gpu_train.py)
NOISE_GRAD = ("linears.1.bias",)


def _close(got, ref, tol, what):
    scale = max(float(np.abs(ref).max()), 1e-30)
    err = float(np.abs(got - ref).max())
    assert err <= tol * scale, "%s: max abs err %.3g vs scale %.3g" % (what, err, scale)


@pytest.mark.parametrize("backend", ["tc", "lut"])
@pytest.mark.parametrize("name", EVAL)
def test_eval_forward_matches_reference(name, backend):
    from explainn_b200 import networks
    g = load_golden(name)
    m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), num_fc=int(g["num_fc"]))
    m.load_state_dict(state_dict_from(g))
    m = m.to(DEV).eval()
    m.conv_backend = backend
    x = torch.from_numpy(onehot_from_codes(g["codes"])).to(DEV)
    with torch.no_grad():
        logits = m(x)
        unit = m.linears(x.repeat(1, int(g["U"]), 1))          # the reference's calling convention (interpretation.py:94-95)
        pooled = m.linears[:4](x)
    assert pooled.shape == g["pooled"].shape
    _close(pooled.cpu().numpy(), g["pooled"], 1e-5, "pooled")
    _close(unit.cpu().numpy(), g["unit"], 1e-4, "unit outputs")
    _close(logits.cpu().numpy(), g["logits"], 1e-4, "logits")       # north-star fp32 tolerance
    # empty batch
    assert m(x[:0]).shape == (0, int(g["C"]))


def test_num_fc0_trains_raise_cleanly():
    from explainn_b200 import networks
    m = networks.ExplaiNN(8, 80, 2, 19, num_fc=0).to(DEV).train()
    with pytest.raises(NotImplementedError, match="num_fc"):
        m(torch.zeros(4, 4, 80, device=DEV))


@pytest.mark.parametrize("name", TRAIN)
def test_num_fc1_train_forward_backward_vs_reference(name):
    from explainn_b200 import networks
    g = load_golden(name)
    m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), num_fc=1)
    m.load_state_dict(state_dict_from(g, "sd0."))
    m = m.to(DEV).train()
    x = torch.from_numpy(onehot_from_codes(g["codes"])).to(DEV)
    y = torch.from_numpy(g["labels"]).to(DEV)
    crit = torch.nn.BCEWithLogitsLoss() if str(g["loss_kind"]) == "bce" else torch.nn.MSELoss()
    logits = m(x)
    loss = crit(logits, y)
    loss.backward()
    _close(logits.detach().cpu().numpy(), g["logits"], 1e-4, "train logits")
    assert abs(float(loss.detach()) - g["losses"][0]) <= 1e-5 * abs(g["losses"][0])
    for k, p in m.named_parameters():
        got, ref = p.grad.cpu().numpy(), g["grad." + k]
        if k in ZERO_GRAD:
            assert np.abs(got).max() == 0.0 and np.abs(ref).max() < 1e-6, k
        elif k in NOISE_GRAD:
            dgamma = np.abs(g["grad.linears.1.weight"]).max()
            assert np.abs(got).max() < 1e-3 * dgamma and np.abs(got - ref).max() < 5e-5 * dgamma, k
        else:
            # max-pool / ReLU kinks: a pre-activation within rounding of a tie may fall on the other side, which moves
            # single elements; bound the bulk tightly and the worst element loosely (see test_gpu_train.py)
            d = np.abs(got - ref)
            scale = np.abs(ref).max()
            assert np.quantile(d, 0.98) <= 1e-4 * scale, k
            assert d.max() <= 2e-2 * scale, k
    sd1 = state_dict_from(g, "sd1.")
    for k in ("linears.1.running_mean", "linears.1.running_var", "linears.7.running_mean", "linears.7.running_var"):
        _close(m.state_dict()[k].cpu().numpy(), sd1[k].numpy(), 1e-4, k)


@pytest.mark.parametrize("name", TRAIN)
def test_num_fc1_fused_trainer_two_steps(name):
    """train.py:59-67 twice through the flat-buffer trainer (CUDA-graph replay on the second step)."""
    from explainn_b200 import networks, train as xtrain
    g = load_golden(name)
    m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), num_fc=1)
    m.load_state_dict(state_dict_from(g, "sd0."))
    m = m.to(DEV).train()
    x = torch.from_numpy(onehot_from_codes(g["codes"])).to(DEV)
    y = torch.from_numpy(g["labels"]).to(DEV)
    crit = torch.nn.BCEWithLogitsLoss() if str(g["loss_kind"]) == "bce" else torch.nn.MSELoss()
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    assert xtrain.FusedTrainer.supports(m, crit, opt)
    tr = xtrain.FusedTrainer(m, crit, opt)
    losses = [float(tr.step(x, y).item()) for _ in range(2)]
    assert abs(losses[0] - g["losses"][0]) <= 1e-5 * abs(g["losses"][0])
    assert abs(losses[1] - g["losses"][1]) <= 2e-3 * abs(g["losses"][1])
    sd2 = state_dict_from(g, "sd2.")
    sd0 = state_dict_from(g, "sd0.")
    for k, p in m.named_parameters():
        if k in ZERO_GRAD or k in NOISE_GRAD:
            continue
        upd = p.detach().cpu().numpy() - sd0[k].numpy()
        upd_ref = sd2[k].numpy() - sd0[k].numpy()
        # Adam's first steps are ~ lr * sign(g): bound every element by two steps and the bulk tightly
        d = np.abs(upd - upd_ref)
        assert d.max() <= 4.2e-3, k
        assert np.mean(d > 2e-4) <= 0.05, (k, float(np.mean(d > 2e-4)))
