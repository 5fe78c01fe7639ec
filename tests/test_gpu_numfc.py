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


# ---- the remaining constructor variants: num_fc = 0 in training, num_fc > 2, MaxPool1d other than (7, 7) ---------------
VAR_EVAL = ["var_eval_fc3", "var_eval_fc5_pool44", "var_eval_fc2_pool53", "var_eval_fc1_pool29"]
VAR_TRAIN = ["var_train_fc0", "var_train_fc3", "var_train_fc4_pool44", "var_train_fc2_pool53", "var_train_fc1_pool92"]


def _var_model(g, prefix):
    from explainn_b200 import networks
    m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]), num_fc=int(g["num_fc"]),
                          pool_size=int(g["pool_size"]), pool_stride=int(g["pool_stride"]))
    m.load_state_dict(state_dict_from(g, prefix))
    return m.to(DEV)


@pytest.mark.parametrize("name", VAR_EVAL)
def test_variants_eval_forward_matches_reference(name):
    """networks.py:102,117-145 in eval mode: pooled activations, `model.linears(x)` (for num_fc > 2 the reference's
    `linears` ends behind the first hidden layer: shape (B, 100U, 1)) and logits against the reference."""
    g = load_golden(name)
    m = _var_model(g, "sd.").eval()
    x = torch.from_numpy(onehot_from_codes(g["codes"])).to(DEV)
    with torch.no_grad():
        logits = m(x)
        unit = m.linears(x.repeat(1, int(g["U"]), 1))
        pooled = m.linears[:4](x)
    assert pooled.shape == g["pooled"].shape and unit.shape == g["unit"].shape
    _close(pooled.cpu().numpy(), g["pooled"], 1e-5, "pooled")
    _close(unit.cpu().numpy(), g["unit"], 1e-4, "linears(x)")
    _close(logits.cpu().numpy(), g["logits"], 1e-4, "logits")
    assert m(x[:0]).shape == (0, int(g["C"]))


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("name", VAR_TRAIN)
def test_variants_train_step_vs_reference(name, fused):
    """One training step (dropout probability 0): logits, loss, every gradient and the BatchNorm running statistics
    against the reference; with `fused` the same step through the flat-buffer trainer (loss and updated weights)."""
    from explainn_b200 import train as xtrain
    g = load_golden(name)
    m = _var_model(g, "sd0.").train()
    for mod in m.modules():
        if isinstance(mod, torch.nn.Dropout):
            mod.p = 0.0
    x = torch.from_numpy(onehot_from_codes(g["codes"])).to(DEV)
    y = torch.from_numpy(g["labels"]).to(DEV)
    crit = torch.nn.BCEWithLogitsLoss() if str(g["loss_kind"]) == "bce" else torch.nn.MSELoss()
    # a bias in front of a train-mode BatchNorm has an exactly zero gradient (PyTorch returns rounding noise)
    zero_grad = [k for k, _ in m.named_parameters() if k.endswith(".bias") and not k.startswith("final")
                 and isinstance(dict(m.named_modules())[k.rsplit(".", 1)[0]], torch.nn.Conv1d)]
    if int(g["num_fc"]) == 0:
        zero_grad = ["linears.0.bias"]
    if fused:
        opt = torch.optim.Adam(m.parameters(), lr=1e-3)
        assert xtrain.FusedTrainer.supports(m, crit, opt)
        tr = xtrain.FusedTrainer(m, crit, opt)
        losses = [float(tr.step(x, y).item()) for _ in range(2)]
        assert abs(losses[0] - g["losses"][0]) <= 1e-5 * abs(g["losses"][0])
        assert abs(losses[1] - g["losses"][1]) <= 5e-3 * abs(g["losses"][1])
        sd2, sd0 = state_dict_from(g, "sd2."), state_dict_from(g, "sd0.")
        for k, p in m.named_parameters():
            if k in zero_grad or k in NOISE_GRAD:
                continue
            d = np.abs((p.detach().cpu().numpy() - sd0[k].numpy()) - (sd2[k].numpy() - sd0[k].numpy()))
            assert d.max() <= 4.2e-3, k
            assert np.mean(d > 2e-4) <= 0.05, (k, float(np.mean(d > 2e-4)))
        return
    logits = m(x)
    loss = crit(logits, y)
    loss.backward()
    _close(logits.detach().cpu().numpy(), g["logits"], 1e-4, "train logits")
    assert abs(float(loss.detach()) - g["losses"][0]) <= 1e-5 * abs(g["losses"][0])
    dgamma = np.abs(g["grad.linears.1.weight"]).max()
    for k, p in m.named_parameters():
        got, ref = p.grad.cpu().numpy(), g["grad." + k]
        if k in zero_grad:
            assert np.abs(got).max() == 0.0 and np.abs(ref).max() < 1e-5 * max(dgamma, 1e-30), k
        elif k in NOISE_GRAD and int(g["num_fc"]) != 0:
            assert np.abs(got).max() < 1e-3 * dgamma and np.abs(got - ref).max() < 1e-4 * dgamma, k
        else:
            d = np.abs(got - ref)
            scale = np.abs(ref).max()
            assert np.quantile(d, 0.98) <= 1e-4 * scale, (k, float(np.quantile(d, 0.98) / scale))
            assert d.max() <= 2e-2 * scale, (k, float(d.max() / scale))
    sd1 = state_dict_from(g, "sd1.")
    for k, v in m.state_dict().items():
        if "running" in k:
            _close(v.cpu().numpy(), sd1[k].numpy(), 1e-4, k)
        elif "num_batches" in k:
            assert int(v) == int(sd1[k]), k


def test_variant_dropout_is_active_and_replayed():
    """num_fc = 4 with dropout on: two forwards with the same seed agree, different seeds differ, and the backward
    replays the forward's masks (finite-difference-free check: gradient of a linear functional is reproducible)."""
    from explainn_b200 import autograd_fn, networks
    torch.manual_seed(0)
    m = networks.ExplaiNN(6, 90, 3, 19, num_fc=4).to(DEV).train()
    x = torch.from_numpy(onehot_from_codes(np.random.default_rng(1).integers(0, 4, size=(64, 90)).astype(np.uint8))).to(DEV)
    with torch.no_grad():
        a = autograd_fn.train_forward(m, x, seed=11).logits.clone()
        b = autograd_fn.train_forward(m, x, seed=11).logits.clone()
        c = autograd_fn.train_forward(m, x, seed=12).logits.clone()
    assert torch.equal(a, b) and not torch.equal(a, c)
    gout = torch.randn(64, 3, device=DEV, generator=torch.Generator(device=DEV).manual_seed(2))
    ws = autograd_fn.train_forward(m, x, seed=11)
    g1 = [t.clone() for t in autograd_fn.train_backward(m, ws, gout)]
    ws = autograd_fn.train_forward(m, x, seed=11)
    g2 = [t.clone() for t in autograd_fn.train_backward(m, ws, gout)]
    assert all(torch.equal(p, q) for p, q in zip(g1, g2))
    assert all(torch.isfinite(p).all() for p in g1)
