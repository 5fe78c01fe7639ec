"""GPU parity of the training path (train-mode forward, backward, fused loss, fused Adam) against the
reference's golden vectors and the fp32 CPU port.  Dropout is switched off for the parity cases
(`model.linears[9].eval()`): its random stream cannot match PyTorch's Philox (SURVEY.md 7.3 e); it is
checked statistically below."""
import os

import numpy as np
import pytest
import torch

from util import TRAIN_GOLDENS, load_golden, onehot_from_codes, random_codes, state_dict_from

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
# gradients of biases that feed a train-mode BatchNorm are exactly zero; PyTorch returns ~1e-9 noise there
ZERO_GRAD = ("linears.0.bias", "linears.6.bias", "linears.10.bias")
# BatchNorm-1's beta multiplies every pooled activation of a unit by exp(beta); the train-mode BatchNorm-2
# behind the (linear) FC1 removes that scale, so d loss / d beta1 is mathematically zero: both
# implementations return rounding noise (~1e-8) there, which cannot be compared relatively.
NOISE_GRAD = ("linears.1.bias",)


def _model(g, prefix):
    from explainn_b200 import networks
    m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]))
    m.load_state_dict(state_dict_from(g, prefix))
    m = m.to(DEV).train()
    m.linears[9].eval()
    return m


def _crit(g):
    return torch.nn.BCEWithLogitsLoss() if str(g["loss_kind"]) == "bce" else torch.nn.MSELoss()


def _close(a, b, rtol, name=""):
    """|a-b| <= rtol * max|b| elementwise (tensor-scale relative tolerance) and rtol*10 relative per element
    for the large entries."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(np.abs(b).max(), 1e-30)
    err = np.abs(a - b).max() / scale
    if err > rtol and os.environ.get("XB_DUMP"):          # debugging aid: keep the two arrays of a failing comparison
        np.savez(os.environ["XB_DUMP"], a=a, b=b)
    assert err <= rtol, "%s: max err %.3e of scale (tol %.1e)" % (name, err, rtol)


@pytest.mark.parametrize("backend", ["tc", "lut"])
@pytest.mark.parametrize("name", TRAIN_GOLDENS)
def test_train_forward_backward_vs_reference(name, backend):
    g = load_golden(name)
    m = _model(g, "sd0.")
    m.conv_backend = backend
    x = torch.from_numpy(onehot_from_codes(g["codes"])).to(DEV)
    y = torch.from_numpy(g["labels"]).to(DEV)
    out = m(x)
    loss = _crit(g)(out, y)
    loss.backward()
    m.check_inputs()
    _close(out.detach().cpu().numpy(), g["logits"], 1e-4, "logits")
    assert abs(float(loss) - g["losses"][0]) <= 1e-5 * abs(g["losses"][0])
    for k, p in m.named_parameters():
        ref = g["grad." + k]
        got = p.grad.cpu().numpy()
        if k in ZERO_GRAD:
            assert np.abs(got).max() == 0.0 and np.abs(ref).max() < 1e-6
        elif k in NOISE_GRAD:
            dgamma = np.abs(g["grad.linears.1.weight"]).max()
            assert np.abs(got).max() < 1e-4 * dgamma and np.abs(ref).max() < 1e-4 * dgamma
        else:
            _close(got, ref, 1e-4, "grad " + k)
    # running statistics after one training forward (momentum 0.1, unbiased variance)
    sd1 = state_dict_from(g, "sd1.")
    for k, v in m.state_dict().items():
        if "running" in k:
            _close(v.cpu().numpy(), sd1[k].numpy(), 1e-5, k)
        if "num_batches_tracked" in k:
            assert int(v) == int(sd1[k])


@pytest.mark.parametrize("name", TRAIN_GOLDENS)
@pytest.mark.parametrize("fused", [False, True])
def test_two_adam_steps_vs_reference(name, fused):
    from explainn_b200 import train as xtrain
    g = load_golden(name)
    m = _model(g, "sd0.")
    x = torch.from_numpy(onehot_from_codes(g["codes"])).to(DEV)
    y = torch.from_numpy(g["labels"]).to(DEV)
    crit = _crit(g)
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    losses = []
    if fused:
        assert xtrain.FusedTrainer.supports(m, crit, opt)
        tr = xtrain.FusedTrainer(m, crit, opt)
        for _ in range(2):
            losses.append(float(tr.step(x, y)))
    else:
        for _ in range(2):
            opt.zero_grad()
            loss = crit(m(x), y)
            loss.backward()
            opt.step()
            losses.append(float(loss))
    np.testing.assert_allclose(losses, g["losses"], rtol=2e-5)
    sd2 = state_dict_from(g, "sd2.")
    sd0 = state_dict_from(g, "sd0.")
    for k, v in m.state_dict().items():
        if "num_batches_tracked" in k:
            assert int(v) == 2
            continue
        ref = sd2[k].numpy()
        got = v.cpu().numpy()
        # Adam moves every weight by ~lr per step whatever the gradient scale, so compare the UPDATE
        if "running_mean" in k:
            # the reference's biases in front of a BatchNorm random-walk by +-lr per step (Adam normalises their
            # 1e-9 gradient noise), ours stay put; the batch mean that feeds running_mean shifts accordingly
            _close(got, ref, 2e-3, k)
        elif "running_var" in k:
            # same cause: BN-1's beta (zero true gradient) random-walks by +-lr in the reference, which scales a
            # unit's activations by exp(+-1e-3) and the downstream batch variance by ~(1 +- 2e-3)
            _close(got, ref, 2e-3, k)
        else:
            upd_ref = ref - sd0[k].numpy()
            upd = got - sd0[k].numpy()
            if k in ZERO_GRAD:
                # reference: Adam normalises 1e-9 gradient noise into +-lr steps; ours stays put (exact zero grad)
                assert np.abs(upd).max() == 0.0 and np.abs(upd_ref).max() <= 2.1e-3
            elif k in NOISE_GRAD:
                assert np.abs(upd).max() <= 2.1e-3 and np.abs(upd_ref).max() <= 2.1e-3
            else:
                # Adam's first steps are ~ lr * sign(g): an element whose gradient is at rounding-noise level can
                # legitimately step the other way, so bound the outliers instead of every element
                d = np.abs(upd - upd_ref)
                assert d.max() <= 2.1e-3, k
                outlier = d > 2e-2 * 2e-3 + 1e-7
                assert np.mean(outlier) <= 2e-3, k
                # ... and only where the first-step gradient really is at the noise floor of its tensor (a sign bug
                # on elements with a real gradient must not pass as an outlier)
                g0 = np.abs(g["grad." + k])
                assert np.all(g0[outlier] <= 1e-3 * g0.max()), (k, float(g0[outlier].max() / g0.max()))


def _close_per_unit(got, ref, rtol, name, max_outlier_frac=0.02, outlier_rtol=0.1, batch=160):
    """Gradients upstream of a ReLU / max-pool kink, row by row (unit or hidden channel).  A pre-activation within
    fp32 rounding of a kink can fall on either side in two correct implementations (the reference's own CPU and cuDNN
    paths differ the same way); one such flip switches one sample's whole contribution to that row on or off, i.e.
    ~1/(active samples) of the row's gradient.  The number of flips grows with the number of activations, i.e. with
    the batch, so: all but `max_outlier_frac` x max(1, batch/160) of the rows within `rtol`, all but 1e-4 of the rows
    (none for small tensors) within `outlier_rtol`, and no row off by more than its own magnitude."""
    got = np.asarray(got, np.float64).reshape(got.shape[0], -1)
    ref = np.asarray(ref, np.float64).reshape(ref.shape[0], -1)
    # a one-element row (BatchNorm-1 gamma: one number per unit) is a sum of B*P signed terms that cancel to ~1 % of
    # their magnitude, so an element near a zero crossing carries the fp32 accumulation noise of the whole sum on BOTH
    # sides; the floor below which an element is compared absolutely is 1 % of the tensor scale there, 0.1 % otherwise
    floor = 1e-2 if got.shape[1] == 1 else 1e-3
    scale = np.maximum(np.abs(ref).max(axis=1), floor * np.abs(ref).max())
    err = np.abs(got - ref).max(axis=1) / scale
    msg = "%s: %d of %d rows above %.1e, %d above %.1e (worst %.2e, tensor-scale err %.2e)" % (
        name, (err > rtol).sum(), err.size, rtol, (err > outlier_rtol).sum(), outlier_rtol, err.max(),
        np.abs(got - ref).max() / np.abs(ref).max())
    allowed = max(1, int(max_outlier_frac * max(1.0, batch / 160.0) * err.size))
    if os.environ.get("XB_DUMP") and (err.max() > 1.0 or (err > outlier_rtol).sum() > int(1e-4 * err.size) or (err > rtol).sum() > allowed):
        np.savez(os.environ["XB_DUMP"], a=got, b=ref)          # debugging aid: keep the two arrays of a failing comparison
    assert err.max() <= 1.0, msg
    assert (err > outlier_rtol).sum() <= int(1e-4 * err.size), msg
    assert (err > rtol).sum() <= allowed, msg


@pytest.mark.parametrize("backend", ["tc", "lut"])
@pytest.mark.parametrize("U,L,C,F,B,loss", [(300, 251, 81, 19, 160, "bce"), (20, 1000, 3, 19, 130, "bce"), (33, 96, 2, 7, 70, "bce"),
                                            (300, 251, 81, 19, 512, "bce"),      # BASELINE.json configs[1], full batch
                                            (300, 249, 2, 19, 1024, "mse")])     # BASELINE.json configs[2], full batch
def test_train_step_vs_torch_port(U, L, C, F, B, loss, backend):
    """cfg-2 / cfg-3 at their full batch sizes, cfg-5-like and odd shapes: logits, loss, all gradients and the
    BatchNorm running statistics of one training step against the fp32 CPU port (explainn/train/train.py:54-73)."""
    from explainn_b200 import networks
    from oracle import torch_port as tp
    torch.manual_seed(7)
    port = tp.PortExplaiNN(U, L, C, F)
    gen = torch.Generator().manual_seed(8)
    for mod in port.modules():
        if isinstance(mod, torch.nn.BatchNorm1d):
            mod.weight.data.uniform_(0.5, 1.5, generator=gen)
            mod.bias.data.normal_(0, 0.1, generator=gen)
    m = networks.ExplaiNN(U, L, C, F)
    m.load_state_dict(port.state_dict())
    m = m.to(DEV).train()
    m.conv_backend = backend
    m.linears[9].eval()
    port.train()
    port.linears[9].eval()
    x = torch.from_numpy(onehot_from_codes(random_codes(B, L, 10, n_frac=0.01)))
    if loss == "bce":
        y = (torch.rand(B, C, generator=gen) < 0.5).float()
        crit = torch.nn.BCEWithLogitsLoss()
    else:
        y = torch.randn(B, C, generator=gen)
        crit = torch.nn.MSELoss()
    out_p = port(x)
    lp = crit(out_p, y)
    lp.backward()
    out = m(x.to(DEV))
    _close(out.detach().cpu().numpy(), out_p.detach().numpy(), 1e-4, "logits")
    lm = crit(out, y.to(DEV))
    lm.backward()
    assert abs(float(lm) - float(lp)) <= 1e-5 * abs(float(lp))
    for (k, p), (_, q) in zip(port.named_parameters(), m.named_parameters()):
        if k in ZERO_GRAD:
            assert float(q.grad.abs().max()) == 0.0
        elif k in NOISE_GRAD:
            # a sum of B*P signed terms whose exact value is zero: what is left is the accumulation noise of the sum,
            # which grows with the batch and with a unit's activation scale.  The bulk must sit at rounding level; single
            # units may reach a percent of the BatchNorm-1 gain gradient (Adam normalises the magnitude away in any case)
            dg = float(port.linears[1].weight.grad.abs().max())
            assert float(q.grad.abs().median()) < 1e-5 * dg and float(q.grad.abs().max()) < 5e-2 * dg
        elif k in ("linears.0.weight", "linears.1.weight", "linears.6.weight", "linears.7.weight", "linears.7.bias",
                   "linears.10.weight", "linears.11.weight", "linears.11.bias"):
            # parameters upstream of a ReLU / max-pool kink (the unit-output ReLU behind BatchNorm-3 for linears.10 / 11),
            # compared row by row (unit or hidden channel): a pre-activation within rounding of zero may fall on the
            # other side of the kink, which moves that one row's gradient by ~1/B of its norm
            _close_per_unit(q.grad.cpu().numpy(), p.grad.numpy(), 2e-4, "grad " + k, batch=B)
        else:
            _close(q.grad.cpu().numpy(), p.grad.numpy(), 2e-4, "grad " + k)
    for (k, a), (_, b) in zip(port.state_dict().items(), m.state_dict().items()):
        if "running" in k:
            _close(b.cpu().numpy(), a.numpy(), 1e-5, k)


def test_unit_output_gradients_and_frozen_params():
    """model.linears(x) is differentiable in train mode; requires_grad=False is honoured
    (scripts/finetune.py:216-221); gradient hooks on the filter fire (tools.py:248-264)."""
    from explainn_b200 import networks
    from oracle import torch_port as tp
    torch.manual_seed(1)
    port = tp.PortExplaiNN(10, 80, 3, 19)
    m = networks.ExplaiNN(10, 80, 3, 19)
    m.load_state_dict(port.state_dict())
    m = m.to(DEV).train(); m.linears[9].eval()
    port.train(); port.linears[9].eval()
    x = torch.from_numpy(onehot_from_codes(random_codes(24, 80, 3)))
    wts = torch.randn(24, 10, generator=torch.Generator().manual_seed(2))
    (port.unit_outputs(x) * wts).sum().backward()
    (m.linears(x.to(DEV).repeat(1, 10, 1)) * wts.to(DEV)).sum().backward()
    for (k, p), (_, q) in zip(port.named_parameters(), m.named_parameters()):
        if k.startswith("final"):
            assert q.grad is None or float(q.grad.abs().max()) == 0.0
        elif k not in ZERO_GRAD and k not in NOISE_GRAD:
            _close(q.grad.cpu().numpy(), p.grad.numpy(), 2e-4, k)
    # frozen everything but `final`
    m.zero_grad()
    for k, p in m.named_parameters():
        p.requires_grad = k.startswith("final")
    m(x.to(DEV)).sum().backward()
    assert m.linears[0].weight.grad is None and m.final.weight.grad is not None
    # hook
    for p in m.parameters():
        p.requires_grad = True
    m.zero_grad()
    m.linears[0].weight.register_hook(lambda gr: gr * 0.0)
    m(x.to(DEV)).sum().backward()
    assert float(m.linears[0].weight.grad.abs().max()) == 0.0


def test_batch_of_one_fails_like_batchnorm():
    from explainn_b200 import networks
    m = networks.ExplaiNN(4, 60, 2, 19).to(DEV).train()
    with pytest.raises(ValueError, match="Expected more than 1 value per channel"):
        m(torch.zeros(1, 4, 60, device=DEV))


def test_dropout_statistics_and_mask_replay():
    from explainn_b200 import autograd_fn, networks
    torch.manual_seed(0)
    m = networks.ExplaiNN(16, 100, 2, 19).to(DEV).train()
    x = torch.from_numpy(onehot_from_codes(random_codes(256, 100, 4))).to(DEV)
    with torch.no_grad():
        a = autograd_fn.train_forward(m, x, seed=123).unit.clone()
        b = autograd_fn.train_forward(m, x, seed=123).unit.clone()
        c = autograd_fn.train_forward(m, x, seed=124).unit.clone()
    assert torch.equal(a, b) and not torch.equal(a, c)
    # keep probability 0.7 with 1/0.7 scaling: E[dropout(v)] = v.  Probe through the linear FC2 pre-activation.
    m.linears[9].eval()
    with torch.no_grad():
        ws0 = autograd_fn.train_forward(m, x, seed=5)
        h2_nodrop = ws0.h2.clone()
    m.linears[9].train()
    acc = torch.zeros_like(h2_nodrop)
    n = 64
    with torch.no_grad():
        for s in range(n):
            acc += autograd_fn.train_forward(m, x, seed=1000 + s).h2
    rel = float((acc / n - h2_nodrop).abs().mean() / h2_nodrop.abs().mean())
    assert rel < 0.15
    # the backward replays the forward's mask: rebuild the FC stack in float64 torch from the saved keep bits and
    # compare the forward activations and every FC-stack gradient
    ws = autograd_fn.train_forward(m, x, seed=77)
    U, B, Hd = ws.h1.shape
    bits = ws.dropmask.to(torch.int64) & 0xFFFFFFFF                           # (U,B,4) words, channel j = bit j%32 of word j//32
    keep = ((bits.unsqueeze(-1) >> torch.arange(32, device=DEV)) & 1).reshape(U, B, 128)[:, :, :Hd].bool()
    frac = keep.double().mean().item()
    assert abs(frac - 0.7) < 0.01, frac
    per_channel = keep.double().mean(dim=(0, 1))
    assert float(per_channel.min()) > 0.6 and float(per_channel.max()) < 0.8
    gout = torch.randn(ws.logits.shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(3))
    grads = autograd_fn.train_backward(m, ws, gout)
    names = ["W1", "b1", "g2", "be2", "W2", "b2", "g3", "be3", "Wf", "bf"]
    ps = [p.detach().double().clone().requires_grad_(True) for p in autograd_fn._params(m)[4:]]
    W1, b1, g2, be2, W2, b2, g3, be3, Wf, bf = ps
    P = ws.pooled.shape[2]
    pooled = ws.pooled.double()                                                # (B,U,P)
    h1 = torch.einsum("bup,ujp->ubj", pooled, W1.reshape(U, Hd, P)) + b1.reshape(U, 1, Hd)
    assert (h1.float() - ws.h1).abs().max().item() < 1e-4 * ws.h1.abs().max().item()
    mu, var = h1.mean(1, keepdim=True), h1.var(1, unbiased=False, keepdim=True)
    z = (h1 - mu) / torch.sqrt(var + 1e-5) * g2.reshape(U, 1, Hd) + be2.reshape(U, 1, Hd)
    a = torch.relu(z) * keep / 0.7
    h2 = (a * W2.reshape(U, 1, Hd)).sum(-1) + b2.reshape(U, 1)               # (U,B)
    assert (h2.float() - ws.h2).abs().max().item() < 1e-4 * ws.h2.abs().max().item()
    mu3, var3 = h2.mean(1, keepdim=True), h2.var(1, unbiased=False, keepdim=True)
    unit = torch.relu((h2 - mu3) / torch.sqrt(var3 + 1e-5) * g3.reshape(U, 1) + be3.reshape(U, 1)).t()
    logits = unit @ Wf.t() + bf
    assert (logits.float() - ws.logits).abs().max().item() < 1e-4 * max(ws.logits.abs().max().item(), 1.0)
    (logits * gout.double()).sum().backward()
    for name, pref, got in zip(names, ps, grads[4:]):
        ref = pref.grad.float()
        if name in ("b1", "b2"):       # a bias in front of a train-mode BatchNorm: gradient is exactly zero
            assert float(got.abs().max()) == 0.0
            continue
        err = float((got - ref).norm() / ref.norm().clamp_min(1e-12))
        assert err < 2e-3, (name, err)


def test_train_explainn_drop_in(tmp_path):
    """Epoch loop, checkpoint files and return values of train_explainn (train.py:13-142)."""
    import os
    from explainn_b200 import networks, train as xtrain
    torch.manual_seed(0)
    U, L, C = 8, 60, 2
    x = torch.from_numpy(onehot_from_codes(random_codes(96, L, 5)))
    y = (torch.rand(96, C) < 0.5).float()
    ds = torch.utils.data.TensorDataset(x, y)
    tl = torch.utils.data.DataLoader(ds, batch_size=32)
    vl = torch.utils.data.DataLoader(ds, batch_size=48)
    for crit, opt_cls in ((torch.nn.BCEWithLogitsLoss(), torch.optim.Adam), (torch.nn.MSELoss(), torch.optim.SGD)):
        m = networks.ExplaiNN(U, L, C, 19).to(DEV)
        opt = opt_cls(m.parameters(), lr=1e-3)
        d = tmp_path / opt_cls.__name__
        d.mkdir()
        m2, tr, te = xtrain.train_explainn(tl, vl, m, DEV, crit, opt, num_epochs=3, weights_folder=str(d),
                                           name_ind="t", checkpoint=1, patience=0, trim_weights=True)
        assert m2 is m and len(tr) == 3 and len(te) == 3 and all(np.isfinite(tr + te))
        files = sorted(os.listdir(d))
        assert "model_epoch_1_t.pth" in files and "model_epoch_3_t.pth" in files
        assert any(f.startswith("model_epoch_best_") and f.endswith("_t.pth") for f in files)
        assert not m.training                                   # left in eval mode, like the reference
        assert float(m.final.weight.min()) >= 0.0               # trim_weights: clamp after every step (train.py:70-71)
        sd = torch.load(os.path.join(d, [f for f in files if "best" in f][0]), map_location="cpu")
        assert list(sd.keys()) == list(m.state_dict().keys())


def test_train_explainn_epoch_losses_vs_reference(tmp_path):
    """Three epochs of train_explainn on the inputs, initial weights and loaders of the reference run recorded in
    tests/golden/api_train_loop.npz (made by tests/golden/make_golden_api.py with the unmodified reference, dropout
    probability 0): epoch train / validation losses to 1e-4 relative, same checkpoint files, and returned weights
    within the Adam tolerance used above.  Both the fused trainer and the generic autograd path."""
    import os
    from explainn_b200 import networks, train as xtrain
    g = load_golden("api_train_loop")
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    y = torch.from_numpy(g["labels"])
    ds = torch.utils.data.TensorDataset(x, y)
    tl = torch.utils.data.DataLoader(ds, batch_size=32, shuffle=False)
    vl = torch.utils.data.DataLoader(ds, batch_size=48, shuffle=False)
    for fused in (True, False):
        m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]))
        m.load_state_dict(state_dict_from(g, "sd0."))
        m = m.to(DEV)
        m.linears[9].p = 0.0
        crit = torch.nn.BCEWithLogitsLoss()
        opt = torch.optim.Adam(m.parameters(), lr=1e-3)
        if not fused:
            crit = torch.nn.BCEWithLogitsLoss(pos_weight=torch.ones(int(g["C"]), device=DEV))    # not recognised -> generic path
            assert not xtrain.FusedTrainer.supports(m, crit, opt)
        d = tmp_path / ("fused" if fused else "generic")
        d.mkdir()
        m2, tr, te = xtrain.train_explainn(tl, vl, m, DEV, crit, opt, num_epochs=3, weights_folder=str(d), name_ind="g",
                                           checkpoint=1, patience=0)
        np.testing.assert_allclose(tr, g["train_error"], rtol=1e-4)
        np.testing.assert_allclose(te, g["test_error"], rtol=1e-4)
        assert sorted(os.listdir(d)) == [str(f) for f in g["files"]]
        sd_ref, sd0 = state_dict_from(g, "sd_final."), state_dict_from(g, "sd0.")
        for k, v in m2.state_dict().items():
            if "num_batches_tracked" in k:
                assert int(v) == int(sd_ref[k])
            elif "running" in k:
                # the reference's conv / FC biases random-walk by +-lr per step (Adam on 1e-9 gradient noise) and drag
                # the batch means along: nine steps of 1e-3 against running means of a few 0.1
                _close(v.cpu().numpy(), sd_ref[k].numpy(), 3e-2, k)
            elif k not in ZERO_GRAD and k not in NOISE_GRAD:
                # nine Adam steps of ~lr each: compare the accumulated update
                upd, upd_ref = v.cpu().numpy() - sd0[k].numpy(), sd_ref[k].numpy() - sd0[k].numpy()
                assert np.abs(upd - upd_ref).max() <= 9 * 2.1e-3, k
                assert np.mean(np.abs(upd - upd_ref) > 1e-3) <= 5e-3, k
