"""Full-size property checks at the shapes of BASELINE.json (the CPU oracle needs minutes there, so these use
size-independent properties instead of a second implementation): backend agreement, batch-split invariance, strand
symmetry, batch-duplication invariance of the training gradients, linearity of the backward."""
import numpy as np
import pytest
import torch

from util import onehot_from_codes, random_codes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _randomise_bn(m, seed):
    g = torch.Generator().manual_seed(seed)
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm1d):
            mod.running_mean.normal_(0, 0.1, generator=g)
            mod.running_var.uniform_(0.5, 1.5, generator=g)
            mod.weight.data.uniform_(0.5, 1.5, generator=g)
            mod.bias.data.normal_(0, 0.1, generator=g)


def test_cfg5_inference_properties():
    """Large-scale inference shape: num_cnns=500, L=1000, 100 outputs (BASELINE.json config 5), 2048 sequences."""
    from explainn_b200 import networks, predict
    torch.manual_seed(0)
    m = networks.ExplaiNN(500, 1000, 100, 19)
    _randomise_bn(m, 1)
    m = m.to(DEV).eval()
    x = torch.from_numpy(onehot_from_codes(random_codes(2048, 1000, 4, n_frac=0.002))).to(DEV)
    with torch.no_grad():
        m.conv_backend = "tc"
        full = m(x)
        # per-sequence computation: any split of the batch gives the same rows, bit for bit
        parts = torch.cat([m(x[:700]), m(x[700:701]), m(x[701:])], 0)
        assert torch.equal(full, parts)
        # the table kernel (bit-identical to the oracle) and the tensor-core kernel agree to fp32 rounding
        m.conv_backend = "lut"
        ref = m(x[:512])
        m.conv_backend = "tc"
        assert float((full[:512] - ref).abs().max()) <= 1e-4 * float(ref.abs().max())
        # strand symmetry: the reverse strand of x is the forward strand of its reverse complement
        both = predict.predict_both_strands(x[:256], m, DEV)          # (N, C, 4): Fwd, Rev, Mean, Max
        xr = torch.flip(x[:256], dims=(1, 2)).contiguous()
        both_r = predict.predict_both_strands(xr, m, DEV)
        assert np.array_equal(both[:, :, 0], both_r[:, :, 1]) and np.array_equal(both[:, :, 1], both_r[:, :, 0])
        assert np.array_equal(both[:, :, 2:], both_r[:, :, 2:])
    assert torch.isfinite(full).all()


def test_cfg3_training_properties():
    """DeepSTARR-style regression shape: num_cnns=300, L=249, 2 outputs, MSE, batch 1024 (BASELINE.json config 3)."""
    from explainn_b200 import autograd_fn, networks
    torch.manual_seed(0)
    m = networks.ExplaiNN(300, 249, 2, 19)
    _randomise_bn(m, 2)
    m = m.to(DEV).train()
    m.linears[9].eval()                                  # dropout off: gradients are a deterministic function of the batch
    x = torch.from_numpy(onehot_from_codes(random_codes(1024, 249, 6, n_frac=0.005))).to(DEV)
    gout = torch.randn(1024, 2, device=DEV, generator=torch.Generator(device=DEV).manual_seed(1))
    with torch.no_grad():
        ws = autograd_fn.train_forward(m, x)
        g1 = [g.clone() for g in autograd_fn.train_backward(m, ws, gout)]
        # linearity of the backward in the upstream gradient
        ws = autograd_fn.train_forward(m, x)
        g3 = [g.clone() for g in autograd_fn.train_backward(m, ws, 3.0 * gout)]
        # duplicating the batch leaves every BatchNorm statistic unchanged
        x2, gout2 = torch.cat([x, x], 0), torch.cat([gout, gout], 0)
        ws2 = autograd_fn.train_forward(m, x2)
        assert float((ws2.logits[:1024] - ws.logits).abs().max()) <= 2e-5 * float(ws.logits.abs().max())
        g2 = [g.clone() for g in autograd_fn.train_backward(m, ws2, gout2)]
    names = [k for k, _ in m.named_parameters()]
    gscale = max(float(g.abs().max()) for g in g1)
    for k, a, b, c in zip(names, g1, g3, g2):
        if k == "linears.1.bias":      # mathematically zero gradient: rounding noise on all sides
            assert float(a.abs().max()) <= 1e-4 * float(g1[2].abs().max())
            continue
        scale = max(float(a.abs().max()), 1e-6 * gscale)
        assert float((b - 3.0 * a).abs().max()) <= 1e-4 * 3.0 * scale, ("linearity", k)
        # the upstream gradient is per sample (not a mean), so the duplicated batch gives exactly twice the gradient;
        # max-pool ties may resolve differently for single elements, hence the quantile
        d = (c - 2.0 * a).abs()
        assert float(torch.quantile(d.flatten()[:1_000_000].float(), 0.98)) <= 2e-4 * 2.0 * scale, ("duplication", k)
    assert all(torch.isfinite(g).all() for g in g1)
