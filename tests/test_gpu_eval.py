"""GPU parity: eval-mode forward and the interpretation pass, through the C ABI, against the oracle
(bit-exact where the arithmetic order is canonical) and the reference's golden vectors."""
import numpy as np
import pytest
import torch

from util import EVAL_GOLDENS, load_golden, onehot_from_codes, random_codes, state_dict_from

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _model(g, prefix="sd."):
    from explainn_b200 import networks
    m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]))
    m.load_state_dict(state_dict_from(g, prefix))
    return m.to(DEV).eval()


def _canonical(g):
    from oracle import canonical as cn
    sd = state_dict_from(g)
    n = lambda k: sd[k].numpy()
    x = onehot_from_codes(g["codes"])
    codes, nmask, _ = cn.pack_onehot(x)
    wq, shift, scale = cn.fold_bn(n("linears.0.weight"), n("linears.0.bias"), n("linears.1.weight"),
                                  n("linears.1.bias"), n("linears.1.running_mean"), n("linears.1.running_var"))
    z = cn.conv_z(codes, nmask, int(g["L"]), wq, shift)
    return x, codes, nmask, wq, shift, scale, z


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_pack_bit_exact(name):
    from explainn_b200 import functional as Fn
    g = load_golden(name)
    x, codes, nmask, *_ = _canonical(g)
    xt = torch.from_numpy(x).to(DEV)
    pk = Fn.pack_onehot(xt)
    assert np.array_equal(pk.codes.cpu().numpy().view(np.uint32), codes)
    assert np.array_equal(pk.nmask.cpu().numpy().view(np.uint32), nmask)
    assert int(pk.invalid.item()) == 0
    # the U-times tiled input the reference passes around, and float16 input
    U = int(g["U"])
    pk2 = Fn.pack_onehot(xt[:8].repeat(1, U, 1), U)
    assert torch.equal(pk2.codes, pk.codes[:8]) and torch.equal(pk2.nmask, pk.nmask[:8])
    pk3 = Fn.pack_onehot(xt.half())
    assert torch.equal(pk3.codes, pk.codes)
    # invalid columns are counted
    bad = xt[:4].clone()
    bad[0, :, 3] = 0.5
    bad[1, 0, 5] = 1.0
    bad[1, 1, 5] = 1.0
    pk4 = Fn.pack_onehot(bad)
    assert int(pk4.invalid.item()) == 2
    with pytest.raises(ValueError):
        Fn.check_invalid(pk4.invalid)


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_conv_prepare_bit_exact(name):
    g = load_golden(name)
    m = _model(g)
    x, codes, nmask, wq, shift, scale, z = _canonical(g)
    prep = m._conv_prep_eval()
    assert np.array_equal(prep.shift.cpu().numpy(), shift)
    assert np.array_equal(prep.scale.cpu().numpy(), scale)
    U, F = int(g["U"]), int(g["F"])
    tiles = (U + 31) // 32
    wq_t = prep.wq_t.cpu().numpy().reshape(tiles, 4, F, 32)
    for u in range(U):
        assert np.array_equal(wq_t[u // 32, :, :, u % 32], wq[u])


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_activations_and_pooled_bit_exact_vs_oracle(name):
    from oracle import canonical as cn
    g = load_golden(name)
    m = _model(g)
    x, codes, nmask, wq, shift, scale, z = _canonical(g)
    xt = torch.from_numpy(x).to(DEV)
    U = int(g["U"])
    m.conv_backend = "lut"          # the table kernel follows the oracle's arithmetic order exactly
    with torch.no_grad():
        act32 = m.linears[:3](xt.repeat(1, U, 1)).cpu().numpy()
        act16 = m._activations(xt, torch.float16).cpu().numpy()
        pooled = m.linears[:4](xt).cpu().numpy()
    assert np.array_equal(act32, cn.expf(z)), "conv+BN+exp differs from the canonical-order oracle"
    assert np.array_equal(act16, cn.activations_fp16(z))
    ref_pooled, ref_idx = cn.pool_exp(z)
    assert np.array_equal(pooled, ref_pooled)
    # and against the reference itself: fp32 tolerance of the north star (1e-4 relative)
    n_act = g["pooled"].shape[0]
    np.testing.assert_allclose(pooled[:n_act], g["pooled"], rtol=1e-5)


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_tensor_core_conv_matches_oracle_and_reference(name):
    """tcgen05 implicit-GEMM conv (default backend): fp32-rounding accurate, not bit-exact.  Tolerance 1e-5
    relative on the pooled activations (the north-star asks 1e-4 on logits)."""
    from oracle import canonical as cn
    g = load_golden(name)
    m = _model(g)
    assert m.conv_backend == "tc"
    x, codes, nmask, wq, shift, scale, z = _canonical(g)
    xt = torch.from_numpy(x).to(DEV)
    with torch.no_grad():
        pooled = m.linears[:4](xt).cpu().numpy()
    ref_pooled, _ = cn.pool_exp(z)
    np.testing.assert_allclose(pooled, ref_pooled, rtol=1e-5)
    n_act = g["pooled"].shape[0]
    np.testing.assert_allclose(pooled[:n_act], g["pooled"], rtol=1e-5)


@pytest.mark.parametrize("backend", ["tc", "lut"])
@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_logits_and_unit_outputs_match_reference(name, backend):
    g = load_golden(name)
    m = _model(g)
    m.conv_backend = backend
    xt = torch.from_numpy(onehot_from_codes(g["codes"])).to(DEV)
    with torch.no_grad():
        logits = m(xt).cpu().numpy()
        unit = m.linears(xt.repeat(1, int(g["U"]), 1)).cpu().numpy()
    m.check_inputs()
    scale = np.abs(g["unit"]).max()
    np.testing.assert_allclose(unit, g["unit"], rtol=1e-4, atol=1e-5 * scale)
    np.testing.assert_allclose(logits, g["logits"], rtol=1e-4, atol=1e-5 * np.abs(g["logits"]).max())


@pytest.mark.parametrize("backend", ["tc", "lut"])
@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_pwm_counts_bit_exact(name, backend):
    from oracle import canonical as cn
    from explainn_b200 import interpretation
    g = load_golden(name)
    m = _model(g)
    x, codes, nmask, wq, shift, scale, z = _canonical(g)
    F = int(g["F"])
    xt = torch.from_numpy(x)
    # several uneven chunks so that the accumulate path is exercised
    chunks = [xt[:5], xt[5:6], xt[6:]]
    pwm, counts, hits = interpretation.get_pwms_explainn_fused([(c, None) for c in chunks], m, DEV, return_counts=True,
                                                               backend=backend)
    a16 = cn.activations_fp16(z)
    ref_counts, ref_hits, ref_thr = cn.pwm_counts(a16, x, F)
    assert np.array_equal(counts, ref_counts)
    assert np.array_equal(hits, ref_hits)
    assert np.array_equal(pwm, cn.pwm_from_counts_hits(ref_counts, ref_hits))
    # the reference's own output
    assert np.array_equal(pwm, g["pwm"])
    acc = interpretation.PwmAccumulator(m, backend=backend)
    acc.scan_max(xt.to(DEV))
    acc.finish_max()
    assert np.array_equal(acc.zmax.cpu().numpy(), z.max(axis=(0, 2)))
    assert np.array_equal(acc.amax.cpu().numpy().astype(np.float16), g["amax16"])
    assert np.array_equal(acc.thr.cpu().numpy().astype(np.float16), ref_thr)


def test_api_helpers_and_meme(tmp_path):
    from explainn_b200 import interpretation, test as xtest
    from oracle import canonical as cn
    g = load_golden("eval_u40_f20")
    m = _model(g)
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    y = torch.arange(x.shape[0]).float().unsqueeze(1)
    dl = torch.utils.data.DataLoader(torch.utils.data.TensorDataset(x, y), batch_size=25)
    labels, outs = xtest.run_test(m, dl, DEV)
    assert outs.shape == g["logits"].shape and np.array_equal(labels, y.numpy())
    np.testing.assert_allclose(outs, g["logits"], rtol=1e-4, atol=1e-5 * np.abs(g["logits"]).max())
    labels2, outs2 = xtest.run_test(m, dl, DEV, isSigmoid=True)
    np.testing.assert_allclose(outs2, 1 / (1 + np.exp(-outs.astype(np.float64))), rtol=1e-5)
    p, l = interpretation.get_explainn_predictions(dl, m, DEV)
    assert np.array_equal(p, outs) and np.array_equal(l, labels)
    uo = interpretation.get_explainn_unit_outputs(dl, m, DEV)
    np.testing.assert_allclose(uo, g["unit"], rtol=1e-4, atol=1e-5 * np.abs(g["unit"]).max())
    act = interpretation.get_explainn_unit_activations(dl, m, DEV)
    assert act.dtype == np.float16 and act.shape == (x.shape[0], 40, 81)
    n_act = g["act16"].shape[0]
    assert np.mean(act[:n_act] != g["act16"]) < 2e-3
    pwm = interpretation.get_pwms_explainn_fused(dl, m, DEV)
    path = tmp_path / "f.meme"
    interpretation.pwm_to_meme(pwm, str(path), verbose=False)
    assert path.read_text() == cn.pwm_to_meme_text(pwm)


@pytest.mark.parametrize("U,L,C,F,B", [(300, 251, 81, 19, 96), (64, 1000, 10, 19, 40), (33, 50, 2, 5, 70)])
def test_eval_forward_vs_torch_port(U, L, C, F, B):
    """Shapes of BASELINE.json configs 2 and 5 (reduced batch) against the fp32 CPU port."""
    from explainn_b200 import networks
    from oracle import torch_port as tp
    torch.manual_seed(3)
    port = tp.PortExplaiNN(U, L, C, F)
    g = torch.Generator().manual_seed(5)
    for mod in port.modules():
        if isinstance(mod, torch.nn.BatchNorm1d):
            mod.running_mean.normal_(0, 0.1, generator=g)
            mod.running_var.uniform_(0.5, 1.5, generator=g)
            mod.weight.data.uniform_(0.5, 1.5, generator=g)
            mod.bias.data.normal_(0, 0.1, generator=g)
    port.eval()
    m = networks.ExplaiNN(U, L, C, F)
    m.load_state_dict(port.state_dict())
    m = m.to(DEV).eval()
    x = torch.from_numpy(onehot_from_codes(random_codes(B, L, 9, n_frac=0.01)))
    with torch.no_grad():
        ref = port(x).numpy()
        refu = port.unit_outputs(x).numpy()
        out = m(x.to(DEV)).cpu().numpy()
        unit = m.linears(x.to(DEV)).cpu().numpy()
    # north-star fp32 tolerance: 1e-4 relative; for entries near zero "relative" is taken against the
    # largest magnitude of the tensor (a sum over U=300 units cannot be relatively exact at a zero crossing)
    np.testing.assert_allclose(unit, refu, rtol=1e-4, atol=1e-4 * np.abs(refu).max())
    np.testing.assert_allclose(out, ref, rtol=1e-4, atol=1e-4 * np.abs(ref).max())


def test_empty_batch_and_errors():
    from explainn_b200 import networks
    m = networks.ExplaiNN(8, 60, 2, 19).to(DEV).eval()
    with torch.no_grad():
        out = m(torch.zeros(0, 4, 60, device=DEV))
    assert out.shape == (0, 2)
    with pytest.raises(ValueError):
        m(torch.zeros(2, 4, 61, device=DEV))
    with pytest.raises(NotImplementedError):
        networks.ExplaiNN(8, 60, 2, 25)                      # filter_size > 24: limit of the k-mer tables, named in the message
    with pytest.raises(NotImplementedError):
        networks.ExplaiNN(8, 4096, 2, 19)                    # input_length > 2048
    with pytest.raises(ValueError):
        networks.ExplaiNN(8, 20, 2, 19, pool_size=7)         # too short for one pool window


def test_tc_fp16_weight_mode_stated_tolerance():
    """conv_backend = "tc_fp16": one fp16 piece per folded weight (relative rounding 2^-11 per tap), fp32 accumulation.
    Stated tolerance: pooled activations within 2e-3 relative of the fp32-accurate path, logits within 2e-3 of the
    tensor scale; the two-piece default stays at 1e-5 / 1e-4.  Includes folded weights far outside the fp16 range
    (row scaling in explainn_conv_tc_prepare)."""
    from explainn_b200 import networks
    torch.manual_seed(5)
    m = networks.ExplaiNN(300, 251, 81, 19)
    with torch.no_grad():
        m.linears[1].weight.uniform_(0.5, 1.5)
        m.linears[1].running_var.uniform_(0.5, 1.5)
        m.linears[0].weight[:20].mul_(1e-6)            # folded taps ~1e-7: fp16 subnormal range without the row scale
        m.linears[1].bias[:20] = -3.0
        m.linears[0].weight[20:24].mul_(3e5)           # folded taps ~3e4 .. 1e5: fp16 overflow without the row scale
        m.linears[1].bias[20:24] = -2e6                # keeps exp(z) finite (it underflows) for these units
    m = m.to(DEV).eval()
    x = torch.from_numpy(onehot_from_codes(random_codes(512, 251, 9, n_frac=0.01))).to(DEV)
    with torch.no_grad():
        m.conv_backend = "lut"
        p_ref, o_ref = m.linears[:4](x), m(x)
        m.conv_backend = "tc"
        p_tc, o_tc = m.linears[:4](x), m(x)
        m.conv_backend = "tc_fp16"
        p_16, o_16 = m.linears[:4](x), m(x)
    assert torch.isfinite(p_ref).all()
    rel = lambda a, b: float(((a - b).abs() / b.abs().clamp_min(1e-30)).max())
    # units 20..23 have |z| of 1e5: relative error of exp(z) there is |z| * (relative error of z)
    keep = torch.ones(300, dtype=torch.bool, device=DEV)
    keep[20:24] = False
    assert rel(p_tc[:, keep], p_ref[:, keep]) <= 1e-5
    assert rel(p_16[:, keep], p_ref[:, keep]) <= 2e-3
    assert float((p_tc[:, ~keep] - p_ref[:, ~keep]).abs().max()) <= 1e-2 * float(p_ref[:, ~keep].abs().max() + 1e-30)
    assert float((o_tc - o_ref).abs().max()) <= 1e-4 * float(o_ref.abs().max())
    assert float((o_16 - o_ref).abs().max()) <= 2e-3 * float(o_ref.abs().max())
    assert not torch.equal(p_16, p_tc)
