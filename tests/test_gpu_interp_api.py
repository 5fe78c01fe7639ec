"""GPU parity of the interpretation drop-ins that take the caller's activation array
(get_pwms_explainn, get_median_unit_importance, get_specific_unit_importance) and of the per-sequence
max / first-argmax scan, through the C ABI, against the reference's golden vectors and the oracle."""
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


def _port_act16(g):
    """(N,U,S) float16 activations of the reference's own torch graph (the port is bit-equal to the reference:
    tests/test_oracle_golden.py::test_torch_port_matches_reference_eval)."""
    from oracle import torch_port as tp
    port = tp.PortExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]))
    port.load_state_dict(state_dict_from(g))
    port.eval()
    x = onehot_from_codes(g["codes"])
    return tp.unit_activations_fp16(port, torch.from_numpy(x)), x


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_get_pwms_explainn_from_reference_activations(name, monkeypatch):
    """The reference's activation array in, the reference's PWMs out -- exactly (interpretation.py:176-227)."""
    from explainn_b200 import interpretation
    from oracle import canonical as cn
    g = load_golden(name)
    a16, x = _port_act16(g)
    n_act = g["act16"].shape[0]
    assert np.array_equal(a16[:n_act], g["act16"])
    F = int(g["F"])
    pwm, counts, hits = interpretation.get_pwms_explainn(a16, torch.from_numpy(x), F, return_counts=True)
    assert pwm.dtype == np.float64 and np.array_equal(pwm, g["pwm"])
    ref_c, ref_h, _ = cn.pwm_counts(a16, x, F)
    assert np.array_equal(counts, ref_c) and np.array_equal(hits, ref_h)
    # numpy sequences, CUDA sequences, several chunks: same integers
    assert np.array_equal(interpretation.get_pwms_explainn(a16, x, F), pwm)
    assert np.array_equal(interpretation.get_pwms_explainn(a16, torch.from_numpy(x).to(DEV), F), pwm)
    real = interpretation._chunks_on_device
    monkeypatch.setattr(interpretation, "_chunks_on_device", lambda a, d, budget_bytes=0: real(a, d, 7 * a[0].nbytes + 1))
    assert np.array_equal(interpretation.get_pwms_explainn(a16, x, F), pwm)
    monkeypatch.undo()
    # float32 activations: threshold and comparison in float32 (brute-force restatement)
    a32 = a16.astype(np.float32) * np.float32(1.0003)
    thr = np.float32(0.5) * np.amax(a32, axis=(0, 2))
    n_i, u_i, s_i = np.nonzero(a32 > thr[None, :, None])
    ref32 = np.zeros_like(ref_c)
    for k in range(F):
        np.add.at(ref32[:, :, k], u_i, (x[n_i, :, s_i + k] != 0).astype(np.int64))
    _, c32, h32 = interpretation.get_pwms_explainn(a32, x, F, return_counts=True)
    assert np.array_equal(c32, ref32) and np.array_equal(h32, np.bincount(u_i, minlength=ref_c.shape[0]))
    # inf and NaN in the caller's array behave as in numpy: inf -> threshold inf -> no hit; NaN -> threshold NaN -> no hit
    b = a16.copy()
    b[0, 0, 1] = np.inf
    b[1, 1, 2] = np.nan
    with np.errstate(all="ignore"):
        rc, rh, _ = cn.pwm_counts(b, x, F)
    _, cb, hb = interpretation.get_pwms_explainn(b, x, F, return_counts=True)
    assert hb[0] == 0 and hb[1] == 0 and np.array_equal(cb, rc) and np.array_equal(hb, rh)
    with pytest.raises(ValueError):
        interpretation.get_pwms_explainn(a16, x[:, :, :-1], F)


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_seq_argmax_bit_exact(name):
    """Per (sequence, unit) max activation and first argmax position == the oracle, bit for bit."""
    from explainn_b200 import interpretation
    from oracle import canonical as cn
    g = load_golden(name)
    m = _model(g)
    sd = state_dict_from(g)
    n = lambda k: sd[k].numpy()
    x = onehot_from_codes(g["codes"])
    codes, nmask, _ = cn.pack_onehot(x)
    wq, shift, _ = cn.fold_bn(n("linears.0.weight"), n("linears.0.bias"), n("linears.1.weight"), n("linears.1.bias"),
                              n("linears.1.running_mean"), n("linears.1.running_var"))
    ref_v, ref_p = cn.seq_argmax(cn.conv_z(codes, nmask, int(g["L"]), wq, shift))
    v, p = interpretation.get_explainn_unit_argmax(torch.from_numpy(x), m, DEV)
    assert v.dtype == np.float32 and p.dtype == np.int32
    assert np.array_equal(v.view(np.uint32), ref_v.view(np.uint32)) and np.array_equal(p, ref_p)
    gv, gn, gp = interpretation.unit_argmax_global(v, p)
    assert np.array_equal(gv, ref_v.max(axis=0)) and np.array_equal(gn, ref_v.argmax(axis=0))


def test_seq_argmax_ties_through_exp():
    """Filters so small that most positions of a sequence differ by less than one float32 ulp of exp(z): the first
    position of the maximal ACTIVATION is not the first position of the maximal z.  Also saturated (inf) units."""
    from explainn_b200 import interpretation, networks
    from oracle import canonical as cn
    torch.manual_seed(3)
    U, L, F = 37, 150, 19
    m = networks.ExplaiNN(U, L, 2, F)
    with torch.no_grad():
        m.linears[0].weight.mul_(2e-7)
        m.linears[0].weight[5:9].mul_(1e3)
        m.linears[1].bias[10:14] = 95.0           # exp overflows: every position ties at inf -> position 0
        m.linears[1].bias[14:16] = -95.0          # subnormal range
        m.linears[1].bias[16:18] = -120.0         # exp underflows to 0 everywhere -> position 0
    m = m.to(DEV).eval()
    x = onehot_from_codes(random_codes(256, L, 12, n_frac=0.02))
    sd = {k: v.cpu().numpy() for k, v in m.state_dict().items()}
    codes, nmask, _ = cn.pack_onehot(x)
    wq, shift, _ = cn.fold_bn(sd["linears.0.weight"], sd["linears.0.bias"], sd["linears.1.weight"], sd["linears.1.bias"],
                              sd["linears.1.running_mean"], sd["linears.1.running_var"])
    z = cn.conv_z(codes, nmask, L, wq, shift)
    ref_v, ref_p = cn.seq_argmax(z)
    assert np.mean(ref_p != z.argmax(axis=2)) > 0.05       # the case the fast path alone would get wrong
    v, p = interpretation.get_explainn_unit_argmax(torch.from_numpy(x), m, DEV)
    assert np.array_equal(v.view(np.uint32), ref_v.view(np.uint32)) and np.array_equal(p, ref_p)
    assert np.all(p[:, 10:14] == 0) and np.all(np.isinf(v[:, 10:14])) and np.all(p[:, 16:18] == 0)


def test_unit_importance_drop_ins_vs_reference():
    """get_median_unit_importance / get_specific_unit_importance on the reference's own inputs == its outputs
    (golden: tests/golden/make_golden_api.py), and the fused variant agrees on the medians."""
    from explainn_b200 import interpretation
    from oracle import torch_port as tp
    g = load_golden("api_importance")
    m = _model(g)
    a16, x = _port_act16(g)
    port = tp.PortExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]))
    port.load_state_dict(state_dict_from(g))
    port.eval()
    with torch.no_grad():
        unit = port.unit_outputs(torch.from_numpy(x)).numpy()
    labels = ["c%d" % i for i in range(int(g["C"]))]
    med = interpretation.get_median_unit_importance(a16, m, unit, labels)
    assert med.shape == g["median"].shape and np.array_equal(med, g["median"], equal_nan=True)
    spec = interpretation.get_specific_unit_importance(a16, m, unit, int(g["specific_filt"]), labels)
    assert np.array_equal(np.stack([np.asarray(spec[k]) for k in labels]), g["specific"])
    _, med_f = interpretation.unit_importance_fused(torch.from_numpy(x), m, DEV, labels)
    np.testing.assert_allclose(med_f, g["median"], rtol=2e-4, atol=1e-6)
