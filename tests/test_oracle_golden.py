"""Pins the oracle (oracle/) against outputs of the reference itself (tests/golden/*.npz, made by
tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import canonical as cn
from oracle import torch_port as tp
from util import EVAL_GOLDENS, TRAIN_GOLDENS, load_golden, onehot_from_codes, state_dict_from


def _port(g, prefix="sd."):
    m = tp.PortExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]))
    m.load_state_dict(state_dict_from(g, prefix))
    return m


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_torch_port_matches_reference_eval(name):
    g = load_golden(name)
    m = _port(g).eval()
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    with torch.no_grad():
        assert torch.equal(m(x), torch.from_numpy(g["logits"]))
        assert torch.equal(m.unit_outputs(x), torch.from_numpy(g["unit"]))
    a16 = tp.unit_activations_fp16(m, x)
    n_act = g["act16"].shape[0]
    assert np.array_equal(a16[:n_act], g["act16"])
    assert np.array_equal(np.amax(a16, axis=(0, 2)), g["amax16"])


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_canonical_conv_pool_matches_reference(name):
    g = load_golden(name)
    sd = state_dict_from(g)
    L, F = int(g["L"]), int(g["F"])
    x = onehot_from_codes(g["codes"])
    codes, nmask, bad = cn.pack_onehot(x)
    assert bad == 0
    assert np.array_equal(cn.decode_onehot(codes, nmask, L), x)
    n = lambda k: sd[k].numpy()
    wq, shift, _ = cn.fold_bn(n("linears.0.weight"), n("linears.0.bias"), n("linears.1.weight"),
                              n("linears.1.bias"), n("linears.1.running_mean"), n("linears.1.running_var"))
    z = cn.conv_z(codes, nmask, L, wq, shift)
    pooled, idx = cn.pool_exp(z)
    n_act = g["pooled"].shape[0]
    # fp32 tolerance of the north star: 1e-4 relative
    np.testing.assert_allclose(pooled[:n_act], g["pooled"], rtol=1e-5, atol=0)
    # first-max positions: bit-exact
    assert np.array_equal(idx, g["argmax"])
    # float16 activations: equal up to float16 rounding of a <= 2-ulp fp32 difference
    a16 = cn.activations_fp16(z[:n_act]).astype(np.float32)
    ref = g["act16"].astype(np.float32)
    assert np.mean(a16 != ref) < 1e-3
    np.testing.assert_allclose(a16, ref, rtol=1.1e-3, atol=1e-7)


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_pwm_counts_match_reference(name):
    """Integer count restatement == get_pwms_explainn, both from the canonical activations and
    (count stage alone) from activations produced by the reference's own ATen path."""
    g = load_golden(name)
    sd = state_dict_from(g)
    L, F = int(g["L"]), int(g["F"])
    x = onehot_from_codes(g["codes"])
    m = _port(g).eval()
    a16_ref = tp.unit_activations_fp16(m, torch.from_numpy(x))
    cnt, hits, _ = cn.pwm_counts(a16_ref, x, F)
    assert np.array_equal(cn.pwm_from_counts_hits(cnt, hits), g["pwm"])
    codes, nmask, _ = cn.pack_onehot(x)
    n = lambda k: sd[k].numpy()
    wq, shift, _ = cn.fold_bn(n("linears.0.weight"), n("linears.0.bias"), n("linears.1.weight"),
                              n("linears.1.bias"), n("linears.1.running_mean"), n("linears.1.running_var"))
    a16 = cn.activations_fp16(cn.conv_z(codes, nmask, L, wq, shift))
    cnt2, hits2, _ = cn.pwm_counts(a16, x, F)
    assert np.array_equal(cnt2, cnt), "canonical-order activations changed a PWM count"
    assert np.array_equal(cn.pwm_from_counts_hits(cnt2, hits2), g["pwm"])


@pytest.mark.parametrize("name", TRAIN_GOLDENS)
def test_torch_port_matches_reference_training(name):
    g = load_golden(name)
    m = _port(g, "sd0.")
    m.train()
    m.linears[9].eval()
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    y = torch.from_numpy(g["labels"])
    crit = torch.nn.BCEWithLogitsLoss() if str(g["loss_kind"]) == "bce" else torch.nn.MSELoss()
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    for step in range(2):
        opt.zero_grad()
        out = m(x)
        loss = crit(out, y)
        loss.backward()
        if step == 0:
            assert torch.equal(out.detach(), torch.from_numpy(g["logits"]))
            for k, p in m.named_parameters():
                assert torch.equal(p.grad, torch.from_numpy(g["grad." + k])), k
        opt.step()
        assert abs(float(loss) - g["losses"][step]) == 0
        for k, v in m.state_dict().items():
            assert torch.equal(v, torch.from_numpy(np.array(g["sd%d.%s" % (step + 1, k)]))), k


def test_det_expf_accuracy_and_monotone_sample():
    x = np.linspace(-104.5, 89.0, 2_000_001).astype(np.float32)
    y = cn.expf(x)
    ref = np.exp(x.astype(np.float64))
    ok = (ref > 1e-37) & (ref < 3e38)
    rel = np.abs(y[ok].astype(np.float64) - ref[ok]) / ref[ok]
    assert rel.max() < 2.5e-7           # <= 2 ulp
    assert np.all(y[1:] >= y[:-1])
    assert cn.expf(np.array([np.inf, -np.inf, 0.0], np.float32)).tolist() == [np.inf, 0.0, 1.0]
    assert np.isnan(cn.expf(np.array([np.nan], np.float32))[0])
    # dense neighbourhoods: consecutive floats around a few anchors stay monotone at fp16 resolution
    for a in (-9.7, -3.2, -0.7, -0.01, 0.01, 0.3, 1.0, 2.5, 5.1, 10.9):
        base = int(np.array([a], np.float32).view(np.uint32)[0])
        v = np.sort((np.arange(-20000, 20000, dtype=np.int64) + base).astype(np.uint32).view(np.float32))
        e = cn.expf(v)
        assert np.all(e[1:] >= e[:-1]), a
        h = e.astype(np.float16)
        assert np.all(h[1:] >= h[:-1]), a


@pytest.mark.parametrize("name", EVAL_GOLDENS)
def test_c_count_stage_and_argmax_match_reference(name):
    """The C (OpenMP) count stage used for the large shapes equals the numpy restatement and the reference's PWMs;
    the per-sequence max / first-argmax restatement equals torch.max(act, dim=2) (networks.py:228,281) on the
    activations of the reference's own graph wherever the canonical exp does not move the maximum."""
    g = load_golden(name)
    F = int(g["F"])
    x = onehot_from_codes(g["codes"])
    m = _port(g).eval()
    a16_ref = tp.unit_activations_fp16(m, torch.from_numpy(x))
    assert np.array_equal(cn.unit_max_fp16_c(a16_ref).view(np.uint16), cn.unit_max_fp16(a16_ref).view(np.uint16))
    cnt, hits, thr = cn.pwm_counts(a16_ref, x, F)
    cnt_c, hits_c, thr_c, seq_hit = cn.pwm_counts_c(a16_ref, x, F, want_seq_hit=True)
    assert np.array_equal(cnt_c, cnt) and np.array_equal(hits_c, hits) and np.array_equal(thr_c, thr)
    assert np.array_equal(cn.pwm_from_counts_hits(cnt_c, hits_c), g["pwm"])
    assert np.array_equal(seq_hit.astype(bool), (a16_ref > thr[None, :, None]).any(axis=2))
    # chunked accumulation with the global maxima gives the same integers
    amax = cn.unit_max_fp16_c(a16_ref)
    h = a16_ref.shape[0] // 2
    c2, h2, _ = cn.pwm_counts_c(a16_ref[:h], x[:h], F, amax16=amax)
    c2, h2, _ = cn.pwm_counts_c(a16_ref[h:], x[h:], F, amax16=amax, counts=c2, hits=h2)
    assert np.array_equal(c2, cnt) and np.array_equal(h2, hits)
    # NaN and inf handling of the maximum follows numpy
    b = a16_ref[:4].copy()
    b[1, 0, 3] = np.nan
    b[2, 1, 5] = np.inf
    assert np.array_equal(cn.unit_max_fp16_c(b).view(np.uint16)[1:], np.amax(b, axis=(0, 2)).view(np.uint16)[1:])
    assert np.isnan(cn.unit_max_fp16_c(b)[0])
    # per-sequence max and first argmax
    sd = state_dict_from(g)
    n = lambda k: sd[k].numpy()
    codes, nmask, _ = cn.pack_onehot(x)
    wq, shift, _ = cn.fold_bn(n("linears.0.weight"), n("linears.0.bias"), n("linears.1.weight"),
                              n("linears.1.bias"), n("linears.1.running_mean"), n("linears.1.running_var"))
    z = cn.conv_z(codes, nmask, int(g["L"]), wq, shift)
    amax_c, apos_c = cn.seq_argmax(z)
    with torch.no_grad():
        act = m.activations(torch.from_numpy(x))
    vmax, imax = torch.max(act, dim=2)
    np.testing.assert_allclose(amax_c, vmax.numpy(), rtol=1e-5)
    # positions agree except where the two fp32 evaluations order a near-tie differently
    assert np.mean(apos_c != imax.numpy()) < 2e-3
    e = cn.expf(z)
    assert np.array_equal(apos_c, np.argmax(e, axis=2)) and np.array_equal(amax_c, e.max(axis=2))
