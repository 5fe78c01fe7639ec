"""GPU parity of the SURVEY 8(f) rows built so far: strand-aware prediction (scripts/fasta2predictions.py:148-201)
and unit importance (interpretation.py:303-340), against the CPU port / numpy restatements."""
import numpy as np
import pytest
import torch

from util import load_golden, onehot_from_codes, random_codes, state_dict_from

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _pair(name="eval_u40_f20"):
    from explainn_b200 import networks
    from oracle import torch_port as tp
    g = load_golden(name)
    U, L, C, F = int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"])
    sd = state_dict_from(g)
    m = networks.ExplaiNN(U, L, C, F); m.load_state_dict(sd); m = m.to(DEV).eval()
    port = tp.PortExplaiNN(U, L, C, F); port.load_state_dict(sd); port.eval()
    return g, m, port


def test_revcomp_packed_bit_exact():
    from explainn_b200 import functional as Fn
    for L in (16, 31, 32, 33, 100, 251, 1000):
        x = torch.from_numpy(onehot_from_codes(random_codes(7, L, L, n_frac=0.05))).to(DEV)
        pk = Fn.pack_onehot(x)
        rc = Fn.revcomp_packed(pk)
        ref = Fn.pack_onehot(torch.flip(x, dims=(1, 2)).contiguous())      # s[::-1, ::-1]
        assert torch.equal(rc.codes, ref.codes), L
        assert torch.equal(rc.nmask, ref.nmask), L


@pytest.mark.parametrize("sigmoid", [False, True])
def test_predict_both_strands_matches_reference_loop(sigmoid):
    from explainn_b200 import predict
    g, m, port = _pair()
    x = onehot_from_codes(g["codes"])
    preds = predict.predict_both_strands(x, m, DEV, apply_sigmoid=sigmoid, batch_size=25)
    # the reference loop (fasta2predictions.py:160-187) on the CPU port
    xt = torch.from_numpy(x)
    rc = torch.from_numpy(np.array([s[::-1, ::-1] for s in x]).copy())
    with torch.no_grad():
        fwd = np.expand_dims(port(xt).numpy(), axis=2)
        rev = np.expand_dims(port(rc).numpy(), axis=2)
    fr = np.concatenate((fwd, rev), axis=2)
    ref = np.empty((x.shape[0], fwd.shape[1], 4))
    ref[:] = np.concatenate((fwd, rev, np.expand_dims(np.mean(fr, axis=2), 2), np.expand_dims(np.max(fr, axis=2), 2)), axis=2)
    if sigmoid:
        ref = torch.sigmoid(torch.Tensor(ref)).numpy()
    assert preds.shape == ref.shape and preds.dtype == ref.dtype
    np.testing.assert_allclose(preds, ref, rtol=1e-4, atol=1e-4 * np.abs(ref).max())
    df = predict.predictions_to_frame(preds, ["s%d" % i for i in range(x.shape[0])])
    assert list(df.columns) == ["SeqId", "Class", "Fwd", "Rev", "Mean", "Max"] and len(df) == x.shape[0] * fwd.shape[1]


def test_unit_importance_fused_matches_reference_semantics():
    from explainn_b200 import interpretation
    from oracle import torch_port as tp
    g, m, port = _pair()
    x = torch.from_numpy(onehot_from_codes(g["codes"]))
    labels = ["c%d" % i for i in range(int(g["C"]))]
    imp, med = interpretation.unit_importance_fused(x, m, DEV, labels)
    # reference semantics on host arrays (interpretation.py:303-340)
    act16 = tp.unit_activations_fp16(port, x)
    with torch.no_grad():
        uo = port.unit_outputs(x).numpy()
    thr = 0.5 * np.amax(act16, axis=(0, 2))
    w = port.final.weight.detach().numpy()
    for u in (0, 7, 39):
        rows = (act16[:, u, :] > thr[u]).any(axis=1)
        s = imp[u]
        for c, lab in enumerate(labels):
            ref = (uo * w[c])[:, u][rows]
            assert s[lab].shape == ref.shape
            np.testing.assert_allclose(s[lab], ref, rtol=1e-4, atol=1e-5 * max(np.abs(ref).max(), 1e-6))
        if rows.any():
            np.testing.assert_allclose(med[u], np.median((uo[rows, u][:, None] * w[None, :, u]), axis=0), rtol=1e-4, atol=1e-6)
    # drop-in host version
    s2 = interpretation.get_specific_unit_importance(act16, m, uo, 7, labels)
    rows = (act16[:, 7, :] > thr[7]).any(axis=1)
    np.testing.assert_allclose(s2["c1"], (uo * w[1])[:, 7][rows], rtol=1e-5)
