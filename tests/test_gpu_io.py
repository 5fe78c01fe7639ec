"""GPU parity of SURVEY 8(f) row 4: on-device DNA encoding (tools.py:136-181) and PWM scoring (networks.py:393-437),
through the C ABI, against the reference's golden vectors and the CPU oracle."""
import numpy as np
import pytest
import torch

from util import load_golden, onehot_from_codes, random_codes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _strings(g):
    text, offs = g["text"].tobytes(), g["offsets"]
    return [text[offs[i]:offs[i + 1]].decode("ascii") for i in range(len(offs) - 1)]


def test_encode_dna_matches_reference_golden():
    from explainn_b200 import functional as Fn, tools
    g = load_golden("io_encode")
    seqs = _strings(g)
    for L in (1, 31, 32, 33, 64, 70):
        pk = tools.encode_sequences(seqs, L, DEV)
        x = Fn.unpack_onehot(pk, torch.float16)
        assert x.dtype == torch.float16
        assert np.array_equal(x.cpu().numpy().astype(np.uint8), g["onehot_L%d" % L]), L
        # the packed form is the one pack_onehot builds from that matrix (bit-exact, padding bits zero)
        ref = Fn.pack_onehot(x)
        assert torch.equal(pk.codes, ref.codes) and torch.equal(pk.nmask, ref.nmask)


def test_encode_dna_roundtrip_and_edge_cases():
    from explainn_b200 import functional as Fn, tools
    from oracle import canonical as oc
    # empty batch, empty strings, bytes input
    pk = tools.encode_sequences([], 40, DEV)
    assert pk.B == 0 and pk.codes.shape == (0, 3)
    pk = tools.encode_sequences(["", b"ACGT"], 8, DEV)
    x = Fn.unpack_onehot(pk).cpu().numpy()
    assert x[0].sum() == 0
    assert np.array_equal(x[1], oc.dna_one_hot("ACGT", 8).astype(np.float32))
    # full size: 4096 sequences of length 1000 (cfg 5 shape), text -> packed -> one-hot -> packed is the identity
    rng = np.random.default_rng(0)
    letters = np.frombuffer(b"ACGTN", dtype=np.uint8)
    codes = random_codes(4096, 1000, 3, n_frac=0.01)
    seqs = [letters[c].tobytes() for c in codes]
    pk = tools.encode_sequences(seqs, 1000, DEV)
    ref = Fn.pack_onehot(torch.from_numpy(onehot_from_codes(codes)).to(DEV))
    assert torch.equal(pk.codes, ref.codes) and torch.equal(pk.nmask, ref.nmask)
    again = Fn.pack_onehot(Fn.unpack_onehot(pk))
    assert torch.equal(again.codes, pk.codes) and torch.equal(again.nmask, pk.nmask)
    # drop-in single-sequence helper
    y = tools.dna_one_hot("acgtnACGT", 12)
    assert y.dtype == np.float16 and np.array_equal(y, oc.dna_one_hot("acgtnACGT", 12))
    assert tools.dna_one_hot("ACGT", flatten=True).shape == (1, 16)


@pytest.mark.parametrize("name", ["io_pwm_g7_f8", "io_pwm_g40_f19", "io_pwm_g33_f24"])
def test_pwm_scores_match_reference_and_oracle(name):
    from explainn_b200 import networks
    from oracle import canonical as oc
    g = load_golden(name)
    x = onehot_from_codes(g["codes"])
    G, _, F = g["pwms"].shape
    for scoring in ("max", "sum"):
        m = networks.PWM(g["pwms"], x.shape[2], scoring=scoring).to(DEV).eval()
        assert m._options == {"groups": G, "kernel_size": F, "sequence_length": x.shape[2], "scoring": scoring}
        assert not any(p.requires_grad for p in m.parameters())
        got = m(torch.from_numpy(x).to(DEV)).cpu().numpy()
        # bit-exact against the oracle (same canonical summation order) ...
        assert np.array_equal(got, oc.pwm_scores(x, g["pwms"], scoring)), scoring
        # ... and float32-rounding close to the reference module's own output
        ref = g["score_" + scoring]
        assert np.abs(got - ref).max() <= 2e-5 * np.abs(ref).max()


def test_pwm_scores_large_batch_properties():
    """Full-size property check (8192 x 251, 300 PWMs): strand symmetry and max <= sum bounds."""
    from explainn_b200 import functional as Fn, networks
    rng = np.random.default_rng(1)
    pwms = rng.normal(size=(300, 4, 19)).astype(np.float32)
    x = torch.from_numpy(onehot_from_codes(random_codes(8192, 251, 9, n_frac=0.01))).to(DEV)
    mx = networks.PWM(pwms, 251, "max").to(DEV)
    sm = networks.PWM(pwms, 251, "sum").to(DEV)
    a, s = mx(x), sm(x)
    # scoring both strands is invariant under reverse-complementing the input (max exactly; sum up to summation order)
    xr = torch.flip(x, dims=(1, 2)).contiguous()
    assert torch.equal(mx(xr), a)
    assert float((sm(xr) - s).abs().max()) <= 1e-4 * float(s.abs().max())
    S2 = 2 * (251 - 19 + 1)
    assert bool((s <= a * S2 + 1e-2).all())
