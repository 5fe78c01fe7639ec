"""CPU: the oracle's restatements of `dna_one_hot` (tools.py:136-181) and `networks.PWM.forward`
(networks.py:428-437) against golden vectors produced by the unmodified reference
(tests/golden/make_golden_io.py)."""
import numpy as np
import pytest

from oracle import canonical as oc
from util import load_golden, onehot_from_codes


def _strings(g):
    text, offs = g["text"].tobytes(), g["offsets"]
    return [text[offs[i]:offs[i + 1]].decode("ascii") for i in range(len(offs) - 1)]


def test_dna_one_hot_port_matches_reference():
    g = load_golden("io_encode")
    seqs = _strings(g)
    assert "" in seqs and any(len(s) > 70 for s in seqs)
    for L in (1, 31, 32, 33, 64, 70):
        ref = g["onehot_L%d" % L]
        for i, s in enumerate(seqs):
            x = oc.dna_one_hot(s, L)
            assert x.dtype == np.float16 and x.shape == (4, L)
            assert np.array_equal(x.astype(np.uint8), ref[i]), (L, i, s)


def test_dna_one_hot_port_default_length():
    x = oc.dna_one_hot("acgTN3")
    assert x.shape == (4, 6)
    assert x.astype(int).tolist() == [[1, 0, 0, 0, 0, 0], [0, 1, 0, 0, 0, 0], [0, 0, 1, 0, 0, 0], [0, 0, 0, 1, 0, 1]]


@pytest.mark.parametrize("name", ["io_pwm_g7_f8", "io_pwm_g40_f19", "io_pwm_g33_f24"])
def test_pwm_scores_port_matches_reference(name):
    g = load_golden(name)
    x = onehot_from_codes(g["codes"])
    for scoring in ("max", "sum"):
        got = oc.pwm_scores(x, g["pwms"], scoring)
        ref = g["score_" + scoring]
        scale = np.abs(ref).max()
        # different summation order inside the conv (mkldnn) and over positions (torch.sum): float32 rounding only
        assert np.abs(got - ref).max() <= 2e-5 * scale, (scoring, np.abs(got - ref).max(), scale)
