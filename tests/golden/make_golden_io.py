"""Golden vectors for the input-encoding and PWM-scoring rows, produced by the
UNMODIFIED reference: explainn.utils.tools.dna_one_hot (tools.py:136-181) and explainn.models.networks.PWM
(networks.py:393-437).

    python tests/golden/make_golden_io.py          # rewrites tests/golden/io_encode.npz and io_pwm_*.npz
"""
import os
import sys
import types

import numpy as np
import torch

for _m in ("h5py", "matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(_m, types.ModuleType(_m))
sys.path.insert(0, "/root/reference")
from explainn import networks  # noqa: E402
from explainn.utils import tools  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def random_strings(n, seed):
    g = np.random.default_rng(seed)
    alphabet = list("ACGTacgtNnRYKM-*.0123 4x")
    p = np.array([20, 20, 20, 20, 6, 6, 6, 6, 2, 1] + [0.5] * (len(alphabet) - 10))
    p = p / p.sum()
    out = []
    for i in range(n):
        ln = int(g.integers(0, 90))
        out.append("".join(g.choice(alphabet, size=ln, p=p)))
    out += ["", "A", "acgt" * 16, "N" * 50, "ACGT" * 40, "T" * 64, "0123" * 10]
    return out


def golden_encode():
    seqs = random_strings(57, 5)
    out = {"seqs": np.array([s.encode("ascii") for s in seqs], dtype=object)}
    # numpy cannot save object arrays without pickle: store the bytes flat + offsets
    flat = b"".join(s.encode("ascii") for s in seqs)
    offs = np.cumsum([0] + [len(s) for s in seqs]).astype(np.int64)
    out = {"text": np.frombuffer(flat, dtype=np.uint8).copy(), "offsets": offs}
    for L in (1, 31, 32, 33, 64, 70):
        x = np.stack([tools.dna_one_hot(s, seq_len=L) for s in seqs])
        assert x.dtype == np.float16 and x.shape == (len(seqs), 4, L)
        out["onehot_L%d" % L] = x.astype(np.uint8)            # values are 0/1
    np.savez_compressed(os.path.join(HERE, "io_encode.npz"), **out)
    print("io_encode", len(seqs), "sequences")


def golden_pwm(name, G, F, L, B, seed, n_frac):
    g = np.random.default_rng(seed)
    pwms = g.normal(size=(G, 4, F)).astype(np.float32)
    codes = g.integers(0, 4, size=(B, L)).astype(np.uint8)
    codes[g.random((B, L)) < n_frac] = 4
    x = np.zeros((B, 4, L), np.float32)
    for c in range(4):
        x[:, c, :] = codes == c
    out = {"pwms": pwms, "codes": codes}
    for scoring in ("max", "sum"):
        m = networks.PWM(pwms, L, scoring=scoring).eval()
        with torch.no_grad():
            out["score_" + scoring] = m(torch.from_numpy(x)).numpy().copy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, out["score_max"].shape)


if __name__ == "__main__":
    golden_encode()
    golden_pwm("io_pwm_g7_f8", 7, 8, 50, 24, 3, 0.03)
    golden_pwm("io_pwm_g40_f19", 40, 19, 200, 16, 4, 0.01)
    golden_pwm("io_pwm_g33_f24", 33, 24, 101, 12, 5, 0.0)
